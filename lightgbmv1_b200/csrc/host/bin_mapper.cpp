#include "bin_mapper.h"

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <iomanip>

namespace lgb {
namespace {

const double kInf = std::numeric_limits<double>::infinity();

// GreedyFindBin (src/io/bin.cpp:78-155): equal-count bins over distinct values; values whose count reaches the
// mean bin size get a bin of their own.
std::vector<double> GreedyBounds(const double* vals, const int* cnts, int n_distinct, int max_bin, size_t total_cnt,
                                 int min_data_in_bin) {
  std::vector<double> bounds;
  if (max_bin <= 0) Log::Fatal("Check failed: (max_bin) > (0)");
  if (n_distinct <= max_bin) {
    int acc = 0;
    for (int i = 0; i + 1 < n_distinct; ++i) {
      acc += cnts[i];
      if (acc >= min_data_in_bin) {
        const double mid = NextUp((vals[i] + vals[i + 1]) / 2.0);
        if (bounds.empty() || !EqualOrdered(bounds.back(), mid)) {
          bounds.push_back(mid);
          acc = 0;
        }
      }
    }
    bounds.push_back(kInf);
    return bounds;
  }
  if (min_data_in_bin > 0) {
    max_bin = std::min(max_bin, static_cast<int>(total_cnt / min_data_in_bin));
    max_bin = std::max(max_bin, 1);
  }
  double mean_size = static_cast<double>(total_cnt) / max_bin;
  int rest_bins = max_bin;
  int rest_cnt = static_cast<int>(total_cnt);
  std::vector<bool> is_big(n_distinct, false);
  for (int i = 0; i < n_distinct; ++i) {
    if (cnts[i] >= mean_size) {
      is_big[i] = true;
      --rest_bins;
      rest_cnt -= cnts[i];
    }
  }
  mean_size = static_cast<double>(rest_cnt) / rest_bins;
  std::vector<double> hi(max_bin, kInf), lo(max_bin, kInf);
  int nb = 0;
  lo[0] = vals[0];
  int acc = 0;
  for (int i = 0; i + 1 < n_distinct; ++i) {
    if (!is_big[i]) rest_cnt -= cnts[i];
    acc += cnts[i];
    if (is_big[i] || acc >= mean_size || (is_big[i + 1] && acc >= std::max(1.0, mean_size * 0.5f))) {
      hi[nb] = vals[i];
      ++nb;
      lo[nb] = vals[i + 1];
      if (nb >= max_bin - 1) break;
      acc = 0;
      if (!is_big[i]) {
        --rest_bins;
        mean_size = rest_cnt / static_cast<double>(rest_bins);
      }
    }
  }
  ++nb;
  for (int i = 0; i + 1 < nb; ++i) {
    const double mid = NextUp((hi[i] + lo[i + 1]) / 2.0);
    if (bounds.empty() || !EqualOrdered(bounds.back(), mid)) bounds.push_back(mid);
  }
  bounds.push_back(kInf);
  return bounds;
}

// FindBinWithZeroAsOneBin (src/io/bin.cpp:256-312): negative side, the zero bin (-1e-35, 1e-35], positive side.
std::vector<double> BoundsWithZeroBin(const double* vals, const int* cnts, int n_distinct, int max_bin,
                                      size_t total_sample_cnt, int min_data_in_bin) {
  std::vector<double> bounds;
  int left_cnt_data = 0, cnt_zero = 0, right_cnt_data = 0;
  for (int i = 0; i < n_distinct; ++i) {
    if (vals[i] <= -kZeroThreshold) left_cnt_data += cnts[i];
    else if (vals[i] > kZeroThreshold) right_cnt_data += cnts[i];
    else cnt_zero += cnts[i];
  }
  int left_cnt = -1;
  for (int i = 0; i < n_distinct; ++i) {
    if (vals[i] > -kZeroThreshold) { left_cnt = i; break; }
  }
  if (left_cnt < 0) left_cnt = n_distinct;
  if (left_cnt > 0 && max_bin > 1) {
    int left_max_bin = static_cast<int>(static_cast<double>(left_cnt_data) / (total_sample_cnt - cnt_zero) * (max_bin - 1));
    left_max_bin = std::max(1, left_max_bin);
    bounds = GreedyBounds(vals, cnts, left_cnt, left_max_bin, left_cnt_data, min_data_in_bin);
    if (!bounds.empty()) bounds.back() = -kZeroThreshold;
  }
  int right_start = -1;
  for (int i = left_cnt; i < n_distinct; ++i) {
    if (vals[i] > kZeroThreshold) { right_start = i; break; }
  }
  const int right_max_bin = max_bin - 1 - static_cast<int>(bounds.size());
  if (right_start >= 0 && right_max_bin > 0) {
    auto right = GreedyBounds(vals + right_start, cnts + right_start, n_distinct - right_start, right_max_bin,
                              right_cnt_data, min_data_in_bin);
    bounds.push_back(kZeroThreshold);
    bounds.insert(bounds.end(), right.begin(), right.end());
  } else {
    bounds.push_back(kInf);
  }
  if (bounds.size() > static_cast<size_t>(max_bin)) Log::Fatal("Check failed: bin_upper_bound.size() <= max_bin");
  return bounds;
}

// NeedFilter (src/io/bin.cpp:54-76), numerical branch: no split leaves >= filter_cnt rows on both sides
bool NoUsefulSplit(const std::vector<int>& cnt_in_bin, int total_cnt, int filter_cnt) {
  int left = 0;
  for (size_t i = 0; i + 1 < cnt_in_bin.size(); ++i) {
    left += cnt_in_bin[i];
    if (left >= filter_cnt && total_cnt - left >= filter_cnt) return false;
  }
  return true;
}

// Ascending sort of finite doubles (NaNs are removed before; zeros never reach FindBin, so -0.0 / +0.0 cannot be
// told apart by it).  Equal doubles are indistinguishable, so any correct sort gives the array the reference's
// std::stable_sort gives; an LSD radix sort over the order-preserving integer image is ~6x faster on the 200K-value
// samples this runs on, once per feature.
void SortDoubles(std::vector<double>* values) {
  const size_t n = values->size();
  if (n < 8192) { std::sort(values->begin(), values->end()); return; }
  // A float32 matrix arrives here as doubles that are all exactly representable as floats: those sort by their 32-bit
  // images in two 16-bit passes over half the bytes (3.9 instead of 7.7 ms per 200K values).
  bool all_f32 = true;
  for (size_t i = 0; i < n && all_f32; ++i) all_f32 = static_cast<double>(static_cast<float>((*values)[i])) == (*values)[i];
  if (all_f32) {
    std::vector<uint32_t> a(n), b(n);
    for (size_t i = 0; i < n; ++i) {
      const float f = static_cast<float>((*values)[i]);
      uint32_t u; std::memcpy(&u, &f, 4);
      a[i] = u ^ ((u >> 31) ? ~0u : 0x80000000u);
    }
    std::vector<uint32_t> count(65536);
    for (int pass = 0; pass < 2; ++pass) {
      const int shift = 16 * pass;
      std::fill(count.begin(), count.end(), 0u);
      for (size_t i = 0; i < n; ++i) ++count[(a[i] >> shift) & 0xffffu];
      uint32_t run = 0;
      for (uint32_t& c : count) { const uint32_t t = c; c = run; run += t; }
      for (size_t i = 0; i < n; ++i) b[count[(a[i] >> shift) & 0xffffu]++] = a[i];
      a.swap(b);
    }
    for (size_t i = 0; i < n; ++i) {
      const uint32_t u = a[i] ^ ((a[i] >> 31) ? 0x80000000u : ~0u);
      float f; std::memcpy(&f, &u, 4);
      (*values)[i] = static_cast<double>(f);
    }
    return;
  }
  std::vector<uint64_t> a(n), b(n);
  for (size_t i = 0; i < n; ++i) {
    uint64_t u; std::memcpy(&u, &(*values)[i], 8);
    a[i] = u ^ ((u >> 63) ? ~0ull : 0x8000000000000000ull);
  }
  std::vector<uint32_t> count(65536);
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 16 * pass;
    std::fill(count.begin(), count.end(), 0u);
    for (size_t i = 0; i < n; ++i) ++count[(a[i] >> shift) & 0xffffu];
    uint32_t run = 0;
    for (uint32_t& c : count) { const uint32_t t = c; c = run; run += t; }
    for (size_t i = 0; i < n; ++i) b[count[(a[i] >> shift) & 0xffffu]++] = a[i];
    a.swap(b);
  }
  for (size_t i = 0; i < n; ++i) {
    const uint64_t u = a[i] ^ ((a[i] >> 63) ? 0x8000000000000000ull : ~0ull);
    std::memcpy(&(*values)[i], &u, 8);
  }
}

}  // namespace

void BinMapper::FindBin(std::vector<double> values, size_t total_sample_cnt, int max_bin, int min_data_in_bin,
                        int min_split_data, bool pre_filter, bool use_missing, bool zero_as_missing, bool categorical) {
  categorical_ = categorical;
  int num_sample = static_cast<int>(values.size());
  // drop NaNs, remember how many
  int kept = 0;
  for (int i = 0; i < num_sample; ++i) if (!std::isnan(values[i])) values[kept++] = values[i];
  int na_cnt = 0;
  if (!use_missing) missing_type_ = kMissingNone;
  else if (zero_as_missing) missing_type_ = kMissingZero;
  else if (kept == num_sample) missing_type_ = kMissingNone;
  else { missing_type_ = kMissingNaN; na_cnt = num_sample - kept; }
  num_sample = kept;
  values.resize(kept);
  default_bin_ = 0;
  const int zero_cnt = static_cast<int>(total_sample_cnt - num_sample - na_cnt);

  SortDoubles(&values);
  // distinct values with counts; the implied zeros are spliced in at their sorted position
  std::vector<double> dv;
  std::vector<int> dc;
  dv.reserve(static_cast<size_t>(num_sample) + 2); dc.reserve(static_cast<size_t>(num_sample) + 2);   // (no regrowth: most sampled values are distinct)
  if (num_sample == 0 || (values[0] > 0.0f && zero_cnt > 0)) { dv.push_back(0.0f); dc.push_back(zero_cnt); }
  if (num_sample > 0) { dv.push_back(values[0]); dc.push_back(1); }
  for (int i = 1; i < num_sample; ++i) {
    if (!EqualOrdered(values[i - 1], values[i])) {
      if (values[i - 1] < 0.0f && values[i] > 0.0f) { dv.push_back(0.0f); dc.push_back(zero_cnt); }
      dv.push_back(values[i]);
      dc.push_back(1);
    } else {
      dv.back() = values[i];  // keep the larger of two values one ulp apart
      ++dc.back();
    }
  }
  if (num_sample > 0 && values[num_sample - 1] < 0.0f && zero_cnt > 0) { dv.push_back(0.0f); dc.push_back(zero_cnt); }
  min_val_ = dv.front();
  max_val_ = dv.back();
  const int nd = static_cast<int>(dv.size());

  std::vector<int> cnt_in_bin;
  if (categorical_) {
    // src/io/bin.cpp:424-494: categories as ints (negative = NaN), by descending count; categories are kept until 99 % of
    // the rows are covered AND max_bin bins exist (the reference's loop condition is an OR), rare ones share bin 0
    std::vector<int> dvi, dci;
    for (int i = 0; i < nd; ++i) {
      const int val = static_cast<int>(dv[i]);
      if (val < 0) { na_cnt += dc[i]; Log::Warning("Met negative value in categorical features, will convert it to NaN"); }
      else if (dvi.empty() || val != dvi.back()) { dvi.push_back(val); dci.push_back(dc[i]); }
      else dci.back() += dc[i];
    }
    upper_.clear(); bin_2_cat_.clear(); cat_2_bin_.clear();
    num_bin_ = 0;
    const int rest_cnt = static_cast<int>(total_sample_cnt - na_cnt);
    if (rest_cnt > 0) {
      if (dvi.back() / 100 > static_cast<int>(dvi.size()))
        Log::Warning("Met categorical feature which contains sparse values. Consider renumbering to consecutive integers started from zero");
      std::vector<std::pair<int, int>> arr;   // Common::SortForPair(counts, values, 0, true): stable, by count descending
      for (size_t i = 0; i < dvi.size(); ++i) arr.emplace_back(dci[i], dvi[i]);
      std::stable_sort(arr.begin(), arr.end(), [](const std::pair<int, int>& a, const std::pair<int, int>& b) { return a.first > b.first; });
      const int cut_cnt = static_cast<int>(static_cast<double>((total_sample_cnt - na_cnt) * 0.99f) + 0.5f);   // RoundInt of a float product
      size_t cur_cat = 0;
      int used_cnt = 0;
      int distinct_cnt = static_cast<int>(arr.size());
      if (na_cnt > 0) ++distinct_cnt;
      max_bin = std::min(distinct_cnt, max_bin);
      bin_2_cat_.push_back(-1); cat_2_bin_[-1] = 0; cnt_in_bin.push_back(0);   // the dummy bin for NaN
      num_bin_ = 1;
      while (cur_cat < arr.size() && (used_cnt < cut_cnt || num_bin_ < max_bin)) {
        if (arr[cur_cat].first < min_data_in_bin && cur_cat > 1) break;
        bin_2_cat_.push_back(arr[cur_cat].second);
        cat_2_bin_[arr[cur_cat].second] = static_cast<uint32_t>(num_bin_);
        used_cnt += arr[cur_cat].first;
        cnt_in_bin.push_back(arr[cur_cat].first);
        ++num_bin_;
        ++cur_cat;
      }
      missing_type_ = (cur_cat == arr.size() && na_cnt == 0) ? kMissingNone : kMissingNaN;
      cnt_in_bin[0] = static_cast<int>(total_sample_cnt - used_cnt);
    }
    trivial_ = num_bin_ <= 1;
    // NeedFilter, categorical branch (bin.cpp:63-73): only a two-bin feature can be filtered
    if (!trivial_ && pre_filter && cnt_in_bin.size() <= 2) {
      bool useful = false;
      for (size_t i = 0; i + 1 < cnt_in_bin.size(); ++i) {
        const int left = cnt_in_bin[i];
        if (left >= min_split_data && static_cast<int>(total_sample_cnt) - left >= min_split_data) useful = true;
      }
      if (!useful) trivial_ = true;
    }
  } else {
  if (missing_type_ == kMissingZero) {
    upper_ = BoundsWithZeroBin(dv.data(), dc.data(), nd, max_bin, total_sample_cnt, min_data_in_bin);
    if (upper_.size() == 2) missing_type_ = kMissingNone;
  } else if (missing_type_ == kMissingNone) {
    upper_ = BoundsWithZeroBin(dv.data(), dc.data(), nd, max_bin, total_sample_cnt, min_data_in_bin);
  } else {
    upper_ = BoundsWithZeroBin(dv.data(), dc.data(), nd, max_bin - 1, total_sample_cnt - na_cnt, min_data_in_bin);
    // src/io/bin.cpp:407 writes `bin_upper_bound_.push_back(NaN)` where `NaN` is the MissingType enumerator (= 2),
    // so the reference's NaN bin carries the upper bound 2.0.  Never searched (ValueToBin stops one bin earlier)
    // nor used as a threshold; kept identical so that mapper dumps compare equal.
    upper_.push_back(static_cast<double>(kMissingNaN));
  }
  num_bin_ = static_cast<int>(upper_.size());
  cnt_in_bin.assign(num_bin_, 0);
  {
    int b = 0;
    for (int i = 0; i < nd; ++i) {
      if (dv[i] > upper_[b]) ++b;
      cnt_in_bin[b] += dc[i];
    }
    if (missing_type_ == kMissingNaN) cnt_in_bin[num_bin_ - 1] = na_cnt;
  }
  if (num_bin_ > max_bin) Log::Fatal("Check failed: (num_bin_) <= (max_bin)");

  trivial_ = num_bin_ <= 1;
  if (!trivial_ && pre_filter && NoUsefulSplit(cnt_in_bin, static_cast<int>(total_sample_cnt), min_split_data)) trivial_ = true;
  }
  if (!trivial_) {
    default_bin_ = ValueToBin(0);
    // ArrayArgs<int>::ArgMax: first maximum
    most_freq_bin_ = static_cast<uint32_t>(std::max_element(cnt_in_bin.begin(), cnt_in_bin.end()) - cnt_in_bin.begin());
    const double max_rate = static_cast<double>(cnt_in_bin[most_freq_bin_]) / total_sample_cnt;
    if (most_freq_bin_ != default_bin_ && max_rate < kSparseThreshold) most_freq_bin_ = default_bin_;
    sparse_rate_ = static_cast<double>(cnt_in_bin[most_freq_bin_]) / total_sample_cnt;
  } else {
    sparse_rate_ = 1.0f;
  }
}

uint32_t BinMapper::ValueToBin(double value) const {
  if (categorical_) {   // bin.h:457-460, 483-494
    if (std::isnan(value)) return 0;
    // (the reference casts whatever it gets; a value outside the int range is undefined there and "unseen" here)
    if (!(value > -2147483649.0 && value < 2147483648.0)) return 0;
    const int iv = static_cast<int>(value);
    if (iv < 0) return 0;
    const auto it = cat_2_bin_.find(iv);
    return it == cat_2_bin_.end() ? 0 : it->second;
  }
  if (std::isnan(value)) {
    if (missing_type_ == kMissingNaN) return num_bin_ - 1;
    value = 0.0f;
  }
  int l = 0, r = num_bin_ - 1;
  if (missing_type_ == kMissingNaN) r -= 1;
  while (l < r) {
    const int m = (r + l - 1) / 2;
    if (value <= upper_[m]) r = m;
    else l = m + 1;
  }
  return l;
}

std::string BinMapper::info_string() const {
  if (categorical_) {   // Common::Join(bin_2_categorical_, ":")
    std::ostringstream c;
    for (size_t i = 0; i < bin_2_cat_.size(); ++i) { if (i) c << ':'; c << bin_2_cat_[i]; }
    return c.str();
  }
  std::ostringstream o;
  o << std::setprecision(std::numeric_limits<double>::digits10 + 2) << '[' << min_val_ << ':' << max_val_ << ']';
  return o.str();
}

}  // namespace lgb
