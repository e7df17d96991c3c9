#include "dataset.h"

#include <omp.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "lgbcu.h"
#include "random.h"

namespace lgb {
namespace {

// LGBM_B200_TIMING=1: wall-clock of the dataset-construction stages on stderr
struct DsTimer {
  bool on; std::chrono::steady_clock::time_point t;
  DsTimer() : on(std::getenv("LGBM_B200_TIMING") != nullptr), t(std::chrono::steady_clock::now()) {}
  void tick(const char* what) {
    if (!on) return;
    const auto n = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[b200 timing] dataset: %-22s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
    t = n;
  }
};

// RowFunctionFromDenseMatric (src/c_api.cpp:2296-2341): element accessor for f32/f64 x row/col-major
struct MatView {
  const void* data; int type; int32_t nrow, ncol; bool row_major;
  double at(int32_t r, int32_t c) const {
    const size_t i = row_major ? static_cast<size_t>(r) * ncol + c : static_cast<size_t>(c) * nrow + r;
    return type == 0 ? static_cast<double>(static_cast<const float*>(data)[i]) : static_cast<const double*>(data)[i];
  }
};

// ---- Exclusive-feature-bundling decision (src/io/dataset.cpp:50-313).  The grouping DECISION and the resulting
// feature order are reproduced; the storage is not: this library keeps every feature in its own column(s).  What a
// shared column changes in the reference's results -- in a row where two features of a bundle are both off their
// most-frequent bin, the later push overwrites the earlier one -- is applied to the binned matrix instead
// (BinAllRows / lgbcu_bin_matrix: LgbCuBinFeature::bundle), so bundled data trains to the same trees.
int ConflictCount(const std::vector<bool>& mark, const std::vector<int>& idx, int max_cnt) {
  int ret = 0;
  for (int i : idx) {
    if (mark[i]) ++ret;
    if (ret > max_cnt) return -1;
  }
  return ret;
}

// (The conflict marks of a group are only needed when a later feature is tried against it.  With dense features no group
// is ever available to a later feature, so the marks -- one bit per sampled row and feature -- are built on first
// use instead of up front: 17 ms less at 28 dense features x 200K samples.)
struct GroupMarks {
  std::vector<std::vector<bool>> marks;      // empty vector = not materialised yet
  const std::vector<const std::vector<int>*>* nz;
  const std::vector<std::vector<int>>* groups;
  int total;
  std::vector<bool>& get(int g) {
    if (marks[g].empty()) {
      marks[g].assign(total, false);
      for (int f : (*groups)[g]) for (int i : *(*nz)[f]) marks[g][i] = true;
    }
    return marks[g];
  }
  void add(int g, int f) { if (!marks[g].empty()) for (int i : *(*nz)[f]) marks[g][i] = true; }
};

std::vector<std::vector<int>> FindGroups(const std::vector<std::shared_ptr<BinMapper>>& mappers,
                                         const std::vector<int>& order, const std::vector<const std::vector<int>*>& nz_idx,
                                         int total_sample_cnt, data_size_t num_data, bool is_sparse,
                                         std::vector<int8_t>* multi_val) {
  const int max_search_group = 100;
  const int max_conflict = total_sample_cnt / 10000;
  multi_val->clear();
  Random rand(num_data);
  std::vector<std::vector<int>> groups;
  GroupMarks gm; gm.nz = &nz_idx; gm.groups = &groups; gm.total = total_sample_cnt;
  std::vector<int> used_rows, total_cnt, group_bins;
  for (int f : order) {
    const int nz = static_cast<int>(nz_idx[f]->size());
    std::vector<int> avail;
    for (int g = 0; g < static_cast<int>(groups.size()); ++g) {
      if (total_cnt[g] + nz <= total_sample_cnt + max_conflict) avail.push_back(g);
    }
    std::vector<int> search;
    if (!avail.empty()) {
      const int last = static_cast<int>(avail.size()) - 1;
      auto pick = rand.Sample(last, std::min(last, max_search_group - 1));
      search.push_back(avail.back());
      for (int i : pick) search.push_back(avail[i]);
    }
    int best = -1, best_conf = -1;
    for (int g : search) {
      const int rest = max_conflict - total_cnt[g] + used_rows[g];
      const int cnt = ConflictCount(gm.get(g), *nz_idx[f], rest);
      if (cnt >= 0 && cnt <= rest && cnt <= nz / 2) { best = g; best_conf = cnt; break; }
    }
    if (best >= 0) {
      gm.add(best, f);   // (before the push: a materialised mark set is complete up to here)
      groups[best].push_back(f);
      total_cnt[best] += nz;
      used_rows[best] += nz - best_conf;
      group_bins[best] += mappers[f]->num_bin() + (mappers[f]->default_bin() == 0 ? -1 : 0);
    } else {
      groups.emplace_back(1, f);
      gm.marks.emplace_back();
      total_cnt.push_back(nz);
      used_rows.push_back(nz);
      group_bins.push_back(1 + mappers[f]->num_bin() + (mappers[f]->default_bin() == 0 ? -1 : 0));
    }
  }
  if (!is_sparse) { multi_val->assign(groups.size(), 0); return groups; }
  std::vector<int> second;
  std::vector<std::vector<int>> groups2;
  std::vector<std::vector<bool>> marks2;   // (the second round never reads the kept groups' marks)
  for (size_t g = 0; g < groups.size(); ++g) {
    const double dense_rate = static_cast<double>(used_rows[g]) / total_sample_cnt;
    if (dense_rate >= 0.4) { groups2.push_back(std::move(groups[g])); }
    else for (int f : groups[g]) second.push_back(f);
  }
  groups = groups2;
  multi_val->assign(groups.size(), 0);
  if (!second.empty()) {
    groups.emplace_back();
    std::vector<bool> mark(total_sample_cnt, false);
    bool is_multi = false;
    int conflicts = 0;
    for (int f : second) {
      groups.back().push_back(f);
      if (!is_multi) {
        const int rest = max_conflict - conflicts;
        const int cnt = ConflictCount(mark, *nz_idx[f], rest);
        conflicts += cnt;
        if (cnt < 0 || conflicts > max_conflict) { is_multi = true; continue; }
        for (int i : *nz_idx[f]) mark[i] = true;
      }
    }
    multi_val->push_back(is_multi ? 1 : 0);
  }
  return groups;
}

// FastFeatureBundling (src/io/dataset.cpp:236-313) -> inner feature order
// `bundle_of_real`: per raw column, 0 or the 1-based id of the single-column bundle it belongs to
std::vector<int> BundledFeatureOrder(const std::vector<std::shared_ptr<BinMapper>>& mappers,
                                     const std::vector<int>& used, const std::vector<std::vector<double>>& sample_vals,
                                     const std::vector<std::vector<int>>& sample_idx, int total_sample_cnt, data_size_t num_data,
                                     std::vector<int>* bundle_of_real) {
  // FixSampleIndices: when the most frequent bin is not the zero bin, "non-zero" means "not in the most frequent bin"
  std::vector<const std::vector<int>*> nz(sample_idx.size(), nullptr);
  std::vector<std::vector<int>> fixed_lists(sample_idx.size());
  for (size_t f = 0; f < sample_idx.size(); ++f) nz[f] = &sample_idx[f];
  for (int f : used) {
    const BinMapper& m = *mappers[f];
    if (m.default_bin() == m.most_freq_bin()) continue;
    std::vector<int>& fixed = fixed_lists[f];
    const auto& idx = sample_idx[f];
    const auto& val = sample_vals[f];
    int i = 0; size_t j = 0;
    while (i < total_sample_cnt) {
      if (j < idx.size() && idx[j] < i) ++j;
      else if (j < idx.size() && idx[j] == i) {
        if (m.ValueToBin(val[j]) != m.most_freq_bin()) fixed.push_back(i);
        ++i;
      } else fixed.push_back(i++);
    }
    // an empty result means "keep the original list" in the reference (it tests ret.empty())
    if (!fixed.empty()) nz[f] = &fixed;
  }
  std::vector<int> by_cnt(used.size());
  for (size_t i = 0; i < used.size(); ++i) by_cnt[i] = static_cast<int>(i);
  // NOTE the reference sorts by the ORIGINAL non-zero counts (num_per_col), not the fixed ones; ordering only
  // matters when it yields fewer groups, which for one-feature groups it never does.
  std::vector<int8_t> mv1, mv2;
  auto g1 = FindGroups(mappers, used, nz, total_sample_cnt, num_data, true, &mv1);
  std::vector<int> order2(used);
  std::stable_sort(order2.begin(), order2.end(), [&](int a, int b) { return sample_vals[a].size() > sample_vals[b].size(); });
  auto g2 = FindGroups(mappers, order2, nz, total_sample_cnt, num_data, true, &mv2);
  if (g1.size() > g2.size()) { g1 = g2; mv1 = mv2; }
  const int ng = static_cast<int>(g1.size());
  Random shuffle(num_data);
  for (int i = 0; i < ng - 1; ++i) {
    const int j = shuffle.NextShort(i + 1, ng);
    std::swap(g1[i], g1[j]);
    std::swap(mv1[i], mv1[j]);
  }
  std::vector<int> order;
  int next_bundle = 0;
  for (int g = 0; g < ng; ++g) {
    // a multi-value group keeps one bin per feature (FeatureGroup::multi_bin_data_): nothing is overwritten there
    const int id = (g1[g].size() > 1 && !mv1[g]) ? ++next_bundle : 0;
    for (int f : g1[g]) { order.push_back(f); (*bundle_of_real)[f] = id; }
  }
  return order;
}

}  // namespace

Dataset* Dataset::FromDenseMatrix(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major,
                                  const Config& config) {
  if (nrow <= 0) Log::Fatal("Check failed: (num_data) > (0)");
  if (data_type != 0 && data_type != 1) Log::Fatal("Unknown data type in RowFunctionFromDenseMatric");
  const MatView M{data, data_type, nrow, ncol, row_major};
  std::unique_ptr<Dataset> ds(new Dataset(nrow));
  ds->num_total_features_ = ncol;
  DsTimer st;
  // sample rows (src/c_api.cpp:1045-1070)
  Random rand(config.data_random_seed);
  const int want = nrow < config.bin_construct_sample_cnt ? nrow : config.bin_construct_sample_cnt;
  const std::vector<int> sample = rand.Sample(nrow, want);
  const int sample_cnt = static_cast<int>(sample.size());
  st.tick("sample indices");
  std::vector<std::vector<double>> sample_vals(ncol);
  std::vector<std::vector<int>> sample_idx(ncol);
  // The sampled rows are first gathered into a compact [sample][column] block (each row of the user's matrix is
  // touched once, by one thread), then every column's lists are cut from that block on the host threads -- in sample
  // order, as the reference's row-by-row loop produces them.
  std::unique_ptr<double[]> block;   // (not value-initialised: its pages are first touched by the threads that fill them)
  if (row_major) {
    block.reset(new double[static_cast<size_t>(sample_cnt) * ncol]);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < sample_cnt; ++i) {
      double* dst = block.get() + static_cast<size_t>(i) * ncol;
      const size_t base = static_cast<size_t>(sample[i]) * ncol;
      if (data_type == 0) { const float* src = static_cast<const float*>(data) + base; for (int c = 0; c < ncol; ++c) dst[c] = static_cast<double>(src[c]); }
      else { const double* src = static_cast<const double*>(data) + base; for (int c = 0; c < ncol; ++c) dst[c] = src[c]; }
    }
  }
  st.tick("sample gather");
  if (row_major) {
    // Every thread takes a contiguous range of the samples and walks its rows of the block once, all columns (a column
    // at a time over the whole block reads every cache line of it once per column: 30 ms instead of 4 at 200K x 28):
    // first the non-zero counts per (thread, column), then each thread writes its entries behind those of the
    // threads before it -- the lists come out in sample order, as the reference's row-by-row loop produces them.
    const int T = std::max(1, omp_get_max_threads());
    std::vector<int> cnt(static_cast<size_t>(T + 1) * ncol, 0);
    auto keep = [](double v) { return std::fabs(v) > kZeroThreshold || std::isnan(v); };
#pragma omp parallel num_threads(T)
    {
      const int t = omp_get_thread_num(), nt = omp_get_num_threads();   // (the runtime may grant fewer than T threads)
      const int i0 = static_cast<int>(static_cast<int64_t>(sample_cnt) * t / nt), i1 = static_cast<int>(static_cast<int64_t>(sample_cnt) * (t + 1) / nt);
      int* my = cnt.data() + static_cast<size_t>(t + 1) * ncol;
      for (int i = i0; i < i1; ++i) {
        const double* row = block.get() + static_cast<size_t>(i) * ncol;
        for (int c = 0; c < ncol; ++c) my[c] += keep(row[c]) ? 1 : 0;
      }
#pragma omp barrier
#pragma omp for schedule(static)
      for (int c = 0; c < ncol; ++c) {   // exclusive prefix over the threads, per column; the lists get their final size
        int run = 0;
        for (int k = 1; k <= nt; ++k) { const int v = cnt[static_cast<size_t>(k) * ncol + c]; cnt[static_cast<size_t>(k) * ncol + c] = run; run += v; }
        sample_vals[c].resize(run); sample_idx[c].resize(run);
      }
      // (implicit barrier of the omp for)
      std::vector<int> pos(my, my + ncol);
      for (int i = i0; i < i1; ++i) {
        const double* row = block.get() + static_cast<size_t>(i) * ncol;
        for (int c = 0; c < ncol; ++c) {
          if (keep(row[c])) { sample_vals[c][pos[c]] = row[c]; sample_idx[c][pos[c]] = i; ++pos[c]; }
        }
      }
    }
  } else {
#pragma omp parallel for schedule(dynamic)
    for (int c = 0; c < ncol; ++c) {
      sample_vals[c].reserve(sample_cnt); sample_idx[c].reserve(sample_cnt);
      for (int i = 0; i < sample_cnt; ++i) {
        const double v = M.at(sample[i], c);
        if (std::fabs(v) > kZeroThreshold || std::isnan(v)) {
          sample_vals[c].push_back(v);
          sample_idx[c].push_back(i);
        }
      }
    }
  }
  st.tick("sample rows");
  // bin mappers (DatasetLoader::ConstructFromSampleData, src/io/dataset_loader.cpp:527-585)
  const data_size_t filter_cnt = static_cast<data_size_t>(static_cast<double>(config.min_data_in_leaf * static_cast<size_t>(sample_cnt)) / nrow);
  if (!config.max_bin_by_feature.empty() && static_cast<int>(config.max_bin_by_feature.size()) != ncol)
    Log::Fatal("Check failed: (static_cast<size_t>(num_col)) == (config_.max_bin_by_feature.size())");   // dataset_loader.cpp:534-537
  std::vector<std::shared_ptr<BinMapper>> all(ncol);
#pragma omp parallel for schedule(guided)
  for (int c = 0; c < ncol; ++c) {
    all[c].reset(new BinMapper());
    const bool categorical = std::find(config.categorical_feature.begin(), config.categorical_feature.end(), c) != config.categorical_feature.end();
    all[c]->FindBin(sample_vals[c], static_cast<size_t>(sample_cnt), config.max_bin_by_feature.empty() ? config.max_bin : config.max_bin_by_feature[c],
                    config.min_data_in_bin, filter_cnt,
                    config.feature_pre_filter, config.use_missing, config.zero_as_missing, categorical);
  }
  st.tick("FindBin");
  std::vector<int> used;
  for (int c = 0; c < ncol; ++c) if (!all[c]->is_trivial()) used.push_back(c);
  if (used.empty()) Log::Warning("There are no meaningful features, as all feature values are constant.");
  std::vector<int> order = used;
  std::vector<int> bundle_of_real(ncol, 0);
  if (config.enable_bundle && !used.empty()) {
    order = BundledFeatureOrder(all, used, sample_vals, sample_idx, sample_cnt, nrow, &bundle_of_real);
  }
  ds->real_to_inner_.assign(ncol, -1);
  for (int f : order) {
    if (all[f]->num_bin() + (all[f]->most_freq_bin() == 0 ? 0 : 1) > 65536) {
      Log::Fatal("feature %d needs more than 65536 stored values; at most 16-bit bins are built", f);
    }
    ds->real_to_inner_[f] = static_cast<int>(ds->inner_to_real_.size());
    ds->inner_to_real_.push_back(f);
    ds->bundle_of_.push_back(bundle_of_real[f]);
    ds->mappers_.push_back(all[f]);
  }
  for (int c = 0; c < ncol; ++c) ds->feature_names_.push_back("Column_" + std::to_string(c));
  st.tick("feature order");
  ds->BinAllRows(data, data_type, nrow, ncol, row_major, config.gpu_device_id);
  st.tick("bin all rows");
  return ds.release();
}

Dataset* Dataset::FromDenseMatrixAligned(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major,
                                         const Dataset& ref, const Config& config) {
  if (ncol != ref.num_total_features_) Log::Fatal("The number of features in data (%d) is not the same as it was in training data (%d).", ncol, ref.num_total_features_);
  std::unique_ptr<Dataset> ds(new Dataset(nrow));
  ds->num_total_features_ = ncol;
  ds->mappers_ = ref.mappers_;
  ds->inner_to_real_ = ref.inner_to_real_;
  // (no bundles: Dataset::CreateValid, src/io/dataset.cpp:710-728, gives every feature of an aligned dataset a group of
  // its own, so nothing is overwritten there)
  ds->real_to_inner_ = ref.real_to_inner_;
  ds->feature_names_ = ref.feature_names_;
  ds->BinAllRows(data, data_type, nrow, ncol, row_major, config.gpu_device_id);
  return ds.release();
}

Dataset* Dataset::Subset(const Dataset& full, const int32_t* idx, int32_t n) {
  if (n <= 0) Log::Fatal("Check failed: (num_used_row_indices) > (0)");
  for (int32_t i = 0; i < n; ++i) {
    if (idx[i] < 0 || idx[i] >= full.num_data_) Log::Fatal("Used indices of subset: element [#%d = %d] outside [0, %d]", i, idx[i], full.num_data_ - 1);
    if (i > 0 && idx[i] < idx[i - 1]) Log::Fatal("used_row_indices should be sorted in Subset");
  }
  std::unique_ptr<Dataset> ds(new Dataset(n));
  ds->num_total_features_ = full.num_total_features_;
  ds->mappers_ = full.mappers_;
  ds->inner_to_real_ = full.inner_to_real_;
  ds->real_to_inner_ = full.real_to_inner_;
  ds->bundle_of_ = full.bundle_of_;
  ds->feature_names_ = full.feature_names_;
  // the stored values travel as they are (a bundle's overwritten rows included): one row = FS bytes
  const size_t FS = static_cast<size_t>(std::max(full.num_storage_columns(), 1));
  const std::vector<uint8_t>& src = full.bins();
  ds->bins_.resize(static_cast<size_t>(n) * FS);
#pragma omp parallel for schedule(static)
  for (int32_t i = 0; i < n; ++i) std::memcpy(ds->bins_.data() + static_cast<size_t>(i) * FS, src.data() + static_cast<size_t>(idx[i]) * FS, FS);
  ds->bins_pinned_ = lgbcu_pin_host_buffer(ds->bins_.data(), ds->bins_.size()) == 0;
  // Metadata::Init(fullset, used_indices) (src/io/metadata.cpp:58-125).  (Its query loop runs over the NEW object's
  // query count, which is still zero there: a subset carries no queries until the caller sets them -- kept.)
  const Metadata& fm = full.meta_;
  if (!fm.label.empty()) { ds->meta_.label.resize(n); for (int32_t i = 0; i < n; ++i) ds->meta_.label[i] = fm.label[idx[i]]; }
  if (!fm.weights.empty()) { ds->meta_.weights.resize(n); for (int32_t i = 0; i < n; ++i) ds->meta_.weights[i] = fm.weights[idx[i]]; }
  if (!fm.init_score.empty()) {
    const int K = static_cast<int>(fm.init_score.size() / static_cast<size_t>(full.num_data_));
    ds->meta_.init_score.resize(static_cast<size_t>(n) * K);
    for (int k = 0; k < K; ++k)
      for (int32_t i = 0; i < n; ++i) ds->meta_.init_score[static_cast<size_t>(k) * n + i] = fm.init_score[static_cast<size_t>(k) * full.num_data_ + idx[i]];
  }
  return ds.release();
}

uint32_t Dataset::stored_value(int inner, double v) const {
  const BinMapper& m = *mappers_[inner];
  uint32_t bin = m.ValueToBin(v);
  if (bin == m.most_freq_bin()) return 0;       // FeatureGroup::PushData (feature_group.h:153-165)
  if (m.most_freq_bin() == 0) bin -= 1;
  return bin + 1;
}

// The row push loop of LGBM_DatasetCreateFromMat (src/c_api.cpp:1082-1097).  With a CUDA device the N x F map runs
// there (lgbcu_bin_matrix: the reference's ValueToBin search per element, bit-identical) and the binned matrix stays
// in HBM for the learner to adopt; the host copy is fetched only if somebody asks for it (bins()).  Without any
// device a Dataset can still be built -- the same map on the host threads -- but nothing can be trained with it:
// LGBM_BoosterCreate fails ("no CUDA device available").  LGBM_B200_HOST_BINNING=1 forces the host loop (tests
// compare the two).
void Dataset::BinAllRows(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major, int device) {
  const int F = num_features();
  int ndev = 0;
  lgbcu_device_count(&ndev);
  const char* force_host = std::getenv("LGBM_B200_HOST_BINNING");
  if (ndev > 0 && F > 0 && !(force_host && std::atoi(force_host) != 0)) {
    std::vector<LgbCuBinFeature> feats(F);
    std::vector<double> bounds;
    for (int f = 0; f < F; ++f) {
      const BinMapper& m = *mappers_[f];
      feats[f].column = inner_to_real_[f]; feats[f].num_bin = m.num_bin();
      feats[f].most_freq_bin = static_cast<int>(m.most_freq_bin()); feats[f].missing_type = m.missing_type();
      feats[f].bound_offset = static_cast<int>(bounds.size()); feats[f].bundle = bundle_of(f);
      feats[f].num_categories = -1; feats[f].pad = 0;
      if (m.is_categorical()) {   // the categories in ascending order, then their bins (bin 0's pseudo-category -1 is left out)
        std::vector<std::pair<int, int>> cb;
        for (int b = 1; b < m.num_bin(); ++b) cb.emplace_back(m.bin_2_categorical()[b], b);
        std::sort(cb.begin(), cb.end());
        feats[f].num_categories = static_cast<int>(cb.size());
        for (auto& e : cb) bounds.push_back(static_cast<double>(e.first));
        for (auto& e : cb) bounds.push_back(static_cast<double>(e.second));
        continue;
      }
      bounds.insert(bounds.end(), m.upper_bounds().begin(), m.upper_bounds().end());
    }
    int dev = device >= 0 ? device : 0;
    if (dev >= ndev) Log::Fatal("gpu_device_id %d out of range (have %d devices)", dev, ndev);
    if (lgbcu_bin_matrix(dev, data, data_type, nrow, ncol, row_major ? 1 : 0, F, feats.data(), bounds.data(),
                         static_cast<int>(bounds.size()), &dev_bins_) != 0) {
      Log::Fatal("%s", lgbcu_last_error());
    }
    return;
  }
  const MatView M{data, data_type, nrow, ncol, row_major};
  // byte matrix: one column per feature, two (low byte first) per feature with more than 256 stored values -- the
  // layout lgbcu_set_dataset / lgbcu_bin_matrix use
  std::vector<int> scol(F), wide(F);
  int FS = 0;
  for (int f = 0; f < F; ++f) { wide[f] = stored_values(f) > 256; scol[f] = FS; FS += wide[f] ? 2 : 1; }
  bins_.assign(static_cast<size_t>(nrow) * std::max(FS, 1), 0);
  std::vector<std::vector<int>> bundles;
  for (int f = 0; f < F; ++f) {
    const int b = bundle_of(f);
    if (b <= 0) continue;
    if (static_cast<int>(bundles.size()) < b) bundles.resize(b);
    bundles[b - 1].push_back(f);
  }
  for (auto& g : bundles) std::sort(g.begin(), g.end(), [&](int a, int b) { return inner_to_real_[a] > inner_to_real_[b]; });
#pragma omp parallel for schedule(static)
  for (int32_t r = 0; r < nrow; ++r) {
    uint8_t* out = bins_.data() + static_cast<size_t>(r) * FS;
    for (int f = 0; f < F; ++f) {
      const uint32_t v = stored_value(f, M.at(r, inner_to_real_[f]));
      out[scol[f]] = static_cast<uint8_t>(v & 255u);
      if (wide[f]) out[scol[f] + 1] = static_cast<uint8_t>(v >> 8);
    }
    for (const auto& g : bundles) {   // members by descending raw column: the last push of the reference's row loop wins
      bool seen = false;
      for (int f : g) {
        const int v = out[scol[f]] | (wide[f] ? out[scol[f] + 1] << 8 : 0);
        if (v == 0) continue;
        if (seen) { out[scol[f]] = 0; if (wide[f]) out[scol[f] + 1] = 0; }
        seen = true;
      }
    }
  }
  // the matrix only exists to be uploaded: page-lock it (a no-op without a device)
  bins_pinned_ = lgbcu_pin_host_buffer(bins_.data(), bins_.size()) == 0;
}

const std::vector<uint8_t>& Dataset::bins() const {
  if (bins_.empty() && dev_bins_ != nullptr) {
    bins_.assign(static_cast<size_t>(num_data_) * std::max(num_storage_columns(), 1), 0);
    if (lgbcu_binned_copy_to_host(dev_bins_, bins_.data()) != 0) Log::Fatal("%s", lgbcu_last_error());
  }
  return bins_;
}

Dataset::~Dataset() {
  if (label_pinned_) lgbcu_unpin_host_buffer(meta_.label.data());
  if (bins_pinned_) lgbcu_unpin_host_buffer(bins_.data());
  if (dev_bins_) lgbcu_binned_free(dev_bins_);
}

void Dataset::set_feature_names(const std::vector<std::string>& names) {
  if (static_cast<int>(names.size()) != num_total_features_) Log::Fatal("Size of feature_names error, should equal with total number of features");
  feature_names_ = names;
  for (std::string& n : feature_names_) {   // spaces would break the model text (the reference replaces them and warns)
    for (char& ch : n) if (ch == ' ') ch = '_';
  }
}

uint64_t Dataset::num_total_bin() const {
  uint64_t t = 0;
  for (auto& m : mappers_) t += 1 + m->num_bin() - (m->most_freq_bin() == 0 ? 1 : 0);
  return t;
}

std::vector<std::string> Dataset::feature_infos() const {
  std::vector<std::string> out;
  for (int c = 0; c < num_total_features_; ++c) {
    const int f = real_to_inner_[c];
    out.push_back(f < 0 ? "none" : mappers_[f]->info_string());
  }
  return out;
}

// LGBM_DatasetSetField (src/c_api.cpp:1422-1439 -> Metadata::Set*)
void Dataset::SetFloatField(const char* name, const float* data, int n) {
  const std::string k(name);
  if (k == "label" || k == "target") {
    if (n != num_data_) Log::Fatal("Length of label is not same with #data");
    if (label_pinned_) { lgbcu_unpin_host_buffer(meta_.label.data()); label_pinned_ = false; }
    meta_.label.assign(data, data + n);
    // uploaded at every LGBM_BoosterCreate: page-lock it like the bin matrix (a no-op without a device)
    label_pinned_ = lgbcu_pin_host_buffer(meta_.label.data(), meta_.label.size() * sizeof(label_t)) == 0;
  } else if (k == "weight" || k == "weights") {
    if (data == nullptr || n == 0) { meta_.weights.clear(); return; }
    if (n != num_data_) Log::Fatal("Length of weights is not same with #data");
    meta_.weights.assign(data, data + n);
  } else {
    Log::Fatal("Input field %s is not a float field", name);
  }
}

void Dataset::SetDoubleField(const char* name, const double* data, int n) {
  const std::string k(name);
  if (k == "init_score") {
    if (data == nullptr || n == 0) { meta_.init_score.clear(); return; }
    if (n != num_data_) Log::Fatal("Initial score size doesn't match data size");
    meta_.init_score.assign(data, data + n);
  } else {
    Log::Fatal("Input field %s is not a double field", name);
  }
}

void Dataset::SetIntField(const char* name, const int32_t* data, int n) {
  const std::string k(name);
  if (k == "query" || k == "group") {
    if (data == nullptr || n == 0) { meta_.query_boundaries.clear(); return; }
    long long sum = 0;
    for (int i = 0; i < n; ++i) sum += data[i];
    if (sum != num_data_) Log::Fatal("Sum of query counts is not same with #data");
    meta_.query_boundaries.assign(n + 1, 0);
    for (int i = 0; i < n; ++i) meta_.query_boundaries[i + 1] = meta_.query_boundaries[i] + data[i];
  } else {
    Log::Fatal("Input field %s is not an int field", name);
  }
}

}  // namespace lgb
