#include "tree.h"

#include <cstdlib>
#include <memory>
#include <unordered_map>

namespace lgb {

static inline double MaybeRoundToZero(double v) { return (v >= -kZeroThreshold && v <= kZeroThreshold) ? 0 : v; }

Tree::Tree(int max_leaves) : max_leaves_(max_leaves), num_leaves_(1), shrinkage_(1.0f) {
  const int n = std::max(max_leaves, 2);
  left_child_.resize(n - 1); right_child_.resize(n - 1);
  split_feature_inner_.resize(n - 1); split_feature_.resize(n - 1);
  threshold_in_bin_.resize(n - 1); threshold_.resize(n - 1);
  decision_type_.assign(n - 1, 0); split_gain_.resize(n - 1);
  leaf_parent_.resize(n); leaf_value_.resize(n); leaf_weight_.resize(n); leaf_count_.resize(n);
  internal_value_.resize(n - 1); internal_weight_.resize(n - 1); internal_count_.resize(n - 1);
  leaf_depth_.resize(n);
  leaf_depth_[0] = 0; leaf_value_[0] = 0.0f; leaf_weight_[0] = 0.0f; leaf_parent_[0] = -1;
}

int Tree::Split(int leaf, int feature_inner, int feature_real, uint32_t threshold_bin, double threshold_double,
                double left_value, double right_value, int left_cnt, int right_cnt, double left_weight,
                double right_weight, float gain, int missing_type, bool default_left) {
  const int node = num_leaves_ - 1;
  const int parent = leaf_parent_[leaf];
  if (parent >= 0) {
    if (left_child_[parent] == ~leaf) left_child_[parent] = node;
    else right_child_[parent] = node;
  }
  split_feature_inner_[node] = feature_inner;
  split_feature_[node] = feature_real;
  split_gain_[node] = gain;
  left_child_[node] = ~leaf;
  right_child_[node] = ~num_leaves_;
  leaf_parent_[leaf] = node;
  leaf_parent_[num_leaves_] = node;
  internal_weight_[node] = leaf_weight_[leaf];
  internal_value_[node] = leaf_value_[leaf];
  internal_count_[node] = left_cnt + right_cnt;
  leaf_value_[leaf] = std::isnan(left_value) ? 0.0f : left_value;
  leaf_weight_[leaf] = left_weight;
  leaf_count_[leaf] = left_cnt;
  leaf_value_[num_leaves_] = std::isnan(right_value) ? 0.0f : right_value;
  leaf_weight_[num_leaves_] = right_weight;
  leaf_count_[num_leaves_] = right_cnt;
  leaf_depth_[num_leaves_] = leaf_depth_[leaf] + 1;
  leaf_depth_[leaf]++;
  // decision_type: bit0 categorical, bit1 default_left, bits 2-3 missing type (tree.h SetMissingType)
  int8_t dt = 0;
  if (default_left) dt |= kDefaultLeftMask;
  dt = static_cast<int8_t>((dt & 3) | (missing_type << 2));
  decision_type_[node] = dt;
  threshold_in_bin_[node] = threshold_bin;
  threshold_[node] = threshold_double;
  ++num_leaves_;
  return num_leaves_ - 1;
}

int Tree::SplitCategorical(int leaf, int feature_inner, int feature_real, const uint32_t* bin_set, const std::vector<uint32_t>& cat_set,
                           double left_value, double right_value, int left_cnt, int right_cnt, double left_weight,
                           double right_weight, float gain, int missing_type) {
  const int node = num_leaves_ - 1;
  const int right = Split(leaf, feature_inner, feature_real, static_cast<uint32_t>(num_cat_), static_cast<double>(num_cat_),
                          left_value, right_value, left_cnt, right_cnt, left_weight, right_weight, gain, missing_type, false);
  decision_type_[node] = static_cast<int8_t>(kCategoricalMask | (missing_type << 2));
  ++num_cat_;
  cat_boundaries_.push_back(cat_boundaries_.back() + static_cast<int>(cat_set.size()));
  cat_threshold_.insert(cat_threshold_.end(), cat_set.begin(), cat_set.end());
  cat_bin_sets_.insert(cat_bin_sets_.end(), bin_set, bin_set + 8);
  return right;
}

void Tree::Shrinkage(double rate) {
  for (int i = 0; i < num_leaves_ - 1; ++i) {
    leaf_value_[i] = MaybeRoundToZero(leaf_value_[i] * rate);
    internal_value_[i] = MaybeRoundToZero(internal_value_[i] * rate);
  }
  leaf_value_[num_leaves_ - 1] = MaybeRoundToZero(leaf_value_[num_leaves_ - 1] * rate);
  shrinkage_ *= rate;
}

void Tree::AddBias(double val) {
  for (int i = 0; i < num_leaves_ - 1; ++i) {
    leaf_value_[i] = MaybeRoundToZero(leaf_value_[i] + val);
    internal_value_[i] = MaybeRoundToZero(internal_value_[i] + val);
  }
  leaf_value_[num_leaves_ - 1] = MaybeRoundToZero(leaf_value_[num_leaves_ - 1] + val);
  shrinkage_ = 1.0f;
}

std::string Tree::ToString() const {
  std::ostringstream o;
  const size_t ni = static_cast<size_t>(num_leaves_ - 1), nl = static_cast<size_t>(num_leaves_);
  std::vector<int> dt(decision_type_.begin(), decision_type_.end());
  o << "num_leaves=" << num_leaves_ << '\n';
  o << "num_cat=" << num_cat_ << '\n';
  o << "split_feature=" << JoinInts(split_feature_, ni) << '\n';
  o << "split_gain=" << JoinG(split_gain_, ni) << '\n';
  o << "threshold=" << Join17(threshold_, ni) << '\n';
  o << "decision_type=" << JoinInts(dt, ni) << '\n';
  o << "left_child=" << JoinInts(left_child_, ni) << '\n';
  o << "right_child=" << JoinInts(right_child_, ni) << '\n';
  o << "leaf_value=" << Join17(leaf_value_, nl) << '\n';
  o << "leaf_weight=" << Join17(leaf_weight_, nl) << '\n';
  o << "leaf_count=" << JoinInts(leaf_count_, nl) << '\n';
  o << "internal_value=" << JoinG(internal_value_, ni) << '\n';
  o << "internal_weight=" << JoinG(internal_weight_, ni) << '\n';
  o << "internal_count=" << JoinInts(internal_count_, ni) << '\n';
  if (num_cat_ > 0) {
    std::vector<int> ct(cat_threshold_.size());
    o << "cat_boundaries=" << JoinInts(cat_boundaries_, static_cast<size_t>(num_cat_) + 1) << '\n';
    o << "cat_threshold=";
    for (size_t i = 0; i < cat_threshold_.size(); ++i) { if (i) o << ' '; o << cat_threshold_[i]; }
    o << '\n';
  }
  o << "shrinkage=" << shrinkage_ << '\n';
  o << '\n';
  return o.str();
}

Tree* Tree::FromString(const std::string& text) {
  std::unordered_map<std::string, std::string> kv;
  size_t pos = 0;
  while (pos < text.size()) {
    size_t e = text.find('\n', pos);
    if (e == std::string::npos) e = text.size();
    const std::string line = text.substr(pos, e - pos);
    pos = e + 1;
    const size_t eq = line.find('=');
    if (eq != std::string::npos) kv[line.substr(0, eq)] = line.substr(eq + 1);
  }
  auto need = [&](const char* k) -> const std::string& {
    auto it = kv.find(k);
    if (it == kv.end()) Log::Fatal("Tree model string format error, should contain %s field", k);
    return it->second;
  };
  const int nl = std::atoi(need("num_leaves").c_str());
  std::unique_ptr<Tree> t(new Tree(std::max(nl, 2)));
  t->num_leaves_ = nl;
  t->num_cat_ = kv.count("num_cat") ? std::atoi(kv["num_cat"].c_str()) : 0;
  if (t->num_cat_ > 0) {   // src/io/tree.cpp:608-623
    const std::string& b = need("cat_boundaries");
    const std::string& c = need("cat_threshold");
    t->cat_boundaries_.assign(static_cast<size_t>(t->num_cat_) + 1, 0);
    const char* p = b.c_str(); char* end = nullptr;
    for (int i = 0; i <= t->num_cat_; ++i) { t->cat_boundaries_[i] = static_cast<int>(std::strtol(p, &end, 10)); p = end; }
    t->cat_threshold_.assign(static_cast<size_t>(t->cat_boundaries_.back()), 0u);
    p = c.c_str();
    for (size_t i = 0; i < t->cat_threshold_.size(); ++i) { t->cat_threshold_[i] = static_cast<uint32_t>(std::strtoul(p, &end, 10)); p = end; }
  }
  auto doubles = [](const std::string& s, size_t n, std::vector<double>* out) {
    const char* p = s.c_str(); char* end = nullptr;
    for (size_t i = 0; i < n; ++i) { (*out)[i] = std::strtod(p, &end); p = end; }
  };
  auto ints = [](const std::string& s, size_t n, std::vector<int>* out) {
    const char* p = s.c_str(); char* end = nullptr;
    for (size_t i = 0; i < n; ++i) { (*out)[i] = static_cast<int>(std::strtol(p, &end, 10)); p = end; }
  };
  const size_t ni = static_cast<size_t>(std::max(nl - 1, 0)), nlz = static_cast<size_t>(nl);
  doubles(need("leaf_value"), nlz, &t->leaf_value_);
  if (nl > 1) {
    ints(need("split_feature"), ni, &t->split_feature_);
    t->split_feature_inner_ = t->split_feature_;
    doubles(need("threshold"), ni, &t->threshold_);
    std::vector<int> dt(ni); ints(need("decision_type"), ni, &dt);
    for (size_t i = 0; i < ni; ++i) t->decision_type_[i] = static_cast<int8_t>(dt[i]);
    ints(need("left_child"), ni, &t->left_child_);
    ints(need("right_child"), ni, &t->right_child_);
    std::vector<double> tmp(ni);
    if (kv.count("split_gain")) { doubles(kv["split_gain"], ni, &tmp); for (size_t i = 0; i < ni; ++i) t->split_gain_[i] = static_cast<float>(tmp[i]); }
    if (kv.count("leaf_weight")) doubles(kv["leaf_weight"], nlz, &t->leaf_weight_);
    if (kv.count("leaf_count")) ints(kv["leaf_count"], nlz, &t->leaf_count_);
    if (kv.count("internal_value")) doubles(kv["internal_value"], ni, &t->internal_value_);
    if (kv.count("internal_weight")) doubles(kv["internal_weight"], ni, &t->internal_weight_);
    if (kv.count("internal_count")) ints(kv["internal_count"], ni, &t->internal_count_);
  }
  if (kv.count("shrinkage")) t->shrinkage_ = std::strtod(kv["shrinkage"].c_str(), nullptr);
  return t.release();
}

// Tree::Predict -> GetLeaf -> NumericalDecision (include/LightGBM/tree.h:250-283)
double Tree::Predict(const double* x) const { return leaf_value_[PredictLeaf(x)]; }

int Tree::PredictLeaf(const double* x) const {
  if (num_leaves_ <= 1) return 0;
  int node = 0;
  while (node >= 0) {
    double fval = x[split_feature_[node]];
    const int missing_type = (decision_type_[node] >> 2) & 3;
    if (decision_type_[node] & kCategoricalMask) {   // Tree::CategoricalDecision (tree.h:302-320)
      // (static_cast<int> of a NaN or of a value outside the int range is undefined in the reference; on x86-64 it
      // yields INT_MIN, i.e. "negative -> right", which is what is done here)
      int iv = (fval > -2147483649.0 && fval < 2147483648.0) ? static_cast<int>(fval) : -1;
      bool left = false;
      if (iv >= 0) {
        const int ci = static_cast<int>(threshold_[node]);
        const int n = cat_boundaries_[ci + 1] - cat_boundaries_[ci];
        const int i1 = iv / 32;
        if (i1 < n) left = ((cat_threshold_[cat_boundaries_[ci] + i1] >> (iv % 32)) & 1u) != 0;
      }
      node = left ? left_child_[node] : right_child_[node];
      continue;
    }
    const bool default_left = (decision_type_[node] & kDefaultLeftMask) != 0;
    if (std::isnan(fval) && missing_type != kMissingNaN) fval = 0.0f;
    if ((missing_type == kMissingZero && (fval >= -kZeroThreshold && fval <= kZeroThreshold)) ||
        (missing_type == kMissingNaN && std::isnan(fval))) {
      node = default_left ? left_child_[node] : right_child_[node];
    } else {
      node = fval <= threshold_[node] ? left_child_[node] : right_child_[node];
    }
  }
  return ~node;
}

}  // namespace lgb
