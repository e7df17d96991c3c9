#include "objective.h"

#include <algorithm>
#include <cmath>
#include <functional>
#include <limits>
#include <memory>
#include <sstream>
#include <vector>

namespace lgb {
namespace {

void Check(int rc) {
  if (rc != 0) Log::Fatal("%s", lgbcu_last_error());
}

// src/objective/binary_objective.hpp
class BinaryLogloss : public Objective {
 public:
  // pos_class < 0: a row is positive when label > 0; otherwise when (int)label == pos_class (MulticlassOVA)
  explicit BinaryLogloss(const Config& c, int pos_class = -1) : sigmoid_(c.sigmoid), is_unbalance_(c.is_unbalance), scale_pos_weight_(c.scale_pos_weight), pos_class_(pos_class) {
    if (sigmoid_ <= 0.0) Log::Fatal("Sigmoid parameter %f should be greater than zero", sigmoid_);
    if (is_unbalance_ && std::fabs(scale_pos_weight_ - 1.0f) > 1e-6) Log::Fatal("Cannot set is_unbalance and scale_pos_weight at the same time");
  }
  data_size_t NumPositiveData() const override { return num_pos_; }
  void Init(const Metadata& meta, data_size_t n, const SumSync& sum_over_ranks) override {
    num_data_ = n; label_ = meta.label.data(); weights_ = meta.weights.empty() ? nullptr : meta.weights.data();
    if (meta.label.empty()) Log::Fatal("Label should not be NULL");
    long long pos = 0, neg = 0;
#pragma omp parallel for schedule(static) reduction(+:pos, neg)
    for (data_size_t i = 0; i < n; ++i) { if (IsPos(label_[i])) ++pos; else ++neg; }
    num_pos_ = static_cast<data_size_t>(pos);
    {  // binary_objective.hpp:72-75: the class counts of ALL machines decide need_train_ and the label weights
      double cnt[2] = {static_cast<double>(pos), static_cast<double>(neg)};  // counts < 2^53: exact as doubles
      sum_over_ranks(cnt, 2);
      pos = static_cast<long long>(cnt[0]); neg = static_cast<long long>(cnt[1]);
    }
    need_train_ = !(neg == 0 || pos == 0);
    if (!need_train_) Log::Warning("Contains only one class");
    Log::Info("Number of positive: %lld, number of negative: %lld", pos, neg);
    label_weights_[0] = label_weights_[1] = 1.0f;
    if (is_unbalance_ && pos > 0 && neg > 0) {
      if (pos > neg) { label_weights_[1] = 1.0f; label_weights_[0] = static_cast<double>(pos) / neg; }
      else { label_weights_[1] = static_cast<double>(neg) / pos; label_weights_[0] = 1.0f; }
    }
    label_weights_[1] *= scale_pos_weight_;
  }
  void Configure(LgbCuContext* ctx) const override {
    Check(lgbcu_set_objective(ctx, LGBCU_OBJ_BINARY, sigmoid_, label_weights_[0], label_weights_[1]));
  }
  double BoostFromScore(int) const override {
    double suml = 0.0f, sumw = 0.0f;
    if (weights_) {  // (the reference reduces over OpenMP threads too: binary_objective.hpp BoostFromScore)
#pragma omp parallel for schedule(static) reduction(+:suml, sumw)
      for (data_size_t i = 0; i < num_data_; ++i) { suml += IsPos(label_[i]) * weights_[i]; sumw += weights_[i]; }
    } else {
      sumw = static_cast<double>(num_data_);
      suml = static_cast<double>(num_pos_);  // a count: exact in any order
    }
    double pavg = suml / sumw;
    pavg = std::min(pavg, 1.0 - kEpsilon);
    pavg = std::max<double>(pavg, kEpsilon);
    const double init = std::log(pavg / (1.0f - pavg)) / sigmoid_;
    Log::Info("[binary:BoostFromScore]: pavg=%f -> initscore=%f", pavg, init);
    return init;
  }
  bool NeedTrain(int) const override { return need_train_; }
  double label_weight(int pos) const { return label_weights_[pos]; }
  void ConvertOutput(const double* in, double* out) const override { out[0] = 1.0f / (1.0f + std::exp(-sigmoid_ * in[0])); }
  std::string ToString() const override { std::ostringstream o; o << "binary sigmoid:" << sigmoid_; return o.str(); }
  const char* Name() const override { return "binary"; }

 private:
  bool IsPos(label_t l) const { return pos_class_ < 0 ? l > 0 : static_cast<int>(l) == pos_class_; }
  double sigmoid_; bool is_unbalance_; double scale_pos_weight_; int pos_class_;
  data_size_t num_data_ = 0, num_pos_ = 0; const label_t* label_ = nullptr; const label_t* weights_ = nullptr;
  bool need_train_ = true; double label_weights_[2] = {1, 1};
};

// src/objective/regression_objective.hpp (RegressionL2loss)
class RegressionL2 : public Objective {
 public:
  explicit RegressionL2(const Config& c) : sqrt_(c.reg_sqrt) {}
  void Init(const Metadata& meta, data_size_t n, const SumSync&) override {
    num_data_ = n; label_ = meta.label.data(); weights_ = meta.weights.empty() ? nullptr : meta.weights.data();
    if (meta.label.empty()) Log::Fatal("Label should not be NULL");
    if (sqrt_) {   // regression_objective.hpp:110-118 (the device's gradient kernel applies the same transform)
      trans_label_.resize(n);
      for (data_size_t i = 0; i < n; ++i) trans_label_[i] = ((label_[i] > 0.0f) - (label_[i] < 0.0f)) * std::sqrt(std::fabs(label_[i]));
      label_ = trans_label_.data();
    }
  }
  void Configure(LgbCuContext* ctx) const override { Check(lgbcu_set_objective(ctx, LGBCU_OBJ_L2, sqrt_ ? 1 : 0, 0, 0)); }
  void ConvertOutput(const double* in, double* out) const override {
    out[0] = sqrt_ ? ((in[0] > 0.0) - (in[0] < 0.0)) * in[0] * in[0] : in[0];
  }
  double BoostFromScore(int) const override {
    double suml = 0.0f, sumw = 0.0f;
    if (weights_) {  // (OpenMP reductions as in regression_objective.hpp BoostFromScore)
#pragma omp parallel for schedule(static) reduction(+:suml, sumw)
      for (data_size_t i = 0; i < num_data_; ++i) { suml += label_[i] * weights_[i]; sumw += weights_[i]; }
    } else {
      sumw = static_cast<double>(num_data_);
#pragma omp parallel for schedule(static) reduction(+:suml)
      for (data_size_t i = 0; i < num_data_; ++i) suml += label_[i];
    }
    return suml / sumw;
  }
  bool IsConstantHessian() const override { return weights_ == nullptr; }
  std::string ToString() const override { return sqrt_ ? "regression sqrt" : "regression"; }
  const char* Name() const override { return "regression"; }

 private:
  bool sqrt_;
  std::vector<label_t> trans_label_;
  data_size_t num_data_ = 0; const label_t* label_ = nullptr; const label_t* weights_ = nullptr;
};

// src/objective/rank_objective.hpp (LambdarankNDCG); label checks / max DCG / sigmoid table happen in lgbcu
class Lambdarank : public Objective {
 public:
  explicit Lambdarank(const Config& c)
      : sigmoid_(c.sigmoid), norm_(c.lambdarank_norm), trunc_(c.lambdarank_truncation_level), label_gain_(c.label_gain) {
    if (sigmoid_ <= 0.0) Log::Fatal("Sigmoid param %f should be greater than zero", sigmoid_);
  }
  void Init(const Metadata& meta, data_size_t, const SumSync&) override {
    if (meta.query_boundaries.empty()) Log::Fatal("Ranking tasks require query information");
  }
  void Configure(LgbCuContext* ctx) const override {
    Check(lgbcu_set_objective(ctx, LGBCU_OBJ_LAMBDARANK, sigmoid_, trunc_, norm_ ? 1 : 0));
    if (!label_gain_.empty()) Check(lgbcu_set_label_gain(ctx, label_gain_.data(), static_cast<int>(label_gain_.size())));
  }
  double BoostFromScore(int) const override { return 0.0; }
  std::string ToString() const override { return "lambdarank"; }
  const char* Name() const override { return "lambdarank"; }

 private:
  double sigmoid_; bool norm_; int trunc_; std::vector<double> label_gain_;
};

// src/objective/rank_objective.hpp:288-366 (RankXENDCG): cross-entropy ranking with randomised gains
class RankXendcg : public Objective {
 public:
  explicit RankXendcg(const Config& c) : seed_(c.objective_seed) {}
  void Init(const Metadata& meta, data_size_t, const SumSync&) override {
    if (meta.query_boundaries.empty()) Log::Fatal("Ranking tasks require query information");
  }
  void Configure(LgbCuContext* ctx) const override { Check(lgbcu_set_objective(ctx, LGBCU_OBJ_XENDCG, seed_, 0, 0)); }
  double BoostFromScore(int) const override { return 0.0; }
  std::string ToString() const override { return "rank_xendcg"; }
  const char* Name() const override { return "rank_xendcg"; }

 private:
  int seed_;
};

}  // namespace

// PercentileFun / WeightedPercentileFun (src/objective/regression_objective.hpp:18-88) over the labels, for the
// BoostFromScore of the percentile objectives; T = label_t as there.  `one_minus_alpha_times_cnt` is formed by the
// caller in the arithmetic its objective uses (double for L1's 0.5, float for the quantile objective's float alpha).
static double LabelPercentile(const label_t* label, data_size_t n, double float_pos) {
  if (n <= 1) return label[0];
  std::vector<label_t> v(label, label + n);
  const data_size_t pos = static_cast<data_size_t>(float_pos);
  if (pos < 1) return *std::max_element(v.begin(), v.end());
  if (pos >= n) return *std::min_element(v.begin(), v.end());
  const double bias = float_pos - pos;
  // the pos-th and (pos+1)-th largest values, whichever selection the reference's two branches run
  std::nth_element(v.begin(), v.begin() + (pos - 1), v.end(), std::greater<label_t>());
  const label_t v1 = v[pos - 1];
  const label_t v2 = *std::max_element(v.begin() + pos, v.end());
  return static_cast<label_t>(v1 - (v1 - v2) * bias);
}
static double LabelWeightedPercentile(const label_t* label, const label_t* weight, data_size_t n, double alpha) {
  if (n <= 1) return label[0];
  std::vector<data_size_t> idx(n);
  for (data_size_t i = 0; i < n; ++i) idx[i] = i;
  std::stable_sort(idx.begin(), idx.end(), [&](data_size_t a, data_size_t b) { return label[a] < label[b]; });
  std::vector<double> cdf(n);
  cdf[0] = weight[idx[0]];
  for (data_size_t i = 1; i < n; ++i) cdf[i] = cdf[i - 1] + weight[idx[i]];
  const double threshold = cdf[n - 1] * alpha;
  size_t pos = std::upper_bound(cdf.begin(), cdf.end(), threshold) - cdf.begin();
  pos = std::min(pos, static_cast<size_t>(n - 1));
  if (pos == 0 || pos == static_cast<size_t>(n - 1)) return label[idx[pos]];
  const label_t v1 = label[idx[pos - 1]], v2 = label[idx[pos]];
  if (cdf[pos + 1] - cdf[pos] >= 1.0f) return static_cast<label_t>((threshold - cdf[pos]) / (cdf[pos + 1] - cdf[pos]) * (v2 - v1) + v1);
  return static_cast<label_t>(v2);
}

// src/objective/regression_objective.hpp (RegressionL1loss :204-283, RegressionQuantileloss :479-571)
class RegressionPercentile : public Objective {
 public:
  RegressionPercentile(const Config& c, bool quantile) : quantile_(quantile), alpha_f_(static_cast<float>(c.alpha)) {
    if (c.reg_sqrt) Log::Fatal("reg_sqrt is outside this library's path");
    if (quantile_ && !(alpha_f_ > 0 && alpha_f_ < 1)) Log::Fatal("Check failed: alpha_ > 0 && alpha_ < 1");
  }
  void Init(const Metadata& meta, data_size_t n, const SumSync&) override {
    num_data_ = n; label_ = meta.label.data(); weights_ = meta.weights.empty() ? nullptr : meta.weights.data();
    if (meta.label.empty()) Log::Fatal("Label should not be NULL");
  }
  void Configure(LgbCuContext* ctx) const override {
    Check(lgbcu_set_objective(ctx, quantile_ ? LGBCU_OBJ_QUANTILE : LGBCU_OBJ_L1, quantile_ ? alpha_f_ : 0.5, 0, 0));
  }
  double BoostFromScore(int) const override {
    if (weights_) return LabelWeightedPercentile(label_, weights_, num_data_, quantile_ ? static_cast<double>(alpha_f_) : 0.5);
    const double float_pos = quantile_ ? static_cast<double>((1.0f - alpha_f_) * num_data_) : (1.0f - 0.5) * num_data_;
    return LabelPercentile(label_, num_data_, float_pos);
  }
  bool IsRenewTreeOutput() const override { return true; }
  double RenewAlpha() const override { return quantile_ ? static_cast<double>(alpha_f_) : 0.5; }
  bool RenewAlphaIsFloat() const override { return quantile_; }
  std::string ToString() const override { return quantile_ ? "quantile" : "regression_l1"; }
  const char* Name() const override { return quantile_ ? "quantile" : "regression_l1"; }

 private:
  bool quantile_; float alpha_f_;
  data_size_t num_data_ = 0; const label_t* label_ = nullptr; const label_t* weights_ = nullptr;
};

// src/objective/xentropy_objective.hpp:150-283 (CrossEntropyLambda): labels in [0, 1], weights inside the link
class CrossEntropyLambda : public Objective {
 public:
  explicit CrossEntropyLambda(const Config&) {}
  void Init(const Metadata& meta, data_size_t n, const SumSync&) override {
    num_data_ = n; label_ = meta.label.data(); weights_ = meta.weights.empty() ? nullptr : meta.weights.data();
    if (meta.label.empty()) Log::Fatal("Label should not be NULL");
    for (data_size_t i = 0; i < n; ++i)
      if (!(label_[i] >= 0.0f && label_[i] <= 1.0f)) Log::Fatal("[%s]: does not tolerate element [#%d = %f] outside [0, 1]", Name(), i, label_[i]);
    if (weights_) {
      label_t minw = weights_[0];
      for (data_size_t i = 0; i < n; ++i) minw = std::min(minw, weights_[i]);
      if (minw <= 0.0f) Log::Fatal("[%s]: at least one weight is non-positive", Name());
    }
  }
  void Configure(LgbCuContext* ctx) const override { Check(lgbcu_set_objective(ctx, LGBCU_OBJ_XENTLAMBDA, 0, 0, 0)); }
  double BoostFromScore(int) const override {
    double suml = 0.0f, sumw = 0.0f;
    if (weights_) {
#pragma omp parallel for schedule(static) reduction(+:suml, sumw)
      for (data_size_t i = 0; i < num_data_; ++i) { suml += label_[i] * weights_[i]; sumw += weights_[i]; }
    } else {
      sumw = static_cast<double>(num_data_);
#pragma omp parallel for schedule(static) reduction(+:suml)
      for (data_size_t i = 0; i < num_data_; ++i) suml += label_[i];
    }
    const double havg = suml / sumw;
    return std::log(std::exp(havg) - 1.0f);
  }
  void ConvertOutput(const double* in, double* out) const override { out[0] = std::log(1.0f + std::exp(in[0])); }
  std::string ToString() const override { return Name(); }
  const char* Name() const override { return "cross_entropy_lambda"; }

 private:
  data_size_t num_data_ = 0; const label_t* label_ = nullptr; const label_t* weights_ = nullptr;
};

// RegressionMAPELOSS (src/objective/regression_objective.hpp:577-669): L1 with label_weight_ = 1 / max(1, |label|)
class RegressionMAPE : public Objective {
 public:
  explicit RegressionMAPE(const Config& c) { if (c.reg_sqrt) Log::Fatal("reg_sqrt is outside this library's path"); }
  void Init(const Metadata& meta, data_size_t n, const SumSync&) override {
    num_data_ = n; label_ = meta.label.data();
    if (meta.label.empty()) Log::Fatal("Label should not be NULL");
    // with weights the reference calls the hessian constant although it holds the weights: not imitated
    if (!meta.weights.empty()) Log::Fatal("objective=mape with weights is outside this library's path");
    for (data_size_t i = 0; i < n; ++i) {
      if (std::fabs(label_[i]) < 1) { Log::Warning("Met 'abs(label) < 1', will convert them to '1' in MAPE objective and metric"); break; }
    }
    label_weight_.resize(n);
    for (data_size_t i = 0; i < n; ++i) label_weight_[i] = 1.0f / std::max(1.0f, std::fabs(label_[i]));
  }
  void Configure(LgbCuContext* ctx) const override { Check(lgbcu_set_objective(ctx, LGBCU_OBJ_MAPE, 0, 0, 0)); }
  double BoostFromScore(int) const override { return LabelWeightedPercentile(label_, label_weight_.data(), num_data_, 0.5); }
  bool IsRenewTreeOutput() const override { return true; }
  double RenewAlpha() const override { return 0.5; }
  bool RenewAlphaIsFloat() const override { return false; }
  std::string ToString() const override { return "mape"; }
  const char* Name() const override { return "mape"; }

 private:
  data_size_t num_data_ = 0; const label_t* label_ = nullptr; std::vector<label_t> label_weight_;
};

// src/objective/multiclass_objective.hpp (MulticlassSoftmax)
class MulticlassSoftmax : public Objective {
 public:
  explicit MulticlassSoftmax(const Config& c) : num_class_(c.num_class) {
    factor_ = static_cast<double>(num_class_) / (num_class_ - 1.0f);
  }
  void Init(const Metadata& meta, data_size_t n, const SumSync& sum_over_ranks) override {
    num_data_ = n;
    if (meta.label.empty()) Log::Fatal("Label should not be NULL");
    const label_t* label = meta.label.data();
    const label_t* weights = meta.weights.empty() ? nullptr : meta.weights.data();
    class_init_probs_.assign(num_class_, 0.0);
    double sum_weight = 0.0;
    for (data_size_t i = 0; i < n; ++i) {   // sequential like the reference: the order of the double sums is part of the result
      const int li = static_cast<int>(label[i]);
      if (li < 0 || li >= num_class_) Log::Fatal("Label must be in [0, %d), but found %d in label", num_class_, li);
      if (weights == nullptr) class_init_probs_[li] += 1.0;
      else { class_init_probs_[li] += weights[i]; sum_weight += weights[i]; }
    }
    if (weights == nullptr) sum_weight = n;
    {  // Network::GlobalSyncUpBySum of the total and of every class
      std::vector<double> v(class_init_probs_);
      v.push_back(sum_weight);
      sum_over_ranks(v.data(), static_cast<int>(v.size()));
      sum_weight = v.back();
      for (int k = 0; k < num_class_; ++k) class_init_probs_[k] = v[k];
    }
    for (int k = 0; k < num_class_; ++k) class_init_probs_[k] /= sum_weight;
  }
  void Configure(LgbCuContext* ctx) const override {
    Check(lgbcu_set_objective(ctx, LGBCU_OBJ_MULTICLASS, num_class_, factor_, 0));
  }
  double BoostFromScore(int class_id) const override {
    return std::log(std::max<double>(kEpsilon, class_init_probs_[class_id]));
  }
  bool NeedTrain(int class_id) const override {
    const double p = std::fabs(class_init_probs_[class_id]);
    return !(p <= kEpsilon || p >= 1.0 - kEpsilon);
  }
  int NumModelPerIteration() const override { return num_class_; }
  void ConvertOutput(const double* in, double* out) const override {  // Common::Softmax
    double wmax = in[0];
    for (int i = 1; i < num_class_; ++i) wmax = std::max(in[i], wmax);
    double wsum = 0.0f;
    for (int i = 0; i < num_class_; ++i) { out[i] = std::exp(in[i] - wmax); wsum += out[i]; }
    for (int i = 0; i < num_class_; ++i) out[i] /= static_cast<double>(wsum);
  }
  std::string ToString() const override { std::ostringstream o; o << "multiclass num_class:" << num_class_; return o.str(); }
  const char* Name() const override { return "multiclass"; }

 private:
  int num_class_; double factor_; data_size_t num_data_ = 0;
  std::vector<double> class_init_probs_;
};

// src/objective/regression_objective.hpp: RegressionHuberLoss (:290-348), RegressionFairLoss (:352-396),
// RegressionPoissonLoss (:399-477).  All start from RegressionL2loss's (weighted) label mean.
class RegressionPointwise : public Objective {
 public:
  RegressionPointwise(const Config& c, int kind) : kind_(kind) {
    if (kind_ == LGBCU_OBJ_HUBER) p0_ = static_cast<double>(c.alpha);
    else if (kind_ == LGBCU_OBJ_FAIR) p0_ = static_cast<double>(c.fair_c);
    else if (kind_ == LGBCU_OBJ_TWEEDIE) p0_ = c.tweedie_variance_power;
    else p0_ = static_cast<double>(c.poisson_max_delta_step);
    if (c.reg_sqrt) Log::Warning("Cannot use sqrt transform in %s Regression, will auto disable it", Name());
  }
  void Init(const Metadata& meta, data_size_t n, const SumSync&) override {
    num_data_ = n; label_ = meta.label.data(); weights_ = meta.weights.empty() ? nullptr : meta.weights.data();
    if (meta.label.empty()) Log::Fatal("Label should not be NULL");
    if (LogLink()) {   // RegressionPoissonLoss::Init, inherited by gamma and tweedie
      label_t miny = label_[0]; double sumy = 0.0;
      for (data_size_t i = 0; i < n; ++i) { miny = std::min(miny, label_[i]); sumy += label_[i]; }
      if (miny < 0.0f) Log::Fatal("[%s]: at least one target label is negative", Name());
      if (sumy == 0.0f) Log::Fatal("[%s]: sum of labels is zero", Name());
    }
  }
  void Configure(LgbCuContext* ctx) const override { Check(lgbcu_set_objective(ctx, kind_, p0_, 0, 0)); }
  double BoostFromScore(int) const override {
    double suml = 0.0f, sumw = 0.0f;
    if (weights_) {
#pragma omp parallel for schedule(static) reduction(+:suml, sumw)
      for (data_size_t i = 0; i < num_data_; ++i) { suml += label_[i] * weights_[i]; sumw += weights_[i]; }
    } else {
      sumw = static_cast<double>(num_data_);
#pragma omp parallel for schedule(static) reduction(+:suml)
      for (data_size_t i = 0; i < num_data_; ++i) suml += label_[i];
    }
    const double mean = suml / sumw;
    if (LogLink()) return mean > 0 ? std::log(mean) : -std::numeric_limits<double>::infinity();  // Common::SafeLog
    return mean;
  }
  void ConvertOutput(const double* in, double* out) const override { out[0] = LogLink() ? std::exp(in[0]) : in[0]; }
  std::string ToString() const override { return Name(); }
  const char* Name() const override {
    return kind_ == LGBCU_OBJ_HUBER ? "huber" : kind_ == LGBCU_OBJ_FAIR ? "fair" : kind_ == LGBCU_OBJ_GAMMA ? "gamma" :
           kind_ == LGBCU_OBJ_TWEEDIE ? "tweedie" : "poisson";
  }

 private:
  bool LogLink() const { return kind_ == LGBCU_OBJ_POISSON || kind_ == LGBCU_OBJ_GAMMA || kind_ == LGBCU_OBJ_TWEEDIE; }
  int kind_; double p0_ = 0.0;
  data_size_t num_data_ = 0; const label_t* label_ = nullptr; const label_t* weights_ = nullptr;
};

// src/objective/xentropy_objective.hpp:44-147 (CrossEntropy): labels are probabilities in [0, 1]
class CrossEntropy : public Objective {
 public:
  explicit CrossEntropy(const Config&) {}
  void Init(const Metadata& meta, data_size_t n, const SumSync&) override {
    num_data_ = n; label_ = meta.label.data(); weights_ = meta.weights.empty() ? nullptr : meta.weights.data();
    if (meta.label.empty()) Log::Fatal("Label should not be NULL");
    for (data_size_t i = 0; i < n; ++i) {   // Common::CheckElementsIntervalClosed
      if (!(label_[i] >= 0.0f && label_[i] <= 1.0f)) Log::Fatal("[%s]: does not tolerate element [#%d = %f] outside [0, 1]", Name(), i, label_[i]);
    }
    if (weights_) {
      label_t minw = weights_[0]; double sumw = 0.0;
      for (data_size_t i = 0; i < n; ++i) { minw = std::min(minw, weights_[i]); sumw += weights_[i]; }
      if (minw < 0.0f) Log::Fatal("[%s]: at least one weight is negative", Name());
      if (sumw == 0.0f) Log::Fatal("[%s]: sum of weights is zero", Name());
    }
  }
  void Configure(LgbCuContext* ctx) const override { Check(lgbcu_set_objective(ctx, LGBCU_OBJ_XENTROPY, 0, 0, 0)); }
  double BoostFromScore(int) const override {
    double suml = 0.0f, sumw = 0.0f;
    if (weights_) {
#pragma omp parallel for schedule(static) reduction(+:suml, sumw)
      for (data_size_t i = 0; i < num_data_; ++i) { suml += label_[i] * weights_[i]; sumw += weights_[i]; }
    } else {
      sumw = static_cast<double>(num_data_);
#pragma omp parallel for schedule(static) reduction(+:suml)
      for (data_size_t i = 0; i < num_data_; ++i) suml += label_[i];
    }
    double pavg = suml / sumw;
    pavg = std::min(pavg, 1.0 - kEpsilon);
    pavg = std::max<double>(pavg, kEpsilon);
    return std::log(pavg / (1.0f - pavg));
  }
  void ConvertOutput(const double* in, double* out) const override { out[0] = 1.0f / (1.0f + std::exp(-in[0])); }
  std::string ToString() const override { return Name(); }
  const char* Name() const override { return "cross_entropy"; }

 private:
  data_size_t num_data_ = 0; const label_t* label_ = nullptr; const label_t* weights_ = nullptr;
};

// src/objective/multiclass_objective.hpp:186-276 (MulticlassOVA): one BinaryLogloss per class
class MulticlassOVA : public Objective {
 public:
  explicit MulticlassOVA(const Config& c) : num_class_(c.num_class), sigmoid_(c.sigmoid) {
    for (int k = 0; k < num_class_; ++k) binary_.emplace_back(new BinaryLogloss(c, k));
  }
  void Init(const Metadata& meta, data_size_t n, const SumSync& sum_over_ranks) override {
    for (auto& b : binary_) b->Init(meta, n, sum_over_ranks);
  }
  void Configure(LgbCuContext* ctx) const override {
    Check(lgbcu_set_objective(ctx, LGBCU_OBJ_MULTICLASSOVA, num_class_, sigmoid_, 0));
    std::vector<double> w;
    for (auto& b : binary_) { w.push_back(b->label_weight(0)); w.push_back(b->label_weight(1)); }
    Check(lgbcu_set_ova_label_weights(ctx, w.data(), num_class_));
  }
  double BoostFromScore(int k) const override { return binary_[k]->BoostFromScore(0); }
  bool NeedTrain(int k) const override { return binary_[k]->NeedTrain(0); }
  int NumModelPerIteration() const override { return num_class_; }
  void ConvertOutput(const double* in, double* out) const override {
    for (int i = 0; i < num_class_; ++i) out[i] = 1.0f / (1.0f + std::exp(-sigmoid_ * in[i]));
  }
  std::string ToString() const override { std::ostringstream o; o << "multiclassova num_class:" << num_class_ << " sigmoid:" << sigmoid_; return o.str(); }
  const char* Name() const override { return "multiclassova"; }

 private:
  int num_class_; double sigmoid_;
  std::vector<std::unique_ptr<BinaryLogloss>> binary_;
};

Objective* Objective::Create(const Config& c) {
  if (c.objective == "multiclass") return new MulticlassSoftmax(c);
  if (c.objective == "multiclassova") return new MulticlassOVA(c);
  if (c.objective == "huber") return new RegressionPointwise(c, LGBCU_OBJ_HUBER);
  if (c.objective == "fair") return new RegressionPointwise(c, LGBCU_OBJ_FAIR);
  if (c.objective == "poisson") return new RegressionPointwise(c, LGBCU_OBJ_POISSON);
  if (c.objective == "gamma") return new RegressionPointwise(c, LGBCU_OBJ_GAMMA);
  if (c.objective == "tweedie") return new RegressionPointwise(c, LGBCU_OBJ_TWEEDIE);
  if (c.objective == "regression_l1") return new RegressionPercentile(c, false);
  if (c.objective == "quantile") return new RegressionPercentile(c, true);
  if (c.objective == "cross_entropy") return new CrossEntropy(c);
  if (c.objective == "cross_entropy_lambda") return new CrossEntropyLambda(c);
  if (c.objective == "mape") return new RegressionMAPE(c);
  if (c.objective == "binary") return new BinaryLogloss(c);
  if (c.objective == "regression") return new RegressionL2(c);
  if (c.objective == "lambdarank") return new Lambdarank(c);
  if (c.objective == "rank_xendcg") return new RankXendcg(c);
  if (c.objective == "custom") return nullptr;
  Log::Fatal("objective=%s is outside this library's path (binary, regression/l2, regression_l1, huber, fair, poisson, gamma, tweedie, quantile, lambdarank, multiclass, multiclassova, custom)", c.objective.c_str());
}

}  // namespace lgb
