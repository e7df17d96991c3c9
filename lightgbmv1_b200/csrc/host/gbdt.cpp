#include "gbdt.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>

#include <omp.h>

#include <algorithm>
#include <functional>
#include <cmath>
#include <cstring>

namespace lgb {

NetworkState& Net() {
  static thread_local NetworkState s;
  return s;
}

namespace {
void Check(int rc) {
  if (rc != 0) Log::Fatal("%s", lgbcu_last_error());
}
// LGBM_B200_TIMING=1: wall-clock of the host-side stages on stderr (where does an end-to-end call spend its time)
struct StageTimer {
  bool on; std::chrono::steady_clock::time_point t;
  StageTimer() : on(std::getenv("LGBM_B200_TIMING") != nullptr), t(std::chrono::steady_clock::now()) {}
  void tick(const char* what) {
    if (!on) return;
    const auto n = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[b200 timing] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
    t = n;
  }
};
}  // namespace

GBDT::GBDT(const Dataset* train, const Config& config)
    : train_(train), config_(config), shrinkage_rate_(config.learning_rate), bag_cnt_(train->num_data()),
      feature_rand_(config.feature_fraction_seed) {
  const int N = train_->num_data(), F = train_->num_features();
  if (F <= 0) Log::Fatal("the CUDA learner needs at least one non-constant feature");
  is_dart_ = config_.boosting == "dart";
  random_for_drop_ = Random(config_.drop_seed);
  StageTimer st;
  objective_.reset(Objective::Create(config_));
  // ---- device learner (TreeLearner::Init + ScoreUpdater ctor, gbdt.cpp:78-91)
  int ndev = 0;
  lgbcu_device_count(&ndev);
  int dev = config_.gpu_device_id >= 0 ? config_.gpu_device_id : 0;
  if (Net().active && config_.gpu_device_id < 0 && ndev > 0) dev = Net().rank % ndev;
  Check(lgbcu_create(dev, &ctx_));
  num_tree_per_iteration_ = objective_ ? objective_->NumModelPerIteration() : 1;
  Check(lgbcu_set_num_class(ctx_, num_tree_per_iteration_));   // before the dataset: the score / gradient planes
  class_need_train_.assign(num_tree_per_iteration_, 1);
  st.tick("lgbcu_create");
  // the communicator comes first: the objective's Init sums class counts over the machines
  data_parallel_ = Net().active && Net().num_machines > 1 && config_.tree_learner == "data";
  if (data_parallel_) Check(lgbcu_comm_init(ctx_, Net().unique_id, Net().rank, Net().num_machines));
  if (objective_) {
    LgbCuContext* ctx = ctx_;
    objective_->Init(train_->metadata(), N, [ctx](double* v, int n) { Check(lgbcu_comm_allreduce_sum_f64(ctx, v, n)); });
    for (int k = 0; k < num_tree_per_iteration_; ++k) class_need_train_[k] = objective_->NeedTrain(k) ? 1 : 0;
  }
  st.tick("comm + objective init");
  std::vector<LgbCuFeatureMeta> meta(F);
  for (int f = 0; f < F; ++f) {
    const BinMapper& m = train_->mapper(f);
    meta[f].num_bin = m.num_bin(); meta[f].most_freq_bin = static_cast<int>(m.most_freq_bin());
    meta[f].default_bin = static_cast<int>(m.default_bin()); meta[f].missing_type = m.missing_type();
    meta[f].real_index = train_->real_feature_index(f);
    meta[f].bin_type = m.is_categorical() ? LGBCU_BIN_CATEGORICAL : LGBCU_BIN_NUMERICAL;
    has_categorical_ = has_categorical_ || m.is_categorical();
    meta[f].penalty = config_.feature_contri.empty() ? 1.0 : config_.feature_contri[meta[f].real_index];
  }
  if (!config_.feature_contri.empty() && static_cast<int>(config_.feature_contri.size()) != train_->num_total_features())
    Log::Fatal("Check failed: feature_contri size");
  {
    // the matrix is usually already in HBM (binned there by LGBM_DatasetCreateFromMat): adopt it without a host trip
    LgbCuBinnedMatrix* dm = train_->device_bins();
    int dm_dev = -1;
    if (dm) lgbcu_binned_device(dm, &dm_dev);
    if (dm && dm_dev == dev) Check(lgbcu_set_dataset_binned(ctx_, dm, meta.data()));
    else Check(lgbcu_set_dataset(ctx_, N, F, train_->bins().data(), meta.data()));
  }
  st.tick("lgbcu_set_dataset");
  LgbCuTreeParams tp;
  tp.num_leaves = config_.num_leaves; tp.max_depth = config_.max_depth; tp.min_data_in_leaf = config_.min_data_in_leaf;
  tp.reserved = 0; tp.min_sum_hessian_in_leaf = config_.min_sum_hessian_in_leaf; tp.lambda_l1 = config_.lambda_l1;
  tp.lambda_l2 = config_.lambda_l2; tp.min_gain_to_split = config_.min_gain_to_split;
  tp.max_delta_step = config_.max_delta_step; tp.path_smooth = config_.path_smooth;
  Check(lgbcu_set_tree_params(ctx_, &tp));
  if (has_categorical_) {
    Check(lgbcu_set_cat_params(ctx_, config_.cat_smooth, config_.cat_l2, config_.max_cat_threshold, config_.min_data_per_group,
                               config_.max_cat_to_onehot));
  }
  if (!config_.monotone_constraints.empty()) {
    if (static_cast<int>(config_.monotone_constraints.size()) != train_->num_total_features())
      Log::Fatal("Check failed: (num_total_features_) == (io_config.monotone_constraints.size())");
    std::vector<int8_t> types(F);
    for (int f = 0; f < F; ++f) {
      types[f] = config_.monotone_constraints[train_->real_feature_index(f)];
      if (types[f] != 0 && train_->mapper(f).is_categorical()) Log::Fatal("The output cannot be monotone with respect to categorical features");
    }
    Check(lgbcu_set_monotone(ctx_, types.data(), F, config_.monotone_penalty));
  }
  Check(lgbcu_set_hist_mode(ctx_, config_.hist_mode));
  st.tick("lgbcu_set_tree_params");
  const Metadata& md = train_->metadata();
  if (!md.label.empty()) Check(lgbcu_set_label(ctx_, md.label.data(), N));
  if (!md.weights.empty()) Check(lgbcu_set_weight(ctx_, md.weights.data(), N));
  if (!md.query_boundaries.empty())
    Check(lgbcu_set_query(ctx_, md.query_boundaries.data(), static_cast<int>(md.query_boundaries.size()) - 1));
  if (objective_) objective_->Configure(ctx_);
  if (!md.init_score.empty()) {   // [class][row] like the reference's ScoreUpdater
    if (md.init_score.size() != static_cast<size_t>(N) * num_tree_per_iteration_) Log::Fatal("Initial score size doesn't match data size");
    for (int k = 0; k < num_tree_per_iteration_; ++k) {
      Check(lgbcu_set_class(ctx_, k));
      Check(lgbcu_set_score(ctx_, md.init_score.data() + static_cast<size_t>(k) * N, N));
    }
    Check(lgbcu_set_class(ctx_, 0));
    has_init_score_ = true;
  }
  records_.resize(std::max(config_.num_leaves - 1, 1));
  st.tick("labels/objective/comm");
  // ---- metrics (Metric::CreateMetric, src/metric/metric.cpp:16-63; "custom" yields none)
  eval_at_ = config_.eval_at;
  if (eval_at_.empty()) { for (int i = 1; i <= 5; ++i) eval_at_.push_back(i); }  // DCGCalculator::DefaultEvalAt
  for (int k : eval_at_) if (k <= 0) Log::Fatal("Check failed: eval_at > 0");
  for (const std::string& m : config_.metric) {
    MetricSpec spec;
    if (m == "l2") { spec.kind = LGBCU_METRIC_L2; spec.names = {"l2"}; }
    else if (m == "rmse") { spec.kind = LGBCU_METRIC_RMSE; spec.names = {"rmse"}; }
    else if (m == "binary_logloss") { spec.kind = LGBCU_METRIC_BINARY_LOGLOSS; spec.names = {"binary_logloss"}; }
    else if (m == "binary_error") { spec.kind = LGBCU_METRIC_BINARY_ERROR; spec.names = {"binary_error"}; }
    else if (m == "auc") { spec.kind = LGBCU_METRIC_AUC; spec.names = {"auc"}; }
    else if (m == "ndcg") {
      spec.kind = LGBCU_METRIC_NDCG;
      for (int k : eval_at_) spec.names.push_back("ndcg@" + std::to_string(k));
    }
    else if (m == "map") {
      spec.kind = LGBCU_METRIC_MAP;
      for (int k : eval_at_) spec.names.push_back("map@" + std::to_string(k));
    }
    else if (m == "gamma_deviance") { spec.kind = LGBCU_METRIC_GAMMA_DEVIANCE; spec.names = {"gamma_deviance"}; }
    else if (m == "l1") { spec.kind = LGBCU_METRIC_L1; spec.names = {"l1"}; }
    else if (m == "huber") { spec.kind = LGBCU_METRIC_HUBER; spec.names = {"huber"}; }
    else if (m == "fair") { spec.kind = LGBCU_METRIC_FAIR; spec.names = {"fair"}; }
    else if (m == "poisson") { spec.kind = LGBCU_METRIC_POISSON; spec.names = {"poisson"}; }
    else if (m == "gamma") { spec.kind = LGBCU_METRIC_GAMMA; spec.names = {"gamma"}; }
    else if (m == "tweedie") { spec.kind = LGBCU_METRIC_TWEEDIE; spec.names = {"tweedie"}; }
    else if (m == "quantile") { spec.kind = LGBCU_METRIC_QUANTILE; spec.names = {"quantile"}; }
    else if (m == "multi_logloss") { spec.kind = LGBCU_METRIC_MULTI_LOGLOSS; spec.names = {"multi_logloss"}; }
    else if (m == "multi_error") {
      spec.kind = LGBCU_METRIC_MULTI_ERROR;
      spec.names = {config_.multi_error_top_k == 1 ? std::string("multi_error") : "multi_error@" + std::to_string(config_.multi_error_top_k)};
    }
    else if (m == "mape") { spec.kind = LGBCU_METRIC_MAPE; spec.names = {"mape"}; }
    else if (m == "cross_entropy_lambda") { spec.kind = LGBCU_METRIC_XENTLAMBDA; spec.names = {"cross_entropy_lambda"}; }
    else if (m == "cross_entropy") { spec.kind = LGBCU_METRIC_XENTROPY; spec.names = {"cross_entropy"}; }
    else if (m == "kullback_leibler") { spec.kind = LGBCU_METRIC_XENTROPY; spec.names = {"kullback_leibler"}; spec.kl = true; }
    else if (m == "custom") continue;
    else Log::Fatal("metric=%s is outside this library's path (l2, rmse, l1, huber, fair, poisson, gamma, tweedie, quantile, mape, gamma_deviance, map, binary_logloss, binary_error, cross_entropy, cross_entropy_lambda, kullback_leibler, auc, ndcg, multi_logloss, multi_error)", m.c_str());
    if ((spec.kind == LGBCU_METRIC_MULTI_LOGLOSS || spec.kind == LGBCU_METRIC_MULTI_ERROR) != (num_tree_per_iteration_ > 1))
      Log::Fatal("Multiclass objective and metrics don't match");
    metrics_.push_back(spec);
  }

  // ---- GOSS::ResetGoss (goss.hpp:77-101)
  is_goss_ = config_.boosting == "goss";
  if (is_goss_) {
    if (num_tree_per_iteration_ > 1) Log::Fatal("boosting=goss with a multiclass objective is outside this library's path");
    if (!(config_.top_rate + config_.other_rate <= 1.0)) Log::Fatal("Check failed: config_->top_rate + config_->other_rate <= 1.0f");
    if (!(config_.top_rate > 0.0 && config_.other_rate > 0.0)) Log::Fatal("Check failed: config_->top_rate > 0.0f && config_->other_rate > 0.0f");
    if (config_.bagging_freq > 0 && config_.bagging_fraction != 1.0) Log::Fatal("Cannot use bagging in GOSS");
    Log::Info("Using GOSS");
    const int nblock = (N + 1023) / 1024;
    goss_rands_.clear();
    for (int i = 0; i < nblock; ++i) goss_rands_.emplace_back(config_.bagging_seed + i);
    goss_score_.resize(static_cast<size_t>(N));
    goss_inbag_.assign(static_cast<size_t>(nblock) * 32, 0u);
    goss_amp_.assign(static_cast<size_t>(nblock) * 32, 0u);
    Check(lgbcu_bagging_reset(ctx_, config_.bagging_seed));  // allocates the device-side bag
    bag_cnt_ = N;  // "flag to not bagging first"
  }
  // ---- ResetBaggingConfig (gbdt.cpp:739-797)
  const data_size_t num_pos = objective_ ? objective_->NumPositiveData() : 0;
  const bool balance_cond = (config_.pos_bagging_fraction < 1.0 || config_.neg_bagging_fraction < 1.0) && num_pos > 0;
  if (!is_goss_ && (config_.bagging_fraction < 1.0 || balance_cond) && config_.bagging_freq > 0) {
    if (balance_cond) {
      balanced_bagging_ = true;
      bag_cnt_ = static_cast<data_size_t>(num_pos * config_.pos_bagging_fraction) +
                 static_cast<data_size_t>((N - num_pos) * config_.neg_bagging_fraction);
    } else {
      bag_cnt_ = static_cast<data_size_t>(config_.bagging_fraction * N);
    }
    Check(lgbcu_bagging_reset(ctx_, config_.bagging_seed));  // one generator per 1024-row block, seeded bagging_seed + block
    use_bagging_ = true;
    need_re_bagging_ = true;
  }
  // ---- ColSampler::SetTrainingData (col_sampler.hpp:39-53): note the sample drawn here is discarded by the
  // ResetByTree of the first BeforeTrain, exactly as in the reference
  for (int c = 0; c < train_->num_total_features(); ++c) if (train_->inner_feature_index(c) >= 0) valid_features_.push_back(c);
  feature_used_.assign(F, 1);
  if (config_.feature_fraction >= 1.0f) {
    need_reset_bytree_ = false;
    used_cnt_bytree_ = static_cast<int>(valid_features_.size());
  } else {
    need_reset_bytree_ = true;
    const int total = static_cast<int>(valid_features_.size());
    used_cnt_bytree_ = std::max(RoundInt(total * config_.feature_fraction), std::min(1, total));
  }
  ResetFeatureMask();
  is_rf_ = config_.boosting == "rf";
  if (is_rf_) RFInit();
}

GBDT::~GBDT() {
  if (ctx_) lgbcu_destroy(ctx_);
}

int64_t GBDT::total_device_launches() const { return lgbcu_launch_count(); }

// ColSampler::ResetByTree (col_sampler.hpp:74-89)
void GBDT::ResetFeatureMask() {
  if (!need_reset_bytree_) return;
  std::fill(feature_used_.begin(), feature_used_.end(), 0);
  const std::vector<int> pick = feature_rand_.Sample(static_cast<int>(valid_features_.size()), used_cnt_bytree_);
  for (int i : pick) feature_used_[train_->inner_feature_index(valid_features_[i])] = 1;
}

// GBDT::Bagging (gbdt.cpp:209-243): the draw itself (BaggingHelper, one generator per 1024-row block) runs on the
// device and the bag stays there; the host only learns its size
void GBDT::Bagging(int iter) {
  const int N = train_->num_data();
  if (is_goss_) { GossSample(iter); return; }
  if (!use_bagging_) return;
  if (!((bag_cnt_ < N && config_.bagging_freq > 0 && iter % config_.bagging_freq == 0) || need_re_bagging_)) return;
  need_re_bagging_ = false;
  int32_t cnt = 0;
  if (balanced_bagging_) Check(lgbcu_bagging_balanced(ctx_, config_.pos_bagging_fraction, config_.neg_bagging_fraction, &cnt));
  else Check(lgbcu_bagging(ctx_, config_.bagging_fraction, &cnt));
  bag_cnt_ = cnt;
  Log::Debug("Re-bagging, using %d data to train", bag_cnt_);
}

// GOSS::Bagging + GOSS::BaggingHelper (goss.hpp:103-177).  The walk over a chunk is sequential (the probability of
// keeping a small-gradient row depends on how many were kept before it) and the chunks are the reference's: one per
// thread, a multiple of 1024 rows (Threading::BlockInfoForceSize, threading.h:42-56), so like the reference the sample
// depends on num_threads.  The device supplies |g * h| and applies the outcome (lgbcu_goss_scores / lgbcu_goss_apply).
void GBDT::GossSample(int iter) {
  const int N = train_->num_data();
  bag_cnt_ = N;
  if (iter < static_cast<int>(1.0f / config_.learning_rate)) return;  // not subsample for first iterations
  Check(lgbcu_goss_scores(ctx_, goss_score_.data(), N));
  const int num_threads = config_.num_threads > 0 ? config_.num_threads : omp_get_max_threads();
  const int min_block = 1024;
  int nblock = std::min<int>(num_threads, (N + min_block - 1) / min_block);
  int inner = N;
  if (nblock > 1) {
    inner = (N + nblock - 1) / nblock;
    inner = (inner + min_block - 1) / min_block * min_block;
  }
  std::fill(goss_inbag_.begin(), goss_inbag_.end(), 0u);
  std::fill(goss_amp_.begin(), goss_amp_.end(), 0u);
  std::vector<float> mult(static_cast<size_t>(nblock), 1.0f);
  const float* score = goss_score_.data();
#pragma omp parallel for schedule(static, 1) num_threads(num_threads)
  for (int b = 0; b < nblock; ++b) {
    const int start = b * inner;
    const int cnt = std::min(inner, N - start);
    if (cnt <= 0) continue;
    int top_k = static_cast<int>(cnt * config_.top_rate);
    const int other_k = static_cast<int>(cnt * config_.other_rate);
    top_k = std::max(1, top_k);
    // ArgMaxAtK(..., top_k - 1): the top_k-th largest value
    std::vector<float> tmp(score + start, score + start + cnt);
    std::nth_element(tmp.begin(), tmp.begin() + (top_k - 1), tmp.end(), std::greater<float>());
    const float threshold = tmp[top_k - 1];
    const float multiply = static_cast<float>(cnt - top_k) / other_k;
    mult[b] = multiply;
    int left = 0, big = 0;
    for (int i = 0; i < cnt; ++i) {
      const int idx = start + i;
      bool in = false;
      if (score[idx] >= threshold) {
        in = true; ++big;
      } else {
        const int sampled = left - big;
        const int rest_need = other_k - sampled;
        const int rest_all = (cnt - i) - (top_k - big);
        const double prob = rest_need / static_cast<double>(rest_all);
        if (goss_rands_[idx / 1024].NextFloat() < prob) {
          in = true;
          goss_amp_[static_cast<size_t>(idx) >> 5] |= 1u << (idx & 31);  // chunks are multiples of 1024 rows: no word is shared
        }
      }
      if (in) { ++left; goss_inbag_[static_cast<size_t>(idx) >> 5] |= 1u << (idx & 31); }
    }
  }
  int32_t cnt = 0;
  Check(lgbcu_goss_apply(ctx_, goss_inbag_.data(), goss_amp_.data(), inner, mult.data(), nblock, &cnt));
  bag_cnt_ = cnt;
}

// The device's split records -> Tree::Split / Tree::SplitCategorical, in order: node numbering, leaf ids and the model
// text come out as the reference's (SerialTreeLearner::SplitInner, serial_tree_learner.cpp:541-659)
void GBDT::ReplayRecords(Tree* tree, int ns) {
  std::vector<uint32_t> cat_sets;   // the tree's categorical bin sets (8 words each), if it has any
  if (has_categorical_) {
    int32_t nsets = 0;
    Check(lgbcu_tree_cat_sets(ctx_, nullptr, 0, &nsets));
    if (nsets > 0) { cat_sets.resize(static_cast<size_t>(nsets) * 8); Check(lgbcu_tree_cat_sets(ctx_, cat_sets.data(), nsets, &nsets)); }
  }
  for (int i = 0; i < ns; ++i) {
    const LgbCuSplitRecord& r = records_[i];
    const BinMapper& m = train_->mapper(r.feature);
    if (m.is_categorical()) {
      // SerialTreeLearner::SplitInner, categorical branch (serial_tree_learner.cpp:592-621): the record names its set
      // of bins; the model carries the set of the bins' category values
      if (r.threshold < 0 || static_cast<size_t>(r.threshold) * 8 + 8 > cat_sets.size()) Log::Fatal("categorical split names bin set %d of %d", r.threshold, static_cast<int>(cat_sets.size() / 8));
      const uint32_t* bin_set = cat_sets.data() + static_cast<size_t>(r.threshold) * 8;
      std::vector<uint32_t> cat_set;   // Common::ConstructBitset over RealThreshold(bin)
      for (int b = 0; b < m.num_bin(); ++b) {
        if (!((bin_set[b >> 5] >> (b & 31)) & 1u)) continue;
        const int v = static_cast<int>(m.BinToValue(b));
        const int i1 = v / 32, i2 = v % 32;
        if (static_cast<int>(cat_set.size()) < i1 + 1) cat_set.resize(i1 + 1, 0);
        cat_set[i1] |= (1u << i2);
      }
      tree->SplitCategorical(r.leaf, r.feature, r.real_feature, bin_set, cat_set, r.left_output, r.right_output, r.left_count,
                             r.right_count, r.left_sum_hessian, r.right_sum_hessian, r.gain, m.missing_type());
      continue;
    }
    tree->Split(r.leaf, r.feature, r.real_feature, static_cast<uint32_t>(r.threshold), m.BinToValue(r.threshold),
                r.left_output, r.right_output, r.left_count, r.right_count, r.left_sum_hessian, r.right_sum_hessian,
                r.gain, m.missing_type(), r.default_left != 0);
  }

}

// ---- boosting=rf (src/boosting/rf.hpp) -----------------------------------------------------------------------------

void GBDT::MultiplyScore(int class_id, double val) {   // RF::MultiplyScore: training and validation scores
  Check(lgbcu_set_class(ctx_, class_id));
  Check(lgbcu_multiply_score(ctx_, val));
  for (const ValidEntry& v : valid_) Check(lgbcu_valid_multiply_score(ctx_, v.set, val));
}

// RF::Init + RF::Boosting (rf.hpp:34-53, 79-99)
void GBDT::RFInit() {
  if (!(config_.bagging_freq > 0 && config_.bagging_fraction < 1.0 && config_.bagging_fraction > 0.0))
    Log::Fatal("Check failed: config->bagging_freq > 0 && config->bagging_fraction < 1.0f && config->bagging_fraction > 0.0f");
  if (!(config_.feature_fraction <= 1.0 && config_.feature_fraction > 0.0)) Log::Fatal("Check failed: config->feature_fraction <= 1.0f && config->feature_fraction > 0.0f");
  if (has_init_score_) Log::Fatal("Check failed: (train_data->metadata().init_score()) == (nullptr)");
  if (!objective_) Log::Fatal("RF mode do not support custom objective function, please use built-in objectives.");
  if (objective_->IsRenewTreeOutput()) Log::Fatal("boosting=rf with a leaf-renewing objective (regression_l1, quantile, mape) is outside this library's path");
  if (is_goss_) Log::Fatal("boosting=rf cannot be combined with goss");
  shrinkage_rate_ = 1.0f;
  const int K = num_tree_per_iteration_;
  rf_init_scores_.assign(K, 0.0);
  for (int k = 0; k < K; ++k) {   // BoostFromAverage(k, false)
    if (!config_.boost_from_average) continue;
    double s = objective_->BoostFromScore(k);
    if (data_parallel_) { Check(lgbcu_comm_allreduce_sum_f64(ctx_, &s, 1)); s /= Net().num_machines; }
    if (std::fabs(s) > kEpsilon) { rf_init_scores_[k] = s; Log::Info("Start training from score %lf", s); }
  }
  // the gradients of every iteration: the objective at the constant initial scores (the scores themselves stay 0: the
  // constants are added, the gradients taken, the constants subtracted -- 0 + x - x is exact)
  for (int k = 0; k < K; ++k) if (rf_init_scores_[k] != 0.0) { Check(lgbcu_set_class(ctx_, k)); Check(lgbcu_add_score_const(ctx_, rf_init_scores_[k])); }
  Check(lgbcu_set_class(ctx_, 0));
  Check(lgbcu_get_gradients(ctx_));
  for (int k = 0; k < K; ++k) if (rf_init_scores_[k] != 0.0) { Check(lgbcu_set_class(ctx_, k)); Check(lgbcu_add_score_const(ctx_, -rf_init_scores_[k])); }
  Check(lgbcu_set_class(ctx_, 0));
}

// RF::TrainOneIter (rf.hpp:101-159)
bool GBDT::RFTrainOneIter() {
  const int N = train_->num_data();
  const int K = num_tree_per_iteration_;
  Bagging(iter_);
  const bool bagged = use_bagging_ && bag_cnt_ < N;
  for (int k = 0; k < K; ++k) {
    Check(lgbcu_set_class(ctx_, k));
    std::unique_ptr<Tree> tree(new Tree(2));
    if (class_need_train_[k]) {
      ResetFeatureMask();
      Check(lgbcu_begin_tree(ctx_, bagged ? LGBCU_DEVICE_BAG : nullptr, bagged ? bag_cnt_ : 0, need_reset_bytree_ ? feature_used_.data() : nullptr));
      int ns = 0;
      Check(lgbcu_train_tree(ctx_, records_.data(), &ns));
      tree.reset(new Tree(config_.num_leaves));
      ReplayRecords(tree.get(), ns);
    }
    if (tree->num_leaves() > 1) {
      if (std::fabs(rf_init_scores_[k]) > kEpsilon) tree->AddBias(rf_init_scores_[k]);
      MultiplyScore(k, static_cast<double>(iter_));
      Check(lgbcu_set_class(ctx_, k));
      Check(lgbcu_add_score_by_partition(ctx_, tree->leaf_values().data(), tree->num_leaves()));
      if (tree->num_cat() > 0 && (bagged || !valid_.empty())) Check(lgbcu_set_tree_cat_sets(ctx_, tree->cat_bin_sets().data(), tree->num_cat()));
      if (bagged) {
        std::vector<int> thr(tree->threshold_in_bin().begin(), tree->threshold_in_bin().end());
        Check(lgbcu_add_score_by_tree(ctx_, tree->num_leaves(), tree->split_feature_inner().data(), thr.data(), tree->decision_type().data(),
                                      tree->left_child().data(), tree->right_child().data(), tree->leaf_values().data(), LGBCU_DEVICE_OOB, N - bag_cnt_, 0));
      }
      AddTreeToValid(*tree);
      MultiplyScore(k, 1.0 / (iter_ + 1));
    } else if (models_.size() < static_cast<size_t>(K)) {
      double output = 0.0;
      if (!class_need_train_[k]) output = objective_->BoostFromScore(k);
      tree->AsConstantTree(output);
      MultiplyScore(k, static_cast<double>(iter_));
      Check(lgbcu_set_class(ctx_, k));
      Check(lgbcu_add_score_const(ctx_, output));
      for (const ValidEntry& v : valid_) Check(lgbcu_valid_add_const(ctx_, v.set, output));
      MultiplyScore(k, 1.0 / (iter_ + 1));
    }
    models_.push_back(std::move(tree));
  }
  Check(lgbcu_set_class(ctx_, 0));
  ++iter_;
  return false;
}

bool GBDT::TrainOneIter(const float* grad, const float* hess) {
  if (is_rf_) {
    if (grad != nullptr || hess != nullptr) Log::Fatal("RF mode do not support custom objective function, please use built-in objectives.");
    return RFTrainOneIter();
  }
  const int N = train_->num_data();
  const int K = num_tree_per_iteration_;
  std::vector<double> init_scores(K, 0.0);
  StageTimer st;
  if (grad == nullptr || hess == nullptr) {
    if (!objective_) Log::Fatal("No object function provided");
    // BoostFromAverage (gbdt.cpp:312-335), once per class
    if (models_.empty() && !has_init_score_ && config_.boost_from_average) {
      for (int k = 0; k < K; ++k) {
        double s = objective_->BoostFromScore(k);
        if (data_parallel_) {  // Network::GlobalSyncUpByMean
          Check(lgbcu_comm_allreduce_sum_f64(ctx_, &s, 1));
          s /= Net().num_machines;
        }
        if (std::fabs(s) > kEpsilon) {
          init_scores[k] = s;
          Check(lgbcu_set_class(ctx_, k));
          Check(lgbcu_add_score_const(ctx_, s));
          for (const ValidEntry& v : valid_) Check(lgbcu_valid_add_const(ctx_, v.set, s));
          Log::Info("Start training from score %lf", s);
        }
      }
    }
    st.tick("boost from average");
    if (is_dart_) DroppingTrees();   // DART::GetTrainingScore: the gradients are taken without the dropped trees
    Check(lgbcu_get_gradients(ctx_));  // GBDT::Boosting: every class's gradients from the scores as they are now
  } else {
    if (is_dart_) Log::Fatal("boosting=dart with gradients from the caller is outside this library's path");
    for (int k = 0; k < K; ++k) {
      Check(lgbcu_set_class(ctx_, k));
      Check(lgbcu_set_gradients(ctx_, grad + static_cast<size_t>(k) * N, hess + static_cast<size_t>(k) * N, N));
    }
  }
  Check(lgbcu_set_class(ctx_, 0));
  Bagging(iter_);   // one bag per iteration, shared by the iteration's trees

  bool should_continue = false;
  const bool bagged = (use_bagging_ || is_goss_) && bag_cnt_ < N;
  for (int k = 0; k < K; ++k) {   // GBDT::TrainOneIter's loop over cur_tree_id (gbdt.cpp:353-412)
    Check(lgbcu_set_class(ctx_, k));
    std::unique_ptr<Tree> tree(new Tree(2));
    if (class_need_train_[k]) {
      ResetFeatureMask();  // BeforeTrain: col_sampler_.ResetByTree()
      Check(lgbcu_begin_tree(ctx_, bagged ? LGBCU_DEVICE_BAG : nullptr, bagged ? bag_cnt_ : 0,
                             need_reset_bytree_ ? feature_used_.data() : nullptr));
      int ns = 0;
      st.tick("gradients/bagging/begin_tree");
      Check(lgbcu_train_tree(ctx_, records_.data(), &ns));
      st.tick("lgbcu_train_tree");
      tree.reset(new Tree(config_.num_leaves));
      ReplayRecords(tree.get(), ns);
    }
    if (tree->num_leaves() > 1) {
      should_continue = true;
      // RenewTreeOutput (serial_tree_learner.cpp:661-699): percentile objectives replace every leaf's Newton step by
      // the (weighted) percentile of label - score over the rows the tree grew on, before shrinkage
      if (objective_ && objective_->IsRenewTreeOutput()) {
        const int nl = tree->num_leaves();
        std::vector<double> out(nl); std::vector<int32_t> has(nl);
        Check(lgbcu_renew_leaf_outputs(ctx_, nl, objective_->RenewAlpha(), objective_->RenewAlphaIsFloat() ? 1 : 0, out.data(), has.data()));
        if (data_parallel_) {   // Network::GlobalSum of the outputs and of "this machine holds rows of the leaf"
          std::vector<double> v(2 * static_cast<size_t>(nl));
          for (int i = 0; i < nl; ++i) { v[i] = has[i] ? out[i] : 0.0; v[nl + i] = has[i] ? 1.0 : 0.0; }
          for (size_t o = 0; o < v.size(); o += 64) Check(lgbcu_comm_allreduce_sum_f64(ctx_, v.data() + o, static_cast<int>(std::min<size_t>(64, v.size() - o))));
          for (int i = 0; i < nl; ++i) tree->SetLeafOutput(i, v[i] / v[nl + i]);
        } else {
          for (int i = 0; i < nl; ++i) tree->SetLeafOutput(i, out[i]);
        }
      }
      tree->Shrinkage(shrinkage_rate_);
      // UpdateScore (gbdt.cpp:458-478): in-bag rows through the partition, out-of-bag rows by walking the tree
      Check(lgbcu_add_score_by_partition(ctx_, tree->leaf_values().data(), tree->num_leaves()));
      if (tree->num_cat() > 0 && (bagged || !valid_.empty()))
        Check(lgbcu_set_tree_cat_sets(ctx_, tree->cat_bin_sets().data(), tree->num_cat()));
      if (bagged) {
        std::vector<int> thr(tree->threshold_in_bin().begin(), tree->threshold_in_bin().end());
        Check(lgbcu_add_score_by_tree(ctx_, tree->num_leaves(), tree->split_feature_inner().data(), thr.data(),
                                      tree->decision_type().data(), tree->left_child().data(), tree->right_child().data(),
                                      tree->leaf_values().data(), LGBCU_DEVICE_OOB, N - bag_cnt_, 0));
      }
      AddTreeToValid(*tree);
      if (std::fabs(init_scores[k]) > kEpsilon) tree->AddBias(init_scores[k]);
    } else {
      if (models_.size() < static_cast<size_t>(K)) {  // only add default score one-time (gbdt.cpp:385-401)
        double output = 0.0;
        if (!class_need_train_[k]) { if (objective_) output = objective_->BoostFromScore(k); }
        else output = init_scores[k];
        tree->AsConstantTree(output);
        Check(lgbcu_add_score_const(ctx_, output));
        for (const ValidEntry& v : valid_) Check(lgbcu_valid_add_const(ctx_, v.set, output));
      }
    }
    st.tick("tree + score update");
    models_.push_back(std::move(tree));
  }
  Check(lgbcu_set_class(ctx_, 0));
  if (!should_continue) {
    Log::Warning("Stopped training because there are no more leaves that meet the split requirements");
    if (models_.size() > static_cast<size_t>(K)) for (int k = 0; k < K; ++k) models_.pop_back();
    return true;
  }
  ++iter_;
  if (is_dart_) {   // DART::TrainOneIter after GBDT::TrainOneIter
    Normalize();
    if (!config_.uniform_drop) { tree_weight_.push_back(shrinkage_rate_); sum_weight_ += shrinkage_rate_; }
  }
  return false;
}

// ---- boosting=dart (src/boosting/dart.hpp) -------------------------------------------------------------------------

void GBDT::AddTreeToTrain(const Tree& t, int class_id) {
  Check(lgbcu_set_class(ctx_, class_id));
  if (t.num_cat() > 0) Check(lgbcu_set_tree_cat_sets(ctx_, t.cat_bin_sets().data(), t.num_cat()));
  std::vector<int> thr(t.threshold_in_bin().begin(), t.threshold_in_bin().end());
  Check(lgbcu_add_score_by_tree(ctx_, t.num_leaves(), t.split_feature_inner().data(), thr.data(), t.decision_type().data(),
                                t.left_child().data(), t.right_child().data(), t.leaf_values().data(), nullptr, 0, 0));
}

// DART::DroppingTrees (dart.hpp:87-142)
void GBDT::DroppingTrees() {
  const int K = num_tree_per_iteration_;
  drop_index_.clear();
  const bool is_skip = random_for_drop_.NextFloat() < config_.skip_drop;
  if (!is_skip) {
    double drop_rate = config_.drop_rate;
    if (!config_.uniform_drop) {
      const double inv_average_weight = static_cast<double>(tree_weight_.size()) / sum_weight_;
      if (config_.max_drop > 0) drop_rate = std::min(drop_rate, config_.max_drop * inv_average_weight / sum_weight_);
      for (int i = 0; i < iter_; ++i) {
        if (random_for_drop_.NextFloat() < drop_rate * tree_weight_[i] * inv_average_weight) {
          drop_index_.push_back(i);
          if (drop_index_.size() >= static_cast<size_t>(config_.max_drop)) break;
        }
      }
    } else {
      if (config_.max_drop > 0) drop_rate = std::min(drop_rate, config_.max_drop / static_cast<double>(iter_));
      for (int i = 0; i < iter_; ++i) {
        if (random_for_drop_.NextFloat() < drop_rate) {
          drop_index_.push_back(i);
          if (drop_index_.size() >= static_cast<size_t>(config_.max_drop)) break;
        }
      }
    }
  }
  for (int i : drop_index_) {
    for (int k = 0; k < K; ++k) {
      Tree& t = *models_[static_cast<size_t>(i) * K + k];
      t.Shrinkage(-1.0);
      AddTreeToTrain(t, k);
    }
  }
  Check(lgbcu_set_class(ctx_, 0));
  if (!config_.xgboost_dart_mode) {
    shrinkage_rate_ = config_.learning_rate / (1.0f + static_cast<double>(drop_index_.size()));
  } else {
    if (drop_index_.empty()) shrinkage_rate_ = config_.learning_rate;
    else shrinkage_rate_ = config_.learning_rate / (config_.learning_rate + static_cast<double>(drop_index_.size()));
  }
}

// DART::Normalize (dart.hpp:144-196)
void GBDT::Normalize() {
  const int K = num_tree_per_iteration_;
  const double k = static_cast<double>(drop_index_.size());
  for (int i : drop_index_) {
    for (int c = 0; c < K; ++c) {
      Tree& t = *models_[static_cast<size_t>(i) * K + c];
      // the tree is at -1 x itself: first its share leaves the validation scores, then it returns to the training scores
      t.Shrinkage(!config_.xgboost_dart_mode ? 1.0f / (k + 1.0f) : shrinkage_rate_);
      Check(lgbcu_set_class(ctx_, c));
      AddTreeToValid(t);
      t.Shrinkage(!config_.xgboost_dart_mode ? -k : -k / config_.learning_rate);
      AddTreeToTrain(t, c);
    }
    if (!config_.uniform_drop) {
      if (!config_.xgboost_dart_mode) {
        sum_weight_ -= tree_weight_[i] * (1.0f / (k + 1.0f));
        tree_weight_[i] *= (k / (k + 1.0f));
      } else {
        sum_weight_ -= tree_weight_[i] * (1.0f / (k + config_.learning_rate));
        tree_weight_[i] *= (k / (k + config_.learning_rate));
      }
    }
  }
  Check(lgbcu_set_class(ctx_, 0));
}

void GBDT::GetTrainScore(double* out, bool converted) const {
  const int N = train_->num_data();
  StageTimer st;
  const int K = num_tree_per_iteration_;
  for (int k = 0; k < K; ++k) {   // [class][row], the reference's GetPredictAt layout before conversion
    Check(lgbcu_set_class(ctx_, k));
    Check(lgbcu_copy_score(ctx_, out + static_cast<size_t>(k) * N, N));
  }
  Check(lgbcu_set_class(ctx_, 0));
  st.tick("copy scores to host");
  if (converted && objective_) ConvertScores(out, N);
  st.tick("convert output");
}

// ---- validation data and metrics ----------------------------------------------------------------------------------

void GBDT::AddTreeToValid(const Tree& tree) {
  if (valid_.empty()) return;
  if (tree.num_cat() > 0) Check(lgbcu_set_tree_cat_sets(ctx_, tree.cat_bin_sets().data(), tree.num_cat()));
  std::vector<int> thr(tree.threshold_in_bin().begin(), tree.threshold_in_bin().end());
  for (const ValidEntry& v : valid_) {
    Check(lgbcu_valid_add_tree(ctx_, v.set, tree.num_leaves(), tree.split_feature_inner().data(), thr.data(),
                               tree.decision_type().data(), tree.left_child().data(), tree.right_child().data(),
                               tree.leaf_values().data()));
  }
}

// GBDT::AddValidDataset (gbdt.cpp:121-149)
void GBDT::AddValidData(const Dataset* valid) {
  // Dataset::CheckAlign: same features, same bin mappers
  bool aligned = valid->num_features() == train_->num_features() && valid->num_total_features() == train_->num_total_features();
  if (aligned) aligned = valid->feature_infos() == train_->feature_infos();
  for (int f = 0; aligned && f < train_->num_features(); ++f) aligned = valid->real_feature_index(f) == train_->real_feature_index(f);
  if (!aligned) Log::Fatal("Cannot add validation data, since it has different bin mappers with training data");
  ValidEntry e;
  e.data = valid;
  Check(lgbcu_valid_create(ctx_, valid->num_data(), valid->bins().data(), &e.set));
  const Metadata& md = valid->metadata();
  const int n = valid->num_data();
  if (!md.label.empty()) Check(lgbcu_valid_set_label(ctx_, e.set, md.label.data(), n));
  if (!md.weights.empty()) Check(lgbcu_valid_set_weight(ctx_, e.set, md.weights.data(), n));
  if (!md.query_boundaries.empty())
    Check(lgbcu_valid_set_query(ctx_, e.set, md.query_boundaries.data(), static_cast<int>(md.query_boundaries.size()) - 1));
  if (!md.init_score.empty()) {  // ScoreUpdater ctor
    if (md.init_score.size() != static_cast<size_t>(n) * num_tree_per_iteration_) Log::Fatal("Initial score size doesn't match data size");
    for (int k = 0; k < num_tree_per_iteration_; ++k) {
      Check(lgbcu_set_class(ctx_, k));
      Check(lgbcu_valid_set_score(ctx_, e.set, md.init_score.data() + static_cast<size_t>(k) * n, n));
    }
    Check(lgbcu_set_class(ctx_, 0));
  }
  valid_.push_back(e);
  // the trees grown so far (tree i belongs to class i % K)
  for (size_t i = 0; i < models_.size(); ++i) {
    const auto& t = models_[i];
    std::vector<int> thr(t->threshold_in_bin().begin(), t->threshold_in_bin().end());
    Check(lgbcu_set_class(ctx_, static_cast<int>(i % num_tree_per_iteration_)));
    if (t->num_cat() > 0) Check(lgbcu_set_tree_cat_sets(ctx_, t->cat_bin_sets().data(), t->num_cat()));
    Check(lgbcu_valid_add_tree(ctx_, e.set, t->num_leaves(), t->split_feature_inner().data(), thr.data(),
                               t->decision_type().data(), t->left_child().data(), t->right_child().data(),
                               t->leaf_values().data()));
  }
  Check(lgbcu_set_class(ctx_, 0));
  for (const MetricSpec& m : metrics_) {
    if (m.kind == LGBCU_METRIC_NDCG || m.kind == LGBCU_METRIC_MAP) SetupNdcg(e.set, md, n);
    if (md.label.empty()) Log::Fatal("Label should not be NULL");
  }
}

// NDCGMetric::Init (rank_metric.hpp:33-75) + DCGCalculator::{Init,CheckLabel,CalMaxDCG} (dcg_calculator.cpp:43-107)
void GBDT::SetupNdcg(int set, const Metadata& md, int num_data) {
  if (md.query_boundaries.empty()) Log::Fatal("The NDCG metric requires query information");
  std::vector<double> gain = config_.label_gain;
  if (gain.empty()) {  // DefaultLabelGain: 2^i - 1
    gain.push_back(0.0);
    for (int i = 1; i < 31; ++i) gain.push_back(static_cast<double>((1 << i) - 1));
  }
  const int kMaxPosition = 10000;
  int max_k = 0;
  for (int k : eval_at_) max_k = std::max(max_k, k);
  std::vector<double> discount(std::min(std::max(max_k, 1), kMaxPosition));
  for (size_t i = 0; i < discount.size(); ++i) discount[i] = 1.0 / std::log2(2.0 + i);
  for (int i = 0; i < num_data; ++i) {  // CheckLabel
    const label_t l = md.label[i];
    if (std::fabs(l - static_cast<int>(l)) > kEpsilon)
      Log::Fatal("label should be int type (met %f) for ranking task,\nfor the gain of label, please set the label_gain parameter", l);
    if (l < 0) Log::Fatal("Label should be non-negative (met %f) for ranking task", l);
    if (static_cast<size_t>(l) >= gain.size())
      Log::Fatal("Label %zu is not less than the number of label mappings (%zu)", static_cast<size_t>(l), gain.size());
  }
  const int nq = static_cast<int>(md.query_boundaries.size()) - 1;
  const int ne = static_cast<int>(eval_at_.size());
  std::vector<double> inv(static_cast<size_t>(nq) * ne, 0.0);
#pragma omp parallel for schedule(static)
  for (int q = 0; q < nq; ++q) {
    const label_t* lab = md.label.data() + md.query_boundaries[q];
    const int cnt = md.query_boundaries[q + 1] - md.query_boundaries[q];
    std::vector<int> label_cnt(gain.size(), 0);
    for (int i = 0; i < cnt; ++i) ++label_cnt[static_cast<int>(lab[i])];
    double cur = 0.0;
    int left = 0, top = static_cast<int>(gain.size()) - 1;
    for (int e = 0; e < ne; ++e) {
      int k = eval_at_[e];
      if (k > cnt) k = cnt;
      for (int j = left; j < k; ++j) {
        while (top > 0 && label_cnt[top] <= 0) top -= 1;
        if (top < 0) break;
        cur += discount[j] * gain[top];
        label_cnt[top] -= 1;
      }
      inv[static_cast<size_t>(q) * ne + e] = cur > 0.0 ? 1.0f / cur : -1.0f;
      left = k;
    }
  }
  std::vector<float> qw;
  if (!md.weights.empty()) {  // Metadata::LoadQueryWeights (metadata.cpp:455-469): float mean of the row weights
    qw.resize(nq);
    for (int q = 0; q < nq; ++q) {
      float w = 0.0f;
      for (int j = md.query_boundaries[q]; j < md.query_boundaries[q + 1]; ++j) w += md.weights[j];
      qw[q] = w / (md.query_boundaries[q + 1] - md.query_boundaries[q]);
    }
  }
  Check(lgbcu_eval_set_ndcg(ctx_, set, eval_at_.data(), ne, inv.data(), qw.empty() ? nullptr : qw.data(), gain.data(),
                            static_cast<int>(gain.size()), discount.data(), static_cast<int>(discount.size())));
}

std::vector<std::string> GBDT::EvalNames() const {
  std::vector<std::string> out;
  for (const MetricSpec& m : metrics_) for (const std::string& n : m.names) out.push_back(n);
  return out;
}

// The cross-entropy metrics divide by Common::ObtainMinMaxSum's sum of the weights (utils/common.h:843-883), which is
// accumulated pairwise IN FLOAT; lgbcu_eval divides by the double sum.  Returns {float-accumulated sum, double sum}.
static std::pair<double, double> XentWeightSums(const Metadata& md) {
  const int n = static_cast<int>(md.weights.size());
  const label_t* w = md.weights.data();
  label_t sumw; int i;
  if (n & 1) { sumw = w[0]; i = 2; } else { sumw = w[0] + w[1]; i = 3; }
  for (; i < n; i += 2) sumw += w[i - 1] + w[i];
  double exact = 0.0;
  for (int k = 0; k < n; ++k) exact += w[k];
  return {static_cast<double>(sumw), exact};
}

// GBDT::GetEvalAt (gbdt.cpp:545-565)
std::vector<double> GBDT::GetEvalAt(int data_idx) {
  if (data_idx < 0 || data_idx > static_cast<int>(valid_.size())) Log::Fatal("Check failed: data_idx >= 0 && data_idx <= valid size");
  const int set = data_idx == 0 ? -1 : valid_[data_idx - 1].set;
  std::vector<double> ret;
  for (const MetricSpec& m : metrics_) {
    if ((m.kind == LGBCU_METRIC_NDCG || m.kind == LGBCU_METRIC_MAP) && data_idx == 0 && !train_ndcg_ready_) {
      SetupNdcg(-1, train_->metadata(), train_->num_data());
      train_ndcg_ready_ = true;
    }
    double buf[64];
    // metrics convert the raw score through the objective (ObjectiveFunction::ConvertOutput): a sigmoid for binary
    double sig = (objective_ && std::string(objective_->Name()) == "binary") ? config_.sigmoid : 0.0;
    if (objective_ && std::string(objective_->Name()) == "cross_entropy") sig = 1.0;   // CrossEntropy::ConvertOutput
    if (m.kind == LGBCU_METRIC_QUANTILE || m.kind == LGBCU_METRIC_HUBER) sig = config_.alpha;
    if (m.kind == LGBCU_METRIC_FAIR) sig = config_.fair_c;
    if (m.kind == LGBCU_METRIC_TWEEDIE) sig = config_.tweedie_variance_power;
    if (m.kind == LGBCU_METRIC_MULTI_LOGLOSS) sig = objective_ ? 1.0 : 0.0;                                       // softmax first
    if (m.kind == LGBCU_METRIC_MULTI_ERROR) sig = (objective_ ? 1.0 : -1.0) * config_.multi_error_top_k;
    Check(lgbcu_eval(ctx_, set, m.kind, sig, buf, 64));
    if (m.kind == LGBCU_METRIC_XENTROPY) {
      const Metadata& md = data_idx == 0 ? train_->metadata() : valid_[data_idx - 1].data->metadata();
      if (!md.weights.empty()) { const auto s = XentWeightSums(md); buf[0] *= s.second / s.first; }
      if (m.kl) buf[0] += LabelEntropyOffset(data_idx);
    }
    for (size_t i = 0; i < m.names.size(); ++i) ret.push_back(buf[i]);
  }
  return ret;
}

double GBDT::LabelEntropyOffset(int data_idx) {
  if (static_cast<int>(kl_offset_.size()) <= data_idx) { kl_offset_.resize(data_idx + 1, 0.0); kl_offset_ready_.resize(data_idx + 1, 0); }
  if (kl_offset_ready_[data_idx]) return kl_offset_[data_idx];
  const Metadata& md = data_idx == 0 ? train_->metadata() : valid_[data_idx - 1].data->metadata();
  const int n = static_cast<int>(md.label.size());
  auto yent = [](double p) {   // YentLoss (xentropy_metric.hpp:60-67)
    double hp = 0.0;
    if (p > 0) hp += p * std::log(p);
    const double q = 1.0f - p;
    if (q > 0) hp += q * std::log(q);
    return hp;
  };
  const double sum_w = md.weights.empty() ? static_cast<double>(n) : XentWeightSums(md).first;
  double presum = 0.0f;
  for (int i = 0; i < n; ++i) presum += md.weights.empty() ? yent(md.label[i]) : yent(md.label[i]) * md.weights[i];
  kl_offset_[data_idx] = presum / sum_w;
  kl_offset_ready_[data_idx] = 1;
  return kl_offset_[data_idx];
}

void GBDT::GetValidScore(int idx, double* out, bool converted) const {
  const int N = valid_[idx].data->num_data();
  for (int k = 0; k < num_tree_per_iteration_; ++k) {
    Check(lgbcu_set_class(ctx_, k));
    Check(lgbcu_valid_copy_score(ctx_, valid_[idx].set, out + static_cast<size_t>(k) * N, N));
  }
  Check(lgbcu_set_class(ctx_, 0));
  if (converted && objective_) ConvertScores(out, N);
}

// ObjectiveFunction::ConvertOutput over [class][row] planes (GBDT::GetPredictAt, gbdt.cpp:612-624)
void GBDT::ConvertScores(double* out, int N) const {
  const int K = num_tree_per_iteration_;
  if (K == 1) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) { double v = out[i]; objective_->ConvertOutput(&v, &out[i]); }
    return;
  }
#pragma omp parallel
  {
    std::vector<double> raw(K), conv(K);
#pragma omp for schedule(static)
    for (int i = 0; i < N; ++i) {
      for (int k = 0; k < K; ++k) raw[k] = out[static_cast<size_t>(k) * N + i];
      objective_->ConvertOutput(raw.data(), conv.data());
      for (int k = 0; k < K; ++k) out[static_cast<size_t>(k) * N + i] = conv[k];
    }
  }
}

void GBDT::PredictMat(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major, bool raw_score,
                      int num_iteration, double* out) const {
  if (ncol != train_->num_total_features())
    Log::Fatal("The number of features in data (%d) is not the same as it was in training data (%d).", ncol, train_->num_total_features());
  const int K = num_tree_per_iteration_;
  int n_iter = static_cast<int>(models_.size()) / K;
  if (num_iteration > 0) n_iter = std::min(n_iter, num_iteration);
#pragma omp parallel
  {
    std::vector<double> row(ncol), s(K), conv(K);
#pragma omp for schedule(static)
    for (int32_t r = 0; r < nrow; ++r) {
      for (int c = 0; c < ncol; ++c) {
        const size_t i = row_major ? static_cast<size_t>(r) * ncol + c : static_cast<size_t>(c) * nrow + r;
        row[c] = data_type == 0 ? static_cast<double>(static_cast<const float*>(data)[i]) : static_cast<const double*>(data)[i];
      }
      // GBDT::PredictRaw (gbdt_prediction.cpp:13-31): tree i * K + k adds to class k; a row's K outputs are adjacent
      for (int k = 0; k < K; ++k) s[k] = 0.0;
      for (int it = 0; it < n_iter; ++it) for (int k = 0; k < K; ++k) s[k] += models_[static_cast<size_t>(it) * K + k]->Predict(row.data());
      if (is_rf_ && !raw_score) for (int k = 0; k < K; ++k) s[k] /= n_iter;   // average_output_: GBDT::Predict divides, PredictRaw does not (gbdt_prediction.cpp:13-31, 55-60)
      if (!raw_score && objective_) { objective_->ConvertOutput(s.data(), conv.data()); for (int k = 0; k < K; ++k) s[k] = conv[k]; }
      for (int k = 0; k < K; ++k) out[static_cast<size_t>(r) * K + k] = s[k];
    }
  }
}

void GBDT::RollbackOneIter() {
  const int K = num_tree_per_iteration_;
  if (iter_ <= 0 || models_.size() < static_cast<size_t>(K)) return;
  for (int k = 0; k < K; ++k) {
    Tree& t = *models_[models_.size() - K + k];
    t.ScaleLeafValues(-1.0);   // Tree::Shrinkage(-1): subtracting the tree = adding its negative
    std::vector<int> thr(t.threshold_in_bin().begin(), t.threshold_in_bin().end());
    Check(lgbcu_set_class(ctx_, k));
    if (t.num_cat() > 0) Check(lgbcu_set_tree_cat_sets(ctx_, t.cat_bin_sets().data(), t.num_cat()));
    Check(lgbcu_add_score_by_tree(ctx_, t.num_leaves(), t.split_feature_inner().data(), thr.data(), t.decision_type().data(),
                                  t.left_child().data(), t.right_child().data(), t.leaf_values().data(), nullptr, 0, 0));
    AddTreeToValid(t);
  }
  Check(lgbcu_set_class(ctx_, 0));
  for (int k = 0; k < K; ++k) models_.pop_back();
  --iter_;
}

void GBDT::ResetParameter(const Config& merged) {
  if (merged.objective != config_.objective || merged.boosting != config_.boosting || merged.tree_learner != config_.tree_learner ||
      merged.max_bin != config_.max_bin || merged.num_leaves > config_.num_leaves)
    Log::Fatal("LGBM_BoosterResetParameter: only the learning rate and the tree-growing parameters (num_leaves not above the "
               "value the booster was created with) can be reset on this booster");
  config_ = merged;
  shrinkage_rate_ = config_.learning_rate;
  LgbCuTreeParams tp;
  tp.num_leaves = config_.num_leaves; tp.max_depth = config_.max_depth; tp.min_data_in_leaf = config_.min_data_in_leaf;
  tp.reserved = 0; tp.min_sum_hessian_in_leaf = config_.min_sum_hessian_in_leaf; tp.lambda_l1 = config_.lambda_l1;
  tp.lambda_l2 = config_.lambda_l2; tp.min_gain_to_split = config_.min_gain_to_split;
  tp.max_delta_step = config_.max_delta_step; tp.path_smooth = config_.path_smooth;
  Check(lgbcu_set_tree_params(ctx_, &tp));
}

std::vector<double> GBDT::FeatureImportance(int num_iteration, int importance_type) const {
  std::vector<double> imp(train_->num_total_features(), 0.0);
  int n_models = static_cast<int>(models_.size());
  if (num_iteration > 0) n_models = std::min(n_models, num_iteration * num_tree_per_iteration_);
  for (int m = 0; m < n_models; ++m) {
    const Tree& t = *models_[m];
    for (int i = 0; i < t.num_leaves() - 1; ++i) {
      if (t.split_gain()[i] > 0) imp[t.split_feature()[i]] += importance_type == 0 ? 1.0 : t.split_gain()[i];
    }
  }
  return imp;
}

std::vector<double> GBDT::FeatureImportanceSplit(int num_iteration) const {
  std::vector<double> imp(train_->num_total_features(), 0.0);
  int n_models = static_cast<int>(models_.size());
  if (num_iteration > 0) n_models = std::min(n_models, num_iteration * num_tree_per_iteration_);
  for (int m = 0; m < n_models; ++m) {
    const Tree& t = *models_[m];
    for (int i = 0; i < t.num_leaves() - 1; ++i) if (t.split_gain()[i] > 0) imp[t.split_feature()[i]] += 1.0;
  }
  return imp;
}

// GBDT::SaveModelToString (src/boosting/gbdt_model_text.cpp:306-397): same text, so the reference can load it
std::string GBDT::SaveModelToString(int start_iteration, int num_iteration) const {
  std::ostringstream ss;
  ss << "tree" << '\n';
  ss << "version=v3" << '\n';
  ss << "num_class=" << num_tree_per_iteration_ << '\n';
  ss << "num_tree_per_iteration=" << num_tree_per_iteration_ << '\n';
  ss << "label_index=0" << '\n';
  ss << "max_feature_idx=" << train_->num_total_features() - 1 << '\n';
  if (objective_) ss << "objective=" << objective_->ToString() << '\n';
  if (is_rf_) ss << "average_output" << '\n';
  {
    std::string names;
    for (size_t i = 0; i < train_->feature_names().size(); ++i) { if (i) names += ' '; names += train_->feature_names()[i]; }
    ss << "feature_names=" << names << '\n';
    if (!config_.monotone_constraints.empty()) {   // gbdt_model_text.cpp:329-332
      ss << "monotone_constraints=";
      for (size_t i = 0; i < config_.monotone_constraints.size(); ++i) { if (i) ss << ' '; ss << static_cast<int>(config_.monotone_constraints[i]); }
      ss << '\n';
    }
    std::string infos; const auto fi = train_->feature_infos();
    for (size_t i = 0; i < fi.size(); ++i) { if (i) infos += ' '; infos += fi[i]; }
    ss << "feature_infos=" << infos << '\n';
  }
  int num_used = static_cast<int>(models_.size());   // gbdt_model_text.cpp:336-347: iterations -> models
  const int total_iter = num_used / num_tree_per_iteration_;
  start_iteration = std::min(std::max(start_iteration, 0), total_iter);
  if (num_iteration > 0) num_used = std::min((start_iteration + num_iteration) * num_tree_per_iteration_, num_used);
  const int start_model = start_iteration * num_tree_per_iteration_;
  std::vector<std::string> strs;
  std::string sizes;
  for (int i = start_model; i < num_used; ++i) {
    strs.push_back("Tree=" + std::to_string(i - start_model) + '\n' + models_[i]->ToString() + '\n');
    if (!sizes.empty()) sizes += ' ';
    sizes += std::to_string(strs.back().size());
  }
  ss << "tree_sizes=" << sizes << '\n' << '\n';
  for (auto& s : strs) ss << s;
  ss << "end of trees" << "\n";
  const std::vector<double> imp = FeatureImportanceSplit(num_iteration);
  std::vector<std::pair<size_t, std::string>> pairs;
  for (size_t i = 0; i < imp.size(); ++i) {
    const size_t v = static_cast<size_t>(imp[i]);
    if (v > 0) pairs.emplace_back(v, train_->feature_names()[i]);
  }
  std::stable_sort(pairs.begin(), pairs.end(), [](const std::pair<size_t, std::string>& a, const std::pair<size_t, std::string>& b) { return a.first > b.first; });
  ss << '\n' << "feature_importances:" << '\n';
  for (auto& p : pairs) ss << p.second << "=" << std::to_string(p.first) << '\n';
  ss << "\nparameters:" << '\n' << config_.ToString() << "\n" << "end of parameters" << '\n';
  return ss.str();
}

}  // namespace lgb
