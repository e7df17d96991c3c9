#include "gbdt.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>

#include <omp.h>

#include <algorithm>
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
const int kBaggingRandBlock = 1024;  // gbdt.h:519
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
  StageTimer st;
  objective_.reset(Objective::Create(config_));
  if (objective_) {
    objective_->Init(train_->metadata(), N);
    class_need_train_ = objective_->NeedTrain();
  }
  // ---- device learner (TreeLearner::Init + ScoreUpdater ctor, gbdt.cpp:78-91)
  int ndev = 0;
  lgbcu_device_count(&ndev);
  int dev = config_.gpu_device_id >= 0 ? config_.gpu_device_id : 0;
  if (Net().active && config_.gpu_device_id < 0 && ndev > 0) dev = Net().rank % ndev;
  st.tick("objective init");
  Check(lgbcu_create(dev, &ctx_));
  st.tick("lgbcu_create");
  std::vector<LgbCuFeatureMeta> meta(F);
  for (int f = 0; f < F; ++f) {
    const BinMapper& m = train_->mapper(f);
    meta[f].num_bin = m.num_bin(); meta[f].most_freq_bin = static_cast<int>(m.most_freq_bin());
    meta[f].default_bin = static_cast<int>(m.default_bin()); meta[f].missing_type = m.missing_type();
    meta[f].real_index = train_->real_feature_index(f); meta[f].reserved = 0;
    meta[f].penalty = config_.feature_contri.empty() ? 1.0 : config_.feature_contri[meta[f].real_index];
  }
  if (!config_.feature_contri.empty() && static_cast<int>(config_.feature_contri.size()) != train_->num_total_features())
    Log::Fatal("Check failed: feature_contri size");
  Check(lgbcu_set_dataset(ctx_, N, F, train_->bins().data(), meta.data()));
  st.tick("lgbcu_set_dataset");
  LgbCuTreeParams tp;
  tp.num_leaves = config_.num_leaves; tp.max_depth = config_.max_depth; tp.min_data_in_leaf = config_.min_data_in_leaf;
  tp.reserved = 0; tp.min_sum_hessian_in_leaf = config_.min_sum_hessian_in_leaf; tp.lambda_l1 = config_.lambda_l1;
  tp.lambda_l2 = config_.lambda_l2; tp.min_gain_to_split = config_.min_gain_to_split;
  tp.max_delta_step = config_.max_delta_step; tp.path_smooth = config_.path_smooth;
  Check(lgbcu_set_tree_params(ctx_, &tp));
  Check(lgbcu_set_hist_mode(ctx_, config_.hist_mode));
  st.tick("lgbcu_set_tree_params");
  const Metadata& md = train_->metadata();
  if (!md.label.empty()) Check(lgbcu_set_label(ctx_, md.label.data(), N));
  if (!md.weights.empty()) Check(lgbcu_set_weight(ctx_, md.weights.data(), N));
  if (!md.query_boundaries.empty())
    Check(lgbcu_set_query(ctx_, md.query_boundaries.data(), static_cast<int>(md.query_boundaries.size()) - 1));
  if (objective_) objective_->Configure(ctx_);
  if (!md.init_score.empty()) { Check(lgbcu_set_score(ctx_, md.init_score.data(), N)); has_init_score_ = true; }
  if (Net().active && Net().num_machines > 1 && config_.tree_learner == "data") {
    Check(lgbcu_comm_init(ctx_, Net().unique_id, Net().rank, Net().num_machines));
  }
  records_.resize(std::max(config_.num_leaves - 1, 1));
  st.tick("labels/objective/comm");

  // ---- ResetBaggingConfig (gbdt.cpp:739-797)
  if (config_.bagging_fraction < 1.0 && config_.bagging_freq > 0) {
    bag_cnt_ = static_cast<data_size_t>(config_.bagging_fraction * N);
    bag_indices_.resize(N);
    for (int i = 0; i < (N + kBaggingRandBlock - 1) / kBaggingRandBlock; ++i) bagging_rands_.emplace_back(config_.bagging_seed + i);
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

// GBDT::Bagging + BaggingHelper (gbdt.cpp:162-181, 209-243): one Random per 1024-row block, rows kept in order
void GBDT::Bagging(int iter) {
  const int N = train_->num_data();
  if (!((bag_cnt_ < N && config_.bagging_freq > 0 && iter % config_.bagging_freq == 0) || need_re_bagging_)) return;
  if (bagging_rands_.empty()) return;
  need_re_bagging_ = false;
  const int nblock = static_cast<int>(bagging_rands_.size());
  std::vector<uint8_t> in_bag(N);
#pragma omp parallel for schedule(static)
  for (int b = 0; b < nblock; ++b) {
    const int s = b * kBaggingRandBlock, e = std::min(N, s + kBaggingRandBlock);
    for (int i = s; i < e; ++i) in_bag[i] = bagging_rands_[b].NextFloat() < config_.bagging_fraction;
  }
  data_size_t l = 0;
  for (int i = 0; i < N; ++i) if (in_bag[i]) bag_indices_[l++] = i;
  bag_cnt_ = l;
  for (int i = 0; i < N; ++i) if (!in_bag[i]) bag_indices_[l++] = i;
  ++bag_version_;
  Log::Debug("Re-bagging, using %d data to train", bag_cnt_);
}

bool GBDT::TrainOneIter(const float* grad, const float* hess) {
  const int N = train_->num_data();
  double init_score = 0.0;
  StageTimer st;
  if (grad == nullptr || hess == nullptr) {
    if (!objective_) Log::Fatal("No object function provided");
    // BoostFromAverage (gbdt.cpp:312-335)
    if (models_.empty() && !has_init_score_ && config_.boost_from_average) {
      double s = objective_->BoostFromScore();
      if (Net().active && Net().num_machines > 1) {  // Network::GlobalSyncUpByMean
        Check(lgbcu_comm_allreduce_sum_f64(ctx_, &s, 1));
        s /= Net().num_machines;
      }
      if (std::fabs(s) > kEpsilon) {
        init_score = s;
        Check(lgbcu_add_score_const(ctx_, init_score));
        Log::Info("Start training from score %lf", init_score);
      }
    }
    st.tick("boost from average");
    Check(lgbcu_get_gradients(ctx_));  // GBDT::Boosting
  } else {
    Check(lgbcu_set_gradients(ctx_, grad, hess, N));
  }
  Bagging(iter_);

  std::unique_ptr<Tree> tree(new Tree(2));
  bool should_continue = false;
  const bool bagged = !bag_indices_.empty() && bag_cnt_ < N;
  if (class_need_train_) {
    ResetFeatureMask();  // BeforeTrain: col_sampler_.ResetByTree()
    Check(lgbcu_begin_tree(ctx_, bagged ? bag_indices_.data() : nullptr, bagged ? bag_cnt_ : 0,
                           need_reset_bytree_ ? feature_used_.data() : nullptr));
    int ns = 0;
    st.tick("gradients/bagging/begin_tree");
    Check(lgbcu_train_tree(ctx_, records_.data(), &ns));
    st.tick("lgbcu_train_tree");
    tree.reset(new Tree(config_.num_leaves));
    for (int i = 0; i < ns; ++i) {
      const LgbCuSplitRecord& r = records_[i];
      const BinMapper& m = train_->mapper(r.feature);
      tree->Split(r.leaf, r.feature, r.real_feature, static_cast<uint32_t>(r.threshold), m.BinToValue(r.threshold),
                  r.left_output, r.right_output, r.left_count, r.right_count, r.left_sum_hessian, r.right_sum_hessian,
                  r.gain, m.missing_type(), r.default_left != 0);
    }
  }
  if (tree->num_leaves() > 1) {
    should_continue = true;
    tree->Shrinkage(shrinkage_rate_);
    // UpdateScore (gbdt.cpp:458-478): in-bag rows through the partition, out-of-bag rows by walking the tree
    Check(lgbcu_add_score_by_partition(ctx_, tree->leaf_values().data(), tree->num_leaves()));
    if (bagged) {
      std::vector<int> thr(tree->threshold_in_bin().begin(), tree->threshold_in_bin().end());
      Check(lgbcu_add_score_by_tree(ctx_, tree->num_leaves(), tree->split_feature_inner().data(), thr.data(),
                                    tree->decision_type().data(), tree->left_child().data(), tree->right_child().data(),
                                    tree->leaf_values().data(), bag_indices_.data() + bag_cnt_, N - bag_cnt_, bag_version_));
    }
    if (std::fabs(init_score) > kEpsilon) tree->AddBias(init_score);
  } else {
    if (models_.empty()) {  // only add default score one-time (gbdt.cpp:385-401)
      double output = 0.0;
      if (!class_need_train_) { if (objective_) output = objective_->BoostFromScore(); }
      else output = init_score;
      tree->AsConstantTree(output);
      Check(lgbcu_add_score_const(ctx_, output));
    }
  }
  st.tick("tree + score update");
  models_.push_back(std::move(tree));
  if (!should_continue) {
    Log::Warning("Stopped training because there are no more leaves that meet the split requirements");
    if (models_.size() > 1) models_.pop_back();
    return true;
  }
  ++iter_;
  return false;
}

void GBDT::GetTrainScore(double* out, bool converted) const {
  const int N = train_->num_data();
  Check(lgbcu_copy_score(ctx_, out, N));
  if (converted && objective_) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) { double v = out[i]; objective_->ConvertOutput(&v, &out[i]); }
  }
}

void GBDT::PredictMat(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major, bool raw_score,
                      int num_iteration, double* out) const {
  if (ncol != train_->num_total_features())
    Log::Fatal("The number of features in data (%d) is not the same as it was in training data (%d).", ncol, train_->num_total_features());
  int n_models = static_cast<int>(models_.size());
  if (num_iteration > 0) n_models = std::min(n_models, num_iteration);
#pragma omp parallel
  {
    std::vector<double> row(ncol);
#pragma omp for schedule(static)
    for (int32_t r = 0; r < nrow; ++r) {
      for (int c = 0; c < ncol; ++c) {
        const size_t i = row_major ? static_cast<size_t>(r) * ncol + c : static_cast<size_t>(c) * nrow + r;
        row[c] = data_type == 0 ? static_cast<double>(static_cast<const float*>(data)[i]) : static_cast<const double*>(data)[i];
      }
      double s = 0.0;
      for (int m = 0; m < n_models; ++m) s += models_[m]->Predict(row.data());
      if (!raw_score && objective_) objective_->ConvertOutput(&s, &s);
      out[r] = s;
    }
  }
}

std::vector<double> GBDT::FeatureImportanceSplit(int num_iteration) const {
  std::vector<double> imp(train_->num_total_features(), 0.0);
  int n_models = static_cast<int>(models_.size());
  if (num_iteration > 0) n_models = std::min(n_models, num_iteration);
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
  ss << "num_class=1" << '\n';
  ss << "num_tree_per_iteration=1" << '\n';
  ss << "label_index=0" << '\n';
  ss << "max_feature_idx=" << train_->num_total_features() - 1 << '\n';
  if (objective_) ss << "objective=" << objective_->ToString() << '\n';
  {
    std::string names;
    for (size_t i = 0; i < train_->feature_names().size(); ++i) { if (i) names += ' '; names += train_->feature_names()[i]; }
    ss << "feature_names=" << names << '\n';
    std::string infos; const auto fi = train_->feature_infos();
    for (size_t i = 0; i < fi.size(); ++i) { if (i) infos += ' '; infos += fi[i]; }
    ss << "feature_infos=" << infos << '\n';
  }
  int num_used = static_cast<int>(models_.size());
  const int total_iter = num_used;
  start_iteration = std::min(std::max(start_iteration, 0), total_iter);
  if (num_iteration > 0) num_used = std::min(start_iteration + num_iteration, num_used);
  std::vector<std::string> strs;
  std::string sizes;
  for (int i = start_iteration; i < num_used; ++i) {
    strs.push_back("Tree=" + std::to_string(i - start_iteration) + '\n' + models_[i]->ToString() + '\n');
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
