// Boosting driver for the CUDA learner: the part of GBDT (src/boosting/gbdt.cpp:42-119,151-160,209-243,312-419,
// 458-478,739-797) that calls the hot path, with gradients, scores, partition and histograms resident on the
// device.  The host keeps the O(num_leaves) Tree objects, the bagging / column-sampling random streams and the
// model text.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "config.h"
#include "dataset.h"
#include "lgbcu.h"
#include "objective.h"
#include "random.h"
#include "tree.h"

namespace lgb {

// process-wide (thread-local, like the reference's Network statics, src/network/network.cpp:17-27) NCCL rendezvous
struct NetworkState {
  bool active = false;
  uint8_t unique_id[128];
  int rank = 0, num_machines = 1;
};
NetworkState& Net();

class GBDT {
 public:
  GBDT(const Dataset* train, const Config& config);
  ~GBDT();
  bool TrainOneIter(const float* grad, const float* hess);  // true = finished (cannot split any more)
  int current_iteration() const { return static_cast<int>(models_.size()) / num_tree_per_iteration_; }   // GBDT::GetCurrentIteration
  int num_models() const { return static_cast<int>(models_.size()); }
  const Tree& model(int i) const { return *models_[i]; }
  int max_feature_idx() const { return train_->num_total_features() - 1; }
  int num_tree_per_iteration() const { return num_tree_per_iteration_; }
  int num_data() const { return train_->num_data(); }
  void GetTrainScore(double* out, bool converted) const;      // LGBM_BoosterGetPredict(data_idx = 0)
  // validation data and metrics (GBDT::AddValidDataset / GetEvalAt, src/boosting/gbdt.cpp:121-149, 545-565; the metric
  // objects of src/c_api.cpp:160-172, 332-344 are the MetricSpec list here, evaluated on the device)
  void AddValidData(const Dataset* valid);
  int num_valid() const { return static_cast<int>(valid_.size()); }
  int valid_num_data(int idx) const { return valid_[idx].data->num_data(); }
  void GetValidScore(int idx, double* out, bool converted) const;  // LGBM_BoosterGetPredict(data_idx = idx + 1)
  std::vector<double> GetEvalAt(int data_idx);
  bool average_output() const { return is_rf_; }
  std::vector<std::string> EvalNames() const;
  void PredictMat(const void* data, int data_type, int32_t nrow, int32_t ncol, bool row_major, bool raw_score,
                  int num_iteration, double* out) const;
  std::string SaveModelToString(int start_iteration, int num_iteration) const;
  // GBDT::RollbackOneIter (src/boosting/gbdt.cpp:421-444): the last tree leaves the scores and the model
  void RollbackOneIter();
  // Booster::ResetConfig (src/c_api.cpp:232-283) for the parameters that can change while training: learning rate
  // and the tree-growing parameters; anything that would change the data or the objective is refused
  void ResetParameter(const Config& merged);
  // GBDT::FeatureImportance (src/boosting/gbdt_model_text.cpp:576-620): 0 = split counts, 1 = total gains
  std::vector<double> FeatureImportance(int num_iteration, int importance_type) const;
  const Dataset* train_data() const { return train_; }
  const Config& config() const { return config_; }
  LgbCuContext* device() const { return ctx_; }
  int64_t total_device_launches() const;

 private:
  struct MetricSpec { int kind; std::vector<std::string> names; bool kl = false; };
  // KullbackLeiblerDivergence's presummed label entropy (xentropy_metric.hpp:281-294) of the training set / a validation set
  double LabelEntropyOffset(int data_idx);
  std::vector<double> kl_offset_; std::vector<char> kl_offset_ready_;
  struct ValidEntry { const Dataset* data; int set; };
  void SetupNdcg(int set, const Metadata& md, int num_data);
  void AddTreeToValid(const Tree& tree);
  // boosting=dart (src/boosting/dart.hpp:20-205): trees dropped before the gradients are taken, put back scaled after
  bool is_dart_ = false;
  Random random_for_drop_;
  std::vector<double> tree_weight_;
  double sum_weight_ = 0.0;
  std::vector<int> drop_index_;
  void DroppingTrees();
  void Normalize();
  void AddTreeToTrain(const Tree& tree, int class_id);   // ScoreUpdater::AddScore(tree, class) over every training row
  // boosting=rf (src/boosting/rf.hpp:25-215): gradients taken once at the initial scores, every tree on a fresh bag at
  // shrinkage 1, the scores kept as the running mean of the trees, predictions averaged (average_output)
  bool is_rf_ = false;
  std::vector<double> rf_init_scores_;
  void ReplayRecords(Tree* tree, int num_splits);
  void RFInit();
  bool RFTrainOneIter();
  void MultiplyScore(int class_id, double val);
  bool has_categorical_ = false;   // some feature of the training data is categorical
  std::vector<MetricSpec> metrics_;
  std::vector<int> eval_at_;
  std::vector<ValidEntry> valid_;
  bool train_ndcg_ready_ = false;
  void Bagging(int iter);
  void ConvertScores(double* out, int N) const;
  void ResetFeatureMask();
  std::vector<double> FeatureImportanceSplit(int num_iteration) const;

  const Dataset* train_;
  Config config_;
  std::unique_ptr<Objective> objective_;
  LgbCuContext* ctx_ = nullptr;
  std::vector<std::unique_ptr<Tree>> models_;
  int iter_ = 0;
  double shrinkage_rate_;
  bool has_init_score_ = false;
  bool data_parallel_ = false;  // this booster trains over the NCCL communicator (tree_learner=data, num_machines > 1)
  int num_tree_per_iteration_ = 1;            // = num_class for multiclass (one tree and one score plane per class)
  std::vector<char> class_need_train_;        // ObjectiveFunction::ClassNeedTrain per class
  // bagging (gbdt.cpp:209-243, 739-797)
  data_size_t bag_cnt_;     // the bag itself (in-bag rows ascending, then out-of-bag rows) lives on the device
  bool use_bagging_ = false;
  bool need_re_bagging_ = false;
  bool balanced_bagging_ = false;  // pos_/neg_bagging_fraction by label (BalancedBaggingHelper, gbdt.cpp:183-207)
  // GOSS (goss.hpp): host-side sampling state
  bool is_goss_ = false;
  void GossSample(int iter);
  std::vector<Random> goss_rands_;          // one generator per 1024-row block, seeded bagging_seed + block
  std::vector<float> goss_score_;           // |g * h| per row, fetched from the device
  std::vector<uint32_t> goss_inbag_, goss_amp_;
  // column sampling (col_sampler.hpp)
  Random feature_rand_;
  std::vector<int> valid_features_;  // real indices of used features
  std::vector<int8_t> feature_used_;
  int used_cnt_bytree_ = 0;
  bool need_reset_bytree_ = false;
  std::vector<LgbCuSplitRecord> records_;
};

}  // namespace lgb
