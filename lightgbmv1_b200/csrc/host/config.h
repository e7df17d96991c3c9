// The slice of the reference's Config (include/LightGBM/config.h, src/io/config.cpp, config_auto.cpp) that the
// hot path reads.  Same keys, aliases and defaults; parameters outside the path are accepted and ignored with a
// warning where they cannot change the result, and rejected with an error where they would.
#pragma once
#include <string>
#include <unordered_map>
#include <vector>

namespace lgb {

struct Config {
  // core
  std::string objective = "regression";
  std::string boosting = "gbdt";
  std::string tree_learner = "serial";
  std::string device_type = "cuda";
  std::vector<std::string> metric;   // normalised names (ParseMetricAlias), in the order given
  std::vector<int> eval_at;          // NDCG positions; empty = 1..5 (DCGCalculator::DefaultEvalAt)
  int num_iterations = 100;
  double learning_rate = 0.1;
  int num_leaves = 31;
  int num_threads = 0;
  int seed = 0;
  int num_machines = 1;
  int gpu_device_id = -1;
  int verbosity = 1;
  int num_class = 1;
  int multi_error_top_k = 1;
  // tree
  int max_depth = -1;
  int min_data_in_leaf = 20;
  double min_sum_hessian_in_leaf = 1e-3;
  double bagging_fraction = 1.0;
  double top_rate = 0.2, other_rate = 0.1;  // GOSS (config.h:399-405)
  double pos_bagging_fraction = 1.0, neg_bagging_fraction = 1.0;
  int bagging_freq = 0;
  int bagging_seed = 3;
  double feature_fraction = 1.0;
  double feature_fraction_bynode = 1.0;
  int feature_fraction_seed = 2;
  double lambda_l1 = 0.0, lambda_l2 = 0.0;
  double min_gain_to_split = 0.0;
  double max_delta_step = 0.0;
  double path_smooth = 0.0;
  bool extra_trees = false;
  std::vector<double> feature_contri;
  // dataset
  int max_bin = 255;
  int min_data_in_bin = 3;
  int bin_construct_sample_cnt = 200000;
  int data_random_seed = 1;
  bool use_missing = true;
  bool zero_as_missing = false;
  bool feature_pre_filter = true;
  bool enable_bundle = true;
  // objective
  double sigmoid = 1.0;
  double alpha = 0.9;   // huber / quantile objectives and metrics (config.h:815)
  double fair_c = 1.0, poisson_max_delta_step = 0.7, tweedie_variance_power = 1.5;
  // categorical features: raw column indices (config.h categorical_feature; "name:" lists need a data file's header)
  std::vector<int> categorical_feature;
  int min_data_per_group = 100, max_cat_threshold = 32, max_cat_to_onehot = 4;
  double cat_l2 = 10.0, cat_smooth = 10.0;
  std::vector<int> max_bin_by_feature;   // per raw column; empty = max_bin everywhere
  // boosting=dart (src/boosting/dart.hpp)
  double drop_rate = 0.1, skip_drop = 0.5;
  int max_drop = 50, drop_seed = 4;
  bool xgboost_dart_mode = false, uniform_drop = false;
  // monotone constraints (basic method): -1 / 0 / +1 per raw column; monotone_penalty
  std::vector<int8_t> monotone_constraints;
  double monotone_penalty = 0.0;
  bool is_unbalance = false;
  double scale_pos_weight = 1.0;
  bool boost_from_average = true;
  bool reg_sqrt = false;
  int lambdarank_truncation_level = 20;
  bool lambdarank_norm = true;
  std::vector<double> label_gain;
  int objective_seed = 5;
  // device
  int hist_mode = 0;  // ours: lgbcu histogram kernel selection (0 auto)

  std::unordered_map<std::string, std::string> raw;  // everything as given (after alias resolution)

  static std::unordered_map<std::string, std::string> Str2Map(const char* parameters);
  void Set(const std::unordered_map<std::string, std::string>& params);
  std::string ToString() const;  // "[key: value]" lines for the model file's parameters block
};

}  // namespace lgb
