#include "config.h"

#include <algorithm>
#include <cstdlib>
#include <limits>

#include "common.h"
#include "random.h"

namespace lgb {
namespace {

// alias -> canonical key (src/io/config_auto.cpp alias table, the entries for keys this library reads)
const std::unordered_map<std::string, std::string>& Aliases() {
  static const std::unordered_map<std::string, std::string> t = {
      {"objective_type", "objective"}, {"app", "objective"}, {"application", "objective"},
      {"boosting_type", "boosting"}, {"boost", "boosting"},
      {"num_iteration", "num_iterations"}, {"n_iter", "num_iterations"}, {"num_tree", "num_iterations"},
      {"num_trees", "num_iterations"}, {"num_round", "num_iterations"}, {"num_rounds", "num_iterations"},
      {"num_boost_round", "num_iterations"}, {"n_estimators", "num_iterations"},
      {"shrinkage_rate", "learning_rate"}, {"eta", "learning_rate"},
      {"num_leaf", "num_leaves"}, {"max_leaves", "num_leaves"}, {"max_leaf", "num_leaves"},
      {"tree", "tree_learner"}, {"tree_type", "tree_learner"}, {"tree_learner_type", "tree_learner"},
      {"num_thread", "num_threads"}, {"nthread", "num_threads"}, {"nthreads", "num_threads"}, {"n_jobs", "num_threads"},
      {"device", "device_type"}, {"random_seed", "seed"}, {"random_state", "seed"},
      {"min_data_per_leaf", "min_data_in_leaf"}, {"min_data", "min_data_in_leaf"}, {"min_child_samples", "min_data_in_leaf"},
      {"min_sum_hessian_per_leaf", "min_sum_hessian_in_leaf"}, {"min_sum_hessian", "min_sum_hessian_in_leaf"},
      {"min_hessian", "min_sum_hessian_in_leaf"}, {"min_child_weight", "min_sum_hessian_in_leaf"},
      {"sub_row", "bagging_fraction"}, {"subsample", "bagging_fraction"}, {"bagging", "bagging_fraction"},
      {"pos_sub_row", "pos_bagging_fraction"}, {"pos_subsample", "pos_bagging_fraction"}, {"pos_bagging", "pos_bagging_fraction"},
      {"neg_sub_row", "neg_bagging_fraction"}, {"neg_subsample", "neg_bagging_fraction"}, {"neg_bagging", "neg_bagging_fraction"},
      {"subsample_freq", "bagging_freq"}, {"bagging_fraction_seed", "bagging_seed"},
      {"sub_feature", "feature_fraction"}, {"colsample_bytree", "feature_fraction"},
      {"sub_feature_bynode", "feature_fraction_bynode"}, {"colsample_bynode", "feature_fraction_bynode"},
      {"max_tree_output", "max_delta_step"}, {"max_leaf_output", "max_delta_step"},
      {"reg_alpha", "lambda_l1"}, {"reg_lambda", "lambda_l2"}, {"lambda", "lambda_l2"},
      {"min_split_gain", "min_gain_to_split"}, {"verbose", "verbosity"},
      {"subsample_for_bin", "bin_construct_sample_cnt"}, {"data_seed", "data_random_seed"},
      {"is_enable_bundle", "enable_bundle"}, {"bundle", "enable_bundle"},
      {"unbalance", "is_unbalance"}, {"unbalanced_sets", "is_unbalance"},
      {"metrics", "metric"}, {"metric_types", "metric"}, {"ndcg_eval_at", "eval_at"}, {"ndcg_at", "eval_at"},
      {"map_eval_at", "eval_at"}, {"map_at", "eval_at"}, {"num_machine", "num_machines"},
      {"feature_contrib", "feature_contri"}, {"fc", "feature_contri"}, {"fp", "feature_contri"}, {"feature_penalty", "feature_contri"},
      {"num_classes", "num_class"},
      {"mc", "monotone_constraints"}, {"monotone_constraint", "monotone_constraints"},
      {"monotone_constraining_method", "monotone_constraints_method"}, {"mc_method", "monotone_constraints_method"},
      {"monotone_splits_penalty", "monotone_penalty"}, {"ms_penalty", "monotone_penalty"}, {"mc_penalty", "monotone_penalty"},
      {"rate_drop", "drop_rate"},
      {"cat_feature", "categorical_feature"}, {"categorical_column", "categorical_feature"}, {"cat_column", "categorical_feature"},
  };
  return t;
}

bool GetStr(const std::unordered_map<std::string, std::string>& m, const char* k, std::string* out) {
  auto it = m.find(k);
  if (it == m.end() || it->second.empty()) return false;
  *out = it->second;
  return true;
}
bool GetInt(const std::unordered_map<std::string, std::string>& m, const char* k, int* out) {
  std::string s;
  if (!GetStr(m, k, &s)) return false;
  char* end = nullptr;
  const double v = std::strtod(s.c_str(), &end);
  if (end == s.c_str()) Log::Fatal("Parameter %s should be of type int, got \"%s\"", k, s.c_str());
  *out = static_cast<int>(v);
  return true;
}
bool GetDouble(const std::unordered_map<std::string, std::string>& m, const char* k, double* out) {
  std::string s;
  if (!GetStr(m, k, &s)) return false;
  char* end = nullptr;
  const double v = std::strtod(s.c_str(), &end);
  if (end == s.c_str()) Log::Fatal("Parameter %s should be of type double, got \"%s\"", k, s.c_str());
  *out = v;
  return true;
}
bool GetBool(const std::unordered_map<std::string, std::string>& m, const char* k, bool* out) {
  std::string s;
  if (!GetStr(m, k, &s)) return false;
  std::transform(s.begin(), s.end(), s.begin(), ::tolower);
  if (s == "false" || s == "-" || s == "0") *out = false;
  else if (s == "true" || s == "+" || s == "1") *out = true;
  else Log::Fatal("Parameter %s should be \"true\"/\"+\" or \"false\"/\"-\", got \"%s\"", k, s.c_str());
  return true;
}
// ParseMetricAlias (include/LightGBM/config.h:1116-1146)
std::string MetricAlias(const std::string& t) {
  auto in = [&](std::initializer_list<const char*> l) { for (auto* x : l) if (t == x) return true; return false; };
  if (in({"regression", "regression_l2", "l2", "mean_squared_error", "mse"})) return "l2";
  if (in({"l2_root", "root_mean_squared_error", "rmse"})) return "rmse";
  if (in({"regression_l1", "l1", "mean_absolute_error", "mae"})) return "l1";
  if (in({"binary_logloss", "binary"})) return "binary_logloss";
  if (in({"ndcg", "lambdarank", "rank_xendcg", "xendcg", "xe_ndcg", "xe_ndcg_mart", "xendcg_mart"})) return "ndcg";
  if (in({"map", "mean_average_precision"})) return "map";
  if (in({"multi_logloss", "multiclass", "softmax", "multiclassova", "multiclass_ova", "ova", "ovr"})) return "multi_logloss";
  if (in({"xentropy", "cross_entropy"})) return "cross_entropy";
  if (in({"xentlambda", "cross_entropy_lambda"})) return "cross_entropy_lambda";
  if (in({"kldiv", "kullback_leibler"})) return "kullback_leibler";
  if (in({"mean_absolute_percentage_error", "mape"})) return "mape";
  if (in({"none", "null", "custom", "na"})) return "custom";
  return t;
}

std::vector<double> ParseDoubles(const std::string& s) {
  std::vector<double> v;
  for (auto& t : SplitAny(s, ",")) v.push_back(std::strtod(t.c_str(), nullptr));
  return v;
}

// ParseObjectiveAlias (include/LightGBM/config.h:1089-1115)
std::string ObjectiveAlias(std::string t) {
  std::transform(t.begin(), t.end(), t.begin(), ::tolower);
  if (t == "regression" || t == "regression_l2" || t == "mean_squared_error" || t == "mse" || t == "l2" ||
      t == "l2_root" || t == "root_mean_squared_error" || t == "rmse") return "regression";
  if (t == "none" || t == "null" || t == "custom" || t == "na") return "custom";
  if (t == "multiclass" || t == "softmax") return "multiclass";
  if (t == "multiclassova" || t == "multiclass_ova" || t == "ova" || t == "ovr") return "multiclassova";
  if (t == "regression_l1" || t == "mean_absolute_error" || t == "l1" || t == "mae") return "regression_l1";
  if (t == "xentropy" || t == "cross_entropy") return "cross_entropy";
  if (t == "xentlambda" || t == "cross_entropy_lambda") return "cross_entropy_lambda";
  if (t == "mean_absolute_percentage_error" || t == "mape") return "mape";
  if (t == "rank_xendcg" || t == "xendcg" || t == "xe_ndcg" || t == "xe_ndcg_mart" || t == "xendcg_mart") return "rank_xendcg";
  return t;
}

}  // namespace

std::unordered_map<std::string, std::string> Config::Str2Map(const char* parameters) {
  std::unordered_map<std::string, std::string> params;
  if (!parameters) return params;
  for (auto& arg : SplitAny(parameters, " \t\n\r")) {
    const std::string kv = Trim(arg);
    const size_t eq = kv.find('=');
    if (eq == std::string::npos || kv.find('=', eq + 1) != std::string::npos) {
      if (!kv.empty()) Log::Warning("Unknown parameter %s", kv.c_str());
      continue;
    }
    std::string key = Trim(kv.substr(0, eq)), value = Trim(kv.substr(eq + 1));
    if (value.size() >= 2 && ((value.front() == '"' && value.back() == '"') || (value.front() == '\'' && value.back() == '\'')))
      value = value.substr(1, value.size() - 2);
    if (key.empty()) continue;
    if (!params.emplace(key, value).second)
      Log::Warning("%s is set=%s, %s=%s will be ignored.", key.c_str(), params[key].c_str(), key.c_str(), value.c_str());
  }
  // ParameterAlias::KeyAliasTransform: a canonical key given explicitly wins over its aliases
  std::unordered_map<std::string, std::string> out;
  for (auto& kv : params) if (!Aliases().count(kv.first)) out[kv.first] = kv.second;
  for (auto& kv : params) {
    auto it = Aliases().find(kv.first);
    if (it != Aliases().end() && !out.count(it->second)) out[it->second] = kv.second;
  }
  return out;
}

void Config::Set(const std::unordered_map<std::string, std::string>& p) {
  raw = p;
  // seed fan-out (src/io/config.cpp:190-200)
  if (GetInt(p, "seed", &seed)) {
    Random rand(seed);
    const int int_max = std::numeric_limits<int16_t>::max();
    data_random_seed = rand.NextShort(0, int_max);
    bagging_seed = rand.NextShort(0, int_max);
    drop_seed = rand.NextShort(0, int_max);
    feature_fraction_seed = rand.NextShort(0, int_max);
    objective_seed = rand.NextShort(0, int_max);
    (void)rand.NextShort(0, int_max);  // extra_seed
  }
  std::string s;
  if (GetStr(p, "boosting", &s)) {
    std::transform(s.begin(), s.end(), s.begin(), ::tolower);
    if (s == "gbdt" || s == "gbrt") boosting = "gbdt";
    else if (s == "goss") boosting = "goss";
    else if (s == "dart") boosting = "dart";
    else if (s == "rf" || s == "random_forest") boosting = "rf";
    else Log::Fatal("boosting=%s is outside this library's path (gbdt, goss, dart and rf are built)", s.c_str());
  }
  if (GetStr(p, "objective", &s)) objective = ObjectiveAlias(s);
  // GetMetricType (src/io/config.cpp:87-100): lower-case, split on ',', ParseMetricAlias, drop repeats; with no
  // `metric` given at all the objective names the metric.
  {
    std::string mv;
    const bool given = GetStr(p, "metric", &mv);
    std::transform(mv.begin(), mv.end(), mv.begin(), ::tolower);
    metric.clear();
    auto add = [&](const std::string& raw_name) {
      const std::string t = MetricAlias(raw_name);
      if (std::find(metric.begin(), metric.end(), t) == metric.end()) metric.push_back(t);
    };
    if (given) for (auto& m : SplitAny(mv, ",")) add(m);
    if (metric.empty() && mv.empty()) {
      std::string ov;
      if (GetStr(p, "objective", &ov)) {
        std::transform(ov.begin(), ov.end(), ov.begin(), ::tolower);
        for (auto& m : SplitAny(ov, ",")) add(m);
      }
    }
  }
  if (GetStr(p, "eval_at", &s)) {
    eval_at.clear();
    for (double d : ParseDoubles(s)) eval_at.push_back(static_cast<int>(d));
  }
  if (GetStr(p, "device_type", &s)) {
    std::transform(s.begin(), s.end(), s.begin(), ::tolower);
    // the reference accepts cpu|gpu (src/io/config.cpp:121-133); here every value selects the CUDA learner,
    // "cpu" included: there is no CPU path in this library.
    if (s == "cpu" || s == "gpu" || s == "cuda") device_type = "cuda";
    else Log::Fatal("Unknown device type %s", s.c_str());
  }
  if (GetStr(p, "tree_learner", &s)) {
    std::transform(s.begin(), s.end(), s.begin(), ::tolower);
    if (s == "serial") tree_learner = "serial";
    else if (s == "data" || s == "data_parallel") tree_learner = "data";
    else if (s == "feature" || s == "feature_parallel" || s == "voting" || s == "voting_parallel")
      Log::Fatal("tree_learner=%s is outside this library's path (serial|data)", s.c_str());
    else Log::Fatal("Unknown tree learner type %s", s.c_str());
  }
  GetInt(p, "num_iterations", &num_iterations);
  GetDouble(p, "learning_rate", &learning_rate);
  GetInt(p, "num_leaves", &num_leaves);
  GetInt(p, "num_threads", &num_threads);
  GetInt(p, "num_machines", &num_machines);
  GetInt(p, "gpu_device_id", &gpu_device_id);
  GetInt(p, "verbosity", &verbosity);
  GetInt(p, "num_class", &num_class);
  GetInt(p, "max_depth", &max_depth);
  GetInt(p, "min_data_in_leaf", &min_data_in_leaf);
  GetDouble(p, "min_sum_hessian_in_leaf", &min_sum_hessian_in_leaf);
  GetDouble(p, "bagging_fraction", &bagging_fraction);
  GetDouble(p, "top_rate", &top_rate);
  GetDouble(p, "other_rate", &other_rate);
  if (top_rate < 0.0 || top_rate > 1.0) Log::Fatal("top_rate should be in [0, 1]");
  if (other_rate < 0.0 || other_rate > 1.0) Log::Fatal("other_rate should be in [0, 1]");
  GetDouble(p, "pos_bagging_fraction", &pos_bagging_fraction);
  GetDouble(p, "neg_bagging_fraction", &neg_bagging_fraction);
  GetInt(p, "bagging_freq", &bagging_freq);
  GetInt(p, "bagging_seed", &bagging_seed);
  GetDouble(p, "feature_fraction", &feature_fraction);
  GetDouble(p, "feature_fraction_bynode", &feature_fraction_bynode);
  GetInt(p, "feature_fraction_seed", &feature_fraction_seed);
  GetDouble(p, "lambda_l1", &lambda_l1);
  GetDouble(p, "lambda_l2", &lambda_l2);
  GetDouble(p, "min_gain_to_split", &min_gain_to_split);
  GetDouble(p, "max_delta_step", &max_delta_step);
  GetDouble(p, "path_smooth", &path_smooth);
  GetBool(p, "extra_trees", &extra_trees);
  if (GetStr(p, "feature_contri", &s)) feature_contri = ParseDoubles(s);
  GetInt(p, "max_bin", &max_bin);
  GetInt(p, "min_data_in_bin", &min_data_in_bin);
  GetInt(p, "bin_construct_sample_cnt", &bin_construct_sample_cnt);
  GetInt(p, "data_random_seed", &data_random_seed);
  GetBool(p, "use_missing", &use_missing);
  GetBool(p, "zero_as_missing", &zero_as_missing);
  GetBool(p, "feature_pre_filter", &feature_pre_filter);
  GetBool(p, "enable_bundle", &enable_bundle);
  GetDouble(p, "sigmoid", &sigmoid);
  GetDouble(p, "alpha", &alpha);
  GetDouble(p, "fair_c", &fair_c);
  GetDouble(p, "poisson_max_delta_step", &poisson_max_delta_step);
  if (GetStr(p, "categorical_feature", &s) && !s.empty()) {   // DatasetLoader::SetHeader (dataset_loader.cpp:142-164)
    if (s.rfind("name:", 0) == 0) Log::Fatal("categorical_feature by column name needs a data file with a header; pass column indices");
    categorical_feature.clear();
    for (auto& t : SplitAny(s, ",")) {
      char* end = nullptr;
      const long v = std::strtol(t.c_str(), &end, 10);
      if (end == t.c_str() || *end != 0) Log::Fatal("categorical_feature is not a number,\nif you want to use a column name,\nplease add the prefix \"name:\" to the column name");
      categorical_feature.push_back(static_cast<int>(v));
    }
  }
  if (GetStr(p, "max_bin_by_feature", &s) && !s.empty()) {
    max_bin_by_feature.clear();
    for (double d : ParseDoubles(s)) max_bin_by_feature.push_back(static_cast<int>(d));
    for (int v : max_bin_by_feature) {
      if (v <= 1) Log::Fatal("Check failed: (*(std::min_element(max_bin_by_feature.begin(), max_bin_by_feature.end()))) > (1)");
      if (v > 65535) Log::Fatal("max_bin_by_feature should be <= 65535 (8- and 16-bit bins are built), got %d", v);
    }
  }
  GetDouble(p, "drop_rate", &drop_rate);
  GetDouble(p, "skip_drop", &skip_drop);
  GetInt(p, "max_drop", &max_drop);
  GetInt(p, "drop_seed", &drop_seed);
  GetBool(p, "xgboost_dart_mode", &xgboost_dart_mode);
  GetBool(p, "uniform_drop", &uniform_drop);
  if (drop_rate < 0.0 || drop_rate > 1.0) Log::Fatal("drop_rate should be in [0, 1]");
  if (skip_drop < 0.0 || skip_drop > 1.0) Log::Fatal("skip_drop should be in [0, 1]");
  GetInt(p, "min_data_per_group", &min_data_per_group);
  GetInt(p, "max_cat_threshold", &max_cat_threshold);
  GetInt(p, "max_cat_to_onehot", &max_cat_to_onehot);
  GetDouble(p, "cat_l2", &cat_l2);
  GetDouble(p, "cat_smooth", &cat_smooth);
  if (min_data_per_group <= 0) Log::Fatal("min_data_per_group should be > 0");
  if (max_cat_threshold <= 0) Log::Fatal("max_cat_threshold should be > 0");
  if (max_cat_to_onehot <= 0) Log::Fatal("max_cat_to_onehot should be > 0");
  if (cat_l2 < 0.0) Log::Fatal("cat_l2 should be >= 0");
  if (cat_smooth < 0.0) Log::Fatal("cat_smooth should be >= 0");
  GetDouble(p, "tweedie_variance_power", &tweedie_variance_power);
  if (tweedie_variance_power < 1.0 || tweedie_variance_power >= 2.0) Log::Fatal("tweedie_variance_power should be in [1, 2)");
  if (fair_c <= 0.0) Log::Fatal("fair_c should be > 0");
  if (poisson_max_delta_step <= 0.0) Log::Fatal("poisson_max_delta_step should be > 0");
  if (alpha <= 0.0) Log::Fatal("alpha should be > 0");
  GetBool(p, "is_unbalance", &is_unbalance);
  GetDouble(p, "scale_pos_weight", &scale_pos_weight);
  GetBool(p, "boost_from_average", &boost_from_average);
  GetBool(p, "reg_sqrt", &reg_sqrt);
  GetInt(p, "lambdarank_truncation_level", &lambdarank_truncation_level);
  GetBool(p, "lambdarank_norm", &lambdarank_norm);
  if (GetStr(p, "label_gain", &s)) label_gain = ParseDoubles(s);
  GetInt(p, "objective_seed", &objective_seed);
  GetInt(p, "hist_mode", &hist_mode);

  // range checks the reference's config_auto.cpp performs for these keys
  if (num_leaves <= 1 || num_leaves > 131072) Log::Fatal("num_leaves should be in (1, 131072], got %d", num_leaves);
  if (learning_rate <= 0.0) Log::Fatal("learning_rate should be > 0");
  if (max_bin <= 1) Log::Fatal("max_bin should be > 1");
  if (max_bin > 65535) Log::Fatal("max_bin should be <= 65535 (8- and 16-bit bins are built), got %d", max_bin);
  if (min_data_in_leaf < 0) Log::Fatal("min_data_in_leaf should be >= 0");
  if (min_sum_hessian_in_leaf < 0) Log::Fatal("min_sum_hessian_in_leaf should be >= 0");
  if (bagging_fraction <= 0.0 || bagging_fraction > 1.0) Log::Fatal("bagging_fraction should be in (0, 1]");
  if (feature_fraction <= 0.0 || feature_fraction > 1.0) Log::Fatal("feature_fraction should be in (0, 1]");
  if (lambda_l1 < 0 || lambda_l2 < 0) Log::Fatal("lambda_l1 / lambda_l2 should be >= 0");
  if (min_gain_to_split < 0) Log::Fatal("min_gain_to_split should be >= 0");
  if (bin_construct_sample_cnt <= 0) Log::Fatal("bin_construct_sample_cnt should be > 0");
  // CheckParamConflict (src/io/config.cpp:249-273): objective, metric and num_class must agree
  if (objective == "multiclass" || objective == "multiclassova") {
    if (num_class <= 1) Log::Fatal("Number of classes should be specified and greater than 1 for multiclass training");
  } else {
    if (num_class != 1) Log::Fatal("Number of classes must be 1 for non-multiclass training");
  }
  GetInt(p, "multi_error_top_k", &multi_error_top_k);
  if (multi_error_top_k <= 0) Log::Fatal("multi_error_top_k should be > 0");
  // CheckParamConflict (src/io/config.cpp:275-287): one machine => serial learner
  if (num_machines <= 1) tree_learner = "serial";
  if (max_depth > 0) {  // config.cpp:309-320: cap num_leaves by the depth
    const double full = std::pow(2, max_depth);
    if (full < num_leaves) num_leaves = static_cast<int>(full);
  }
  if (path_smooth > kEpsilon && min_data_in_leaf < 2) min_data_in_leaf = 2;  // config.cpp:330-333
  // options that would change the trees but are not built: refuse rather than silently differ
  if (extra_trees) Log::Fatal("extra_trees is outside this library's path");
  if (feature_fraction_bynode < 1.0) Log::Fatal("feature_fraction_bynode < 1 is outside this library's path");
  if (pos_bagging_fraction <= 0.0 || pos_bagging_fraction > 1.0) Log::Fatal("pos_bagging_fraction should be in (0, 1]");
  if (neg_bagging_fraction <= 0.0 || neg_bagging_fraction > 1.0) Log::Fatal("neg_bagging_fraction should be in (0, 1]");
  if (GetStr(p, "monotone_constraints", &s) && !s.empty()) {
    monotone_constraints.clear();
    for (auto& t : SplitAny(s, ",")) monotone_constraints.push_back(static_cast<int8_t>(std::atoi(t.c_str())));
    for (int8_t v : monotone_constraints) if (v < -1 || v > 1) Log::Fatal("monotone_constraints should be -1, 0 or 1");
    // (an all-zero list still selects the constrained code path in the reference -- USE_MC = !empty(), feature_histogram.hpp:
    // 117 -- whose gains are formed from the outputs; kept as given)
  }
  if (GetStr(p, "monotone_constraints_method", &s) && s != "basic")
    Log::Fatal("monotone_constraints_method=%s is outside this library's path (basic is built)", s.c_str());
  GetDouble(p, "monotone_penalty", &monotone_penalty);
  if (monotone_penalty < 0.0) Log::Fatal("monotone_penalty should be >= 0");
  for (const char* k : {"forcedsplits_filename", "forcedbins_filename", "interaction_constraints",
                        "cegb_penalty_feature_lazy", "cegb_penalty_feature_coupled",
                        "linear_tree"}) {
    if (GetStr(p, k, &s)) Log::Fatal("parameter %s is outside this library's path", k);
  }
  Log::level() = verbosity;
}

std::string Config::ToString() const {
  std::ostringstream o;
  o << "[boosting: " << boosting << "]\n[objective: " << objective << "]\n[tree_learner: " << tree_learner << "]\n"
    << "[device_type: " << device_type << "]\n[learning_rate: " << learning_rate << "]\n[num_leaves: " << num_leaves << "]\n"
    << "[max_depth: " << max_depth << "]\n[min_data_in_leaf: " << min_data_in_leaf << "]\n"
    << "[min_sum_hessian_in_leaf: " << min_sum_hessian_in_leaf << "]\n[bagging_fraction: " << bagging_fraction << "]\n"
    << "[bagging_freq: " << bagging_freq << "]\n[bagging_seed: " << bagging_seed << "]\n"
    << "[feature_fraction: " << feature_fraction << "]\n[feature_fraction_seed: " << feature_fraction_seed << "]\n"
    << "[lambda_l1: " << lambda_l1 << "]\n[lambda_l2: " << lambda_l2 << "]\n[min_gain_to_split: " << min_gain_to_split << "]\n"
    << "[max_bin: " << max_bin << "]\n[min_data_in_bin: " << min_data_in_bin << "]\n"
    << "[bin_construct_sample_cnt: " << bin_construct_sample_cnt << "]\n[data_random_seed: " << data_random_seed << "]";
  return o.str();
}

}  // namespace lgb
