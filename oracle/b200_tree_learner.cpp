// TEST INFRASTRUCTURE ONLY (like ref_shim.cpp): the inner boundary of INTEGRATION.md section 2, compiled for real.
//
// `CudaB200TreeLearner` is a subclass of the REFERENCE's own `TreeLearner` interface (include/LightGBM/tree_learner.h)
// that forwards the hot path -- histogram construction, split search, row partition -- to this repository's C layer
// (include/lgbcu.h) and nothing else; it is compiled against the reference's headers where they lie under
// /root/reference and linked to the unmodified reference library (oracle/_ref/lib_lightgbm.so) and to the product
// (lightgbmv1_b200/lib_lightgbm_b200.so).  `GBDTWithB200Learner` is the reference's `GBDT` with that learner put in
// the place of the one its factory made (TreeLearner::CreateTreeLearner, src/treelearner/tree_learner.cpp:13-36, is
// where a maintainer would add the branch; it is a static function of the reference library and is not touched here).
// Everything else -- Dataset and bin mappers, objective and gradients on the host, score updater, shrinkage, model
// text -- is the reference's own code.  tests/test_inner_boundary_gpu.py trains the same Dataset through the stock
// reference learner and through this one and wants identical models.
#include <LightGBM/bin.h>
#include <LightGBM/config.h>
#include <LightGBM/dataset.h>
#include <LightGBM/objective_function.h>
#include <LightGBM/tree.h>
#include <LightGBM/tree_learner.h>
#include <LightGBM/utils/common.h>
#include <LightGBM/utils/log.h>

#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "boosting/gbdt.h"
#include "lgbcu.h"

using namespace LightGBM;

namespace {

void Need(int rc) { if (rc != 0) Log::Fatal("%s", lgbcu_last_error()); }

class CudaB200TreeLearner : public TreeLearner {
 public:
  explicit CudaB200TreeLearner(const Config* config) : config_(config) {}
  ~CudaB200TreeLearner() override { if (ctx_) lgbcu_destroy(ctx_); }

  void Init(const Dataset* train_data, bool) override {   // serial_tree_learner.cpp:28-69
    train_data_ = train_data;
    num_data_ = train_data->num_data();
    const int F = train_data->num_features();
    if (ctx_) { lgbcu_destroy(ctx_); ctx_ = nullptr; }
    Need(lgbcu_create(config_->gpu_device_id < 0 ? 0 : config_->gpu_device_id, &ctx_));
    // one stored-value column per inner feature, exactly what the reference's DenseBin holds
    std::vector<std::vector<uint8_t>> cols(F, std::vector<uint8_t>(num_data_));
    std::vector<const uint8_t*> ptrs(F);
    std::vector<LgbCuFeatureMeta> meta(F);
    for (int f = 0; f < F; ++f) {
      std::unique_ptr<BinIterator> it(train_data->FeatureIterator(f));
      it->Reset(0);
      for (data_size_t i = 0; i < num_data_; ++i) cols[f][i] = static_cast<uint8_t>(it->RawGet(i));
      ptrs[f] = cols[f].data();
      const BinMapper* m = train_data->FeatureBinMapper(f);
      if (m->num_bin() + (m->GetMostFreqBin() == 0 ? 0 : 1) > 256)
        Log::Fatal("CudaB200TreeLearner: feature %d needs more than 8-bit bins (this shim hands over byte columns)", f);
      has_categorical_ = has_categorical_ || m->bin_type() == BinType::CategoricalBin;
      meta[f].num_bin = m->num_bin(); meta[f].most_freq_bin = static_cast<int>(m->GetMostFreqBin());
      meta[f].default_bin = static_cast<int>(m->GetDefaultBin()); meta[f].missing_type = static_cast<int>(m->missing_type());
      meta[f].real_index = train_data->RealFeatureIndex(f);
      meta[f].bin_type = m->bin_type() == BinType::CategoricalBin ? LGBCU_BIN_CATEGORICAL : LGBCU_BIN_NUMERICAL; meta[f].penalty = 1.0;
    }
    Need(lgbcu_set_dataset_columns(ctx_, num_data_, F, ptrs.data(), meta.data()));
    PushParams();
    records_.resize(config_->num_leaves > 1 ? config_->num_leaves - 1 : 1);
    bag_ = nullptr; bag_cnt_ = num_data_;
  }
  void ResetIsConstantHessian(bool) override {}
  void ResetTrainingData(const Dataset* train_data, bool is_constant_hessian) override { Init(train_data, is_constant_hessian); }
  void ResetConfig(const Config* config) override { config_ = config; if (ctx_) PushParams(); }
  void SetForcedSplit(const Json*) override {}

  Tree* Train(const score_t* gradients, const score_t* hessians) override {   // serial_tree_learner.cpp:152-202
    Need(lgbcu_set_gradients(ctx_, gradients, hessians, num_data_));   // host objective: one H2D copy per tree
    Need(lgbcu_begin_tree(ctx_, bag_, bag_ ? bag_cnt_ : 0, nullptr));
    int n = 0;
    Need(lgbcu_train_tree(ctx_, records_.data(), &n));
    std::unique_ptr<Tree> tree(new Tree(config_->num_leaves, false));
    std::vector<uint32_t> cat_sets;   // the tree's categorical bin sets (8 words each)
    if (has_categorical_) {
      int32_t nsets = 0;
      Need(lgbcu_tree_cat_sets(ctx_, nullptr, 0, &nsets));
      if (nsets > 0) { cat_sets.resize(static_cast<size_t>(nsets) * 8); Need(lgbcu_tree_cat_sets(ctx_, cat_sets.data(), nsets, &nsets)); }
    }
    for (int i = 0; i < n; ++i) {
      const LgbCuSplitRecord& r = records_[i];
      const BinMapper* m = train_data_->FeatureBinMapper(r.feature);
      if (m->bin_type() == BinType::CategoricalBin) {   // SerialTreeLearner::SplitInner, categorical branch (:592-621)
        const uint32_t* set = cat_sets.data() + static_cast<size_t>(r.threshold) * 8;
        std::vector<uint32_t> bins;
        std::vector<int> cats;
        for (int b = 0; b < m->num_bin(); ++b) {
          if (!((set[b >> 5] >> (b & 31)) & 1u)) continue;
          bins.push_back(static_cast<uint32_t>(b));
          cats.push_back(static_cast<int>(train_data_->RealThreshold(r.feature, static_cast<uint32_t>(b))));
        }
        const std::vector<uint32_t> inner = Common::ConstructBitset(bins.data(), static_cast<int>(bins.size()));
        const std::vector<uint32_t> outer = Common::ConstructBitset(cats.data(), static_cast<int>(cats.size()));
        tree->SplitCategorical(r.leaf, r.feature, r.real_feature, inner.data(), static_cast<int>(inner.size()), outer.data(),
                               static_cast<int>(outer.size()), r.left_output, r.right_output, r.left_count, r.right_count,
                               r.left_sum_hessian, r.right_sum_hessian, static_cast<float>(r.gain), m->missing_type());
        continue;
      }
      tree->Split(r.leaf, r.feature, r.real_feature, static_cast<uint32_t>(r.threshold),
                  train_data_->RealThreshold(r.feature, static_cast<uint32_t>(r.threshold)), r.left_output, r.right_output,
                  r.left_count, r.right_count, r.left_sum_hessian, r.right_sum_hessian, static_cast<float>(r.gain),
                  m->missing_type(), r.default_left != 0);
    }
    return tree.release();
  }
  Tree* FitByExistingTree(const Tree*, const score_t*, const score_t*) const override {
    Log::Fatal("CudaB200TreeLearner: FitByExistingTree (refit) is outside the path");
    return nullptr;
  }
  Tree* FitByExistingTree(const Tree*, const std::vector<int>&, const score_t*, const score_t*) override {
    Log::Fatal("CudaB200TreeLearner: FitByExistingTree (refit) is outside the path");
    return nullptr;
  }
  void SetBaggingData(const Dataset* subset, const data_size_t* used_indices, data_size_t num_data) override {
    if (subset != nullptr) Log::Fatal("CudaB200TreeLearner: bagging by subset copy is outside this shim (use a bagging rate above 0.5)");
    bag_ = used_indices; bag_cnt_ = num_data;
  }
  // SerialTreeLearner::AddPredictionToScore (serial_tree_learner.h:96-111): the reference keeps the scores on the host
  // here, so the leaves' row lists come back from the device partition
  void AddPredictionToScore(const Tree* tree, double* out_score) const override {
    if (tree->num_leaves() <= 1) return;
    std::vector<int32_t> rows(num_data_);
    for (int leaf = 0; leaf < tree->num_leaves(); ++leaf) {
      int32_t cnt = 0;
      Need(lgbcu_leaf_indices(ctx_, leaf, rows.data(), num_data_, &cnt));
      const double out = tree->LeafOutput(leaf);
      for (int32_t i = 0; i < cnt; ++i) out_score[rows[i]] += out;
    }
  }
  void RenewTreeOutput(Tree*, const ObjectiveFunction* obj, std::function<double(const label_t*, int)>, data_size_t,
                       const data_size_t*, data_size_t) const override {
    if (obj != nullptr && obj->IsRenewTreeOutput()) Log::Fatal("CudaB200TreeLearner: RenewTreeOutput objectives are outside this shim");
  }

 private:
  void PushParams() {
    LgbCuTreeParams p;
    p.num_leaves = config_->num_leaves; p.max_depth = config_->max_depth; p.min_data_in_leaf = config_->min_data_in_leaf;
    p.reserved = 0; p.min_sum_hessian_in_leaf = config_->min_sum_hessian_in_leaf; p.lambda_l1 = config_->lambda_l1;
    p.lambda_l2 = config_->lambda_l2; p.min_gain_to_split = config_->min_gain_to_split;
    p.max_delta_step = config_->max_delta_step; p.path_smooth = config_->path_smooth;
    Need(lgbcu_set_tree_params(ctx_, &p));
    if (has_categorical_)
      Need(lgbcu_set_cat_params(ctx_, config_->cat_smooth, config_->cat_l2, config_->max_cat_threshold, config_->min_data_per_group,
                                config_->max_cat_to_onehot));
    if (!config_->monotone_constraints.empty()) {   // FeatureMetainfo::monotone_type (feature_histogram.hpp:1131-1138)
      const int F = train_data_->num_features();
      std::vector<int8_t> types(F);
      for (int f = 0; f < F; ++f) types[f] = config_->monotone_constraints[train_data_->RealFeatureIndex(f)];
      Need(lgbcu_set_monotone(ctx_, types.data(), F, config_->monotone_penalty));
    }
  }
  const Config* config_;
  const Dataset* train_data_ = nullptr;
  data_size_t num_data_ = 0;
  LgbCuContext* ctx_ = nullptr;
  std::vector<LgbCuSplitRecord> records_;
  const data_size_t* bag_ = nullptr; data_size_t bag_cnt_ = 0;
  bool has_categorical_ = false;
};

class GBDTWithB200Learner : public GBDT {
 public:
  void Init(const Config* config, const Dataset* train_data, const ObjectiveFunction* objective_function,
            const std::vector<const Metric*>& training_metrics) override {
    GBDT::Init(config, train_data, objective_function, training_metrics);
    tree_learner_.reset(new CudaB200TreeLearner(config_.get()));
    tree_learner_->Init(train_data_, is_constant_hessian_);
  }
};

}  // namespace

extern "C" {

// `dataset` = a DatasetHandle of the REFERENCE library (LGBM_DatasetCreateFromMat + SetField there).  Trains `iters`
// iterations with the reference's GBDT around CudaB200TreeLearner and returns the reference's model text.
int b200_shim_train(void* dataset, const char* params, int iters, char* out, long long cap, long long* out_len) {
  try {
    const Dataset* ds = reinterpret_cast<const Dataset*>(dataset);
    Config config;
    config.Set(Config::Str2Map(params));
    std::unique_ptr<ObjectiveFunction> obj(ObjectiveFunction::CreateObjectiveFunction(config.objective, config));
    if (obj) obj->Init(ds->metadata(), ds->num_data());
    GBDTWithB200Learner gbdt;
    gbdt.Init(&config, ds, obj.get(), std::vector<const Metric*>());
    for (int i = 0; i < iters; ++i) if (gbdt.TrainOneIter(nullptr, nullptr)) break;
    const std::string text = gbdt.SaveModelToString(0, -1, 0);
    *out_len = static_cast<long long>(text.size()) + 1;
    if (static_cast<long long>(text.size()) + 1 <= cap) std::memcpy(out, text.c_str(), text.size() + 1);
    return 0;
  } catch (const std::exception& e) {
    std::strncpy(out, e.what(), static_cast<size_t>(cap > 0 ? cap - 1 : 0));
    if (cap > 0) out[cap - 1] = 0;
    *out_len = -1;
    return -1;
  }
}

}  // extern "C"
