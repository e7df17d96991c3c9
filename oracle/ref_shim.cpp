// TEST INFRASTRUCTURE ONLY.  A thin C ABI over a few INTERNALS of the unmodified reference (compiled against the
// reference's own headers under /root/reference, linked to oracle/_ref/lib_lightgbm.so) so that the plain-C
// oracle (oracle/oracle.c) and the host-side binning of the product can be pinned against the reference's real
// code, function by function.  Nothing here is reachable from the product.
#include <LightGBM/bin.h>
#include <LightGBM/config.h>
#include <LightGBM/dataset.h>
#include <LightGBM/objective_function.h>
#include <LightGBM/tree.h>

#include <cstring>
#include <memory>
#include <vector>

#include "treelearner/feature_histogram.hpp"
#include "treelearner/split_info.hpp"

using namespace LightGBM;

struct ShimFeature {  // same layout as LgbCuFeatureMeta
  int32_t num_bin, most_freq_bin, default_bin, missing_type, real_index, reserved;
  double penalty;
};
struct ShimParams {  // same layout as LgbCuTreeParams
  int32_t num_leaves, max_depth, min_data_in_leaf, reserved;
  double min_sum_hessian_in_leaf, lambda_l1, lambda_l2, min_gain_to_split, max_delta_step, path_smooth;
};
struct ShimSplit {  // same layout as LgbCuSplitInfo
  int32_t feature, real_feature, threshold, default_left, left_count, right_count;
  double gain, left_output, right_output, left_sum_gradient, left_sum_hessian, right_sum_gradient, right_sum_hessian;
};

extern "C" {

int ref_num_inner_features(void* ds) { return reinterpret_cast<Dataset*>(ds)->num_features(); }

int ref_feature_info(void* ds, int inner, int32_t* meta5, double* bounds, int cap) {
  const Dataset* d = reinterpret_cast<Dataset*>(ds);
  const BinMapper* m = d->FeatureBinMapper(inner);
  meta5[0] = m->num_bin(); meta5[1] = static_cast<int>(m->GetMostFreqBin()); meta5[2] = static_cast<int>(m->GetDefaultBin());
  meta5[3] = static_cast<int>(m->missing_type()); meta5[4] = d->RealFeatureIndex(inner);
  for (int i = 0; i < m->num_bin() && i < cap; ++i) bounds[i] = m->BinToValue(i);
  return 0;
}

// stored (raw) column of one inner feature, exactly what DenseBin holds
int ref_feature_stored_column(void* ds, int inner, uint8_t* out) {
  const Dataset* d = reinterpret_cast<Dataset*>(ds);
  std::unique_ptr<BinIterator> it(d->FeatureIterator(inner));
  it->Reset(0);
  for (data_size_t i = 0; i < d->num_data(); ++i) out[i] = static_cast<uint8_t>(it->RawGet(i));
  return 0;
}

// bin of every row in the FEATURE's space, as the split and the tree walk see it (BinIterator::Get: a stored value
// outside the feature's range of a shared column reads as the most frequent bin), and the group of the feature
int ref_feature_bin_column(void* ds, int inner, int32_t* out) {
  const Dataset* d = reinterpret_cast<Dataset*>(ds);
  std::unique_ptr<BinIterator> it(d->FeatureIterator(inner));
  it->Reset(0);
  for (data_size_t i = 0; i < d->num_data(); ++i) out[i] = static_cast<int32_t>(it->Get(i));
  return 0;
}
int ref_feature_group(void* ds, int inner) { return reinterpret_cast<Dataset*>(ds)->Feature2Group(inner); }
int ref_num_groups(void* ds) { return reinterpret_cast<Dataset*>(ds)->num_feature_groups(); }

static Bin* make_dense(const uint8_t* stored, int n_total, int num_total_bin) {
  Bin* b = Bin::CreateDenseBin(n_total, num_total_bin);
  for (int i = 0; i < n_total; ++i) b->Push(0, i, stored[i]);
  b->FinishLoad();
  return b;
}

// DenseBin<uint8_t>::ConstructHistogram (src/io/dense_bin.hpp:97-170) over one stored column
int ref_dense_hist(const uint8_t* stored, int n_total, int num_total_bin, const int* idx, int n, const float* ordered_grad,
                   const float* ordered_hess, double* out) {
  std::unique_ptr<Bin> b(make_dense(stored, n_total, num_total_bin));
  std::memset(out, 0, sizeof(double) * 2 * num_total_bin);
  if (idx) b->ConstructHistogram(idx, 0, n, ordered_grad, ordered_hess, out);
  else b->ConstructHistogram(0, n, ordered_grad, ordered_hess, out);
  return 0;
}

// DenseBin::Split as FeatureGroup::Split calls it for a single-feature group (feature_group.h:231-250)
int ref_dense_split(const uint8_t* stored, int n_total, const ShimFeature* f, int threshold, int default_left,
                    const int* idx, int n, int* lte, int* gt) {
  const int num_total_bin = 1 + f->num_bin - (f->most_freq_bin == 0 ? 1 : 0);
  std::unique_ptr<Bin> b(make_dense(stored, n_total, num_total_bin));
  const uint32_t max_bin = num_total_bin - 1;
  return b->Split(max_bin, f->default_bin, f->most_freq_bin, static_cast<MissingType>(f->missing_type), default_left != 0,
                  static_cast<uint32_t>(threshold), idx, n, lte, gt);
}

// FeatureHistogram::FindBestThreshold (src/treelearner/feature_histogram.hpp:85-95)
int ref_find_best_threshold(double* data, const ShimFeature* f, const ShimParams* p, double sum_g, double sum_h,
                            int num_data, double parent_output, ShimSplit* out, int* is_splittable) {
  Config cfg;
  cfg.min_data_in_leaf = p->min_data_in_leaf; cfg.min_sum_hessian_in_leaf = p->min_sum_hessian_in_leaf;
  cfg.lambda_l1 = p->lambda_l1; cfg.lambda_l2 = p->lambda_l2; cfg.min_gain_to_split = p->min_gain_to_split;
  cfg.max_delta_step = p->max_delta_step; cfg.path_smooth = p->path_smooth;
  FeatureMetainfo meta;
  meta.num_bin = f->num_bin; meta.missing_type = static_cast<MissingType>(f->missing_type);
  meta.offset = f->most_freq_bin == 0 ? 1 : 0; meta.default_bin = f->default_bin; meta.monotone_type = 0;
  meta.penalty = f->penalty; meta.config = &cfg; meta.bin_type = BinType::NumericalBin; meta.rand = Random(0);
  FeatureHistogram fh;
  fh.Init(data, &meta);
  SplitInfo si;
  ConstraintEntry ce;
  fh.FindBestThreshold(sum_g, sum_h, num_data, ce, parent_output, &si);
  out->feature = -1; out->real_feature = -1; out->threshold = static_cast<int>(si.threshold);
  out->default_left = si.default_left; out->left_count = si.left_count; out->right_count = si.right_count;
  out->gain = si.gain; out->left_output = si.left_output; out->right_output = si.right_output;
  out->left_sum_gradient = si.left_sum_gradient; out->left_sum_hessian = si.left_sum_hessian;
  out->right_sum_gradient = si.right_sum_gradient; out->right_sum_hessian = si.right_sum_hessian;
  *is_splittable = fh.is_splittable() ? 1 : 0;
  return 0;
}

// ObjectiveFunction::GetGradients through the reference's factory (src/objective/objective_function.cpp)
int ref_gradients(const char* params, int n, const float* label, const float* weight, const int* query_sizes,
                  int num_queries, const double* score, float* grad, float* hess) {
  try {
    Config cfg;
    cfg.Set(Config::Str2Map(params));
    Metadata md;
    md.Init(n, weight ? 0 : -1, query_sizes ? 0 : -1);
    md.SetLabel(label, n);
    if (weight) md.SetWeights(weight, n);
    if (query_sizes) md.SetQuery(query_sizes, num_queries);
    std::unique_ptr<ObjectiveFunction> obj(ObjectiveFunction::CreateObjectiveFunction(cfg.objective, cfg));
    obj->Init(md, n);
    obj->GetGradients(score, grad, hess);
  } catch (...) {
    return -1;
  }
  return 0;
}

}  // extern "C"
