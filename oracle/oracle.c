/*
 * oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle.h).  Sequential plain-C restatement of the reference's
 * CPU serial tree learner hot path.  Deliberately simple: one thread, same loop order and the same double /
 * float arithmetic as the reference so that results can be compared number by number.
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* include/LightGBM/meta.h:50-54 : score_t (float) constants used in double expressions */
static const float kEpsilonF = 1e-15f;
#define K_EPS ((double)kEpsilonF)
#define K_MIN_SCORE (-INFINITY)

static int feat_offset(const LgbCuFeatureMeta* f) { return f->most_freq_bin == 0 ? 1 : 0; }
static int feat_nslots(const LgbCuFeatureMeta* f) { return 1 + f->num_bin - feat_offset(f); }

int orc_total_bin(const LgbCuFeatureMeta* feats, int num_features, int* hist_base) {
  int total = 0;
  for (int f = 0; f < num_features; ++f) {
    if (hist_base) hist_base[f] = total;
    total += feat_nslots(&feats[f]);
  }
  return total;
}

/* src/io/dataset.cpp:1282-1383 (col-wise branch) + src/io/dense_bin.hpp:97-140 */
void orc_hist_build(const uint8_t* bins, int num_features, const LgbCuFeatureMeta* feats, const int32_t* indices,
                    int n, const float* grad, const float* hess, double* out) {
  int* base = (int*)malloc(sizeof(int) * (size_t)num_features);
  const int total = orc_total_bin(feats, num_features, base);
  memset(out, 0, sizeof(double) * 2 * (size_t)total);
  /* ordered gradients: dataset.cpp:1318-1334 */
  float* og = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
  float* oh = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));
  for (int i = 0; i < n; ++i) {
    const int r = indices ? indices[i] : i;
    og[i] = grad[r];
    oh[i] = hess[r];
  }
  for (int f = 0; f < num_features; ++f) {
    double* h = out + 2 * (size_t)base[f];
    for (int i = 0; i < n; ++i) {
      const int r = indices ? indices[i] : i;
      const uint32_t ti = ((uint32_t)bins[(size_t)r * num_features + f]) << 1;
      h[ti] += og[i];
      h[ti + 1] += oh[i];
    }
  }
  free(og); free(oh); free(base);
}

/* src/io/dataset.cpp:1410-1428 ; data = hist + 2*(base+1) */
static void fix_one(const LgbCuFeatureMeta* f, double sum_g, double sum_h, double* data) {
  const int mfb = f->most_freq_bin;
  if (mfb > 0) {
    data[2 * mfb] = sum_g;
    data[2 * mfb + 1] = sum_h;
    for (int i = 0; i < f->num_bin; ++i) {
      if (i != mfb) {
        data[2 * mfb] -= data[2 * i];
        data[2 * mfb + 1] -= data[2 * i + 1];
      }
    }
  }
}

void orc_hist_fix(const LgbCuFeatureMeta* feats, int num_features, double sum_g, double sum_h, double* hist) {
  int base = 0;
  for (int f = 0; f < num_features; ++f) {
    fix_one(&feats[f], sum_g, sum_h, hist + 2 * (size_t)(base + 1));
    base += feat_nslots(&feats[f]);
  }
}

/* ---- gain math: feature_histogram.hpp:734-851 ---- */
static int sign_d(double x) { return (x > 0.0) - (x < 0.0); }
static double threshold_l1(double s, double l1) {
  const double reg_s = fmax(0.0, fabs(s) - l1);
  return sign_d(s) * reg_s;
}
typedef struct { int use_l1, use_max_output, use_smoothing; } Flags;
static Flags make_flags(const LgbCuTreeParams* p) {
  Flags f;
  f.use_l1 = p->lambda_l1 > 0;
  f.use_max_output = p->max_delta_step > 0;
  f.use_smoothing = p->path_smooth > K_EPS;
  return f;
}
static double leaf_output(double g, double h, const LgbCuTreeParams* p, Flags fl, int n, double parent_output) {
  double ret;
  if (fl.use_l1) ret = -threshold_l1(g, p->lambda_l1) / (h + p->lambda_l2);
  else ret = -g / (h + p->lambda_l2);
  if (fl.use_max_output) {
    if (p->max_delta_step > 0 && fabs(ret) > p->max_delta_step) ret = sign_d(ret) * p->max_delta_step;
  }
  if (fl.use_smoothing) {
    ret = ret * (n / p->path_smooth) / (n / p->path_smooth + 1) + parent_output / (n / p->path_smooth + 1);
  }
  return ret;
}
static double leaf_gain_given_output(double g, double h, const LgbCuTreeParams* p, Flags fl, double out) {
  if (fl.use_l1) {
    const double sg = threshold_l1(g, p->lambda_l1);
    return -(2.0 * sg * out + (h + p->lambda_l2) * out * out);
  }
  return -(2.0 * g * out + (h + p->lambda_l2) * out * out);
}
static double leaf_gain(double g, double h, const LgbCuTreeParams* p, Flags fl, int n, double parent_output) {
  if (!fl.use_max_output && !fl.use_smoothing) {
    if (fl.use_l1) {
      const double sg = threshold_l1(g, p->lambda_l1);
      return (sg * sg) / (h + p->lambda_l2);
    }
    return (g * g) / (h + p->lambda_l2);
  }
  return leaf_gain_given_output(g, h, p, fl, leaf_output(g, h, p, fl, n, parent_output));
}
static double split_gains(double lg, double lh, double rg, double rh, const LgbCuTreeParams* p, Flags fl, int lc, int rc,
                          double parent_output) {
  return leaf_gain(lg, lh, p, fl, lc, parent_output) + leaf_gain(rg, rh, p, fl, rc, parent_output);
}
static int round_int(double x) { return (int)(x + 0.5f); } /* utils/common.h:985-987 */

typedef struct {
  const double* data; const LgbCuFeatureMeta* meta; const LgbCuTreeParams* p; Flags fl;
  int is_splittable;
} FH;

/* feature_histogram.hpp:853-1049 */
static void scan_sequential(FH* fh, int reverse, int skip_default_bin, int na_as_missing, double sum_gradient,
                            double sum_hessian, int num_data, double min_gain_shift, double parent_output,
                            LgbCuSplitInfo* output) {
  const LgbCuFeatureMeta* m = fh->meta;
  const LgbCuTreeParams* p = fh->p;
  const int offset = feat_offset(m);
  double best_sum_left_gradient = NAN, best_sum_left_hessian = NAN, best_gain = K_MIN_SCORE;
  int best_left_count = 0;
  uint32_t best_threshold = (uint32_t)m->num_bin;
  const double cnt_factor = num_data / sum_hessian;
  const double* d = fh->data;
  if (reverse) {
    double sum_right_gradient = 0.0f, sum_right_hessian = K_EPS;
    int right_count = 0;
    int t = m->num_bin - 1 - offset - na_as_missing;
    const int t_end = 1 - offset;
    for (; t >= t_end; --t) {
      if (skip_default_bin && (t + offset) == m->default_bin) continue;
      const double grad = d[2 * t], hess = d[2 * t + 1];
      const int cnt = round_int(hess * cnt_factor);
      sum_right_gradient += grad;
      sum_right_hessian += hess;
      right_count += cnt;
      if (right_count < p->min_data_in_leaf || sum_right_hessian < p->min_sum_hessian_in_leaf) continue;
      const int left_count = num_data - right_count;
      if (left_count < p->min_data_in_leaf) break;
      const double sum_left_hessian = sum_hessian - sum_right_hessian;
      if (sum_left_hessian < p->min_sum_hessian_in_leaf) break;
      const double sum_left_gradient = sum_gradient - sum_right_gradient;
      const double current_gain = split_gains(sum_left_gradient, sum_left_hessian, sum_right_gradient,
                                              sum_right_hessian, p, fh->fl, left_count, right_count, parent_output);
      if (current_gain <= min_gain_shift) continue;
      fh->is_splittable = 1;
      if (current_gain > best_gain) {
        best_left_count = left_count;
        best_sum_left_gradient = sum_left_gradient;
        best_sum_left_hessian = sum_left_hessian;
        best_threshold = (uint32_t)(t - 1 + offset);
        best_gain = current_gain;
      }
    }
  } else {
    double sum_left_gradient = 0.0f, sum_left_hessian = K_EPS;
    int left_count = 0;
    int t = 0;
    const int t_end = m->num_bin - 2 - offset;
    if (na_as_missing && offset == 1) {
      sum_left_gradient = sum_gradient;
      sum_left_hessian = sum_hessian - K_EPS;
      left_count = num_data;
      for (int i = 0; i < m->num_bin - offset; ++i) {
        const double grad = d[2 * i], hess = d[2 * i + 1];
        const int cnt = round_int(hess * cnt_factor);
        sum_left_gradient -= grad;
        sum_left_hessian -= hess;
        left_count -= cnt;
      }
      t = -1;
    }
    for (; t <= t_end; ++t) {
      if (skip_default_bin && (t + offset) == m->default_bin) continue;
      if (t >= 0) {
        sum_left_gradient += d[2 * t];
        sum_left_hessian += d[2 * t + 1];
        left_count += round_int(d[2 * t + 1] * cnt_factor);
      }
      if (left_count < p->min_data_in_leaf || sum_left_hessian < p->min_sum_hessian_in_leaf) continue;
      const int right_count = num_data - left_count;
      if (right_count < p->min_data_in_leaf) break;
      const double sum_right_hessian = sum_hessian - sum_left_hessian;
      if (sum_right_hessian < p->min_sum_hessian_in_leaf) break;
      const double sum_right_gradient = sum_gradient - sum_left_gradient;
      const double current_gain = split_gains(sum_left_gradient, sum_left_hessian, sum_right_gradient,
                                              sum_right_hessian, p, fh->fl, left_count, right_count, parent_output);
      if (current_gain <= min_gain_shift) continue;
      fh->is_splittable = 1;
      if (current_gain > best_gain) {
        best_left_count = left_count;
        best_sum_left_gradient = sum_left_gradient;
        best_sum_left_hessian = sum_left_hessian;
        best_threshold = (uint32_t)(t + offset);
        best_gain = current_gain;
      }
    }
  }
  if (fh->is_splittable && best_gain > output->gain + min_gain_shift) {
    output->threshold = (int32_t)best_threshold;
    output->left_output = leaf_output(best_sum_left_gradient, best_sum_left_hessian, p, fh->fl, best_left_count, parent_output);
    output->left_count = best_left_count;
    output->left_sum_gradient = best_sum_left_gradient;
    output->left_sum_hessian = best_sum_left_hessian - K_EPS;
    output->right_output = leaf_output(sum_gradient - best_sum_left_gradient, sum_hessian - best_sum_left_hessian, p,
                                       fh->fl, num_data - best_left_count, parent_output);
    output->right_count = num_data - best_left_count;
    output->right_sum_gradient = sum_gradient - best_sum_left_gradient;
    output->right_sum_hessian = sum_hessian - best_sum_left_hessian - K_EPS;
    output->gain = best_gain - min_gain_shift;
    output->default_left = reverse;
  }
}

/* feature_histogram.hpp:85-95 + FuncForNumricalL3 (:156-217) + BeforeNumercal (:97-113) */
void orc_find_best_threshold(const double* data, const LgbCuFeatureMeta* feat, const LgbCuTreeParams* p,
                             double sum_g, double sum_h_in, int num_data, double parent_output,
                             LgbCuSplitInfo* out, int* is_splittable) {
  FH fh;
  fh.data = data; fh.meta = feat; fh.p = p; fh.fl = make_flags(p);
  memset(out, 0, sizeof(*out));
  out->feature = -1; out->real_feature = -1;
  out->default_left = 1;
  out->gain = K_MIN_SCORE;
  const double sum_h = sum_h_in + 2 * kEpsilonF;
  fh.is_splittable = 0;
  const double min_gain_shift = leaf_gain(sum_g, sum_h, p, fh.fl, num_data, parent_output) + p->min_gain_to_split;
  if (feat->num_bin > 2 && feat->missing_type != LGBCU_MISSING_NONE) {
    if (feat->missing_type == LGBCU_MISSING_ZERO) {
      scan_sequential(&fh, 1, 1, 0, sum_g, sum_h, num_data, min_gain_shift, parent_output, out);
      scan_sequential(&fh, 0, 1, 0, sum_g, sum_h, num_data, min_gain_shift, parent_output, out);
    } else {
      scan_sequential(&fh, 1, 0, 1, sum_g, sum_h, num_data, min_gain_shift, parent_output, out);
      scan_sequential(&fh, 0, 0, 1, sum_g, sum_h, num_data, min_gain_shift, parent_output, out);
    }
  } else {
    scan_sequential(&fh, 1, 0, 0, sum_g, sum_h, num_data, min_gain_shift, parent_output, out);
    if (feat->missing_type == LGBCU_MISSING_NAN) out->default_left = 0;
  }
  out->gain *= feat->penalty;
  *is_splittable = fh.is_splittable;
}

/* split_info.hpp:126-153 */
static int split_greater(const LgbCuSplitInfo* a, const LgbCuSplitInfo* b) {
  int fa = a->real_feature, fb = b->real_feature;
  if (fa == -1) fa = INT32_MAX;
  if (fb == -1) fb = INT32_MAX;
  if (a->gain != b->gain) return a->gain > b->gain;
  return fa < fb;
}

static void reset_split(LgbCuSplitInfo* s) {
  memset(s, 0, sizeof(*s));
  s->feature = -1; s->real_feature = -1; s->gain = K_MIN_SCORE; s->default_left = 1;
}

/* serial_tree_learner.cpp:358-425 + :701-736 for one leaf whose histogram still needs FixHistogram.
 * is_root: the reference's test is `leaf_splits->leaf_index() == 0` (:710) -- true for the root AND for every later
 * leaf that still carries id 0 (the left child keeps its parent's id), whose "parent output" is therefore its own
 * unsmoothed output instead of the output its parent's split assigned.  Only matters with path_smooth. */
void orc_find_best_split(double* hist, const LgbCuFeatureMeta* feats, int num_features, const int8_t* feature_used,
                         const LgbCuTreeParams* p, double sum_g, double sum_h, int num_data, int is_root,
                         double leaf_out, LgbCuSplitInfo* out, int8_t* splittable) {
  const Flags fl = make_flags(p);
  LgbCuSplitInfo best;
  reset_split(&best);
  int base = 0;
  for (int f = 0; f < num_features; ++f) {
    double* data = hist + 2 * (size_t)(base + 1);
    base += feat_nslots(&feats[f]);
    if (feature_used && !feature_used[f]) continue;
    fix_one(&feats[f], sum_g, sum_h, data);
    double parent_output;
    if (is_root) {
      Flags f0 = fl; f0.use_smoothing = 0;
      parent_output = leaf_output(sum_g, sum_h, p, f0, num_data, 0);
    } else {
      parent_output = leaf_out;
    }
    LgbCuSplitInfo s; int sp = 0;
    orc_find_best_threshold(data, &feats[f], p, sum_g, sum_h, num_data, parent_output, &s, &sp);
    s.feature = f; s.real_feature = feats[f].real_index;
    if (splittable) splittable[f] = (int8_t)sp;
    if (split_greater(&s, &best)) best = s;
  }
  *out = best;
}

/* feature_group.h:231-277 (num_feature_ == 1 branch) -> dense_bin.hpp:282-307 -> SplitInner<...,USE_MIN_BIN=false> (:173-253) */
int orc_partition(const uint8_t* bins, int num_features, const LgbCuFeatureMeta* feats, int feature, int threshold,
                  int default_left, const int32_t* indices, int n, int32_t* lte, int32_t* gt) {
  const LgbCuFeatureMeta* m = &feats[feature];
  const uint32_t min_bin = 1;
  const uint32_t max_bin = (uint32_t)feat_nslots(m) - 1;
  const uint32_t default_bin = (uint32_t)m->default_bin, most_freq_bin = (uint32_t)m->most_freq_bin;
  int miss_is_zero = 0, miss_is_na = 0, mfb_is_zero = 0, mfb_is_na = 0;
  if (m->missing_type == LGBCU_MISSING_ZERO) {
    miss_is_zero = 1;
    if (default_bin == most_freq_bin) mfb_is_zero = 1;
  } else if (m->missing_type == LGBCU_MISSING_NAN) {
    miss_is_na = 1;
    if (max_bin == most_freq_bin + 1 && most_freq_bin > 0) mfb_is_na = 1;
  }
  uint8_t th = (uint8_t)(threshold + min_bin);
  uint8_t t_zero_bin = (uint8_t)(min_bin + default_bin);
  if (most_freq_bin == 0) { --th; --t_zero_bin; }
  const uint8_t maxb = (uint8_t)max_bin;
  int lte_count = 0, gt_count = 0;
  int32_t* default_indices = gt; int* default_count = &gt_count;
  int32_t* missing_default_indices = gt; int* missing_default_count = &gt_count;
  if (most_freq_bin <= (uint32_t)threshold) { default_indices = lte; default_count = &lte_count; }
  if (miss_is_zero || miss_is_na) {
    if (default_left) { missing_default_indices = lte; missing_default_count = &lte_count; }
  }
  if (min_bin < max_bin) {
    for (int i = 0; i < n; ++i) {
      const int32_t idx = indices[i];
      const uint8_t bin = bins[(size_t)idx * num_features + feature];
      if ((miss_is_zero && !mfb_is_zero && bin == t_zero_bin) || (miss_is_na && !mfb_is_na && bin == maxb)) {
        missing_default_indices[(*missing_default_count)++] = idx;
      } else if (bin == 0) {
        if ((miss_is_na && mfb_is_na) || (miss_is_zero && mfb_is_zero)) missing_default_indices[(*missing_default_count)++] = idx;
        else default_indices[(*default_count)++] = idx;
      } else if (bin > th) {
        gt[gt_count++] = idx;
      } else {
        lte[lte_count++] = idx;
      }
    }
  } else {
    int32_t* max_bin_indices = gt; int* max_bin_count = &gt_count;
    if (maxb <= th) { max_bin_indices = lte; max_bin_count = &lte_count; }
    for (int i = 0; i < n; ++i) {
      const int32_t idx = indices[i];
      const uint8_t bin = bins[(size_t)idx * num_features + feature];
      if (miss_is_zero && !mfb_is_zero && bin == t_zero_bin) {
        missing_default_indices[(*missing_default_count)++] = idx;
      } else if (bin != maxb) {
        if ((miss_is_na && mfb_is_na) || (miss_is_zero && mfb_is_zero)) missing_default_indices[(*missing_default_count)++] = idx;
        else default_indices[(*default_count)++] = idx;
      } else {
        if (miss_is_na && !mfb_is_na) missing_default_indices[(*missing_default_count)++] = idx;
        else max_bin_indices[(*max_bin_count)++] = idx;
      }
    }
  }
  return lte_count;
}

/* ---------------- SerialTreeLearner::Train ---------------- */
typedef struct {
  int begin, count, slot, depth;
  double sum_g, sum_h, output;
} Leaf;

int orc_train_tree(const uint8_t* bins, int num_rows_total, int num_features, const LgbCuFeatureMeta* feats,
                   const LgbCuTreeParams* p, const float* grad, const float* hess, const int32_t* bag_indices,
                   int bag_count, const int8_t* feature_used, LgbCuSplitRecord* records, int32_t* leaf_of_row) {
  const int L = p->num_leaves;
  const int total_bin = orc_total_bin(feats, num_features, NULL);
  const int n_root = bag_indices ? bag_count : num_rows_total;
  int32_t* indices = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_root > 0 ? n_root : 1));
  int32_t* tmp_l = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_root > 0 ? n_root : 1));
  int32_t* tmp_r = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_root > 0 ? n_root : 1));
  for (int i = 0; i < n_root; ++i) indices[i] = bag_indices ? bag_indices[i] : i;
  double* pool = (double*)calloc((size_t)L * 2 * (size_t)total_bin, sizeof(double));
  int8_t* splittable = (int8_t*)calloc((size_t)L * (size_t)num_features, 1);
  Leaf* leaves = (Leaf*)calloc((size_t)L, sizeof(Leaf));
  LgbCuSplitInfo* best = (LgbCuSplitInfo*)malloc(sizeof(LgbCuSplitInfo) * (size_t)L);
  int8_t* used = (int8_t*)malloc((size_t)num_features);
  for (int l = 0; l < L; ++l) { reset_split(&best[l]); leaves[l].slot = l; }
  /* BeforeTrain: root sums (leaf_splits.hpp:66-79 / :84-100 for a bag) */
  double sg = 0.0f, sh = 0.0f;
  for (int i = 0; i < n_root; ++i) { sg += grad[indices[i]]; sh += hess[indices[i]]; }
  leaves[0].begin = 0; leaves[0].count = n_root; leaves[0].sum_g = sg; leaves[0].sum_h = sh; leaves[0].output = 0;
  int num_leaves = 1, num_splits = 0;
  int left_leaf = 0, right_leaf = -1;
  for (int split = 0; split < L - 1; ++split) {
    /* BeforeFindBestSplit (serial_tree_learner.cpp:274-314) */
    int go = 1;
    if (p->max_depth > 0 && leaves[left_leaf].depth >= p->max_depth) go = 0;
    const int nl = leaves[left_leaf].count, nr = right_leaf >= 0 ? leaves[right_leaf].count : 0;
    if (go && nr < p->min_data_in_leaf * 2 && nl < p->min_data_in_leaf * 2) go = 0;
    if (!go) {
      best[left_leaf].gain = K_MIN_SCORE;
      if (right_leaf >= 0) best[right_leaf].gain = K_MIN_SCORE;
    } else {
      int smaller, larger;
      if (right_leaf < 0) { smaller = left_leaf; larger = -1; }
      else if (nl < nr) { smaller = left_leaf; larger = right_leaf; }
      else { smaller = right_leaf; larger = left_leaf; }
      /* FindBestSplits (:316-331): features unsplittable in the parent are skipped in both children.  The parent's
       * histogram object (and its flags) lives on in the larger child's slot. */
      for (int f = 0; f < num_features; ++f) {
        used[f] = feature_used ? (feature_used[f] != 0) : 1;
        if (used[f] && larger >= 0 && !splittable[(size_t)leaves[larger].slot * num_features + f]) {
          splittable[(size_t)leaves[smaller].slot * num_features + f] = 0;
          used[f] = 0;
        }
      }
      /* ConstructHistograms: smaller leaf only (parent is always cached: histogram_pool_size = -1) */
      double* hs = pool + (size_t)leaves[smaller].slot * 2 * total_bin;
      orc_hist_build(bins, num_features, feats, indices + leaves[smaller].begin, leaves[smaller].count, grad, hess, hs);
      int8_t* sp_s = splittable + (size_t)leaves[smaller].slot * num_features;
      {
        int8_t* spt = (int8_t*)calloc((size_t)num_features, 1);
        orc_find_best_split(hs, feats, num_features, used, p, leaves[smaller].sum_g, leaves[smaller].sum_h,
                            leaves[smaller].count, smaller == 0, leaves[smaller].output, &best[smaller], spt);
        for (int f = 0; f < num_features; ++f) if (used[f]) sp_s[f] = spt[f];
        free(spt);
      }
      if (larger >= 0) {
        double* hl = pool + (size_t)leaves[larger].slot * 2 * total_bin;
        int8_t* sp_l = splittable + (size_t)leaves[larger].slot * num_features;
        const Flags fl = make_flags(p);
        LgbCuSplitInfo bl; reset_split(&bl);
        int base = 0;
        for (int f = 0; f < num_features; ++f) {
          const int nsl = feat_nslots(&feats[f]);
          double* dl = hl + 2 * (size_t)(base + 1);
          const double* ds = hs + 2 * (size_t)(base + 1);
          base += nsl;
          if (!used[f]) continue;
          const int cnt = (feats[f].num_bin - feat_offset(&feats[f])) * 2;
          for (int i = 0; i < cnt; ++i) dl[i] -= ds[i];  /* FeatureHistogram::Subtract */
          LgbCuSplitInfo s; int sp = 0;
          double parent_output = leaves[larger].output;
          if (larger == 0) {  /* ComputeBestSplitForFeature tests leaf_index() == 0, not "is the root" (see below) */
            Flags f0 = fl; f0.use_smoothing = 0;
            parent_output = leaf_output(leaves[larger].sum_g, leaves[larger].sum_h, p, f0, leaves[larger].count, 0);
          }
          orc_find_best_threshold(dl, &feats[f], p, leaves[larger].sum_g, leaves[larger].sum_h, leaves[larger].count,
                                  parent_output, &s, &sp);
          s.feature = f; s.real_feature = feats[f].real_index;
          sp_l[f] = (int8_t)sp;
          if (split_greater(&s, &bl)) bl = s;
        }
        best[larger] = bl;
      }
    }
    /* ArgMax over leaves (first maximal) */
    int best_leaf = 0;
    for (int l = 1; l < L; ++l) if (split_greater(&best[l], &best[best_leaf])) best_leaf = l;
    const LgbCuSplitInfo bs = best[best_leaf];
    if (bs.gain <= 0.0) break;
    /* SplitInner (:541-659) */
    Leaf parent = leaves[best_leaf];
    const int lc = orc_partition(bins, num_features, feats, bs.feature, bs.threshold, bs.default_left,
                                 indices + parent.begin, parent.count, tmp_l, tmp_r);
    const int rc = parent.count - lc;
    memcpy(indices + parent.begin, tmp_l, sizeof(int32_t) * (size_t)lc);
    memcpy(indices + parent.begin + lc, tmp_r, sizeof(int32_t) * (size_t)rc);
    const int R = num_leaves;
    LgbCuSplitRecord* rec = &records[num_splits++];
    memset(rec, 0, sizeof(*rec));
    rec->leaf = best_leaf; rec->right_leaf = R; rec->feature = bs.feature; rec->real_feature = bs.real_feature;
    rec->threshold = bs.threshold; rec->default_left = bs.default_left; rec->left_count = lc; rec->right_count = rc;
    rec->gain = (float)(bs.gain + p->min_gain_to_split);
    rec->left_output = bs.left_output; rec->right_output = bs.right_output;
    rec->left_sum_hessian = bs.left_sum_hessian; rec->right_sum_hessian = bs.right_sum_hessian;
    rec->left_sum_gradient = bs.left_sum_gradient; rec->right_sum_gradient = bs.right_sum_gradient;
    Leaf A = parent, B = parent;
    A.count = lc; A.depth = parent.depth + 1; A.sum_g = bs.left_sum_gradient; A.sum_h = bs.left_sum_hessian; A.output = bs.left_output;
    B.begin = parent.begin + lc; B.count = rc; B.depth = parent.depth + 1;
    B.sum_g = bs.right_sum_gradient; B.sum_h = bs.right_sum_hessian; B.output = bs.right_output;
    /* pool hand-over (:298-312): larger child keeps the parent's histogram object, smaller gets a fresh one */
    if (lc < rc) { A.slot = R; B.slot = parent.slot; } else { B.slot = R; A.slot = parent.slot; }
    leaves[best_leaf] = A; leaves[R] = B;
    num_leaves = R + 1;
    left_leaf = best_leaf; right_leaf = R;
  }
  if (leaf_of_row) {
    for (int i = 0; i < num_rows_total; ++i) leaf_of_row[i] = -1;
    for (int l = 0; l < num_leaves; ++l)
      for (int i = 0; i < leaves[l].count; ++i) leaf_of_row[indices[leaves[l].begin + i]] = l;
  }
  free(indices); free(tmp_l); free(tmp_r); free(pool); free(splittable); free(leaves); free(best); free(used);
  return num_splits;
}

/* ---------------- objectives ---------------- */
void orc_grad_binary(int n, const double* score, const float* label, const float* weight, double sigmoid,
                     double w_neg, double w_pos, float* grad, float* hess) {
  for (int i = 0; i < n; ++i) {
    const int is_pos = label[i] > 0;
    const int lab = is_pos ? 1 : -1;
    const double lw = is_pos ? w_pos : w_neg;
    const double response = -lab * sigmoid / (1.0f + exp(lab * sigmoid * score[i]));
    const double abs_response = fabs(response);
    if (!weight) {
      grad[i] = (float)(response * lw);
      hess[i] = (float)(abs_response * (sigmoid - abs_response) * lw);
    } else {
      grad[i] = (float)(response * lw * weight[i]);
      hess[i] = (float)(abs_response * (sigmoid - abs_response) * lw * weight[i]);
    }
  }
}

void orc_grad_l2(int n, const double* score, const float* label, const float* weight, float* grad, float* hess) {
  for (int i = 0; i < n; ++i) {
    if (!weight) { grad[i] = (float)(score[i] - label[i]); hess[i] = 1.0f; }
    else { grad[i] = (float)((score[i] - label[i]) * weight[i]); hess[i] = (float)weight[i]; }
  }
}

/* rank_objective.hpp:100-257 ; dcg_calculator.cpp:43-76 */
typedef struct { const double* score; } SortCtx;
static void stable_sort_desc(int* idx, int n, const double* score, int* tmp) {
  /* bottom-up merge sort: stable, comparator score[a] > score[b] */
  for (int w = 1; w < n; w *= 2) {
    for (int lo = 0; lo < n; lo += 2 * w) {
      int mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
      int i = lo, j = mid, k = lo;
      while (i < mid && j < hi) {
        if (score[idx[j]] > score[idx[i]]) tmp[k++] = idx[j++];
        else tmp[k++] = idx[i++];
      }
      while (i < mid) tmp[k++] = idx[i++];
      while (j < hi) tmp[k++] = idx[j++];
    }
    memcpy(idx, tmp, sizeof(int) * (size_t)n);
  }
}

#define SIGMOID_BINS (1024 * 1024)

void orc_grad_lambdarank(int num_queries, const int32_t* boundaries, const double* score_all, const float* label_all,
                         const float* weight, double sigmoid, int truncation_level, int norm,
                         const double* label_gain, int num_label_gain, float* grad, float* hess) {
  /* ConstructSigmoidTable */
  double min_in = -50.0 / sigmoid / 2, max_in = -min_in;
  const double idx_factor = SIGMOID_BINS / (max_in - min_in);
  double* table = (double*)malloc(sizeof(double) * SIGMOID_BINS);
  for (size_t i = 0; i < SIGMOID_BINS; ++i) {
    const double s = i / idx_factor + min_in;
    table[i] = 1.0f / (1.0f + exp(s * sigmoid));
  }
  int maxq = 0;
  for (int q = 0; q < num_queries; ++q) if (boundaries[q + 1] - boundaries[q] > maxq) maxq = boundaries[q + 1] - boundaries[q];
  int* sorted = (int*)malloc(sizeof(int) * (size_t)(maxq + 1));
  int* tmp = (int*)malloc(sizeof(int) * (size_t)(maxq + 1));
  int* label_cnt = (int*)malloc(sizeof(int) * (size_t)num_label_gain);
  for (int q = 0; q < num_queries; ++q) {
    const int start = boundaries[q], cnt = boundaries[q + 1] - boundaries[q];
    const float* label = label_all + start;
    const double* score = score_all + start;
    float* lambdas = grad + start;
    float* hessians = hess + start;
    /* CalMaxDCGAtK */
    double max_dcg = 0.0f;
    {
      for (int i = 0; i < num_label_gain; ++i) label_cnt[i] = 0;
      for (int i = 0; i < cnt; ++i) ++label_cnt[(int)label[i]];
      int top_label = num_label_gain - 1;
      int k = truncation_level; if (k > cnt) k = cnt;
      for (int j = 0; j < k; ++j) {
        while (top_label > 0 && label_cnt[top_label] <= 0) top_label -= 1;
        if (top_label < 0) break;
        max_dcg += (1.0 / log2(2.0 + j)) * label_gain[top_label];
        label_cnt[top_label] -= 1;
      }
    }
    double inverse_max_dcg = max_dcg;
    if (inverse_max_dcg > 0.0) inverse_max_dcg = 1.0f / inverse_max_dcg;
    for (int i = 0; i < cnt; ++i) { lambdas[i] = 0.0f; hessians[i] = 0.0f; sorted[i] = i; }
    if (cnt == 0) continue;
    stable_sort_desc(sorted, cnt, score, tmp);
    const double best_score = score[sorted[0]];
    int worst_idx = cnt - 1;
    if (worst_idx > 0 && score[sorted[worst_idx]] == K_MIN_SCORE) worst_idx -= 1;
    const double worst_score = score[sorted[worst_idx]];
    double sum_lambdas = 0.0;
    for (int i = 0; i < cnt; ++i) {
      const int high = sorted[i];
      const int high_label = (int)label[high];
      const double high_score = score[high];
      if (high_score == K_MIN_SCORE) continue;
      const double high_label_gain = label_gain[high_label];
      const double high_discount = 1.0 / log2(2.0 + i);
      double high_sum_lambda = 0.0, high_sum_hessian = 0.0;
      for (int j = 0; j < cnt; ++j) {
        if (i == j) continue;
        const int low = sorted[j];
        const int low_label = (int)label[low];
        const double low_score = score[low];
        if (high_label <= low_label || low_score == K_MIN_SCORE) continue;
        const double delta_score = high_score - low_score;
        const double low_label_gain = label_gain[low_label];
        const double low_discount = 1.0 / log2(2.0 + j);
        const double dcg_gap = high_label_gain - low_label_gain;
        const double paired_discount = fabs(high_discount - low_discount);
        double delta_pair_NDCG = dcg_gap * paired_discount * inverse_max_dcg;
        if (norm && best_score != worst_score) delta_pair_NDCG /= (0.01f + fabs(delta_score));
        double p_lambda;
        if (delta_score <= min_in) p_lambda = table[0];
        else if (delta_score >= max_in) p_lambda = table[SIGMOID_BINS - 1];
        else p_lambda = table[(size_t)((delta_score - min_in) * idx_factor)];
        double p_hessian = p_lambda * (1.0f - p_lambda);
        p_lambda *= -sigmoid * delta_pair_NDCG;
        p_hessian *= sigmoid * sigmoid * delta_pair_NDCG;
        high_sum_lambda += p_lambda;
        high_sum_hessian += p_hessian;
        lambdas[low] -= (float)p_lambda;
        hessians[low] += (float)p_hessian;
        sum_lambdas -= 2 * p_lambda;
      }
      lambdas[high] += (float)high_sum_lambda;
      hessians[high] += (float)high_sum_hessian;
    }
    if (norm && sum_lambdas > 0) {
      const double norm_factor = log2(1 + sum_lambdas) / sum_lambdas;
      for (int i = 0; i < cnt; ++i) {
        lambdas[i] = (float)(lambdas[i] * norm_factor);
        hessians[i] = (float)(hessians[i] * norm_factor);
      }
    }
    if (weight) {
      for (int j = 0; j < cnt; ++j) {
        lambdas[j] = (float)(lambdas[j] * weight[start + j]);
        hessians[j] = (float)(hessians[j] * weight[start + j]);
      }
    }
  }
  free(table); free(sorted); free(tmp); free(label_cnt);
}
