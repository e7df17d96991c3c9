// Objective gradients (A9), score update (A10), per-tree preparation (A8) and layout helpers.
#pragma once
#include "dev_types.cuh"

// ---------------------------------------------------------------------------------------------------------
// BinaryLogloss::GetGradients (src/objective/binary_objective.hpp:102-133)
__global__ void k_grad_binary(int n, const double* __restrict__ score, const float* __restrict__ label,
                              const float* __restrict__ weight, double sigmoid, double w_neg, double w_pos,
                              float* __restrict__ grad, float* __restrict__ hess, int pos_class = -1) {
  // pos_class < 0: the positive class is label > 0; otherwise it is (int)label == pos_class (MulticlassOVA's
  // per-class BinaryLogloss, multiclass_objective.hpp:188-195)
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int is_pos = pos_class < 0 ? (label[i] > 0) : ((int)label[i] == pos_class);
    const int lab = is_pos ? 1 : -1;
    const double lw = is_pos ? w_pos : w_neg;
    const double response = -lab * sigmoid / (1.0f + exp(lab * sigmoid * score[i]));
    const double abs_response = fabs(response);
    if (weight == nullptr) {
      grad[i] = (float)(response * lw);
      hess[i] = (float)(abs_response * (sigmoid - abs_response) * lw);
    } else {
      grad[i] = (float)(response * lw * weight[i]);
      hess[i] = (float)(abs_response * (sigmoid - abs_response) * lw * weight[i]);
    }
  }
}

// RegressionL2loss::GetGradients (src/objective/regression_objective.hpp:125-140)
__global__ void k_grad_l2(int n, const double* __restrict__ score, const float* __restrict__ label,
                          const float* __restrict__ weight, float* __restrict__ grad, float* __restrict__ hess, int sqrt_label) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    // reg_sqrt: the objective sees sign(label) * sqrt(|label|) (RegressionL2loss::Init, :110-118; float arithmetic there)
    const float lab = sqrt_label ? (float)((label[i] > 0.0f) - (label[i] < 0.0f)) * sqrtf(fabsf(label[i])) : label[i];
    if (weight == nullptr) {
      grad[i] = (float)(score[i] - lab);
      hess[i] = 1.0f;
    } else {
      grad[i] = (float)((score[i] - lab) * weight[i]);
      hess[i] = (float)(weight[i]);
    }
  }
}

// RegressionHuberLoss (regression_objective.hpp:310-335; p0 = alpha), RegressionFairLoss (:363-380; p0 = c),
// RegressionPoissonLoss (:441-457; p0 = max_delta_step): one pointwise kernel, kind = LGBCU_OBJ_*
__global__ void k_grad_regression(int n, int kind, double p0, const double* __restrict__ score, const float* __restrict__ label,
                                  const float* __restrict__ weight, float* __restrict__ grad, float* __restrict__ hess) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double g, h;
    if (kind == LGBCU_OBJ_HUBER) {
      const double diff = score[i] - label[i];
      if (fabs(diff) <= p0) g = diff; else g = ((diff > 0.0) - (diff < 0.0)) * p0;
      if (weight == nullptr) { grad[i] = (float)g; hess[i] = 1.0f; }
      else {
        grad[i] = fabs(diff) <= p0 ? (float)(diff * weight[i]) : (float)(((diff > 0.0) - (diff < 0.0)) * weight[i] * p0);
        hess[i] = (float)(weight[i]);
      }
      continue;
    } else if (kind == LGBCU_OBJ_FAIR) {
      const double x = score[i] - label[i];
      g = p0 * x / (fabs(x) + p0);
      h = p0 * p0 / ((fabs(x) + p0) * (fabs(x) + p0));
    } else if (kind == LGBCU_OBJ_GAMMA) {   // RegressionGammaLoss (:683-701); the weight multiplies the second term only
      if (weight == nullptr) {
        grad[i] = (float)(1.0 - label[i] / exp(score[i]));
        hess[i] = (float)(label[i] / exp(score[i]));
      } else {
        grad[i] = (float)(1.0 - label[i] / exp(score[i]) * weight[i]);
        hess[i] = (float)(label[i] / exp(score[i]) * weight[i]);
      }
      continue;
    } else if (kind == LGBCU_OBJ_TWEEDIE) { // RegressionTweedieLoss (:720-738); p0 = tweedie_variance_power
      g = -label[i] * exp((1 - p0) * score[i]) + exp((2 - p0) * score[i]);
      h = -label[i] * (1 - p0) * exp((1 - p0) * score[i]) + (2 - p0) * exp((2 - p0) * score[i]);
    } else if (kind == LGBCU_OBJ_XENTLAMBDA && weight != nullptr) {   // CrossEntropyLambda::GetGradients, weighted (xentropy_objective.hpp:199-215)
      const double w = weight[i];
      const double y = label[i];
      const double epf = exp(score[i]);
      const double hhat = log(1.0f + epf);
      const double z = 1.0f - exp(-w * hhat);
      const double enf = 1.0f / epf;
      grad[i] = (float)((1.0f - y / z) * w / (1.0f + enf));
      const double c = 1.0f / (1.0f - z);
      double d = 1.0f + epf;
      const double a = w * epf / (d * d);
      d = c - 1.0f;
      const double b = (c / (d * d)) * (1.0f + w * epf - c);
      hess[i] = (float)(a * (1.0f + y * b));
      continue;
    } else if (kind == LGBCU_OBJ_XENTROPY || kind == LGBCU_OBJ_XENTLAMBDA) { // CrossEntropy::GetGradients (the unweighted lambda form is the same) (xentropy_objective.hpp:77-96)
      const double z = 1.0f / (1.0f + exp(-score[i]));
      g = z - label[i];
      h = z * (1.0f - z);
    } else {  // poisson
      g = exp(score[i]) - label[i];
      h = exp(score[i] + p0);
    }
    if (weight == nullptr) { grad[i] = (float)g; hess[i] = (float)h; }
    else { grad[i] = (float)(g * weight[i]); hess[i] = (float)(h * weight[i]); }
  }
}

__device__ __forceinline__ float mape_label_weight(float label) { return 1.0f / fmaxf(1.0f, fabsf(label)); }

// RegressionL1loss::GetGradients (src/objective/regression_objective.hpp:214-231): sign(score - label) [* w], hessian 1 [w]
__global__ void k_grad_l1(int n, const double* __restrict__ score, const float* __restrict__ label,
                          const float* __restrict__ weight, float* __restrict__ grad, float* __restrict__ hess, int mape) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double diff = score[i] - label[i];
    const int sg = (diff > 0.0) - (diff < 0.0);   // Common::Sign
    // RegressionMAPELOSS (:587-626, unweighted): label_weight_ = 1 / max(1, |label|) scales the gradient only
    if (mape) { grad[i] = (float)(sg * mape_label_weight(label[i])); hess[i] = 1.0f; }
    else if (weight == nullptr) { grad[i] = (float)sg; hess[i] = 1.0f; }
    else { grad[i] = (float)(sg * weight[i]); hess[i] = weight[i]; }
  }
}

// RegressionQuantileloss::GetGradients (:491-516): alpha is held as a float; delta is the float of (score - label)
__global__ void k_grad_quantile(int n, const double* __restrict__ score, const float* __restrict__ label,
                                const float* __restrict__ weight, float alpha, float* __restrict__ grad, float* __restrict__ hess) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float delta = (float)(score[i] - label[i]);
    if (weight == nullptr) {
      grad[i] = delta >= 0 ? (1.0f - alpha) : -alpha;
      hess[i] = 1.0f;
    } else {
      grad[i] = delta >= 0 ? (float)((1.0f - alpha) * weight[i]) : (float)(-alpha * weight[i]);
      hess[i] = (float)(weight[i]);
    }
  }
}

// MulticlassSoftmax::GetGradients (src/objective/multiclass_objective.hpp:86-130): one thread per row over the K score
// planes (class-major, [k][n]); Common::Softmax's order of operations (max, exp(x - max) summed in class order, divide)
__global__ void k_grad_multiclass(int n, int K, const double* __restrict__ score, const float* __restrict__ label,
                                  const float* __restrict__ weight, double factor, float* __restrict__ grad, float* __restrict__ hess) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double wmax = score[i];
    for (int k = 1; k < K; ++k) wmax = fmax(score[(size_t)n * k + i], wmax);
    double wsum = 0.0;
    for (int k = 0; k < K; ++k) wsum += exp(score[(size_t)n * k + i] - wmax);
    const int lab = (int)label[i];
    for (int k = 0; k < K; ++k) {
      const size_t idx = (size_t)n * k + i;
      const double p = exp(score[idx] - wmax) / wsum;
      if (weight == nullptr) {
        grad[idx] = lab == k ? (float)(p - 1.0f) : (float)p;
        hess[idx] = (float)(factor * p * (1.0f - p));
      } else {
        grad[idx] = lab == k ? (float)((p - 1.0f) * weight[i]) : (float)(p * weight[i]);
        hess[idx] = (float)((factor * p * (1.0f - p)) * weight[i]);
      }
    }
  }
}

// max |g|, max |h| as float bit patterns (non-negative floats order like unsigned ints)
// Also writes the interleaved (g, h) copy the histogram kernel gathers from: one 8-byte load per indexed row
// instead of two 4-byte loads in two different sectors.
__global__ void k_absmax(int n, const float* __restrict__ grad, const float* __restrict__ hess, float2* __restrict__ gh,
                         DevCtl* ctl) {
  unsigned mg = 0, mh = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float g = grad[i], h = hess[i];
    gh[i] = make_float2(g, h);
    mg = max(mg, __float_as_uint(fabsf(g)));
    mh = max(mh, __float_as_uint(fabsf(h)));
  }
  for (int o = 16; o > 0; o >>= 1) {
    mg = max(mg, __shfl_xor_sync(0xffffffffu, mg, o));
    mh = max(mh, __shfl_xor_sync(0xffffffffu, mh, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (mg) atomicMax(&ctl->max_g_bits, mg);
    if (mh) atomicMax(&ctl->max_h_bits, mh);
  }
}

// 2^k such that  max|v| * 2^k * rows < 2^62  (sums of `rows` fixed-point values cannot overflow an int64)  and
// max|v| * 2^k < 2^46  (a row's value fits the 32 + 16 bit record the histogram kernel consumes, pack_gh).
__device__ __forceinline__ int fixed_exponent(unsigned max_bits, long long rows) {
  const float m = __uint_as_float(max_bits);
  if (!(m > 0.f) || isinf(m) || isnan(m)) return 0;
  int e; frexpf(m, &e);          // m < 2^e
  int rb = 0; while ((1ll << rb) <= rows) ++rb;  // rows < 2^rb
  int k = 62 - e - rb;
  if (k > 46 - e) k = 46 - e;
  if (k > 900) k = 900;
  if (k < -900) k = -900;
  return k;
}

__global__ void k_reset_ctl_max(DevCtl* ctl, int root_count, int num_data) {
  ctl->max_g_bits = 0; ctl->max_h_bits = 0;
  ctl->global_root = root_count; ctl->global_rows = num_data;
}

// BeforeTrain (src/treelearner/serial_tree_learner.cpp:244-272) on device: reset the control block, the
// root leaf and the per-leaf best splits; derive the fixed-point scales from max|g|, max|h|.
// scale_rows = number of rows a histogram entry can sum over (global row count in data-parallel mode).
__global__ void k_prepare_tree(DevView V, int root_count, int root_identity, int lookahead) {
  DevCtl* c = V.ctl;
  const long long scale_rows = c->global_rows;
  const int groot = (int)c->global_root;
  for (int l = threadIdx.x; l < V.p.num_leaves; l += blockDim.x) {
    DevSplit none; none.gain = LGB_MIN_SCORE; none.feature = -1; none.real_feature = -1;
    none.threshold = 0; none.default_left = 1; none.left_count = 0; none.right_count = 0;
    none.left_output = none.right_output = 0; none.left_sum_g = none.left_sum_h = 0;
    none.right_sum_g = none.right_sum_h = 0;
    V.best[l] = none;
    { SplitKey k; k.gain = LGB_MIN_SCORE; k.real_feature = -1; k.pad = 0; V.leaf_key[l] = k; }
    DevLeaf z; z.begin = 0; z.count = 0; z.slot = l; z.depth = 0; z.buf = 0; z.gcount = 0;
    z.sum_g = z.sum_h = z.output = 0; z.mc_min = -LGB_MC_INF; z.mc_max = LGB_MC_INF;
    if (l == 0) { z.count = root_count; z.gcount = groot; }
    V.leaves[l] = z;
    V.expanded[l] = 0;
  }
  if (threadIdx.x == 0) {
    c->num_leaves = 1; c->done = 0; c->smaller = 0; c->larger = -1;
    c->split_leaf = c->split_feature = c->split_threshold = 0;
    c->split_default_left = 0; c->split_right_leaf = 1; c->split_left_count = 0;
    c->num_splits = 0; c->feat_done = 0; c->cat_count = 0; c->cat_overflow = 0;
    c->round = 0; c->num_items = 0; c->lookahead = lookahead;
    c->root_identity = root_identity; c->num_data = root_count;
    c->kg = fixed_exponent(c->max_g_bits, scale_rows);
    c->kh = fixed_exponent(c->max_h_bits, scale_rows);
    c->scale_g = ldexp(1.0, c->kg); c->inv_g = ldexp(1.0, -c->kg);
    c->scale_h = ldexp(1.0, c->kh); c->inv_h = ldexp(1.0, -c->kh);
    // BeforeFindBestSplit for the root: right child count is 0 (serial_tree_learner.h GetGlobalDataCountInLeaf)
    c->skip_find = (0 < V.p.min_data_in_leaf * 2 && groot < V.p.min_data_in_leaf * 2) ? 1 : 0;
    if (V.p.num_leaves < 2) c->done = 1;
  }
}

__global__ void k_iota(int n, int* out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = i;
}

// step-wise API helpers ------------------------------------------------------------------------------------
__global__ void k_set_split(DevView V, int leaf, int feature, int threshold, int default_left, int right_leaf) {
  DevCtl* c = V.ctl;
  c->split_leaf = leaf; c->split_feature = feature; c->split_threshold = threshold;
  c->split_default_left = default_left; c->split_right_leaf = right_leaf; c->done = 0;
}
__global__ void k_set_leaf_sums(DevView V, int leaf, double sg, double sh, double out) {
  V.leaves[leaf].sum_g = sg; V.leaves[leaf].sum_h = sh; V.leaves[leaf].output = out;
}
// step-wise lgbcu_hist_build: clear the leaf's pool slot and its exact totals before accumulating
__global__ void k_zero_slot(DevView V, int leaf) {
  HistEntry* a = V.pool + (size_t)V.leaves[leaf].slot * V.total_bin;
  HistEntry z; z.g = 0; z.h = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < V.total_bin; i += gridDim.x * blockDim.x) a[i] = z;

}

// FixHistogram (src/io/dataset.cpp:1410-1428) for every feature of one slot, then conversion to hist_t doubles
// in the reference layout.  One block per feature.
__global__ void k_hist_export(DevView V, int leaf, double* out) {
  const int slot = V.leaves[leaf].slot;
  const int f = blockIdx.x;
  const DevFeat F = V.feats[f];
  const HistEntry* h = V.pool + (size_t)slot * V.total_bin + F.hist_base;
  double* o = out + 2 * (size_t)F.hist_base;
  const double ig = V.ctl->inv_g, ih = V.ctl->inv_h;
  for (int s = threadIdx.x; s < F.nslots; s += blockDim.x) {
    HistEntry e = h[s];
    if (F.mfb > 0 && s == 1 + F.mfb) e = h[0];
    o[2 * s] = (double)e.g * ig;
    o[2 * s + 1] = (double)e.h * ih;
  }
}

// larger = parent - smaller over a whole slot (FeatureHistogram::Subtract, feature_histogram.hpp:79-83)
__global__ void k_hist_subtract(DevView V, int parent_leaf, int smaller_leaf) {
  const int parent_slot = V.leaves[parent_leaf].slot, smaller_slot = V.leaves[smaller_leaf].slot;
  HistEntry* a = V.pool + (size_t)parent_slot * V.total_bin;
  const HistEntry* b = V.pool + (size_t)smaller_slot * V.total_bin;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < V.total_bin; i += gridDim.x * blockDim.x) {
    HistEntry e = a[i]; const HistEntry s = b[i];
    e.g -= s.g; e.h -= s.h; a[i] = e;
  }
}

// SerialTreeLearner::AddPredictionToScore (src/treelearner/serial_tree_learner.h:96-111):
// score[idx] += leaf_value[leaf] for every row of every leaf range.  Leaves differ in size by orders of magnitude, so
// the work is cut into chunks of ADD_SCORE_CHUNK positions: every block derives the chunk -> (leaf, offset) map from
// the leaf table (a 1024-entry block scan), then handles the chunk with its own index.
#define ADD_SCORE_CHUNK 4096
#define ADD_SCORE_THREADS 256
__global__ void __launch_bounds__(ADD_SCORE_THREADS) k_add_score_partition(DevView V, int num_leaves,
                                                                           const double* __restrict__ leaf_value,
                                                                           double* __restrict__ score) {
  __shared__ int s_first[1025];  // first chunk of leaf l (exclusive prefix of the chunk counts)
  __shared__ int s_warp[ADD_SCORE_THREADS / 32];
  __shared__ int s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int target = blockIdx.x;
  for (int base = 0; base < num_leaves; base += 1024) {  // trees of more than 1024 leaves: window by window
    if (tid == 0) s_carry = 0;
    __syncthreads();
    const int nwin = min(1024, num_leaves - base);
    for (int off = 0; off < nwin; off += ADD_SCORE_THREADS) {
      const int l = off + tid;
      const int v = l < nwin ? (V.leaves[base + l].count + ADD_SCORE_CHUNK - 1) / ADD_SCORE_CHUNK : 0;
      int inc = v;
      for (int o = 1; o < 32; o <<= 1) { const int x = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += x; }
      if (lane == 31) s_warp[warp] = inc;
      __syncthreads();
      int woff = 0;
      for (int w = 0; w < warp; ++w) woff += s_warp[w];
      const int excl = s_carry + woff + inc - v;
      if (l < nwin) s_first[l] = excl;
      __syncthreads();
      if (tid == ADD_SCORE_THREADS - 1) s_carry = excl + v;
      __syncthreads();
    }
    const int total = s_carry;
    if (tid == 0) s_first[nwin] = total;
    __syncthreads();
    if (target < total) {
      int lo = 0, hi = nwin - 1;  // last leaf whose first chunk is <= target (it cannot be an empty one)
      while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (s_first[mid] <= target) lo = mid; else hi = mid - 1; }
      const int leaf = base + lo;
      const DevLeaf L = V.leaves[leaf];
      const double v = leaf_value[leaf];
      const int* idx = V.indices[L.buf] + L.begin;
      const int c0 = (target - s_first[lo]) * ADD_SCORE_CHUNK;
      const int c1 = min(L.count, c0 + ADD_SCORE_CHUNK);
      for (int i = c0 + tid; i < c1; i += ADD_SCORE_THREADS) score[idx[i]] += v;
      return;
    }
    target -= total;
    __syncthreads();
  }
}

// ScoreUpdater::AddScore(double) (src/boosting/score_updater.hpp:54-61)
__global__ void k_add_const(int n, double* score, double v) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) score[i] += v;
}

// ScoreUpdater::MultiplyScore (score_updater.hpp:84-91)
__global__ void k_mul_const(int n, double* score, double v) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) score[i] *= v;
}

// packed [n][F] rows -> padded [n][row_stride] rows (one thread per 4 output bytes)
__global__ void k_pad_rows(long long n, int F, int row_stride, const uint8_t* __restrict__ src, uint8_t* __restrict__ dst) {
  const long long words = n * (row_stride / 4);
  for (long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x; w < words; w += (long long)gridDim.x * blockDim.x) {
    const long long r = w / (row_stride / 4);
    const int c = (int)(w - r * (row_stride / 4)) * 4;
    unsigned v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) if (c + k < F) v |= (unsigned)src[r * F + c + k] << (8 * k);
    reinterpret_cast<unsigned*>(dst)[w] = v;
  }
}

// column vectors -> row-major matrix (one warp-wide 32-row strip per step; coalesced on the column side)
__global__ void k_scatter_column(int n, const uint8_t* __restrict__ col, uint8_t* __restrict__ bins, int row_stride,
                                 int f) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    bins[(size_t)i * row_stride + f] = col[i];
  }
}

// row-major [n][row_stride] -> column-major [F][col_stride] copy (one thread per row: 16-byte loads on the row side,
// byte stores that are contiguous across the warp on the column side)
__global__ void k_transpose_bins(int n, int F, int row_stride, const uint8_t* __restrict__ bins, uint8_t* __restrict__ cols,
                                 long long col_stride) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint4* row = reinterpret_cast<const uint4*>(bins + (size_t)i * row_stride);
    for (int c16 = 0; c16 * 16 < F; ++c16) {
      const uint4 q = row[c16];
      const unsigned w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int f = c16 * 16 + k;
        if (f < F) cols[(size_t)f * col_stride + i] = (uint8_t)((w[k >> 2] >> ((k & 3) * 8)) & 0xffu);
      }
    }
  }
}

// Tree::AddPredictionToScore on binned data (src/io/tree.cpp:109-199 with NumericalDecisionInner,
// include/LightGBM/tree.h:285-300): walk the tree for each listed row (out-of-bag rows, or all rows).
struct TreeDev {
  const int* split_feature; const int* threshold_bin; const signed char* decision_type;
  const int* left_child; const int* right_child; const double* leaf_value;
  int num_leaves;
  const unsigned* cat_sets;   // [num_cat][LGB_CAT_WORDS] bin sets of the categorical nodes (threshold_bin = index), or NULL
};
#define TREE_SMEM_NODES 1024
__global__ void __launch_bounds__(256) k_add_score_tree(DevView V, TreeDev T, const int* __restrict__ rows, int n,
                                                        double* __restrict__ score) {
  __shared__ int s_feat[TREE_SMEM_NODES], s_thr[TREE_SMEM_NODES], s_left[TREE_SMEM_NODES], s_right[TREE_SMEM_NODES];
  __shared__ signed char s_dt[TREE_SMEM_NODES];
  const int ni = T.num_leaves - 1;
  const bool cached = ni <= TREE_SMEM_NODES;
  if (cached) {
    for (int i = threadIdx.x; i < ni; i += blockDim.x) {
      s_feat[i] = T.split_feature[i]; s_thr[i] = T.threshold_bin[i]; s_left[i] = T.left_child[i];
      s_right[i] = T.right_child[i]; s_dt[i] = T.decision_type[i];
    }
    __syncthreads();
  }
  const int* feat = cached ? s_feat : T.split_feature;
  const int* thr = cached ? s_thr : T.threshold_bin;
  const int* lc = cached ? s_left : T.left_child;
  const int* rc = cached ? s_right : T.right_child;
  const signed char* dt = cached ? s_dt : T.decision_type;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int row = rows ? rows[i] : i;
    const uint8_t* r = V.bins + (size_t)row * V.row_stride;
    int node = 0;
    while (node >= 0) {
      const int f = feat[node];
      const DevFeat F = V.feats[f];
      const int stored = stored_bin_row(r, F);
      // DenseBinIterator::Get: stored value -> bin
      const int bin = stored == 0 ? F.mfb : stored - 1 + F.offset;
      const int missing_type = (dt[node] >> 2) & 3;
      bool left;
      if (dt[node] & 1) {   // Tree::CategoricalDecisionInner (tree.h:330-341): left = the bin is in the node's set
        left = ((T.cat_sets[(size_t)thr[node] * LGB_CAT_WORDS + (bin >> 5)] >> (bin & 31)) & 1u) != 0;
      } else if ((missing_type == LGBCU_MISSING_ZERO && bin == F.default_bin) ||
          (missing_type == LGBCU_MISSING_NAN && bin == F.num_bin - 1)) {
        left = (dt[node] & 2) != 0;
      } else {
        left = bin <= thr[node];
      }
      node = left ? lc[node] : rc[node];
    }
    score[row] += T.leaf_value[~node];
  }
}
