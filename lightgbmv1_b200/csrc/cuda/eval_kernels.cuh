// Validation-set scoring and evaluation metrics on the device.
//
// Replaces, for data sets other than the training rows: ScoreUpdater (src/boosting/score_updater.hpp:21-125) with
// Tree::AddPredictionToScore on binned data (src/io/tree.cpp:109-199), and the metrics the path's objectives default
// to -- L2 / RMSE (src/metric/regression_metric.hpp), binary logloss and AUC (src/metric/binary_metric.hpp:20-270),
// NDCG@k (src/metric/rank_metric.hpp:20-160, src/metric/dcg_calculator.cpp:43-153) -- so that a run with
// metric_freq = 1 never copies scores to the host.
//
// Sums are formed in double by a fixed two-stage tree (per-block partials, then one block), so a metric is bit-
// reproducible from run to run; the reference reduces over OpenMP threads, i.e. in a thread-count-dependent order,
// and agreement with it is to rounding (tests: 1e-9 relative; north_star asks for 1e-5).
#pragma once
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>
#include <cub/device/device_scan.cuh>

namespace {

#define EVAL_THREADS 256
#define EVAL_MAX_BLOCKS 1024

__device__ __forceinline__ double block_sum_f64(double v, double* s_warp) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0) for (int w = 0; w < EVAL_THREADS / 32; ++w) t += s_warp[w];
  __syncthreads();
  return t;  // valid in thread 0
}

// kind 0: (score - label)^2 [* w]   (RegressionMetric<L2Metric>::LossOnPoint)
// kind 1: binary logloss of prob = 1 / (1 + exp(-sigmoid * score))   (BinaryLoglossMetric::LossOnPoint, kEpsilon clip)
__device__ __forceinline__ double safe_log(double x) { return x > 0 ? log(x) : -INFINITY; }   // Common::SafeLog
__global__ void __launch_bounds__(EVAL_THREADS) k_eval_pointwise(int n, int kind, double sigmoid, int conv_exp, const double* __restrict__ score,
                                                                 const float* __restrict__ label, const float* __restrict__ weight,
                                                                 double* __restrict__ partial) {
  __shared__ double s_warp[EVAL_THREADS / 32];
  double acc = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double s = score[i];
    if (conv_exp == 2) s = (double)((s > 0.0) - (s < 0.0)) * s * s;   // RegressionL2loss::ConvertOutput with reg_sqrt
    const float y = label[i];
    double loss;
    if (kind == 0) {
      loss = (s - y) * (s - y);
    } else if (kind == 2) {          // L1Metric::LossOnPoint (regression_metric.hpp:177)
      loss = fabs(s - y);
    } else if (kind == 4) {          // HuberLossMetric::LossOnPoint (:191-198); `sigmoid` carries config.alpha
      const double diff = s - y;
      loss = fabs(diff) <= sigmoid ? 0.5f * diff * diff : sigmoid * (fabs(diff) - 0.5f * sigmoid);
    } else if (kind == 5) {          // FairLossMetric::LossOnPoint (:212-216); `sigmoid` carries config.fair_c
      const double x = fabs(s - y);
      loss = sigmoid * x - sigmoid * sigmoid * log(1.0f + x / sigmoid);
    } else if (kind == 6) {          // PoissonMetric::LossOnPoint (:229-235); conv_exp: the objective's exp first
      double t = conv_exp == 1 ? exp(s) : s;
      const double eps = 1e-10f;
      if (t < eps) t = eps;
      loss = t - y * log(t);
    } else if (kind == 9) {          // BinaryErrorMetric::LossOnPoint (binary_metric.hpp:143-149) of the converted score
      const double prob = sigmoid > 0.0 ? 1.0f / (1.0f + exp(-sigmoid * s)) : s;
      loss = prob <= 0.5f ? (y > 0 ? 1.0 : 0.0) : (y <= 0 ? 1.0 : 0.0);
    } else if (kind == 12) {         // XentLambdaLoss (xentropy_metric.hpp:53-55) of hhat = log(1 + exp(score)); the weight is INSIDE
      const double hhat = log(1.0f + exp(s));
      const float w = weight ? weight[i] : 1.0f;
      const double prob = 1.0f - exp(-w * hhat);
      const double log_arg_epsilon = 1.0e-12;
      double a = y, b = 1.0f - y;
      a *= prob > log_arg_epsilon ? log(prob) : log(log_arg_epsilon);
      b *= 1.0f - prob > log_arg_epsilon ? log(1.0f - prob) : log(log_arg_epsilon);
      acc += -(a + b);
      continue;
    } else if (kind == 13) {         // GammaDevianceMetric::LossOnPoint (regression_metric.hpp:284-288)
      const double t = conv_exp == 1 ? exp(s) : s;
      const double tmp = y / (t + 1.0e-9);
      loss = tmp - safe_log(tmp) - 1;
    } else if (kind == 11) {         // MAPEMetric::LossOnPoint (regression_metric.hpp:248-250)
      loss = fabs(y - s) / fmaxf(1.0f, fabsf(y));
    } else if (kind == 10) {         // XentLoss (xentropy_metric.hpp:34-50) of the converted score
      const double prob = sigmoid > 0.0 ? 1.0f / (1.0f + exp(-sigmoid * s)) : s;
      const double log_arg_epsilon = 1.0e-12;
      double a = y, b = 1.0f - y;
      a *= prob > log_arg_epsilon ? log(prob) : log(log_arg_epsilon);
      b *= 1.0f - prob > log_arg_epsilon ? log(1.0f - prob) : log(log_arg_epsilon);
      loss = -(a + b);
    } else if (kind == 7) {          // GammaMetric::LossOnPoint (:261-268), written out as there (psi = 1)
      const double t = conv_exp == 1 ? exp(s) : s;
      const double psi = 1.0;
      const double theta = -1.0 / t;
      const double a = psi;
      const double b = -safe_log(-theta);
      const double c = 1. / psi * safe_log(y / psi) - safe_log(y) - 0;
      loss = -((y * theta - b) / a + c);
    } else if (kind == 8) {          // TweedieMetric::LossOnPoint (:305-314); `sigmoid` carries tweedie_variance_power
      double t = conv_exp == 1 ? exp(s) : s;
      const double eps = 1e-10f;
      if (t < eps) t = eps;
      const double a = y * exp((1 - sigmoid) * log(t)) / (1 - sigmoid);
      const double b = exp((2 - sigmoid) * log(t)) / (2 - sigmoid);
      loss = -a + b;
    } else if (kind == 3) {          // QuantileMetric::LossOnPoint (:157-164); `sigmoid` carries config.alpha
      const double delta = y - s;
      loss = delta < 0 ? (sigmoid - 1.0f) * delta : sigmoid * delta;
    } else {
      // BinaryLogloss::ConvertOutput; sigmoid <= 0 = no objective (custom): the score is taken as the probability
      const double prob = sigmoid > 0.0 ? 1.0f / (1.0f + exp(-sigmoid * s)) : s;
      if (y <= 0) loss = (1.0f - prob > LGB_EPS) ? -log(1.0f - prob) : -log(LGB_EPS);
      else loss = (prob > LGB_EPS) ? -log(prob) : -log(LGB_EPS);
    }
    acc += weight ? loss * weight[i] : loss;
  }
  const double t = block_sum_f64(acc, s_warp);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// Multiclass metrics (src/metric/multiclass_metric.hpp:60-120): the K raw scores of a row ([class][row] planes) are
// converted by the objective's softmax (softmax = 0: no objective, the scores are taken as they are), then
// kind 0: multi_logloss  -log(max(p[label], kEpsilon))          (MultiSoftmaxLoglossMetric::LossOnPoint, :167-175)
// kind 1: multi_error@top_k: 1 if more than top_k classes score >= the label's class   (MultiErrorMetric, :142-151)
__global__ void __launch_bounds__(EVAL_THREADS) k_eval_multiclass(int n, int K, int kind, int top_k, int softmax, double ova_sigmoid,
                                                                  const double* __restrict__ score, const float* __restrict__ label,
                                                                  const float* __restrict__ weight, double* __restrict__ partial) {
  __shared__ double s_warp[EVAL_THREADS / 32];
  double acc = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int lab = (int)label[i];
    double wmax = score[i];
    for (int k = 1; k < K; ++k) wmax = fmax(score[(size_t)n * k + i], wmax);
    double wsum = 0.0;
    if (softmax) for (int k = 0; k < K; ++k) wsum += exp(score[(size_t)n * k + i] - wmax);
    // ova_sigmoid > 0: MulticlassOVA::ConvertOutput, a sigmoid per class (multiclass_objective.hpp:239-243)
    const double raw_l = score[(size_t)n * lab + i];
    const double pl = softmax ? exp(raw_l - wmax) / wsum : ova_sigmoid > 0.0 ? 1.0f / (1.0f + exp(-ova_sigmoid * raw_l)) : raw_l;
    double loss;
    if (kind == 0) {
      loss = pl > LGB_EPS ? -log(pl) : -log(LGB_EPS);
    } else {
      int num_larger = 0;
      for (int k = 0; k < K; ++k) {
        const double raw = score[(size_t)n * k + i];
        const double pk = softmax ? exp(raw - wmax) / wsum : ova_sigmoid > 0.0 ? 1.0f / (1.0f + exp(-ova_sigmoid * raw)) : raw;
        if (pk >= pl) ++num_larger;
      }
      loss = num_larger > top_k ? 1.0 : 0.0;
    }
    acc += weight ? loss * weight[i] : loss;
  }
  const double t = block_sum_f64(acc, s_warp);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

__global__ void __launch_bounds__(EVAL_THREADS) k_sum_f32(int n, const float* __restrict__ v, double* __restrict__ partial) {
  __shared__ double s_warp[EVAL_THREADS / 32];
  double acc = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) acc += v[i];
  const double t = block_sum_f64(acc, s_warp);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// second stage: one block sums `m` partials per output column in index order
__global__ void __launch_bounds__(EVAL_THREADS) k_eval_final(int m, int ncol, const double* __restrict__ partial, double* __restrict__ out) {
  __shared__ double s_warp[EVAL_THREADS / 32];
  for (int c = 0; c < ncol; ++c) {
    double acc = 0.0;
    for (int i = threadIdx.x; i < m; i += blockDim.x) acc += partial[(size_t)i * ncol + c];
    const double t = block_sum_f64(acc, s_warp);
    if (threadIdx.x == 0) out[c] = t;
  }
}

// ---- AUC (AUCMetric::Eval): rows sorted by score descending; per group of equal scores (pos, neg) weights; then
//      accum = sum_g neg_g * (0.5 * pos_g + pos before g)
__global__ void k_auc_keys(int n, const double* __restrict__ score, const float* __restrict__ label, const float* __restrict__ weight,
                           unsigned long long* __restrict__ keys, double2* __restrict__ vals) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double s = score[i];
    if (s == 0.0) s = 0.0;  // -0.0 and +0.0 compare equal in the reference: give them one key
    unsigned long long u = (unsigned long long)__double_as_longlong(s);
    u = (u >> 63) ? ~u : (u | 0x8000000000000000ull);  // order-preserving
    keys[i] = u;
    const double w = weight ? (double)weight[i] : 1.0;
    vals[i] = label[i] > 0 ? make_double2(w, 0.0) : make_double2(0.0, w);
  }
}
struct PairSum {
  __device__ __forceinline__ double2 operator()(const double2& a, const double2& b) const { return make_double2(a.x + b.x, a.y + b.y); }
};
// groups arrive in ASCENDING key order (cub's ReduceByKey keeps input order; the sort is ascending): "positives with a
// higher score" = total positives - inclusive prefix.
__global__ void __launch_bounds__(EVAL_THREADS) k_auc_accum(const int* __restrict__ num_groups, const double2* __restrict__ grp,
                                                            const double* __restrict__ pos_incl, double* __restrict__ partial) {
  __shared__ double s_warp[EVAL_THREADS / 32];
  const int m = *num_groups;
  const double total_pos = m > 0 ? pos_incl[m - 1] : 0.0;
  double acc = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
    const double2 g = grp[i];
    acc += g.y * (g.x * 0.5f + (total_pos - pos_incl[i]));
  }
  const double t = block_sum_f64(acc, s_warp);
  if (threadIdx.x == 0) { partial[2 * blockIdx.x] = t; partial[2 * blockIdx.x + 1] = blockIdx.x == 0 ? total_pos : 0.0; }
}
__global__ void k_auc_pos(int n, const double2* __restrict__ grp, double* __restrict__ pos) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) pos[i] = grp[i].x;
}

// ---- NDCG@k (NDCGMetric::Eval + DCGCalculator::CalDCG): one block per query.  rank of a document = number of
//      documents that a stable descending sort puts before it; the top-K contributions are summed in rank order by one
//      thread, exactly as the reference's sequential loop does.
#define NDCG_MAX_K 1024
__global__ void __launch_bounds__(EVAL_THREADS) k_ndcg(int num_queries, const int* __restrict__ qb, const double* __restrict__ score,
                                                       const float* __restrict__ label, const float* __restrict__ query_weight,
                                                       const double* __restrict__ label_gain, const double* __restrict__ discount,
                                                       const int* __restrict__ eval_at, int num_eval, int max_k,
                                                       const double* __restrict__ inv_max_dcg, double* __restrict__ per_query) {
  __shared__ double s_contrib[NDCG_MAX_K];
  for (int q = blockIdx.x; q < num_queries; q += gridDim.x) {
    const int b = qb[q], cnt = qb[q + 1] - b;
    double* out = per_query + (size_t)q * num_eval;
    const double w = query_weight ? (double)query_weight[q] : 1.0;
    if (inv_max_dcg[(size_t)q * num_eval] <= 0.0) {  // all documents irrelevant: counts as 1 (unweighted, as the reference adds it)
      if (threadIdx.x < num_eval) out[threadIdx.x] = 1.0;
      continue;
    }
    const int kk = max_k < cnt ? max_k : cnt;
    for (int r = threadIdx.x; r < kk; r += blockDim.x) s_contrib[r] = 0.0;
    __syncthreads();
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      const double si = score[b + i];
      int rank = 0;
      for (int j = 0; j < cnt; ++j) {
        const double sj = score[b + j];
        rank += (sj > si) || (sj == si && j < i);
      }
      if (rank < kk) s_contrib[rank] = label_gain[(int)label[b + i]] * discount[rank];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double cur = 0.0; int left = 0;
      for (int e = 0; e < num_eval; ++e) {
        int k = eval_at[e]; if (k > cnt) k = cnt;
        for (int j = left; j < k; ++j) cur += s_contrib[j];
        out[e] = cur * inv_max_dcg[(size_t)q * num_eval + e] * w;
        left = k;
      }
    }
    __syncthreads();
  }
}

// ---- MAP@k (MapMetric::CalMapAtK, map_metric.hpp:72-103): one block per query, ranks as for NDCG (a stable descending
//      sort), then one thread walks the top documents: every hit adds (hits so far) / (position + 1).
__global__ void __launch_bounds__(EVAL_THREADS) k_map(int num_queries, const int* __restrict__ qb, const double* __restrict__ score,
                                                      const float* __restrict__ label, const float* __restrict__ query_weight,
                                                      const int* __restrict__ eval_at, int num_eval, int max_k,
                                                      double* __restrict__ per_query) {
  __shared__ unsigned char s_hit[NDCG_MAX_K];
  __shared__ int s_npos;
  for (int q = blockIdx.x; q < num_queries; q += gridDim.x) {
    const int b = qb[q], cnt = qb[q + 1] - b;
    double* out = per_query + (size_t)q * num_eval;
    const double w = query_weight ? (double)query_weight[q] : 1.0;
    const int kk = max_k < cnt ? max_k : cnt;
    if (threadIdx.x == 0) s_npos = 0;
    for (int r = threadIdx.x; r < kk; r += blockDim.x) s_hit[r] = 0;
    __syncthreads();
    int my_pos = 0;
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      const double si = score[b + i];
      int rank = 0;
      for (int j = 0; j < cnt; ++j) {
        const double sj = score[b + j];
        rank += (sj > si) || (sj == si && j < i);
      }
      const bool hit = label[b + i] > 0.5f;
      my_pos += hit ? 1 : 0;
      if (rank < kk && hit) s_hit[rank] = 1;
    }
    if (my_pos) atomicAdd(&s_npos, my_pos);
    __syncthreads();
    if (threadIdx.x == 0) {
      const int npos = s_npos;
      int num_hit = 0, left = 0;
      double sum_ap = 0.0f;
      for (int e = 0; e < num_eval; ++e) {
        int k = eval_at[e]; if (k > cnt) k = cnt;
        for (int j = left; j < k; ++j) if (s_hit[j]) { ++num_hit; sum_ap += (double)num_hit / (j + 1.0f); }
        out[e] = (npos > 0 ? sum_ap / (npos < k ? npos : k) : 1.0) * w;
        left = k;
      }
    }
    __syncthreads();
  }
}

}  // namespace
