// LambdarankNDCG::GetGradients on device (src/objective/rank_objective.hpp:48-65,100-257;
// src/metric/dcg_calculator.cpp:43-76).
//
// One block per query.  The reference walks pairs (i, j) in score-rank order and accumulates per-document
// lambdas in FLOAT; to reproduce its rounding each document is owned by one thread that replays, in the
// reference's order of i, exactly the updates that document receives: for i != rank(doc) the pair term where
// doc is the lower-labelled one, and at i == rank(doc) the double-accumulated sum over all j where doc is the
// higher-labelled one.  The 2^20-entry sigmoid table and the position discounts are built on the host with
// libm (same values as the reference) and uploaded once.  Only sum_lambdas (a double accumulated over all
// pairs) is summed in a different order (block reduction): a <= 1e-16 relative change of the normalisation
// factor, i.e. at most a one-ulp flip of a rounded float -- the test tolerance says so.
#pragma once

#define RANK_SIGMOID_BINS (1024 * 1024)
#define RANK_MAX_POSITION 10000
#define RANK_THREADS 128

struct RankDev {
  const int* boundaries; const double* score; const float* label; const float* weight;
  const double* label_gain; const double* inv_max_dcg; const double* sigmoid_table; const double* discount;
  float* grad; float* hess;
  double sigmoid, min_in, max_in, idx_factor;
  int norm;
};

namespace {

__device__ __forceinline__ double rank_sigmoid(const RankDev& R, double s) {
  if (s <= R.min_in) return R.sigmoid_table[0];
  if (s >= R.max_in) return R.sigmoid_table[RANK_SIGMOID_BINS - 1];
  return R.sigmoid_table[(size_t)((s - R.min_in) * R.idx_factor)];
}

// smem: rank -> doc (sorted_idx), then per-rank copies of score / label so the pair loops read shared memory
__global__ void __launch_bounds__(RANK_THREADS) k_lambdarank(RankDev R, int max_query) {
  extern __shared__ __align__(16) unsigned char rsm[];
  double* s_score = reinterpret_cast<double*>(rsm);                       // [cnt] by rank
  int* s_label = reinterpret_cast<int*>(s_score + max_query);             // [cnt] by rank
  int* s_doc = s_label + max_query;                                       // [cnt] rank -> doc
  int* s_own = s_doc + max_query;                                         // [cnt] thread position -> rank it owns
  __shared__ double s_red[RANK_THREADS / 32];
  __shared__ double s_sum;
  const int q = blockIdx.x;
  const int start = R.boundaries[q], cnt = R.boundaries[q + 1] - start;
  if (cnt <= 0) return;
  const double* score = R.score + start;
  const float* label = R.label + start;
  // stable descending sort by counting: rank(d) = #{score > score_d} + #{score == score_d, index < d}
  for (int d = threadIdx.x; d < cnt; d += blockDim.x) {
    const double sd = score[d];
    int r = 0;
    for (int e = 0; e < cnt; ++e) {
      const double se = score[e];
      r += (se > sd) || (se == sd && e < d);
    }
    s_doc[r] = d;
    s_score[r] = sd;
    s_label[r] = (int)label[d];
  }
  __syncthreads();
  const double inverse_max_dcg = R.inv_max_dcg[q];
  const double best_score = s_score[0];
  int worst_idx = cnt - 1;
  if (worst_idx > 0 && s_score[worst_idx] == LGB_MIN_SCORE) worst_idx -= 1;
  const double worst_score = s_score[worst_idx];
  const bool use_norm_dist = R.norm && best_score != worst_score;
  // Which document a thread owns: position t of the order (label, rank).  The lanes of a warp then mostly share a
  // label, so "this pair contributes" (the other document's label is higher / lower than mine) is the same for the whole
  // warp and the ~70 % of pairs that contribute nothing are skipped by a uniform branch.  Every document still receives
  // its terms in the reference's order of i, so the float sums are bit for bit what they were.
  for (int r = threadIdx.x; r < cnt; r += blockDim.x) {
    const int L = s_label[r];
    int pos = 0;
    for (int e = 0; e < cnt; ++e) { const int le = s_label[e]; pos += (le < L) || (le == L && e < r); }
    s_own[pos] = r;
  }
  __syncthreads();
  double local_sum = 0.0;
  const int cnt_pad = (cnt + 31) / 32 * 32;
  for (int t = threadIdx.x; t < cnt_pad; t += blockDim.x) {
    // this thread owns the document at rank r (lanes past the end of the query run the loops with nothing to add)
    const int r = t < cnt ? s_own[t] : -1;
    const double my_score = r >= 0 ? s_score[r] : LGB_MIN_SCORE;
    const int my_label = r >= 0 ? s_label[r] : 0;
    const double my_disc = r >= 0 ? R.discount[r] : 0.0;
    float lam = 0.0f, hes = 0.0f;
    // One pair term.  A pair that the reference skips contributes exactly +0 (x - 0 == x and x + 0 == x bit for bit).
    auto pair_term = [&](double high_score, int high_label, double high_disc, double low_score, int low_label,
                         double low_disc, bool ok, double* pl, double* ph) {
      const double delta_score = ok ? high_score - low_score : 0.0;
      const double dcg_gap = R.label_gain[ok ? high_label : 0] - R.label_gain[ok ? low_label : 0];
      const double paired_discount = fabs(high_disc - low_disc);
      double delta_pair_NDCG = dcg_gap * paired_discount * inverse_max_dcg;
      if (use_norm_dist) delta_pair_NDCG /= (0.01f + fabs(delta_score));
      double p_lambda = rank_sigmoid(R, delta_score);
      double p_hessian = p_lambda * (1.0f - p_lambda);
      p_lambda *= -R.sigmoid * delta_pair_NDCG;
      p_hessian *= R.sigmoid * R.sigmoid * delta_pair_NDCG;
      *pl = ok ? p_lambda : 0.0;
      *ph = ok ? p_hessian : 0.0;
    };
    const bool me_valid = r >= 0 && my_score != LGB_MIN_SCORE;
    // doc as the higher-labelled one: the reference sums these in double over all j, in order, when its walk reaches
    // rank r
    double high_sum_lambda = 0.0, high_sum_hessian = 0.0;
    for (int j = 0; j < cnt; ++j) {
      const double ls = s_score[j]; const int ll = s_label[j];
      const bool ok = me_valid && j != r && ls != LGB_MIN_SCORE && my_label > ll;
      if (!__any_sync(0xffffffffu, ok)) continue;
      double pl, ph;
      pair_term(my_score, my_label, my_disc, ls, ll, R.discount[j], ok, &pl, &ph);
      high_sum_lambda += pl; high_sum_hessian += ph;
      local_sum -= 2 * pl;
    }
    // the walk over i: for i != r the pair term where doc is the lower-labelled one; at i == r the sum above
    for (int i = 0; i < cnt; ++i) {
      if (i == r && me_valid) { lam += (float)high_sum_lambda; hes += (float)high_sum_hessian; }
      const double hs = s_score[i]; const int hl = s_label[i];
      const bool ok = me_valid && i != r && hs != LGB_MIN_SCORE && hl > my_label;
      if (!__any_sync(0xffffffffu, ok)) continue;
      double pl, ph;
      pair_term(hs, hl, R.discount[i], my_score, my_label, my_disc, ok, &pl, &ph);
      lam -= (float)pl; hes += (float)ph;
    }
    if (r >= 0) {
      R.grad[start + s_doc[r]] = lam;
      R.hess[start + s_doc[r]] = hes;
    }
  }
  // sum_lambdas over the block
  for (int o = 16; o > 0; o >>= 1) local_sum += __shfl_xor_sync(0xffffffffu, local_sum, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = local_sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < RANK_THREADS / 32; ++w) t += s_red[w];
    s_sum = t;
  }
  __syncthreads();
  const double sum_lambdas = s_sum;
  const bool do_norm = R.norm && sum_lambdas > 0;
  const double norm_factor = do_norm ? log2(1 + sum_lambdas) / sum_lambdas : 1.0;
  for (int d = threadIdx.x; d < cnt; d += blockDim.x) {
    float lam = R.grad[start + d], hes = R.hess[start + d];
    if (do_norm) { lam = (float)(lam * norm_factor); hes = (float)(hes * norm_factor); }
    if (R.weight) { lam = (float)(lam * R.weight[start + d]); hes = (float)(hes * R.weight[start + d]); }
    R.grad[start + d] = lam;
    R.hess[start + d] = hes;
  }
}

// Host part of LambdarankNDCG::Init: label checks, inverse max DCG per query, sigmoid table, discounts.
int rank_prepare(LgbCuContext* c) {
  if (c->rank_ready) return 0;
  NEED(c->num_queries > 0, "Ranking tasks require query information");
  NEED(c->has_label, "lambdarank: label not set");
  std::vector<double> gains;
  if (c->num_label_gain > 0) {
    gains.resize(c->num_label_gain);
    CU(cudaMemcpy(gains.data(), c->d_label_gain, sizeof(double) * gains.size(), cudaMemcpyDeviceToHost));
  } else {  // DCGCalculator::DefaultLabelGain: 2^i - 1, i < 31
    gains.push_back(0.0);
    for (int i = 1; i < 31; ++i) gains.push_back(static_cast<double>((1 << i) - 1));
    CU(dev_alloc(&c->d_label_gain, gains.size()));
    CU(cudaMemcpy(c->d_label_gain, gains.data(), sizeof(double) * gains.size(), cudaMemcpyHostToDevice));
    c->num_label_gain = static_cast<int>(gains.size());
  }
  std::vector<float> label(c->N);
  CU(cudaMemcpy(label.data(), c->d_label, sizeof(float) * c->N, cudaMemcpyDeviceToHost));
  for (int i = 0; i < c->N; ++i) {  // DCGCalculator::CheckLabel
    const float l = label[i];
    NEED(std::fabs(l - static_cast<int>(l)) <= 1e-15f, "label should be int type (met %f) for ranking task,\n"
         "for the gain of label, please set the label_gain parameter", l);
    NEED(l >= 0, "Label should be non-negative (met %f) for ranking task", l);
    NEED(static_cast<size_t>(l) < gains.size(), "Label %zu is not less than the number of label mappings (%zu)",
         static_cast<size_t>(l), gains.size());
  }
  NEED(c->max_query <= RANK_MAX_POSITION, "lambdarank: a query has %d documents; the reference's discount table "
       "stops at %d", c->max_query, RANK_MAX_POSITION);
  std::vector<double> discount(RANK_MAX_POSITION);
  for (int i = 0; i < RANK_MAX_POSITION; ++i) discount[i] = 1.0 / std::log2(2.0 + i);
  const int trunc = static_cast<int>(c->obj_p[1]);
  std::vector<double> inv(c->num_queries);
  std::vector<int> label_cnt(gains.size());
  for (int q = 0; q < c->num_queries; ++q) {  // CalMaxDCGAtK
    const int s = c->h_query[q], n = c->h_query[q + 1] - s;
    std::fill(label_cnt.begin(), label_cnt.end(), 0);
    for (int i = 0; i < n; ++i) ++label_cnt[static_cast<int>(label[s + i])];
    int top = static_cast<int>(gains.size()) - 1;
    int k = trunc; if (k > n) k = n;
    double ret = 0.0f;
    for (int j = 0; j < k; ++j) {
      while (top > 0 && label_cnt[top] <= 0) top -= 1;
      if (top < 0) break;
      ret += discount[j] * gains[top];
      label_cnt[top] -= 1;
    }
    if (ret > 0.0) ret = 1.0f / ret;
    inv[q] = ret;
  }
  // ConstructSigmoidTable
  const double sigmoid = c->obj_p[0];
  const double min_in = -50.0 / sigmoid / 2, max_in = -min_in;
  const double idx_factor = RANK_SIGMOID_BINS / (max_in - min_in);
  std::vector<double> table(RANK_SIGMOID_BINS);
  for (size_t i = 0; i < RANK_SIGMOID_BINS; ++i) {
    const double s = i / idx_factor + min_in;
    table[i] = 1.0f / (1.0f + std::exp(s * sigmoid));
  }
  double* d_table = nullptr; double* d_disc = nullptr;
  CU(dev_alloc(&c->d_inv_max_dcg, inv.size()));
  CU(cudaMemcpy(c->d_inv_max_dcg, inv.data(), sizeof(double) * inv.size(), cudaMemcpyHostToDevice));
  if (c->d_sigmoid_table) { lgb_dev_free(c->d_sigmoid_table); c->d_sigmoid_table = nullptr; }
  CU(lgb_dev_malloc(reinterpret_cast<void**>(&d_table), sizeof(double) * (RANK_SIGMOID_BINS + RANK_MAX_POSITION)));
  d_disc = d_table + RANK_SIGMOID_BINS;
  CU(cudaMemcpy(d_table, table.data(), sizeof(double) * RANK_SIGMOID_BINS, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d_disc, discount.data(), sizeof(double) * RANK_MAX_POSITION, cudaMemcpyHostToDevice));
  c->d_sigmoid_table = reinterpret_cast<float*>(d_table);
  c->rank_params.sigmoid = sigmoid; c->rank_params.truncation_level = trunc; c->rank_params.norm = c->obj_p[2] != 0;
  c->rank_ready = true;
  return 0;
}

int rank_launch(LgbCuContext* c) {
  RankDev R;
  R.boundaries = c->d_query; R.score = c->d_score; R.label = c->d_label;
  R.weight = c->has_weight ? c->d_weight : nullptr;
  R.label_gain = c->d_label_gain; R.inv_max_dcg = c->d_inv_max_dcg;
  R.sigmoid_table = reinterpret_cast<const double*>(c->d_sigmoid_table);
  R.discount = R.sigmoid_table + RANK_SIGMOID_BINS;
  R.grad = c->d_grad; R.hess = c->d_hess;
  R.sigmoid = c->rank_params.sigmoid;
  R.min_in = -50.0 / R.sigmoid / 2; R.max_in = -R.min_in;
  R.idx_factor = RANK_SIGMOID_BINS / (R.max_in - R.min_in);
  R.norm = c->rank_params.norm;
  const size_t smem = (size_t)c->max_query * (sizeof(double) + 3 * sizeof(int));
  NEED(smem <= 200 * 1024, "lambdarank: query of %d documents does not fit the per-query shared memory tile", c->max_query);
  if (!c->rank_attr_set) {  // the attribute is per device: kept per context, not per process
    CU(cudaFuncSetAttribute(k_lambdarank, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    c->rank_attr_set = true;
  }
  LAUNCH(k_lambdarank, c->num_queries, RANK_THREADS, smem, c->stream, R, c->max_query);
  return 0;
}

}  // namespace

// ---- rank_xendcg (RankXENDCG::GetGradientsForOneQuery, src/objective/rank_objective.hpp:304-356) -------------------
// One thread per query walks the reference's loops in their order (softmax of the scores, the randomised gains
// Phi(label, u) with u from the query's own LCG -- utils/random.h, seeded objective_seed + query and advanced one draw
// per document per iteration --, then the three accumulation passes).  rho / params = scratch of one double per
// document each.  The sums are sequential in the reference, so they are here.
__global__ void k_grad_xendcg(int num_queries, const int* __restrict__ qb, const double* __restrict__ score, const float* __restrict__ label,
                              const float* __restrict__ weight, unsigned* __restrict__ rng, double* __restrict__ rho_all,
                              double* __restrict__ params_all, float* __restrict__ grad, float* __restrict__ hess) {
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < num_queries; q += gridDim.x * blockDim.x) {
    const int b = qb[q], cnt = qb[q + 1] - b;
    float* lambdas = grad + b; float* hessians = hess + b;
    if (cnt <= 1) {
      for (int i = 0; i < cnt; ++i) { lambdas[i] = 0.0f; hessians[i] = 0.0f; }
    } else {
      const double* sc = score + b; const float* lab = label + b;
      double* rho = rho_all + b; double* params = params_all + b;
      double wmax = sc[0];   // Common::Softmax
      for (int i = 1; i < cnt; ++i) wmax = fmax(sc[i], wmax);
      double wsum = 0.0f;
      for (int i = 0; i < cnt; ++i) { rho[i] = exp(sc[i] - wmax); wsum += rho[i]; }
      for (int i = 0; i < cnt; ++i) rho[i] /= wsum;
      unsigned x = rng[q];
      double inv_denominator = 0;
      for (int i = 0; i < cnt; ++i) {
        x = 214013u * x + 2531011u;                                         // Random::NextFloat
        const double g = (double)((float)((x >> 16) & 0x7FFF) / 32768.0f);
        params[i] = ldexp(1.0, (int)lab[i]) - g;                            // Phi: 2^label - u
        inv_denominator += params[i];
      }
      rng[q] = x;
      inv_denominator = 1. / fmax(LGB_EPS, inv_denominator);
      double sum_l1 = 0.0;
      for (int i = 0; i < cnt; ++i) {
        const double term = -params[i] * inv_denominator + rho[i];
        lambdas[i] = (float)term;
        params[i] = term / (1. - rho[i]);
        sum_l1 += params[i];
      }
      double sum_l2 = 0.0;
      for (int i = 0; i < cnt; ++i) {
        const double term = rho[i] * (sum_l1 - params[i]);
        lambdas[i] += (float)term;
        params[i] = term / (1. - rho[i]);
        sum_l2 += params[i];
      }
      for (int i = 0; i < cnt; ++i) {
        lambdas[i] += (float)(rho[i] * (sum_l2 - params[i]));
        hessians[i] = (float)(rho[i] * (1.0 - rho[i]));
      }
    }
    if (weight) {   // RankingObjective::GetGradients (:56-62)
      for (int i = 0; i < cnt; ++i) { lambdas[i] = (float)(lambdas[i] * weight[b + i]); hessians[i] = (float)(hessians[i] * weight[b + i]); }
    }
  }
}

