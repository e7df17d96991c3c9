// Host side of the validation-set / metric entry points declared in include/lgbcu.h (kernels: eval_kernels.cuh).
// Included at the end of lgbcu.cu; the entry points get C linkage from their declarations in lgbcu.h.

namespace {

int valid_get(LgbCuContext* c, int32_t set, ValidSet** out) {
  NEED(c, "lgbcu: NULL context");
  NEED(set >= 0 && set < (int)c->valid.size(), "lgbcu: validation set %d does not exist", set);
  *out = &c->valid[set];
  return 0;
}

template <typename T>
int upload_vec(LgbCuContext* c, T** dst, const T* src, size_t n) {
  CU(dev_alloc(dst, n));
  CU(cudaMemcpyAsync(*dst, src, sizeof(T) * n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

void free_ndcg(NdcgTables* t) {
  void* p[] = {t->d_eval_at, t->d_inv_max_dcg, t->d_query_weight, t->d_label_gain, t->d_discount};
  for (void* q : p) if (q) lgb_dev_free(q);
  *t = NdcgTables();
}

void free_valid_sets(LgbCuContext* c) {
  for (ValidSet& v : c->valid) {
    void* p[] = {v.d_bins, v.d_score_base, v.d_label, v.d_weight, v.d_query};
    for (void* q : p) if (q) lgb_dev_free(q);
    free_ndcg(&v.ndcg);
  }
  c->valid.clear();
  free_ndcg(&c->train_ndcg);
  void* p[] = {c->d_eval_partial, c->d_eval_out, c->d_auc_keys[0], c->d_auc_keys[1], c->d_auc_vals[0], c->d_auc_vals[1],
               c->d_auc_pos, c->d_auc_groups, c->d_cub_tmp, c->d_ndcg_per_query};
  for (void* q : p) if (q) lgb_dev_free(q);
  if (c->h_eval_out) lgb_host_free(c->h_eval_out);
  c->d_eval_partial = c->d_eval_out = c->d_auc_pos = c->d_ndcg_per_query = nullptr; c->h_eval_out = nullptr;
  c->d_auc_keys[0] = c->d_auc_keys[1] = nullptr; c->d_auc_vals[0] = c->d_auc_vals[1] = nullptr;
  c->d_auc_groups = nullptr; c->d_cub_tmp = nullptr; c->cub_tmp_bytes = 0; c->auc_cap = 0; c->ndcg_cap = 0;
}

}  // namespace

int lgbcu_valid_create(LgbCuContext* c, int32_t num_data, const uint8_t* bins, int32_t* out_set) {
  NEED(c && bins && out_set, "lgbcu_valid_create: NULL argument");
  NEED(c->N > 0, "lgbcu_valid_create: set the training data first");
  NEED(num_data > 0, "lgbcu_valid_create: empty validation set");
  USE_DEVICE(c->device);
  ValidSet v;
  v.n = num_data;
  CU(dev_alloc(&v.d_bins, (size_t)num_data * c->stride));
  {
    struct Tmp { uint8_t* p = nullptr; ~Tmp() { if (p) lgb_dev_free(p); } } packed;   // released on every way out
    CU(dev_alloc(&packed.p, (size_t)num_data * c->FS));
    CU(cudaMemcpyAsync(packed.p, bins, (size_t)num_data * c->FS, cudaMemcpyHostToDevice, c->stream));
    LAUNCH(k_pad_rows, grid_for(c, (long long)num_data * (c->stride / 4), 256), 256, 0, c->stream, (long long)num_data,
           c->FS, c->stride, packed.p, v.d_bins);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(c->stream));
  }
  CU(dev_alloc(&v.d_score_base, (size_t)num_data * c->num_class));
  CU(cudaMemsetAsync(v.d_score_base, 0, sizeof(double) * num_data * c->num_class, c->stream));
  v.d_score = v.d_score_base + (size_t)c->cur_class * num_data;
  CU(cudaStreamSynchronize(c->stream));
  c->valid.push_back(v);
  *out_set = (int)c->valid.size() - 1;
  return 0;
}

int lgbcu_valid_set_label(LgbCuContext* c, int32_t set, const float* label, int32_t n) {
  ValidSet* v; if (valid_get(c, set, &v)) return -1;
  NEED(label && n == v->n, "lgbcu_valid_set_label: length %d != %d rows", n, v->n);
  USE_DEVICE(c->device);
  return upload_vec(c, &v->d_label, label, (size_t)n);
}

int lgbcu_valid_set_weight(LgbCuContext* c, int32_t set, const float* weight, int32_t n) {
  ValidSet* v; if (valid_get(c, set, &v)) return -1;
  NEED(weight && n == v->n, "lgbcu_valid_set_weight: length %d != %d rows", n, v->n);
  USE_DEVICE(c->device);
  return upload_vec(c, &v->d_weight, weight, (size_t)n);
}

int lgbcu_valid_set_query(LgbCuContext* c, int32_t set, const int32_t* boundaries, int32_t num_queries) {
  ValidSet* v; if (valid_get(c, set, &v)) return -1;
  NEED(boundaries && num_queries > 0, "lgbcu_valid_set_query: bad argument");
  NEED(boundaries[0] == 0 && boundaries[num_queries] == v->n, "lgbcu_valid_set_query: boundaries must span [0, num_data]");
  for (int q = 0; q < num_queries; ++q) NEED(boundaries[q + 1] >= boundaries[q], "lgbcu_valid_set_query: boundaries not monotone");
  USE_DEVICE(c->device);
  v->num_queries = num_queries;
  return upload_vec(c, &v->d_query, boundaries, (size_t)num_queries + 1);
}

int lgbcu_valid_set_score(LgbCuContext* c, int32_t set, const double* score, int32_t n) {
  ValidSet* v; if (valid_get(c, set, &v)) return -1;
  NEED(score && n == v->n, "lgbcu_valid_set_score: length %d != %d rows", n, v->n);
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(v->d_score, score, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_valid_copy_score(LgbCuContext* c, int32_t set, double* out, int32_t n) {
  ValidSet* v; if (valid_get(c, set, &v)) return -1;
  NEED(out && n == v->n, "lgbcu_valid_copy_score: length %d != %d rows", n, v->n);
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(out, v->d_score, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_valid_add_const(LgbCuContext* c, int32_t set, double value) {
  ValidSet* v; if (valid_get(c, set, &v)) return -1;
  USE_DEVICE(c->device);
  LAUNCH(k_add_const, grid_for(c, v->n, 256), 256, 0, c->stream, v->n, v->d_score, value);
  CU(cudaGetLastError());
  return 0;
}

int lgbcu_valid_multiply_score(LgbCuContext* c, int32_t set, double value) {
  ValidSet* v; if (valid_get(c, set, &v)) return -1;
  USE_DEVICE(c->device);
  LAUNCH(k_mul_const, grid_for(c, v->n, 256), 256, 0, c->stream, v->n, v->d_score, value);
  CU(cudaGetLastError());
  return 0;
}

int lgbcu_valid_add_tree(LgbCuContext* c, int32_t set, int32_t num_leaves, const int32_t* split_feature_inner,
                         const int32_t* threshold_bin, const int8_t* decision_type, const int32_t* left_child,
                         const int32_t* right_child, const double* leaf_value) {
  ValidSet* v; if (valid_get(c, set, &v)) return -1;
  NEED(num_leaves >= 1 && leaf_value, "lgbcu_valid_add_tree: bad tree");
  USE_DEVICE(c->device);
  if (num_leaves == 1) {  // constant tree (tree.cpp:110-118)
    if (leaf_value[0] == 0.0) return 0;
    return lgbcu_valid_add_const(c, set, leaf_value[0]);
  }
  TreeDev T;
  if (upload_tree(c, num_leaves, split_feature_inner, threshold_bin, decision_type, left_child, right_child, leaf_value, &T)) return -1;
  DevView V = c->view();
  V.bins = v->d_bins;  // same stride and feature meta as the training matrix
  LAUNCH(k_add_score_tree, grid_for(c, v->n, 256), 256, 0, c->stream, V, T, (const int*)nullptr, v->n, v->d_score);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(c->stream));  // the caller's host arrays may go away
  return 0;
}

int lgbcu_eval_set_ndcg(LgbCuContext* c, int32_t set, const int32_t* eval_at, int32_t num_eval, const double* inv_max_dcg,
                        const float* query_weight, const double* label_gain, int32_t num_label_gain,
                        const double* discount, int32_t num_discount) {
  NEED(c && eval_at && inv_max_dcg && label_gain && discount, "lgbcu_eval_set_ndcg: NULL argument");
  NEED(num_eval >= 1 && num_eval <= 32, "lgbcu_eval_set_ndcg: %d evaluation positions (1..32 supported)", num_eval);
  NdcgTables* t; int nq;
  if (set < 0) { t = &c->train_ndcg; nq = c->num_queries; }
  else { ValidSet* v; if (valid_get(c, set, &v)) return -1; t = &v->ndcg; nq = v->num_queries; }
  NEED(nq > 0, "The NDCG metric requires query information");
  int max_k = 0;
  for (int i = 0; i < num_eval; ++i) { NEED(eval_at[i] > 0, "lgbcu_eval_set_ndcg: eval_at must be positive"); max_k = std::max(max_k, eval_at[i]); }
  NEED(max_k <= NDCG_MAX_K && max_k <= num_discount, "lgbcu_eval_set_ndcg: eval_at %d exceeds the supported %d", max_k, NDCG_MAX_K);
  USE_DEVICE(c->device);
  free_ndcg(t);
  if (upload_vec(c, &t->d_eval_at, eval_at, (size_t)num_eval)) return -1;
  if (upload_vec(c, &t->d_inv_max_dcg, inv_max_dcg, (size_t)nq * num_eval)) return -1;
  if (upload_vec(c, &t->d_label_gain, label_gain, (size_t)num_label_gain)) return -1;
  if (upload_vec(c, &t->d_discount, discount, (size_t)max_k)) return -1;
  t->sum_query_weights = (double)nq;
  if (query_weight) {
    if (upload_vec(c, &t->d_query_weight, query_weight, (size_t)nq)) return -1;
    double sw = 0.0;
    for (int q = 0; q < nq; ++q) sw += query_weight[q];  // NDCGMetric::Init: sequential, in double
    t->sum_query_weights = sw;
  }
  t->num_eval = num_eval; t->max_k = max_k; t->ready = true;
  return 0;
}

namespace {

int eval_scratch(LgbCuContext* c) {
  if (!c->d_eval_partial) {
    CU(dev_alloc(&c->d_eval_partial, (size_t)EVAL_MAX_BLOCKS * 2));
    CU(dev_alloc(&c->d_eval_out, (size_t)64));
    CU(lgb_host_malloc(reinterpret_cast<void**>(&c->h_eval_out), sizeof(double) * 64));
  }
  return 0;
}

// sum of the row weights in double (Metric::Init: sum_weights_ += weights_[i]); n when there are none
int eval_sum_weights(LgbCuContext* c, int n, const float* d_weight, double* out) {
  if (!d_weight) { *out = (double)n; return 0; }
  const int grid = std::min(EVAL_MAX_BLOCKS, (n + EVAL_THREADS - 1) / EVAL_THREADS);
  LAUNCH(k_sum_f32, grid, EVAL_THREADS, 0, c->stream, n, d_weight, c->d_eval_partial);
  LAUNCH(k_eval_final, 1, EVAL_THREADS, 0, c->stream, grid, 1, c->d_eval_partial, c->d_eval_out + 8);
  CU(cudaMemcpyAsync(c->h_eval_out + 8, c->d_eval_out + 8, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  *out = c->h_eval_out[8];
  return 0;
}

int eval_auc(LgbCuContext* c, int n, const double* score, const float* label, const float* weight, double sum_w, double* out) {
  if (c->auc_cap < n) {
    for (int i = 0; i < 2; ++i) { CU(dev_alloc(&c->d_auc_keys[i], (size_t)n)); CU(dev_alloc(&c->d_auc_vals[i], (size_t)n)); }
    CU(dev_alloc(&c->d_auc_pos, (size_t)n));
    CU(dev_alloc(&c->d_auc_groups, (size_t)1));
    c->auc_cap = n;
  }
  size_t b0 = 0, b1 = 0, b2 = 0;
  CU(cub::DeviceRadixSort::SortPairs(nullptr, b0, c->d_auc_keys[0], c->d_auc_keys[1], c->d_auc_vals[0], c->d_auc_vals[1], n, 0, 64, c->stream));
  CU(cub::DeviceReduce::ReduceByKey(nullptr, b1, c->d_auc_keys[1], c->d_auc_keys[0], c->d_auc_vals[1], c->d_auc_vals[0],
                                    c->d_auc_groups, PairSum(), n, c->stream));
  CU(cub::DeviceScan::InclusiveSum(nullptr, b2, c->d_auc_pos, c->d_auc_pos, n, c->stream));
  const size_t need = std::max(b0, std::max(b1, b2));
  if (c->cub_tmp_bytes < need) {
    if (c->d_cub_tmp) lgb_dev_free(c->d_cub_tmp);
    c->d_cub_tmp = nullptr;
    CU(lgb_dev_malloc(&c->d_cub_tmp, need));
    c->cub_tmp_bytes = need;
  }
  size_t tb = c->cub_tmp_bytes;
  LAUNCH(k_auc_keys, grid_for(c, n, 256), 256, 0, c->stream, n, score, label, weight, c->d_auc_keys[0], c->d_auc_vals[0]);
  CU(cub::DeviceRadixSort::SortPairs(c->d_cub_tmp, tb, c->d_auc_keys[0], c->d_auc_keys[1], c->d_auc_vals[0], c->d_auc_vals[1], n, 0, 64, c->stream));
  // groups of equal score, in ascending score order: (sum of positive weights, sum of negative weights)
  CU(cudaMemsetAsync(c->d_auc_vals[0], 0, sizeof(double2) * (size_t)n, c->stream));
  tb = c->cub_tmp_bytes;
  CU(cub::DeviceReduce::ReduceByKey(c->d_cub_tmp, tb, c->d_auc_keys[1], c->d_auc_keys[0], c->d_auc_vals[1], c->d_auc_vals[0],
                                    c->d_auc_groups, PairSum(), n, c->stream));
  tb = c->cub_tmp_bytes;
  LAUNCH(k_auc_pos, grid_for(c, n, 256), 256, 0, c->stream, n, c->d_auc_vals[0], c->d_auc_pos);
  CU(cub::DeviceScan::InclusiveSum(c->d_cub_tmp, tb, c->d_auc_pos, c->d_auc_pos, n, c->stream));
  const int grid = std::min(EVAL_MAX_BLOCKS, (n + EVAL_THREADS - 1) / EVAL_THREADS);
  LAUNCH(k_auc_accum, grid, EVAL_THREADS, 0, c->stream, c->d_auc_groups, c->d_auc_vals[0], c->d_auc_pos, c->d_eval_partial);
  LAUNCH(k_eval_final, 1, EVAL_THREADS, 0, c->stream, grid, 2, c->d_eval_partial, c->d_eval_out);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(c->h_eval_out, c->d_eval_out, sizeof(double) * 2, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  const double accum = c->h_eval_out[0], sum_pos = c->h_eval_out[1];
  double auc = 1.0f;
  if (sum_pos > 0.0f && sum_pos != sum_w) auc = accum / (sum_pos * (sum_w - sum_pos));
  *out = auc;
  return 0;
}

}  // namespace

int lgbcu_eval(LgbCuContext* c, int32_t set, int32_t kind, double sigmoid, double* out, int32_t out_len) {
  NEED(c && out && out_len >= 1, "lgbcu_eval: bad argument");
  NEED(kind >= LGBCU_METRIC_L2 && kind <= LGBCU_METRIC_GAMMA_DEVIANCE, "lgbcu_eval: unknown metric kind %d", kind);
  USE_DEVICE(c->device);
  int n; const double* score; const float* label; const float* weight; const int* query; int nq; NdcgTables* nd;
  if (set < 0) {
    NEED(c->N > 0, "lgbcu_eval: no training data");
    n = c->N; score = c->d_score; label = c->has_label ? c->d_label : nullptr; weight = c->has_weight ? c->d_weight : nullptr;
    if (kind == LGBCU_METRIC_MULTI_LOGLOSS || kind == LGBCU_METRIC_MULTI_ERROR) score = c->d_score_base;
    query = c->d_query; nq = c->num_queries; nd = &c->train_ndcg;
  } else {
    ValidSet* v; if (valid_get(c, set, &v)) return -1;
    n = v->n; score = v->d_score; label = v->d_label; weight = v->d_weight; query = v->d_query; nq = v->num_queries; nd = &v->ndcg;
    if (kind == LGBCU_METRIC_MULTI_LOGLOSS || kind == LGBCU_METRIC_MULTI_ERROR) score = v->d_score_base;
  }
  NEED(label, "lgbcu_eval: the data set has no label");
  if (eval_scratch(c)) return -1;
  if (kind == LGBCU_METRIC_MAP) {
    NEED(nd->ready && query && nq > 0, "The MAP metric requires query information");
    NEED(out_len >= nd->num_eval, "lgbcu_eval: out_len %d < %d evaluation positions", out_len, nd->num_eval);
    const size_t need = (size_t)nq * nd->num_eval;
    if (c->ndcg_cap < need) { CU(dev_alloc(&c->d_ndcg_per_query, need)); c->ndcg_cap = need; }
    LAUNCH(k_map, std::min(nq, c->num_sms * 8), EVAL_THREADS, 0, c->stream, nq, query, score, label, nd->d_query_weight,
           nd->d_eval_at, nd->num_eval, nd->max_k, c->d_ndcg_per_query);
    LAUNCH(k_eval_final, 1, EVAL_THREADS, 0, c->stream, nq, nd->num_eval, c->d_ndcg_per_query, c->d_eval_out);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(c->h_eval_out, c->d_eval_out, sizeof(double) * nd->num_eval, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    for (int e = 0; e < nd->num_eval; ++e) out[e] = c->h_eval_out[e] / nd->sum_query_weights;
    return 0;
  }
  if (kind == LGBCU_METRIC_NDCG) {
    NEED(nd->ready && query && nq > 0, "The NDCG metric requires query information");
    NEED(out_len >= nd->num_eval, "lgbcu_eval: out_len %d < %d evaluation positions", out_len, nd->num_eval);
    const size_t need = (size_t)nq * nd->num_eval;
    if (c->ndcg_cap < need) { CU(dev_alloc(&c->d_ndcg_per_query, need)); c->ndcg_cap = need; }
    LAUNCH(k_ndcg, std::min(nq, c->num_sms * 8), EVAL_THREADS, 0, c->stream, nq, query, score, label, nd->d_query_weight,
           nd->d_label_gain, nd->d_discount, nd->d_eval_at, nd->num_eval, nd->max_k, nd->d_inv_max_dcg, c->d_ndcg_per_query);
    LAUNCH(k_eval_final, 1, EVAL_THREADS, 0, c->stream, nq, nd->num_eval, c->d_ndcg_per_query, c->d_eval_out);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(c->h_eval_out, c->d_eval_out, sizeof(double) * nd->num_eval, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    for (int e = 0; e < nd->num_eval; ++e) out[e] = c->h_eval_out[e] / nd->sum_query_weights;
    return 0;
  }
  double sum_w = 0.0;
  if (eval_sum_weights(c, n, weight, &sum_w)) return -1;
  if (kind == LGBCU_METRIC_AUC) return eval_auc(c, n, score, label, weight, sum_w, out);
  const int grid = std::min(EVAL_MAX_BLOCKS, (n + EVAL_THREADS - 1) / EVAL_THREADS);
  if (kind == LGBCU_METRIC_MULTI_LOGLOSS || kind == LGBCU_METRIC_MULTI_ERROR) {
    // `sigmoid` carries the parameter: multi_logloss > 0 = apply the objective's softmax; multi_error: top_k, negative
    // when the scores are to be taken as they are (no objective)
    const bool ova = c->objective == LGBCU_OBJ_MULTICLASSOVA;   // its ConvertOutput is a sigmoid per class, not a softmax
    const int softmax = (sigmoid > 0.0 && !ova) ? 1 : 0;
    const double ova_sigmoid = (sigmoid > 0.0 && ova) ? c->obj_p[1] : 0.0;
    const int top_k = kind == LGBCU_METRIC_MULTI_ERROR ? (int)std::fabs(sigmoid) : 1;
    LAUNCH(k_eval_multiclass, grid, EVAL_THREADS, 0, c->stream, n, c->num_class, kind == LGBCU_METRIC_MULTI_LOGLOSS ? 0 : 1, top_k,
           softmax, ova_sigmoid, score, label, weight, c->d_eval_partial);
    LAUNCH(k_eval_final, 1, EVAL_THREADS, 0, c->stream, grid, 1, c->d_eval_partial, c->d_eval_out);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(c->h_eval_out, c->d_eval_out, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    out[0] = c->h_eval_out[0] / sum_w;
    return 0;
  }
  const int pk = kind == LGBCU_METRIC_BINARY_LOGLOSS ? 1 : kind == LGBCU_METRIC_L1 ? 2 : kind == LGBCU_METRIC_QUANTILE ? 3 :
                 kind == LGBCU_METRIC_HUBER ? 4 : kind == LGBCU_METRIC_FAIR ? 5 : kind == LGBCU_METRIC_POISSON ? 6 :
                 kind == LGBCU_METRIC_GAMMA ? 7 : kind == LGBCU_METRIC_TWEEDIE ? 8 : kind == LGBCU_METRIC_BINARY_ERROR ? 9 :
                 kind == LGBCU_METRIC_XENTROPY ? 10 : kind == LGBCU_METRIC_MAPE ? 11 : kind == LGBCU_METRIC_XENTLAMBDA ? 12 : kind == LGBCU_METRIC_GAMMA_DEVIANCE ? 13 : 0;
  // metrics see the objective's ConvertOutput: exp for the log-link objectives
  const int conv_exp = (c->objective == LGBCU_OBJ_POISSON || c->objective == LGBCU_OBJ_GAMMA || c->objective == LGBCU_OBJ_TWEEDIE) ? 1 :
                       (c->objective == LGBCU_OBJ_L2 && c->obj_p[0] != 0.0) ? 2 : 0;   // 2: the reg_sqrt objective squares its output
  LAUNCH(k_eval_pointwise, grid, EVAL_THREADS, 0, c->stream, n, pk, sigmoid, conv_exp, score,
         label, weight, c->d_eval_partial);
  LAUNCH(k_eval_final, 1, EVAL_THREADS, 0, c->stream, grid, 1, c->d_eval_partial, c->d_eval_out);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(c->h_eval_out, c->d_eval_out, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  // (cross_entropy_lambda keeps the weight inside the loss and averages over the rows: xentropy_metric.hpp:224)
  double avg = c->h_eval_out[0] / (kind == LGBCU_METRIC_XENTLAMBDA ? (double)n : sum_w);
  if (kind == LGBCU_METRIC_GAMMA_DEVIANCE) avg = c->h_eval_out[0] * 2;   // GammaDevianceMetric::AverageLoss: the sum, doubled
  out[0] = kind == LGBCU_METRIC_RMSE ? std::sqrt(avg) : avg;
  return 0;
}
