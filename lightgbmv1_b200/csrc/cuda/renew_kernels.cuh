// RenewTreeOutput on the device (replaces SerialTreeLearner::RenewTreeOutput,
// src/treelearner/serial_tree_learner.cpp:661-699, and the PercentileFun / WeightedPercentileFun macros of
// src/objective/regression_objective.hpp:18-88 that RegressionL1loss / RegressionQuantileloss call per leaf).
//
// Objectives whose leaf value is a percentile of the residuals (L1: the median, quantile: the alpha-quantile) ignore
// the Newton step of a freshly grown tree: every leaf's output is recomputed from label - score of the rows the tree
// grew on.  The rows of a leaf are a contiguous range of the partition's index buffer, so the leaves are segments:
//   1. gather   residual (double) [and weight] of every in-bag row into its position of that range;
//   2. sort     every segment ascending (cub segmented sort; stable for the weighted form, as the reference's
//               std::stable_sort on indices is);
//   3. pick     unweighted: the two order statistics around (1 - alpha) * count and their interpolation, exactly the
//               values the reference's ArgMaxAtK selection returns; weighted: the weighted CDF (a prefix sum per
//               segment), std::upper_bound of alpha * total, and the reference's interpolation rule.
// The sort is library code (cub), like the AUC metric's; gather and pick are kernels of this file.
#pragma once

__global__ void k_renew_gather(DevView V, int num_leaves, const double* __restrict__ score, const float* __restrict__ label,
                               const float* __restrict__ weight, int mape, double* __restrict__ keys, float* __restrict__ wout,
                               int* __restrict__ seg, int* __restrict__ seg_begin, int* __restrict__ seg_end) {
  const int l = blockIdx.x;
  if (l >= num_leaves) return;
  const DevLeaf L = V.leaves[l];
  if (blockIdx.y == 0 && threadIdx.x == 0) { seg_begin[l] = L.begin; seg_end[l] = L.begin + L.count; }
  const int* idx = V.indices[L.buf] + L.begin;
  for (int p = blockIdx.y * blockDim.x + threadIdx.x; p < L.count; p += gridDim.y * blockDim.x) {
    const int row = idx[p];
    keys[L.begin + p] = (double)label[row] - score[row];   // residual_getter (gbdt.cpp:372)
    if (wout) wout[L.begin + p] = mape ? mape_label_weight(label[row]) : weight[row];   // mape: RegressionMAPELOSS::label_weight_
    if (seg) seg[L.begin + p] = l;
  }
}

// PercentileFun (regression_objective.hpp:18-48) on a segment sorted ASCENDING: the reference's "k-th largest" is
// asc[cnt - 1 - k].  alpha_is_float: RegressionQuantileloss holds alpha as a float and forms (1.0f - alpha) * cnt in
// float; RegressionL1loss uses the double 0.5.
__global__ void k_renew_pick(int num_leaves, const int* __restrict__ seg_begin, const int* __restrict__ seg_end,
                             const double* __restrict__ sorted, double alpha, int alpha_is_float, double* __restrict__ out,
                             int* __restrict__ nonempty) {
  for (int l = blockIdx.x * blockDim.x + threadIdx.x; l < num_leaves; l += gridDim.x * blockDim.x) {
    const int b = seg_begin[l], cnt = seg_end[l] - b;
    nonempty[l] = cnt > 0;
    if (cnt <= 0) { out[l] = 0.0; continue; }
    const double* a = sorted + b;
    if (cnt <= 1) { out[l] = a[0]; continue; }
    double float_pos;
    if (alpha_is_float) { const float fa = (float)alpha; float_pos = (double)((1.0f - fa) * (float)cnt); }
    else float_pos = (1.0f - alpha) * cnt;
    const int pos = (int)float_pos;
    if (pos < 1) { out[l] = a[cnt - 1]; continue; }        // the maximum
    if (pos >= cnt) { out[l] = a[0]; continue; }          // the minimum
    const double bias = float_pos - pos;
    const double v1 = a[cnt - pos], v2 = a[cnt - 1 - pos];  // pos-th and (pos+1)-th largest
    out[l] = v1 - (v1 - v2) * bias;
  }
}

// WeightedPercentileFun (regression_objective.hpp:50-88): `cdf` = inclusive prefix sums of the weights in sorted order
__global__ void k_renew_pick_weighted(int num_leaves, const int* __restrict__ seg_begin, const int* __restrict__ seg_end,
                                      const double* __restrict__ sorted, const double* __restrict__ cdf, double alpha,
                                      double* __restrict__ out, int* __restrict__ nonempty) {
  for (int l = blockIdx.x * blockDim.x + threadIdx.x; l < num_leaves; l += gridDim.x * blockDim.x) {
    const int b = seg_begin[l], cnt = seg_end[l] - b;
    nonempty[l] = cnt > 0;
    if (cnt <= 0) { out[l] = 0.0; continue; }
    const double* a = sorted + b;
    const double* w = cdf + b;
    if (cnt <= 1) { out[l] = a[0]; continue; }
    const double threshold = w[cnt - 1] * alpha;
    int lo = 0, hi = cnt;                                  // std::upper_bound: first cdf > threshold
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (w[mid] > threshold) hi = mid; else lo = mid + 1; }
    int pos = lo < cnt - 1 ? lo : cnt - 1;
    if (pos == 0 || pos == cnt - 1) { out[l] = a[pos]; continue; }
    const double v1 = a[pos - 1], v2 = a[pos];
    if (w[pos + 1] - w[pos] >= 1.0f) out[l] = (threshold - w[pos]) / (w[pos + 1] - w[pos]) * (v2 - v1) + v1;
    else out[l] = v2;
  }
}

__global__ void k_renew_widen(int n, const float* __restrict__ in, double* __restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = (double)in[i];
}

extern "C" int lgbcu_renew_leaf_outputs(LgbCuContext* c, int32_t num_leaves, double alpha, int32_t alpha_is_float,
                                        double* leaf_values, int32_t* leaf_nonempty) {
  NEED(c && leaf_values && leaf_nonempty, "lgbcu_renew_leaf_outputs: NULL argument");
  NEED(c->tree_begun, "lgbcu_renew_leaf_outputs: no tree has been grown");
  NEED(num_leaves >= 1 && num_leaves <= c->alloc_leaves, "lgbcu_renew_leaf_outputs: num_leaves %d out of range", num_leaves);
  NEED(c->has_label, "lgbcu_renew_leaf_outputs: label not set");
  USE_DEVICE(c->device);
  const int n = c->root_count;          // rows the tree grew on (the bag)
  const bool mape = c->objective == LGBCU_OBJ_MAPE;
  const bool weighted = c->has_weight || mape;
  if (c->renew_cap < n) {
    CU(dev_alloc(&c->d_renew_key[0], (size_t)n)); CU(dev_alloc(&c->d_renew_key[1], (size_t)n));
    CU(dev_alloc(&c->d_renew_w[0], (size_t)n)); CU(dev_alloc(&c->d_renew_w[1], (size_t)n));
    CU(dev_alloc(&c->d_renew_cdf, (size_t)n)); CU(dev_alloc(&c->d_renew_wd, (size_t)n)); CU(dev_alloc(&c->d_renew_seg, (size_t)n));
    c->renew_cap = n;
  }
  if (c->renew_leaf_cap < num_leaves) {
    CU(dev_alloc(&c->d_renew_begin, (size_t)num_leaves)); CU(dev_alloc(&c->d_renew_end, (size_t)num_leaves));
    CU(dev_alloc(&c->d_renew_out, (size_t)num_leaves)); CU(dev_alloc(&c->d_renew_nonempty, (size_t)num_leaves));
    c->renew_leaf_cap = num_leaves;
  }
  const DevView V = c->view();
  LAUNCH(k_renew_gather, dim3(num_leaves, 64), 256, 0, c->stream, V, (int)num_leaves, c->d_score, c->d_label,
         c->has_weight ? c->d_weight : nullptr, mape ? 1 : 0, c->d_renew_key[0], weighted ? c->d_renew_w[0] : nullptr,
         weighted ? c->d_renew_seg : nullptr, c->d_renew_begin, c->d_renew_end);
  size_t need = 0;
  if (weighted) {
    CU(cub::DeviceSegmentedSort::StableSortPairs(nullptr, need, c->d_renew_key[0], c->d_renew_key[1], c->d_renew_w[0], c->d_renew_w[1],
                                                 n, num_leaves, c->d_renew_begin, c->d_renew_end, c->stream));
  } else {
    CU(cub::DeviceSegmentedSort::SortKeys(nullptr, need, c->d_renew_key[0], c->d_renew_key[1], n, num_leaves, c->d_renew_begin,
                                          c->d_renew_end, c->stream));
  }
  size_t need2 = 0;
  if (weighted) CU(cub::DeviceScan::InclusiveSumByKey(nullptr, need2, c->d_renew_seg, c->d_renew_wd, c->d_renew_cdf, n, cub::Equality(), c->stream));
  need = std::max(need, need2);
  if (c->cub_tmp_bytes < need) {
    if (c->d_cub_tmp) lgb_dev_free(c->d_cub_tmp);
    c->d_cub_tmp = nullptr; c->cub_tmp_bytes = 0;
    CU(lgb_dev_malloc(&c->d_cub_tmp, need));
    c->cub_tmp_bytes = need;
  }
  size_t bytes = c->cub_tmp_bytes;
  if (weighted) {
    CU(cub::DeviceSegmentedSort::StableSortPairs(c->d_cub_tmp, bytes, c->d_renew_key[0], c->d_renew_key[1], c->d_renew_w[0], c->d_renew_w[1],
                                                 n, num_leaves, c->d_renew_begin, c->d_renew_end, c->stream));
    // the weighted CDF of every segment: double prefix sums of the float weights in sorted order (the segment ids were
    // written position by position, so they are still right after the sort, which never crosses a segment)
    LAUNCH(k_renew_widen, grid_for(c, n, 256), 256, 0, c->stream, n, c->d_renew_w[1], c->d_renew_wd);
    bytes = c->cub_tmp_bytes;
    CU(cub::DeviceScan::InclusiveSumByKey(c->d_cub_tmp, bytes, c->d_renew_seg, c->d_renew_wd, c->d_renew_cdf, n, cub::Equality(), c->stream));
    LAUNCH(k_renew_pick_weighted, grid_for(c, num_leaves, 128), 128, 0, c->stream, (int)num_leaves, c->d_renew_begin, c->d_renew_end,
           c->d_renew_key[1], c->d_renew_cdf, alpha, c->d_renew_out, c->d_renew_nonempty);
  } else {
    CU(cub::DeviceSegmentedSort::SortKeys(c->d_cub_tmp, bytes, c->d_renew_key[0], c->d_renew_key[1], n, num_leaves, c->d_renew_begin,
                                          c->d_renew_end, c->stream));
    LAUNCH(k_renew_pick, grid_for(c, num_leaves, 128), 128, 0, c->stream, (int)num_leaves, c->d_renew_begin, c->d_renew_end,
           c->d_renew_key[1], alpha, (int)alpha_is_float, c->d_renew_out, c->d_renew_nonempty);
  }
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(leaf_values, c->d_renew_out, sizeof(double) * num_leaves, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(leaf_nonempty, c->d_renew_nonempty, sizeof(int) * num_leaves, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}
