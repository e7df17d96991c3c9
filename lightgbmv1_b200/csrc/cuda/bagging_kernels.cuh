// Bagging on the device (replaces GBDT::Bagging + BaggingHelper, src/boosting/gbdt.cpp:162-181,
// 209-243): every 1024-row block draws from its own LCG (include/LightGBM/utils/random.h: x = 214013 x + 2531011,
// NextFloat = ((x >> 16) & 0x7fff) / 32768), which makes the sample independent of the number of threads -- and lets a
// GPU reproduce it exactly: a warp takes one block, lane l jumps the block's generator ahead by 32 l steps (an affine
// map composed on the host) and draws the 32 numbers of rows [32 l, 32 l + 32), which become one ballot-style bit word.
// A scan over the per-block counts and a stable scatter then write the in-bag rows (ascending) followed by the
// out-of-bag rows (ascending) -- the list the tree grows on and the list the tree walk scores afterwards.  Nothing
// crosses the PCIe bus but the bag size.
#pragma once

#define BAG_BLOCK 1024   /* GBDT::bagging_rand_block_ (gbdt.h:519) == PART_TILE: one bit-word row per generator */
static_assert(BAG_BLOCK == PART_TILE, "one partition tile per bagging generator");

struct BagJump { unsigned a[32], c[32]; };   // x_{32 l} = a[l] * x_0 + c[l]  (mod 2^32)

__global__ void k_bag_seed(int nblock, int seed, unsigned* __restrict__ state) {
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < nblock; b += gridDim.x * blockDim.x) state[b] = (unsigned)(seed + b);
}

// one warp per 1024-row block
// label == nullptr: every row is kept with probability `fraction`; otherwise rows with label > 0 with `fraction`, the
// others with `fraction_neg` (BalancedBaggingHelper, gbdt.cpp:183-207)
__global__ void __launch_bounds__(256) k_bag_flags(int n, double fraction, double fraction_neg, const float* __restrict__ label,
                                                   BagJump J, unsigned* __restrict__ state,
                                                   unsigned* __restrict__ bits, int* __restrict__ tile_left) {
  const int lane = threadIdx.x & 31;
  const int nblock = (n + BAG_BLOCK - 1) / BAG_BLOCK;
  for (int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); b < nblock; b += gridDim.x * (blockDim.x >> 5)) {
    const unsigned x0 = state[b];
    unsigned x = J.a[lane] * x0 + J.c[lane];
    unsigned word = 0, x_last = x;
    const int row0 = b * BAG_BLOCK + lane * 32;
#pragma unroll 8
    for (int i = 0; i < 32; ++i) {
      x = 214013u * x + 2531011u;
      const float r = (float)((x >> 16) & 0x7FFFu) / 32768.0f;
      if (row0 + i < n) {
        x_last = x;
        const double fr = (label == nullptr || label[row0 + i] > 0.f) ? fraction : fraction_neg;
        if ((double)r < fr) word |= 1u << i;
      }
    }
    bits[(size_t)b * 32 + lane] = word;
    int cnt = __popc(word);
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    __syncwarp();
    if (lane == 0) tile_left[b] = cnt;
    // the generator has drawn once per EXISTING row of its block (the last block of the data may be partial): the lane
    // that owns the block's last row holds its state
    const int rows_here = min(BAG_BLOCK, n - b * BAG_BLOCK);
    if (lane == (rows_here - 1) / 32) state[b] = x_last;
  }
}

// in-bag rows (ascending) to out[0, bag), out-of-bag rows (ascending) to out[bag, n)
__global__ void __launch_bounds__(256) k_bag_scatter(int n, const unsigned* __restrict__ bits, const int* __restrict__ tile_off,
                                                     const unsigned* __restrict__ ctl, int* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int nblock = (n + BAG_BLOCK - 1) / BAG_BLOCK;
  const int bag = (int)ctl[0];
  for (int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); b < nblock; b += gridDim.x * (blockDim.x >> 5)) {
    const unsigned word = bits[(size_t)b * 32 + lane];
    const int pc = __popc(word);
    int inc = pc;
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    const int in_before = tile_off[b] + inc - pc;                  // in-bag rows before this lane's 32 rows
    const int row0 = b * BAG_BLOCK + lane * 32;
    int li = in_before, ri = bag + (row0 - in_before);
    for (int i = 0; i < 32 && row0 + i < n; ++i) {
      if ((word >> i) & 1u) out[li++] = row0 + i; else out[ri++] = row0 + i;
    }
  }
}

// ---- GOSS (src/boosting/goss.hpp:103-152) -----------------------------------------------------------------------
// The sampling decision of a row depends on how many rows of its thread chunk were sampled before it -- a sequential
// walk per chunk -- so it stays on the host (gbdt.cpp GossSample).  The device supplies |g * h| per row in float, as
// BaggingHelper forms it, and afterwards applies the result: two bit rows per 1024-row block (in the bag / sampled
// small-gradient row whose gradient and hessian are scaled by its chunk's multiplier), from which the same scan +
// stable scatter as plain bagging builds the bag.
__global__ void k_goss_scores(int n, const float* __restrict__ g, const float* __restrict__ h, float* __restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = fabsf(g[i] * h[i]);
}

// one 1024-thread block per 1024-row block: thread = row
__global__ void __launch_bounds__(BAG_BLOCK) k_goss_apply(int n, const unsigned* __restrict__ bits, const unsigned* __restrict__ amp,
                                                         int chunk_rows, const float* __restrict__ chunk_mult,
                                                         float* __restrict__ g, float* __restrict__ h, int* __restrict__ tile_left) {
  const int nblock = (n + BAG_BLOCK - 1) / BAG_BLOCK;
  for (int b = blockIdx.x; b < nblock; b += gridDim.x) {
    const int t = threadIdx.x, row = b * BAG_BLOCK + t;
    const unsigned aw = amp[(size_t)b * 32 + (t >> 5)];
    if (row < n && ((aw >> (t & 31)) & 1u)) {
      const float m = chunk_mult[row / chunk_rows];
      g[row] *= m; h[row] *= m;
    }
    if (t < 32) {
      int cnt = __popc(bits[(size_t)b * 32 + t]);
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      if (t == 0) tile_left[b] = cnt;
    }
  }
}

extern "C" {

int lgbcu_goss_scores(LgbCuContext* c, float* out, int32_t n) {
  NEED(c && out && n == c->N && c->d_grad && c->d_hess, "lgbcu_goss_scores: bad argument / no gradients");
  USE_DEVICE(c->device);
  if (!c->d_goss) CU(dev_alloc(&c->d_goss, (size_t)c->N));
  LAUNCH(k_goss_scores, grid_for(c, c->N, 256), 256, 0, c->stream, c->N, c->d_grad, c->d_hess, c->d_goss);
  CU(cudaGetLastError());
  // straight into the caller's buffer (pinned or pageable); rows are spread over the copy threads by the caller's walk
  CU(cudaMemcpyAsync(out, c->d_goss, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_goss_apply(LgbCuContext* c, const uint32_t* inbag_bits, const uint32_t* amplify_bits, int32_t chunk_rows,
                     const float* chunk_multiply, int32_t nchunks, int32_t* bag_count) {
  NEED(c && inbag_bits && amplify_bits && chunk_multiply && bag_count && chunk_rows > 0 && nchunks > 0, "lgbcu_goss_apply: bad argument");
  NEED((long long)chunk_rows * nchunks >= c->N, "lgbcu_goss_apply: %d chunks of %d rows do not cover %d rows", nchunks, chunk_rows, c->N);
  NEED(c->d_bag_state, "lgbcu_goss_apply: call lgbcu_bagging_reset first");
  USE_DEVICE(c->device);
  const int nblock = (c->N + BAG_BLOCK - 1) / BAG_BLOCK;
  if (!c->d_goss_amp) CU(dev_alloc(&c->d_goss_amp, (size_t)nblock * 32));
  if (!c->d_goss_mult || c->goss_mult_cap < nchunks) {
    if (c->d_goss_mult) { lgb_dev_free(c->d_goss_mult); c->d_goss_mult = nullptr; }
    CU(dev_alloc(&c->d_goss_mult, (size_t)nchunks));
    c->goss_mult_cap = nchunks;
  }
  CU(cudaMemcpyAsync(c->d_bag_bits, inbag_bits, sizeof(unsigned) * (size_t)nblock * 32, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(c->d_goss_amp, amplify_bits, sizeof(unsigned) * (size_t)nblock * 32, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(c->d_goss_mult, chunk_multiply, sizeof(float) * (size_t)nchunks, cudaMemcpyHostToDevice, c->stream));
  const int grid = grid_for(c, (long long)nblock * 32, 256);
  LAUNCH(k_goss_apply, std::min(nblock, c->num_sms * 2), BAG_BLOCK, 0, c->stream, c->N, c->d_bag_bits, c->d_goss_amp, (int)chunk_rows,
         c->d_goss_mult, c->d_grad, c->d_hess, c->d_bag_tile);
  LAUNCH(k_na_scan, 1, 1024, 0, c->stream, c->d_bag_tile, c->N, c->d_bag_ctl, reinterpret_cast<int*>(c->d_bag_ctl + 1));
  LAUNCH(k_bag_scatter, grid, 256, 0, c->stream, c->N, c->d_bag_bits, c->d_bag_tile, c->d_bag_ctl, c->d_bag);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(c->h_ctl, c->d_bag_ctl, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  c->bag_count = (int)*reinterpret_cast<unsigned*>(c->h_ctl);
  c->bag_ready = true;
  *bag_count = c->bag_count;
  return 0;
}

int lgbcu_bagging_reset(LgbCuContext* c, int32_t seed) {
  NEED(c && c->N > 0, "lgbcu_bagging_reset: dataset not set");
  USE_DEVICE(c->device);
  const int nblock = (c->N + BAG_BLOCK - 1) / BAG_BLOCK;
  CU(dev_alloc(&c->d_bag_state, (size_t)nblock));
  CU(dev_alloc(&c->d_bag_bits, (size_t)nblock * 32));
  CU(dev_alloc(&c->d_bag_tile, (size_t)nblock));
  CU(dev_alloc(&c->d_bag_ctl, (size_t)2));
  CU(dev_alloc(&c->d_bag, (size_t)c->N));
  LAUNCH(k_bag_seed, grid_for(c, nblock, 256), 256, 0, c->stream, nblock, (int)seed, c->d_bag_state);
  CU(cudaGetLastError());
  c->bag_count = c->N; c->bag_ready = false;
  return 0;
}

static int bagging_draw(LgbCuContext* c, double fraction, double fraction_neg, const float* label, int32_t* bag_count);

int lgbcu_bagging(LgbCuContext* c, double fraction, int32_t* bag_count) {
  NEED(c && bag_count, "lgbcu_bagging: NULL argument");
  return bagging_draw(c, fraction, fraction, nullptr, bag_count);
}

int lgbcu_bagging_balanced(LgbCuContext* c, double pos_fraction, double neg_fraction, int32_t* bag_count) {
  NEED(c && bag_count, "lgbcu_bagging_balanced: NULL argument");
  NEED(c->d_label, "lgbcu_bagging_balanced: labels not set");
  return bagging_draw(c, pos_fraction, neg_fraction, c->d_label, bag_count);
}

static int bagging_draw(LgbCuContext* c, double fraction, double fraction_neg, const float* label, int32_t* bag_count) {
  NEED(c->d_bag_state, "lgbcu_bagging: call lgbcu_bagging_reset first");
  USE_DEVICE(c->device);
  BagJump J;
  {  // affine maps of the LCG for 32 l steps, l = 0..31
    unsigned a32 = 1, c32 = 0;
    for (int i = 0; i < 32; ++i) { c32 = 214013u * c32 + 2531011u; a32 = 214013u * a32; }
    unsigned a = 1, cc = 0;
    for (int l = 0; l < 32; ++l) { J.a[l] = a; J.c[l] = cc; cc = a32 * cc + c32; a = a32 * a; }
  }
  const int nblock = (c->N + BAG_BLOCK - 1) / BAG_BLOCK;
  const int grid = grid_for(c, (long long)nblock * 32, 256);
  LAUNCH(k_bag_flags, grid, 256, 0, c->stream, c->N, fraction, fraction_neg, label, J, c->d_bag_state, c->d_bag_bits, c->d_bag_tile);
  LAUNCH(k_na_scan, 1, 1024, 0, c->stream, c->d_bag_tile, c->N, c->d_bag_ctl, reinterpret_cast<int*>(c->d_bag_ctl + 1));
  LAUNCH(k_bag_scatter, grid, 256, 0, c->stream, c->N, c->d_bag_bits, c->d_bag_tile, c->d_bag_ctl, c->d_bag);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(c->h_ctl, c->d_bag_ctl, sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  c->bag_count = (int)*reinterpret_cast<unsigned*>(c->h_ctl);
  c->bag_ready = true;
  *bag_count = c->bag_count;
  return 0;
}

int lgbcu_copy_bag(LgbCuContext* c, int32_t* out, int32_t n) {
  NEED(c && out && n == c->N && c->bag_ready, "lgbcu_copy_bag: bad argument / no bag drawn");
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(out, c->d_bag, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

}  // extern "C"
