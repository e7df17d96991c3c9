// Histogram construction (replaces Dataset::ConstructHistograms, src/io/dataset.cpp:1282-1383 and
// DenseBin::ConstructHistogramInner, src/io/dense_bin.hpp:97-140).
//
// Arithmetic: gradients/hessians (float) are converted once per row to 64-bit fixed point with a per-tree
// power-of-two scale (DevCtl::scale_g/scale_h) and accumulated with integer adds.  Integer accumulation is
// exact and order independent, so the result is bit-reproducible for any grid shape, any number of GPUs and
// any atomic ordering; the only rounding is the per-row quantisation (<= 2^-38 of max|g| for 10^7 rows).
//
// Two kernels:
//   k_hist_global       : one thread per (row, 16-feature chunk), REDG.ADD.64 straight into the pool slot.
//                         Launch-latency friendly, used for small leaves.
//   k_hist_warp_private : each warp owns a private 64 KB shared-memory sub-histogram for one tile of 16
//                         features x {g,h}; lane l = (channel, feature) so the 32 lanes of a warp always
//                         hit 32 different columns -> plain LDS/IADD/STS, no atomics, no bank conflicts
//                         (row stride 256 B keeps lane l on bank pair l).  Flushed with REDG.ADD.64.
#pragma once
#include "dev_types.cuh"

#define HIST_WARPS_PER_BLOCK 3
#define HIST_ROWS_SMEM 256                      /* bin rows per private sub-histogram */
#define HIST_SUB_BYTES (HIST_ROWS_SMEM * 32 * 8) /* 64 KB */
#define HIST_STAGE_BYTES 1024                    /* per warp: 32 x 16 B bins + 32 x 16 B (gq,hq) */
#define HIST_SMEM_BYTES (HIST_WARPS_PER_BLOCK * (HIST_SUB_BYTES + HIST_STAGE_BYTES))

__device__ __forceinline__ i64 to_fixed(float v, double scale) {
  return __double2ll_rn((double)v * scale);
}

__device__ __forceinline__ void atomic_add_i64(i64* addr, i64 v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)v);
}

// Resolve which leaf a kernel works on: explicit (step-wise API) or the smaller leaf chosen on device.
__device__ __forceinline__ bool hist_target(const DevView& V, int leaf_arg, int* leaf) {
  if (leaf_arg >= 0) { *leaf = leaf_arg; return true; }
  const DevCtl* c = V.ctl;
  if (c->done || c->skip_find) return false;
  *leaf = c->smaller;
  return *leaf >= 0;
}

// AUTO mode: both kernels are enqueued every split with a `regime`; each looks at the leaf size on device and
// only the one whose regime it is does the work (the other exits after reading a few words), so no host
// round trip is needed to choose.  regime 0 = always run, 1 = only small leaves, 2 = only large leaves.
#define HIST_SMALL_LEAF_ROWS 4096
__device__ __forceinline__ bool hist_regime_ok(int regime, int count) {
  if (regime == 1) return count < HIST_SMALL_LEAF_ROWS;
  if (regime == 2) return count >= HIST_SMALL_LEAF_ROWS;
  return true;
}

__global__ void __launch_bounds__(256) k_hist_global(DevView V, int leaf_arg, int regime) {
  int leaf;
  if (!hist_target(V, leaf_arg, &leaf)) return;
  const DevLeaf L = V.leaves[leaf];
  if (!hist_regime_ok(regime, L.count)) return;
  const DevCtl* c = V.ctl;
  const bool identity = c->root_identity && leaf == 0 && L.count == V.num_data;
  const double sg = c->scale_g, sh = c->scale_h;
  const int* idx = V.indices[L.buf] + L.begin;
  HistEntry* out = V.pool + (size_t)L.slot * V.total_bin;
  const int nchunk = V.num_tiles16;
  const long long total = (long long)L.count * nchunk;
  i64 tg = 0, th = 0;
  for (long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x; p < total;
       p += (long long)gridDim.x * blockDim.x) {
    const int pos = (int)(p / nchunk);
    const int ck = (int)(p - (long long)pos * nchunk);
    const int row = identity ? pos : idx[pos];
    const uint4 q = *reinterpret_cast<const uint4*>(V.bins + (size_t)row * V.row_stride + ck * 16);
    const i64 gq = to_fixed(V.grad[row], sg), hq = to_fixed(V.hess[row], sh);
    if (ck == 0) { tg += gq; th += hq; }
    const unsigned w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const int f = ck * 16 + k;
      if (f < V.num_features) {
        const unsigned b = (w[k >> 2] >> ((k & 3) * 8)) & 0xffu;
        HistEntry* e = out + V.feats[f].hist_base + b;
        atomic_add_i64(&e->g, gq);
        atomic_add_i64(&e->h, hq);
      }
    }
  }
  // leaf totals (LeafSplits::Init for the root, src/treelearner/leaf_splits.hpp:66-79)
  for (int o = 16; o > 0; o >>= 1) {
    tg += __shfl_down_sync(0xffffffffu, tg, o);
    th += __shfl_down_sync(0xffffffffu, th, o);
  }
  if ((threadIdx.x & 31) == 0 && (tg | th)) {
    atomic_add_i64(&V.leaves[leaf].tot_g, tg);
    atomic_add_i64(&V.leaves[leaf].tot_h, th);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&V.ctl->prof_hist_calls, 1ull);
    atomicAdd(identity ? &V.ctl->prof_hist_rows : &V.ctl->prof_hist_rows_indexed, (unsigned long long)L.count);
  }
}

// rows each (tile, part) work item should at least have before another warp is recruited for the tile
#define HIST_MIN_ROWS_PER_WARP 128

__global__ void __launch_bounds__(HIST_WARPS_PER_BLOCK * 32, 1) k_hist_warp_private(DevView V, int leaf_arg, int regime, int min_rows) {
  extern __shared__ __align__(16) unsigned char smem[];
  int leaf;
  if (!hist_target(V, leaf_arg, &leaf)) return;
  const DevLeaf L = V.leaves[leaf];
  if (!hist_regime_ok(regime, L.count)) return;
  const DevCtl* c = V.ctl;
  const bool identity = c->root_identity && leaf == 0 && L.count == V.num_data;
  const double sg = c->scale_g, sh = c->scale_h;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int fl = lane & 15, ch = lane >> 4;

  i64* H = reinterpret_cast<i64*>(smem + (size_t)warp * HIST_SUB_BYTES);
  unsigned char* stage = smem + (size_t)HIST_WARPS_PER_BLOCK * HIST_SUB_BYTES + warp * HIST_STAGE_BYTES;
  uint4* sbins = reinterpret_cast<uint4*>(stage);              // [32] one 16-byte chunk per staged row
  longlong2* svals = reinterpret_cast<longlong2*>(stage + 512);  // [32] (gq, hq) per staged row
  const unsigned char* sb8 = stage;
  const i64* sv64 = reinterpret_cast<const i64*>(stage + 512);

  const int NW = gridDim.x * HIST_WARPS_PER_BLOCK;
  const int gw = blockIdx.x * HIST_WARPS_PER_BLOCK + warp;
  const int NT = V.num_tiles16;
  int P = (L.count + min_rows - 1) / min_rows;
  int cap = NW / NT; if (cap < 1) cap = 1;
  if (P > cap) P = cap;
  if (P < 1) P = 1;
  int per = (L.count + P - 1) / P;
  per = (per + 31) & ~31;

  const int* idx = V.indices[L.buf] + L.begin;
  HistEntry* out = V.pool + (size_t)L.slot * V.total_bin;
  i64 tg = 0, th = 0;

  for (int item = gw; item < NT * P; item += NW) {
    const int tile = item / P, part = item - tile * P;
    const int r0 = part * per;
    int r1 = r0 + per; if (r1 > L.count) r1 = L.count;
    if (r0 >= r1) continue;
    // zero the private sub-histogram
    {
      uint4* z = reinterpret_cast<uint4*>(H);
      const uint4 zero = make_uint4(0, 0, 0, 0);
#pragma unroll 4
      for (int i = lane; i < HIST_SUB_BYTES / 16; i += 32) z[i] = zero;
    }
    __syncwarp();
    const uint8_t* colbase = V.bins + tile * 16;
    // software pipeline: the (index ->) row loads of batch k+1 are in flight while batch k is accumulated
    uint4 nq = make_uint4(0, 0, 0, 0);
    float ng = 0.f, nh = 0.f;
    {
      const int pos = r0 + lane;
      if (pos < r1) {
        const int row = identity ? pos : idx[pos];
        nq = *reinterpret_cast<const uint4*>(colbase + (size_t)row * V.row_stride);
        ng = V.grad[row]; nh = V.hess[row];
      }
    }
    for (int base = r0; base < r1; base += 32) {
      const uint4 q = nq;
      const i64 gq = to_fixed(ng, sg), hq = to_fixed(nh, sh);
      if (tile == 0) { tg += gq; th += hq; }
      nq = make_uint4(0, 0, 0, 0); ng = 0.f; nh = 0.f;
      {
        const int pos = base + 32 + lane;
        if (pos < r1) {
          const int row = identity ? pos : idx[pos];
          nq = *reinterpret_cast<const uint4*>(colbase + (size_t)row * V.row_stride);
          ng = V.grad[row]; nh = V.hess[row];
        }
      }
      sbins[lane] = q;
      svals[lane] = make_longlong2(gq, hq);
      __syncwarp();
      int n = r1 - base; if (n > 32) n = 32;
      if (n == 32) {
        // 4 rows in flight: their 4 read-modify-writes are independent unless two rows of this lane's feature share
        // a bin.  Duplicates are folded forward (the last row of a bin carries the sum, earlier ones add 0), all 4
        // loads are issued before the 4 stores, and stores stay in row order, so the result is the sequential one.
#pragma unroll
        for (int j0 = 0; j0 < 32; j0 += 4) {
          const unsigned b0 = sb8[(j0 + 0) * 16 + fl], b1 = sb8[(j0 + 1) * 16 + fl];
          const unsigned b2 = sb8[(j0 + 2) * 16 + fl], b3 = sb8[(j0 + 3) * 16 + fl];
          i64 v0 = sv64[(j0 + 0) * 2 + ch], v1 = sv64[(j0 + 1) * 2 + ch];
          i64 v2 = sv64[(j0 + 2) * 2 + ch], v3 = sv64[(j0 + 3) * 2 + ch];
          if (b1 == b0) { v1 += v0; v0 = 0; }
          if (b2 == b0) { v2 += v0; v0 = 0; }
          if (b2 == b1) { v2 += v1; v1 = 0; }
          if (b3 == b0) { v3 += v0; v0 = 0; }
          if (b3 == b1) { v3 += v1; v1 = 0; }
          if (b3 == b2) { v3 += v2; v2 = 0; }
          i64* p0 = H + (b0 * 32 + lane); i64* p1 = H + (b1 * 32 + lane);
          i64* p2 = H + (b2 * 32 + lane); i64* p3 = H + (b3 * 32 + lane);
          const i64 a0 = *p0, a1 = *p1, a2 = *p2, a3 = *p3;
          *p0 = a0 + v0; *p1 = a1 + v1; *p2 = a2 + v2; *p3 = a3 + v3;
        }
      } else {
        for (int j = 0; j < n; ++j) {
          const unsigned b = sb8[j * 16 + fl];
          i64* p = H + (b * 32 + lane);
          *p += sv64[j * 2 + ch];
        }
      }
      __syncwarp();
    }
    // flush: lane owns column (ch, feature tile*16+fl)
    const int f = tile * 16 + fl;
    if (f < V.num_features) {
      const DevFeat F = V.feats[f];
      i64* dst = reinterpret_cast<i64*>(out + F.hist_base) + ch;
      for (int b = 0; b < F.nslots; ++b) {
        const i64 v = H[b * 32 + lane];
        if (v) atomic_add_i64(dst + 2 * b, v);
      }
    }
    __syncwarp();
  }
  for (int o = 16; o > 0; o >>= 1) {
    tg += __shfl_down_sync(0xffffffffu, tg, o);
    th += __shfl_down_sync(0xffffffffu, th, o);
  }
  if (lane == 0 && (tg | th)) {
    atomic_add_i64(&V.leaves[leaf].tot_g, tg);
    atomic_add_i64(&V.leaves[leaf].tot_h, th);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&V.ctl->prof_hist_calls, 1ull);
    atomicAdd(identity ? &V.ctl->prof_hist_rows : &V.ctl->prof_hist_rows_indexed, (unsigned long long)L.count);
  }
}

