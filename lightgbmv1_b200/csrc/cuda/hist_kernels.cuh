// Histogram construction (replaces Dataset::ConstructHistograms, src/io/dataset.cpp:1282-1383 and
// DenseBin::ConstructHistogramInner, src/io/dense_bin.hpp:97-140).
//
// Arithmetic: gradients/hessians (float) are converted once per row to 64-bit fixed point with a per-tree
// power-of-two scale (DevCtl::scale_g/scale_h) and accumulated with integer adds.  Integer accumulation is
// exact and order independent, so the result is bit-reproducible for any grid shape, any number of GPUs and
// any atomic ordering; the only rounding is the per-row quantisation (<= 2^-38 of max|g| for 10^7 rows).
//
// Two code paths, chosen ON DEVICE from the leaf size (no host round trip):
//   global atomics : one thread per (row, 16-feature chunk), REDG.ADD.64 straight into the pool slot.  Small
//                    leaves (< HIST_SMALL_LEAF_ROWS rows), where zeroing / flushing a 64 KB sub-histogram per
//                    warp would cost more than the rows themselves.
//   warp private   : warps 0..2 of a block each own a private 64 KB shared-memory sub-histogram for one tile
//                    of 16 features x {g,h}; lane l = (channel, feature) so the 32 lanes of a warp always hit
//                    32 different columns -> plain LDS / IADD / STS, no atomics, and with the layout
//                    [bin][channel][16 features] (256 B per bin) lane l always sits on bank pair l: no bank
//                    conflicts for any bin pattern.  Four rows are kept in flight per warp (duplicate bins are
//                    folded forward).  The three sub-histograms of a block are summed in shared memory and
//                    flushed once with REDG.ADD.64.
#pragma once
#include "dev_types.cuh"

#define HIST_THREADS 256                          /* block size of every kernel that contains a histogram phase */
#define HIST_WARPS_PER_BLOCK 3                    /* warps that accumulate (the private sub-histograms fill smem) */
#define HIST_ROWS_SMEM 256                        /* bin rows per private sub-histogram */
#define HIST_SUB_BYTES (HIST_ROWS_SMEM * 32 * 8)  /* 64 KB */
#define HIST_STAGE_BYTES 1024                     /* per warp: 32 x 16 B bins + 32 x 16 B (gq,hq) */
#define HIST_SMEM_BYTES (HIST_WARPS_PER_BLOCK * (HIST_SUB_BYTES + HIST_STAGE_BYTES))
#define HIST_SMALL_LEAF_ROWS 4096
#define HIST_MIN_ROWS_PER_WARP 128                /* rows a warp should at least get before another block is recruited */

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

__device__ void hist_global_path(const DevView& V, const DevLeaf& L, bool identity, int bid, int nblocks) {
  const DevCtl* c = V.ctl;
  const double sg = c->scale_g, sh = c->scale_h;
  const int* idx = V.indices[L.buf] + L.begin;
  HistEntry* out = V.pool + (size_t)L.slot * V.total_bin;
  const int nchunk = V.num_tiles16;
  const long long total = (long long)L.count * nchunk;
  for (long long p = bid * (long long)blockDim.x + threadIdx.x; p < total; p += (long long)nblocks * blockDim.x) {
    const int pos = (int)(p / nchunk);
    const int ck = (int)(p - (long long)pos * nchunk);
    const int row = identity ? pos : idx[pos];
    const uint4 q = *reinterpret_cast<const uint4*>(V.bins + (size_t)row * V.row_stride + ck * 16);
    const i64 gq = to_fixed(V.grad[row], sg), hq = to_fixed(V.hess[row], sh);
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
}

// smem: HIST_SMEM_BYTES of dynamic shared memory.  Called by all HIST_THREADS threads of every block.
__device__ void hist_private_path(const DevView& V, const DevLeaf& L, bool identity, int bid, int nblocks, int min_rows,
                                  unsigned char* smem) {
  const DevCtl* c = V.ctl;
  const double sg = c->scale_g, sh = c->scale_h;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int fl = lane & 15, ch = lane >> 4;
  const bool worker = warp < HIST_WARPS_PER_BLOCK;
  const int wslot = worker ? warp : 0;
  i64* H = reinterpret_cast<i64*>(smem + (size_t)wslot * HIST_SUB_BYTES);
  i64* Hl = H + lane;  // this lane's column
  const i64* H0 = reinterpret_cast<const i64*>(smem);
  const i64* H1 = reinterpret_cast<const i64*>(smem + (size_t)HIST_SUB_BYTES);
  const i64* H2 = reinterpret_cast<const i64*>(smem + (size_t)2 * HIST_SUB_BYTES);
  unsigned char* stage = smem + (size_t)HIST_WARPS_PER_BLOCK * HIST_SUB_BYTES + wslot * HIST_STAGE_BYTES;
  uint4* sbins = reinterpret_cast<uint4*>(stage);                // [32] one 16-byte chunk per staged row
  longlong2* svals = reinterpret_cast<longlong2*>(stage + 512);  // [32] (gq, hq) per staged row
  const unsigned char* sb8 = stage;
  const i64* sv64 = reinterpret_cast<const i64*>(stage + 512);

  const int NT = V.num_tiles16;
  // block-level work items: (tile, part); a block's rows are split three ways over its accumulating warps
  int PB = (L.count + HIST_WARPS_PER_BLOCK * min_rows - 1) / (HIST_WARPS_PER_BLOCK * min_rows);
  int cap = nblocks / NT; if (cap < 1) cap = 1;
  if (PB > cap) PB = cap;
  if (PB < 1) PB = 1;
  int per_block = (L.count + PB - 1) / PB;
  per_block = (per_block + 95) / 96 * 96;  // three multiples of 32
  const int per_warp = per_block / 3;

  const int* idx = V.indices[L.buf] + L.begin;
  HistEntry* out = V.pool + (size_t)L.slot * V.total_bin;

  for (int item = bid; item < NT * PB; item += nblocks) {
    const int tile = item / PB, part = item - tile * PB;
    const int b0 = part * per_block;
    int b1 = b0 + per_block; if (b1 > L.count) b1 = L.count;
    if (worker) {
      // zero the private sub-histogram
      uint4* z = reinterpret_cast<uint4*>(H);
      const uint4 zero = make_uint4(0, 0, 0, 0);
#pragma unroll 4
      for (int i = lane; i < HIST_SUB_BYTES / 16; i += 32) z[i] = zero;
      __syncwarp();
      const int r0 = b0 + warp * per_warp;
      int r1 = r0 + per_warp; if (r1 > b1) r1 = b1;
      const uint8_t* colbase = V.bins + tile * 16;
      // software pipeline: the (index ->) row loads of batch k+1 are in flight while batch k is accumulated
      uint4 nq = make_uint4(0, 0, 0, 0);
      float ng = 0.f, nh = 0.f;
      if (r0 < r1) {
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
          // 4 rows in flight: their 4 read-modify-writes are independent unless two rows of this lane's feature
          // share a bin; then the later row's (stale) loaded value is patched with the earlier rows' addends.
          // Stores stay in row order, so the result is the sequential one.
#pragma unroll
          for (int j0 = 0; j0 < 32; j0 += 4) {
            const unsigned c0 = sb8[(j0 + 0) * 16 + fl], c1 = sb8[(j0 + 1) * 16 + fl];
            const unsigned c2 = sb8[(j0 + 2) * 16 + fl], c3 = sb8[(j0 + 3) * 16 + fl];
            const i64 v0 = sv64[(j0 + 0) * 2 + ch], v1 = sv64[(j0 + 1) * 2 + ch];
            const i64 v2 = sv64[(j0 + 2) * 2 + ch], v3 = sv64[(j0 + 3) * 2 + ch];
            i64* p0 = Hl + c0 * 32; i64* p1 = Hl + c1 * 32;
            i64* p2 = Hl + c2 * 32; i64* p3 = Hl + c3 * 32;
            i64 a0 = *p0, a1 = *p1, a2 = *p2, a3 = *p3;   // all four loads before any store
            // a later row that shares a bin with an earlier one loaded a stale value: add what the earlier rows add
            if (c1 == c0) a1 += v0;
            if (c2 == c0) a2 += v0;
            if (c2 == c1) a2 += v1;
            if (c3 == c0) a3 += v0;
            if (c3 == c1) a3 += v1;
            if (c3 == c2) a3 += v2;
            *p0 = a0 + v0; *p1 = a1 + v1; *p2 = a2 + v2; *p3 = a3 + v3;  // row order: the last store of a bin wins
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
    }
    __syncthreads();
    // merge the three sub-histograms and flush once: thread tid owns column (tid & 31) of bin rows tid>>5, +8, ...
    {
      const int f = tile * 16 + fl;
      if (f < V.num_features) {
        const DevFeat F = V.feats[f];
        i64* dst = reinterpret_cast<i64*>(out + F.hist_base) + ch;
        for (int b = warp; b < F.nslots; b += HIST_THREADS / 32) {
          const int e = b * 32 + lane;
          const i64 v = H0[e] + H1[e] + H2[e];
          if (v) atomic_add_i64(dst + 2 * b, v);
        }
      }
    }
    __syncthreads();
  }
}

// mode: LGBCU_HIST_AUTO (by leaf size), LGBCU_HIST_GLOBAL_ATOMIC, LGBCU_HIST_WARP_PRIVATE
__device__ void hist_phase(const DevView& V, int leaf, int mode, int min_rows, int bid, int nblocks, unsigned char* smem) {
  const DevLeaf L = V.leaves[leaf];
  const DevCtl* c = V.ctl;
  const bool identity = c->root_identity && leaf == 0 && L.count == V.num_data;
  const bool small = mode == LGBCU_HIST_GLOBAL_ATOMIC || (mode == LGBCU_HIST_AUTO && L.count < HIST_SMALL_LEAF_ROWS);
  if (small) hist_global_path(V, L, identity, bid, nblocks);
  else hist_private_path(V, L, identity, bid, nblocks, min_rows, smem);
  if (bid == 0 && threadIdx.x == 0) {
    atomicAdd(&V.ctl->prof_hist_calls, 1ull);
    atomicAdd(identity ? &V.ctl->prof_hist_rows : &V.ctl->prof_hist_rows_indexed, (unsigned long long)L.count);
  }
}

__global__ void __launch_bounds__(HIST_THREADS, 1) k_hist(DevView V, int leaf_arg, int mode, int min_rows) {
  extern __shared__ __align__(16) unsigned char hist_smem[];
  int leaf;
  STAMP(V.ctl, 0);
  if (!hist_target(V, leaf_arg, &leaf)) return;
#ifdef LGBCU_STAGE_TIMERS
  const int split_ = V.ctl->num_splits;
#endif
  hist_phase(V, leaf, mode, min_rows, blockIdx.x, gridDim.x, hist_smem);
  STAMP_END(split_, 0);
}
