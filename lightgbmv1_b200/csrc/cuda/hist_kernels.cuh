// Histogram construction (replaces Dataset::ConstructHistograms, src/io/dataset.cpp:1282-1383 and
// DenseBin::ConstructHistogramInner, src/io/dense_bin.hpp:97-140).
//
// Arithmetic: gradients/hessians (float) are converted once per row to 64-bit fixed point with a per-tree
// power-of-two scale (DevCtl::scale_g/scale_h) and accumulated with integer adds.  Integer accumulation is
// exact and order independent, so the result is bit-reproducible for any grid shape, any number of GPUs and
// any atomic ordering; the only rounding is the per-row quantisation (<= 2^-38 of max|g| for 10^7 rows).
//
// Two code paths, chosen ON DEVICE from the leaf size (no host round trip):
//   global atomics : one thread per (row, 16-feature chunk), REDG.ADD.64 straight into the pool slot.  Tiny
//                    leaves (< HIST_SMALL_LEAF_ROWS rows), where zeroing / flushing a 64 KB sub-histogram would
//                    cost more than the rows themselves.
//   block shared   : a block (32 warps, one per SM) owns one 64 KB shared-memory sub-histogram per tile of 16 features
//                    x {g,h} -- two tiles at once, so that one 32-byte row read, one (g,h) read and one staging pass
//                    serve 32 features -- and a slice of the leaf's rows.  Lane l of every warp = column (channel,
//                    feature), so the 32 lanes of an instruction always hit 32 different columns, and with the
//                    layout [bin][32 columns] of 32-bit words lane l always sits on bank l: no bank conflicts for
//                    any bin pattern.  sm_100 has native shared-memory atomics only for 32-bit integers, so a 64-bit
//                    entry is kept as two words: ATOMS.ADD on the low word returns the old value, the carry
//                    (old + v wrapped) is added together with the high word by a second ATOMS.ADD.  Each wrap is seen
//                    by exactly the add that caused it and addition commutes, so the pair (hi, lo) read after the
//                    block barrier is the exact 64-bit sum whatever the interleaving.  Measured in isolation
//                    (tools/ubench/atoms.cu): 3.3 cycles per (row, tile) per SM with 32 resident warps against 12.3
//                    for the round's first design (a private LDS/IADD/STS sub-histogram per warp; 3 warps fit).
//                    The blocks of a cluster (2) work on row slices of the same tile pair, sum their sub-histograms
//                    through distributed shared memory and flush once with REDG.ADD.64 (zero entries skipped).
#pragma once
#include <cooperative_groups.h>
#include "dev_types.cuh"

#define HIST_CLUSTER 2                            /* blocks per thread-block cluster: they reduce their sub-histograms through DSMEM
                                                     (2 x 1024 threads x 176 KB: 74 clusters = all 148 SMs; 4 leaves 16 SMs idle) */
#define HIST_THREADS 1024                         /* block size of the histogram kernel: 32 warps */
#define HIST_WARPS (HIST_THREADS / 32)
#define HIST_BLOCKS_PER_SM 1                      /* 176 KB of shared memory */
#define HIST_TILES 2                              /* 16-feature tiles a block accumulates at once */
#define HIST_ROWS_SMEM 256                        /* bin rows of the sub-histogram */
#define HIST_SUB_BYTES (HIST_ROWS_SMEM * 32 * 8)  /* 64 KB: lo words [256][32], then hi words [256][32] */
#define HIST_STAGE_BYTES 1536                     /* per warp: 32 x 32 B bins (two tiles, interleaved) + 2 x 32 x 8 B (gq | hq) */
#define HIST_SMEM_BYTES (HIST_TILES * HIST_SUB_BYTES + HIST_WARPS * HIST_STAGE_BYTES)
#define HIST_SMALL_LEAF_ROWS 256
#define HIST_MIN_ROWS_PER_WARP 16                 /* rows a warp should at least get before another block is recruited */

#ifdef LGBCU_STAGE_TIMERS
__device__ unsigned long long g_hist_cycles[8];
#define HSTAGE(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) { long long t_ = clock64(); g_hist_cycles[i] += (unsigned long long)(t_ - ht_); ht_ = t_; } } while (0)
#else
#define HSTAGE(i) do {} while (0)
#endif

__device__ __forceinline__ i64 to_fixed(float v, double scale) {
  return __double2ll_rn((double)v * scale);
}

__device__ __forceinline__ void atomic_add_i64(i64* addr, i64 v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)v);
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
    const float2 gh = V.gh[row];
    const i64 gq = to_fixed(gh.x, sg), hq = to_fixed(gh.y, sh);
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
__device__ void hist_shared_path(const DevView& V, const DevLeaf& L, bool identity, int bid, int nblocks, int min_rows,
                                 unsigned char* smem) {
  const DevCtl* c = V.ctl;
  const double sg = c->scale_g, sh = c->scale_h;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int fl = lane & 15, ch = lane >> 4;
  unsigned* lo = reinterpret_cast<unsigned*>(smem);   // tile A: lo [256][32], hi [256][32]; tile B follows
  unsigned* lo_l = lo + lane;                          // this lane's column
  unsigned char* stage = smem + HIST_TILES * HIST_SUB_BYTES + warp * HIST_STAGE_BYTES;
  uint4* sbins = reinterpret_cast<uint4*>(stage);      // [32 rows][32 B]: byte 2f = tile A feature f, 2f+1 = tile B
  i64* svals = reinterpret_cast<i64*>(stage + 1024);   // [2][32]: gq of the 32 staged rows, then hq
  const unsigned short* sb16 = reinterpret_cast<const unsigned short*>(stage) + fl;
  const longlong2* sv2 = reinterpret_cast<const longlong2*>(stage + 1024 + ch * 256);

  const int NT = V.num_tiles16;
  const int NP = (NT + HIST_TILES - 1) / HIST_TILES;   // tile pairs
  // Work units: (tile pair, group of HIST_CLUSTER row slices).  A block accumulates BOTH 16-feature tiles of a pair:
  // one 32-byte row read, one (g,h) read and one staging pass serve 32 features.
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int nclusters = nblocks / HIST_CLUSTER, cluster_id = bid / HIST_CLUSTER;
  int PG = (L.count + HIST_CLUSTER * HIST_WARPS * min_rows - 1) / (HIST_CLUSTER * HIST_WARPS * min_rows);
  int cap = nclusters / NP; if (cap < 1) cap = 1;
  if (PG > cap) PG = cap;
  if (PG < 1) PG = 1;
  int per_block = (L.count + PG * HIST_CLUSTER - 1) / (PG * HIST_CLUSTER);
  per_block = (per_block + 31) / 32 * 32;
  const int* idx = V.indices[L.buf] + L.begin;
  HistEntry* out = V.pool + (size_t)L.slot * V.total_bin;
#ifdef LGBCU_STAGE_TIMERS
  long long ht_ = clock64();
#endif
  for (int unit = cluster_id; unit < NP * PG; unit += nclusters) {
    const int group = unit / NP, pair = unit - group * NP, part = group * HIST_CLUSTER + rank;
    const int tile0 = pair * HIST_TILES;
    const bool two = tile0 + 1 < NT;   // the last pair of an odd tile count has one tile only
    int b0 = part * per_block; if (b0 > L.count) b0 = L.count;
    int b1 = b0 + per_block; if (b1 > L.count) b1 = L.count;
    const uint8_t* colbase = V.bins + tile0 * 16;
    const int step = HIST_WARPS * 32;
    uint4 nqa = make_uint4(0, 0, 0, 0), nqb = make_uint4(0, 0, 0, 0);
    float ng = 0.f, nh = 0.f;
    int prow = -1;
    int base = b0 + warp * 32;
    {
      const int pos = base + lane;
      if (pos < b1) {
        const int row = identity ? pos : idx[pos];
        const uint4* rp = reinterpret_cast<const uint4*>(colbase + (size_t)row * V.row_stride);
        nqa = rp[0];
        if (two) nqb = rp[1];
        { const float2 gh = V.gh[row]; ng = gh.x; nh = gh.y; }
      }
      const int pos1 = pos + step;
      if (pos1 < b1) prow = identity ? pos1 : idx[pos1];
    }
    {
      uint4* z = reinterpret_cast<uint4*>(smem);
      const uint4 zero = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int i = tid; i < HIST_TILES * HIST_SUB_BYTES / 16; i += HIST_THREADS) z[i] = zero;
    }
    __syncthreads();
    HSTAGE(0);
    for (; base < b1; base += step) {
      {
        // byte-interleave the two tiles of this lane's row: (a0 b0 a1 b1) (a2 b2 a3 b3) per pair of words
        uint4 w0, w1;
        w0.x = __byte_perm(nqa.x, nqb.x, 0x5140); w0.y = __byte_perm(nqa.x, nqb.x, 0x7362);
        w0.z = __byte_perm(nqa.y, nqb.y, 0x5140); w0.w = __byte_perm(nqa.y, nqb.y, 0x7362);
        w1.x = __byte_perm(nqa.z, nqb.z, 0x5140); w1.y = __byte_perm(nqa.z, nqb.z, 0x7362);
        w1.z = __byte_perm(nqa.w, nqb.w, 0x5140); w1.w = __byte_perm(nqa.w, nqb.w, 0x7362);
        sbins[2 * lane] = w0; sbins[2 * lane + 1] = w1;
      }
      svals[lane] = to_fixed(ng, sg);
      svals[32 + lane] = to_fixed(nh, sh);
      nqa = make_uint4(0, 0, 0, 0); nqb = make_uint4(0, 0, 0, 0); ng = 0.f; nh = 0.f;
      if (prow >= 0) {
        const uint4* rp = reinterpret_cast<const uint4*>(colbase + (size_t)prow * V.row_stride);
        nqa = rp[0];
        if (two) nqb = rp[1];
        { const float2 gh = V.gh[prow]; ng = gh.x; nh = gh.y; }
      }
      {
        const int pos2 = base + 2 * step + lane;
        prow = -1;
        if (pos2 < b1) prow = identity ? pos2 : idx[pos2];
      }
      __syncwarp();
      int n = b1 - base; if (n > 32) n = 32;
      if (n == 32 && two) {
#pragma unroll 8
        for (int j = 0; j < 32; j += 2) {
          const unsigned c0 = sb16[j * 16], c1 = sb16[(j + 1) * 16];   // (tile A bin) | (tile B bin) << 8
          const longlong2 v = sv2[j >> 1];
          const unsigned v0l = (unsigned)v.x, v0h = (unsigned)((unsigned long long)v.x >> 32);
          const unsigned v1l = (unsigned)v.y, v1h = (unsigned)((unsigned long long)v.y >> 32);
          unsigned* p0a = lo_l + (c0 & 0xffu) * 32; unsigned* p0b = lo_l + HIST_SUB_BYTES / 4 + (c0 >> 8) * 32;
          unsigned* p1a = lo_l + (c1 & 0xffu) * 32; unsigned* p1b = lo_l + HIST_SUB_BYTES / 4 + (c1 >> 8) * 32;
          const unsigned o0a = atomicAdd(p0a, v0l);
          const unsigned o0b = atomicAdd(p0b, v0l);
          const unsigned o1a = atomicAdd(p1a, v1l);
          const unsigned o1b = atomicAdd(p1b, v1l);
          atomicAdd(p0a + HIST_ROWS_SMEM * 32, v0h + ((o0a + v0l) < o0a ? 1u : 0u));
          atomicAdd(p0b + HIST_ROWS_SMEM * 32, v0h + ((o0b + v0l) < o0b ? 1u : 0u));
          atomicAdd(p1a + HIST_ROWS_SMEM * 32, v1h + ((o1a + v1l) < o1a ? 1u : 0u));
          atomicAdd(p1b + HIST_ROWS_SMEM * 32, v1h + ((o1b + v1l) < o1b ? 1u : 0u));
        }
      } else if (n == 32) {  // single tile (F <= 16, or the last pair of an odd tile count)
#pragma unroll 8
        for (int j = 0; j < 32; j += 2) {
          const unsigned c0 = sb16[j * 16] & 0xffu, c1 = sb16[(j + 1) * 16] & 0xffu;
          const longlong2 v = sv2[j >> 1];
          const unsigned v0l = (unsigned)v.x, v0h = (unsigned)((unsigned long long)v.x >> 32);
          const unsigned v1l = (unsigned)v.y, v1h = (unsigned)((unsigned long long)v.y >> 32);
          unsigned* p0 = lo_l + c0 * 32; unsigned* p1 = lo_l + c1 * 32;
          const unsigned o0 = atomicAdd(p0, v0l);
          const unsigned o1 = atomicAdd(p1, v1l);
          atomicAdd(p0 + HIST_ROWS_SMEM * 32, v0h + ((o0 + v0l) < o0 ? 1u : 0u));
          atomicAdd(p1 + HIST_ROWS_SMEM * 32, v1h + ((o1 + v1l) < o1 ? 1u : 0u));
        }
      } else {
        const i64* sv = reinterpret_cast<const i64*>(stage + 1024 + ch * 256);
        for (int j = 0; j < n; ++j) {
          const unsigned cc = sb16[j * 16];
          const i64 v = sv[j];
          const unsigned vl = (unsigned)v, vh = (unsigned)((unsigned long long)v >> 32);
          unsigned* pa = lo_l + (cc & 0xffu) * 32;
          const unsigned oa = atomicAdd(pa, vl);
          atomicAdd(pa + HIST_ROWS_SMEM * 32, vh + ((oa + vl) < oa ? 1u : 0u));
          if (two) {
            unsigned* pb = lo_l + HIST_SUB_BYTES / 4 + (cc >> 8) * 32;
            const unsigned ob = atomicAdd(pb, vl);
            atomicAdd(pb + HIST_ROWS_SMEM * 32, vh + ((ob + vl) < ob ? 1u : 0u));
          }
        }
      }
      __syncwarp();
    }
    // Reduce across the cluster, then flush: block `rank` owns bin rows [rank*32, rank*32+32) of the tile, sums
    // that slice over the HIST_CLUSTER sub-histograms (its own and the peers', read through DSMEM) and issues one
    // REDG.ADD.64 per non-zero entry -- 1/HIST_CLUSTER of the global atomics a per-block flush would need.
    HSTAGE(1);
    cluster.sync();
    HSTAGE(2);
    {
      const unsigned* rlo[HIST_CLUSTER];
#pragma unroll
      for (int k = 0; k < HIST_CLUSTER; ++k) rlo[k] = cluster.map_shared_rank(lo, k);
      for (int t = 0; t < (two ? 2 : 1); ++t) {
        const int f = (tile0 + t) * 16 + fl;
        int nslots = 0; i64* dst = nullptr;
        if (f < V.num_features) {
          const DevFeat F = V.feats[f];
          nslots = F.nslots;
          dst = reinterpret_cast<i64*>(out + F.hist_base) + ch;
        }
        for (int e0 = rank * (HIST_ROWS_SMEM / HIST_CLUSTER) * 32; e0 < (rank + 1) * (HIST_ROWS_SMEM / HIST_CLUSTER) * 32; e0 += HIST_THREADS) {
          const int e = e0 + tid;  // column = tid & 31
          const int b = e >> 5;
          const int eo = t * (HIST_SUB_BYTES / 4) + e;
          unsigned l[HIST_CLUSTER], h[HIST_CLUSTER];
#pragma unroll
          for (int k = 0; k < HIST_CLUSTER; ++k) { l[k] = rlo[k][eo]; h[k] = rlo[k][eo + HIST_ROWS_SMEM * 32]; }
          i64 v = 0;
#pragma unroll
          for (int k = 0; k < HIST_CLUSTER; ++k) v += (i64)(((unsigned long long)h[k] << 32) | l[k]);
          if (b < nslots && v) atomic_add_i64(dst + 2 * b, v);
        }
      }
    }
    HSTAGE(3);
    cluster.sync();  // peers may still be reading this block's sub-histogram
    HSTAGE(4);
  }
}

// mode: LGBCU_HIST_AUTO (by leaf size), LGBCU_HIST_GLOBAL_ATOMIC, LGBCU_HIST_WARP_PRIVATE (= the block-shared path)
__device__ void hist_phase(const DevView& V, const DevLeaf& L, bool is_root, int mode, int min_rows, int small_rows,
                           int bid, int nblocks, unsigned char* smem) {
  const DevCtl* c = V.ctl;
  const bool identity = c->root_identity && is_root && L.count == V.num_data;
  const bool small = mode == LGBCU_HIST_GLOBAL_ATOMIC || (mode == LGBCU_HIST_AUTO && L.count < small_rows);
  if (small) hist_global_path(V, L, identity, bid, nblocks);
  else hist_shared_path(V, L, identity, bid, nblocks, min_rows, smem);
  if (bid == 0 && threadIdx.x == 0) {
    atomicAdd(&V.ctl->prof_hist_calls, 1ull);
    atomicAdd(identity ? &V.ctl->prof_hist_rows : &V.ctl->prof_hist_rows_indexed, (unsigned long long)L.count);
  }
}

// leaf_arg >= 0: that leaf (step-wise API).  leaf_arg < 0: training -- round 0 builds the root, every later round the
// smaller child of each leaf being expanded (DevSpec); all of a round's leaves share this one launch.
__global__ void __cluster_dims__(HIST_CLUSTER, 1, 1) __launch_bounds__(HIST_THREADS, HIST_BLOCKS_PER_SM) k_hist(DevView V, int leaf_arg, int mode, int min_rows, int small_rows) {
  extern __shared__ __align__(16) unsigned char hist_smem[];
  const DevCtl* c = V.ctl;
  STAMP(V.ctl, 0);
  if (leaf_arg >= 0) {
    hist_phase(V, V.leaves[leaf_arg], leaf_arg == 0, mode, min_rows, small_rows, blockIdx.x, gridDim.x, hist_smem);
    return;
  }
  if (c->done) return;
  if (c->round == 0) {
    if (!c->skip_find) hist_phase(V, V.leaves[0], true, mode, min_rows, small_rows, blockIdx.x, gridDim.x, hist_smem);
    return;
  }
  // The clusters of the grid are dealt to the round's leaves in proportion to their rows (at least one each), so
  // that the leaves are histogrammed side by side; a cluster works for exactly one leaf.
  const int ni = c->num_items;
  const int nclusters = (int)gridDim.x / HIST_CLUSTER;
  const int my_cluster = (int)blockIdx.x / HIST_CLUSTER, rank = (int)blockIdx.x % HIST_CLUSTER;
  long long rows[LGB_KMAX];
  long long total = 0;
  int live = 0;
#pragma unroll
  for (int e = 0; e < LGB_KMAX; ++e) {
    rows[e] = 0;
    if (e < ni) {
      const DevSpec* S = &V.spec[c->item_leaf[e]];
      if (!S->skip_find) { rows[e] = S->child[S->smaller].count > 0 ? S->child[S->smaller].count : 1; total += rows[e]; ++live; }
    }
  }
  if (live == 0) return;
  const int spare = nclusters - live;
  int off = 0;
#pragma unroll
  for (int e = 0; e < LGB_KMAX; ++e) {
    if (rows[e] == 0) continue;
    const int nc = 1 + (int)(rows[e] * spare / total);
    if (my_cluster >= off && my_cluster < off + nc) {
      const DevSpec* S = &V.spec[c->item_leaf[e]];
      const DevLeaf L = S->child[S->smaller];
      hist_phase(V, L, false, mode, min_rows, small_rows, (my_cluster - off) * HIST_CLUSTER + rank, nc * HIST_CLUSTER, hist_smem);
      return;
    }
    off += nc;
  }
}
