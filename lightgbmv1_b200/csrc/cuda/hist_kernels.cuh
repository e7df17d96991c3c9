// Histogram construction (replaces Dataset::ConstructHistograms, src/io/dataset.cpp:1282-1383 and
// DenseBin::ConstructHistogramInner, src/io/dense_bin.hpp:97-140).
//
// Arithmetic: gradients/hessians (float) are converted once per tree to 64-bit fixed point with a per-tree
// power-of-two scale (DevCtl::scale_g/scale_h; at most 47 value bits per row) by the lane that stages the row, and summed with integer
// adds.  Integer sums are exact and order independent, so a histogram is bit-reproducible for any grid shape, any
// number of GPUs and any atomic ordering, parent - sibling is exact, and the data-parallel all-reduce is exact; the
// only rounding is the per-row quantisation (<= 2^-39 of max|g| for 10^7 rows).
//
// Two code paths, chosen ON DEVICE from the leaf size (no host round trip):
//   global atomics : one thread per (row, 16-feature chunk), REDG.ADD.64 straight into the pool slot.  Tiny
//                    leaves (< HIST_SMALL_LEAF_ROWS rows), where zeroing / flushing a sub-histogram would cost more
//                    than the rows themselves.
//   block shared   : a block (32 warps, one per SM) owns the sub-histogram of one GROUP of 32 features and a slice of
//                    the leaf's rows.  Lane l of every warp = feature l of the group; the layout is
//                    [bin][channel][32 lanes] of 32-bit words, so lane l always sits on bank l: no bank conflicts for
//                    any bin pattern (hot bins of low-cardinality features cost nothing) and any number of warps can
//                    share the sub-histogram.  sm_100 has native shared-memory atomics only for 32-bit integers
//                    (64-bit and floating-point ones compile to ATOMS.CAST.SPIN loops), so a 64-bit entry is kept as
//                    two words: ATOMS.ADD on the low word returns the old value, the carry (`old + v` wrapped) is
//                    added together with the value's high word by a second ATOMS.ADD -- no branch, no predicate.
//                    Each wrap is seen by exactly the add that caused it and addition commutes, so (hi, lo) after
//                    the block barrier is the exact 64-bit sum whatever the interleaving.
//                    Feeding the lanes: a warp stages 32 rows at a time -- lane j loads row index j of the batch
//                    (coalesced) and that row's fixed-point (g, h) record (gathered by row; contiguous for the
//                    full-data root pass) into shared memory;
//                    every lane then reads one 16-byte broadcast per row (row offset, g low, h low, packed high
//                    words) and fetches ITS OWN bin of that row straight from the row-major matrix (32 lanes x 1 byte =
//                    the row's 32-byte sector, one L1 wavefront, 16 loads in flight per warp): no transposition
//                    through shared memory.  L1 / shared-memory data pipe per row (ncu): 4.04 ATOMS + 2.02 LDS (a
//                    16-byte broadcast is two wavefronts) + ~1.3 for the bin bytes and the staging stores = ~7.7
//                    wavefronts at ~1 per clock and SM: this pipe, not HBM, bounds the kernel.
//                    The blocks of a cluster (2) work on row slices of the same feature group, sum their
//                    sub-histograms through distributed shared memory and flush once with REDG.ADD.64 (zero entries
//                    skipped).
#pragma once
#include <cooperative_groups.h>
#include "dev_types.cuh"

#define HIST_CLUSTER 2                            /* blocks per thread-block cluster: they reduce their sub-histograms through DSMEM */
#define HIST_THREADS 1024                         /* block size of the histogram kernel: 32 warps */
#define HIST_WARPS (HIST_THREADS / 32)
#define HIST_BLOCKS_PER_SM 1
#define HIST_GROUP 32                             /* features per sub-histogram = lanes */
#define HIST_ROWS_SMEM 256                        /* bin rows of the sub-histogram */
#define HIST_LO_WORDS (HIST_ROWS_SMEM * 2 * 32)   /* low words [256 bins][2 channels][32 lanes] = 64 KB; the high words follow */
#define HIST_STAGE_BYTES 512                      /* per warp, for 32 staged rows: row offsets (128 B), (g low, h low) (256 B), packed high words (128 B) */
#define HIST_SMEM_BYTES (2 * HIST_LO_WORDS * 4 + HIST_WARPS * HIST_STAGE_BYTES)
#define HIST_SMALL_LEAF_ROWS 256
#define HIST_MIN_ROWS_PER_WARP 16                 /* rows a warp should at least get before another block is recruited */

#ifdef LGBCU_STAGE_TIMERS
__device__ unsigned long long g_hist_cycles[8];
__device__ long long g_hist_t0;
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

// A row's record as the histogram kernel consumes it: x = g low word, y = h low word, z = high words packed
// (g high in bits 0-15, h high in bits 16-31, both signed: values are at most 47 bits wide), w = free.
__device__ __forceinline__ int4 pack_gh(i64 g, i64 h) {
  const unsigned gh = (unsigned)((unsigned long long)g >> 32), hh = (unsigned)((unsigned long long)h >> 32);
  return make_int4((int)(unsigned)g, (int)(unsigned)h, (int)((gh & 0xffffu) | (hh << 16)), 0);
}
__device__ __forceinline__ i64 unpack_g(const int4& r) { return ((i64)(short)(r.z & 0xffff) << 32) | (unsigned)r.x; }
__device__ __forceinline__ i64 unpack_h(const int4& r) { return ((i64)(r.z >> 16) << 32) | (unsigned)r.y; }

__device__ void hist_global_path(const DevView& V, const DevLeaf& L, bool identity, int bid, int nblocks) {
  const int* idx = V.indices[L.buf] + L.begin;
  HistEntry* out = V.hist_pool + (size_t)L.slot * V.total_bin;
  const int nchunk = V.num_tiles16;
  const long long total = (long long)L.count * nchunk;
  for (long long p = bid * (long long)blockDim.x + threadIdx.x; p < total; p += (long long)nblocks * blockDim.x) {
    const int pos = (int)(p / nchunk);
    const int ck = (int)(p - (long long)pos * nchunk);
    const int row = identity ? pos : idx[pos];
    const uint4 q = *reinterpret_cast<const uint4*>(V.bins + (size_t)row * V.row_stride + ck * 16);
    const float2 gh = V.gh[row];
    const i64 gq = to_fixed(gh.x, V.ctl->scale_g), hq = to_fixed(gh.y, V.ctl->scale_h);
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

// Data with a wide feature (more than 256 stored values: two byte columns, dev_types.cuh): one thread per (row, feature),
// two REDG.ADD.64 per pair straight into the pool slot.  The general path -- any number of bins per feature -- and a slow
// one (global atomics; the 8-bit sub-histogram path above it in hist_phase is what the benchmarks run).
__device__ void hist_global_wide_path(const DevView& V, const DevLeaf& L, bool identity, int bid, int nblocks) {
  const int* idx = V.indices[L.buf] + L.begin;
  HistEntry* out = V.hist_pool + (size_t)L.slot * V.total_bin;
  const int F = V.num_features;
  const long long total = (long long)L.count * F;
  for (long long p = bid * (long long)blockDim.x + threadIdx.x; p < total; p += (long long)nblocks * blockDim.x) {
    const int pos = (int)(p / F);
    const int f = (int)(p - (long long)pos * F);
    const int row = identity ? pos : idx[pos];
    const DevFeat Ft = V.feats[f];
    const int b = stored_bin_row(V.bins + (size_t)row * V.row_stride, Ft);
    const float2 gh = V.gh[row];
    HistEntry* e = out + Ft.hist_base + b;
    atomic_add_i64(&e->g, to_fixed(gh.x, V.ctl->scale_g));
    atomic_add_i64(&e->h, to_fixed(gh.y, V.ctl->scale_h));
  }
}

// One 64-bit add into a (low word, high word) pair of shared memory, in two steps so that the low-word adds of
// several rows are in flight before any of their results is needed (an ATOMS result takes > 100 cycles to come back):
//   hist_atom_lo: low += v_lo, returns the old low word.  `e` = 32-bit shared address of the low word.
//   hist_add_hi : high += v_hi + carry, carry = (old + v_lo wrapped).  The high word lies HIST_LO_WORDS words further.
__device__ __forceinline__ unsigned hist_atom_lo(unsigned e, int lo) {
  unsigned old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(e), "r"(lo) : "memory");
  return old;
}
__device__ __forceinline__ void hist_add_hi(unsigned e, unsigned old, int lo, int hi) {
  asm volatile("{\n"
               " .reg .u32 t, h;\n"
               " add.cc.u32 t, %1, %2;\n"
               " addc.u32 h, %3, 0;\n"
               " red.shared.add.u32 [%0+65536], h;\n"
               "}" :: "r"(e), "r"(old), "r"(lo), "r"(hi) : "memory");
}
static_assert(HIST_LO_WORDS * 4 == 65536, "hist_add_hi addresses the high word at +64 KB");

// one bin of one row, read around L1 (every sector is used by exactly one instruction of one warp)
__device__ __forceinline__ unsigned load_bin(const uint8_t* p) {
  unsigned v;
  asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

// smem: HIST_SMEM_BYTES of dynamic shared memory.  Called by all HIST_THREADS threads of every block.
__device__ void hist_shared_path(const DevView& V, const DevLeaf& L, bool identity, int bid, int nblocks, int min_rows,
                                 unsigned char* smem) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned* lo = reinterpret_cast<unsigned*>(smem);   // [256][2][32] low words, then [256][2][32] high words
  const unsigned lo_s = (unsigned)__cvta_generic_to_shared(lo + lane);   // this lane's column: low word of (bin 0, g)
  unsigned char* stage = smem + 2 * HIST_LO_WORDS * 4 + warp * HIST_STAGE_BYTES;
  unsigned* soff = reinterpret_cast<unsigned*>(stage);       // [32] row offsets in 16-byte units
  int2* slo = reinterpret_cast<int2*>(stage + 128);          // [32] (g low word, h low word)
  int* shi = reinterpret_cast<int*>(stage + 384);            // [32] g high | h high << 16
  const uint4* soff4 = reinterpret_cast<const uint4*>(stage);
  const int4* slo4 = reinterpret_cast<const int4*>(stage + 128);
  const int4* shi4 = reinterpret_cast<const int4*>(stage + 384);

  const int NG = (V.num_features + HIST_GROUP - 1) / HIST_GROUP;   // feature groups
  // Work units: (feature group, group of HIST_CLUSTER row slices).
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int nclusters = nblocks / HIST_CLUSTER, cluster_id = bid / HIST_CLUSTER;
  int PG = (L.count + HIST_CLUSTER * HIST_WARPS * min_rows - 1) / (HIST_CLUSTER * HIST_WARPS * min_rows);
  int cap = nclusters / NG; if (cap < 1) cap = 1;
  if (PG > cap) PG = cap;
  if (PG < 1) PG = 1;
  int per_block = (L.count + PG * HIST_CLUSTER - 1) / (PG * HIST_CLUSTER);
  per_block = (per_block + 31) / 32 * 32;
  const int* idx = V.indices[L.buf] + L.begin;
  HistEntry* out = V.hist_pool + (size_t)L.slot * V.total_bin;
  const unsigned stride16 = (unsigned)V.row_stride / 16u;  // row offsets travel as 32-bit counts of 16 bytes (64 GB)
  const double sg = V.ctl->scale_g, sh = V.ctl->scale_h;
#ifdef LGBCU_STAGE_TIMERS
  long long ht_ = clock64();
  if (blockIdx.x == 0 && threadIdx.x == 0) g_hist_cycles[5] += (unsigned long long)(ht_ - g_hist_t0);
#endif
  for (int unit = cluster_id; unit < NG * PG; unit += nclusters) {
    const int group = unit / NG, fg = unit - group * NG, part = group * HIST_CLUSTER + rank;
    const int f = fg * HIST_GROUP + lane;            // this lane's feature
    const bool act = f < V.num_features;
    int b0 = part * per_block; if (b0 > L.count) b0 = L.count;
    int b1 = b0 + per_block; if (b1 > L.count) b1 = L.count;
    unsigned long long col = (unsigned long long)(V.bins + (act ? f : 0));
    asm volatile("" : "+l"(col));   // one 64-bit base register: a row's address is then a single multiply-add
    const int step = HIST_WARPS * 32;
    // prefetch the first batch's (row, gq, hq) of this lane's position
    int4 nst = make_int4(0, 0, 0, 0);
    int base = b0 + warp * 32;
    {
      const int pos = base + lane;
      if (pos < b1) {
        const int row = identity ? pos : idx[pos];
        const float2 gh = V.gh[row];
        const int4 r = pack_gh(to_fixed(gh.x, sg), to_fixed(gh.y, sh));
        nst = make_int4((int)((unsigned)row * stride16), r.x, r.y, r.z);
      }
    }
    {
      uint4* z = reinterpret_cast<uint4*>(smem);
      const uint4 zero = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int i = tid; i < 2 * HIST_LO_WORDS / 4; i += HIST_THREADS) z[i] = zero;
    }
    __syncthreads();
    HSTAGE(0);
    for (; base < b1; base += step) {
      // positions past the end of the slice are staged as (row 0, 0, 0): they add zero to some entry -- exact no-ops
      soff[lane] = (unsigned)nst.x; slo[lane] = make_int2(nst.y, nst.z); shi[lane] = nst.w;
      nst = make_int4(0, 0, 0, 0);
      {
        const int pos1 = base + step + lane;
        if (pos1 < b1) {
          const int row = identity ? pos1 : idx[pos1];
          const float2 gh = V.gh[row];
          const int4 r = pack_gh(to_fixed(gh.x, sg), to_fixed(gh.y, sh));
          nst = make_int4((int)((unsigned)row * stride16), r.x, r.y, r.z);
        }
      }
      __syncwarp();
      if (act) {
        // 16 rows at a time: their 16 bin loads are in flight per lane (x 32 warps = 512 rows per SM) before the first
        // of their 64 shared-memory adds is issued
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          unsigned bin[16];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint4 o = soff4[4 * half + q];
            bin[4 * q + 0] = load_bin(reinterpret_cast<const uint8_t*>(col + (unsigned long long)o.x * 16ull));
            bin[4 * q + 1] = load_bin(reinterpret_cast<const uint8_t*>(col + (unsigned long long)o.y * 16ull));
            bin[4 * q + 2] = load_bin(reinterpret_cast<const uint8_t*>(col + (unsigned long long)o.z * 16ull));
            bin[4 * q + 3] = load_bin(reinterpret_cast<const uint8_t*>(col + (unsigned long long)o.w * 16ull));
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {   // four rows per step: 8 low-word atomics in flight, then their high words
            const int4 la = slo4[8 * half + 2 * q], lb = slo4[8 * half + 2 * q + 1];   // rows 4q, 4q+1 | 4q+2, 4q+3
            const int4 hp = shi4[4 * half + q];
            const unsigned e0 = lo_s + bin[4 * q] * 256u, e1 = lo_s + bin[4 * q + 1] * 256u;
            const unsigned e2 = lo_s + bin[4 * q + 2] * 256u, e3 = lo_s + bin[4 * q + 3] * 256u;
            const unsigned g0 = hist_atom_lo(e0, la.x), h0 = hist_atom_lo(e0 + 128u, la.y);
            const unsigned g1 = hist_atom_lo(e1, la.z), h1 = hist_atom_lo(e1 + 128u, la.w);
            const unsigned g2 = hist_atom_lo(e2, lb.x), h2 = hist_atom_lo(e2 + 128u, lb.y);
            const unsigned g3 = hist_atom_lo(e3, lb.z), h3 = hist_atom_lo(e3 + 128u, lb.w);
            hist_add_hi(e0, g0, la.x, (int)(short)(hp.x & 0xffff)); hist_add_hi(e0 + 128u, h0, la.y, hp.x >> 16);
            hist_add_hi(e1, g1, la.z, (int)(short)(hp.y & 0xffff)); hist_add_hi(e1 + 128u, h1, la.w, hp.y >> 16);
            hist_add_hi(e2, g2, lb.x, (int)(short)(hp.z & 0xffff)); hist_add_hi(e2 + 128u, h2, lb.y, hp.z >> 16);
            hist_add_hi(e3, g3, lb.z, (int)(short)(hp.w & 0xffff)); hist_add_hi(e3 + 128u, h3, lb.w, hp.w >> 16);
          }
        }
      }
      __syncwarp();
    }
    // Reduce across the cluster, then flush: block `rank` owns a 1/HIST_CLUSTER slice of the (bin, channel) rows, sums
    // it over the HIST_CLUSTER sub-histograms (its own and the peers', read through DSMEM) and issues one
    // REDG.ADD.64 per non-zero entry -- 1/HIST_CLUSTER of the global atomics a per-block flush would need.
    HSTAGE(1);
    cluster.sync();
    HSTAGE(2);
    {
      const unsigned* rlo[HIST_CLUSTER];
#pragma unroll
      for (int k = 0; k < HIST_CLUSTER; ++k) rlo[k] = cluster.map_shared_rank(lo, k);
      int nslots = 0; i64* dst = nullptr;
      if (act) {
        const DevFeat F = V.feats[f];
        nslots = F.nslots;
        dst = reinterpret_cast<i64*>(out + F.hist_base);
      }
      const int slice = HIST_LO_WORDS / HIST_CLUSTER;
      for (int e0 = rank * slice; e0 < (rank + 1) * slice; e0 += HIST_THREADS) {
        const int e = e0 + tid;          // column (= lane = feature) is tid & 31
        const int bc = e >> 5, b = bc >> 1, ch = bc & 1;
        i64 v = 0;
#pragma unroll
        for (int k = 0; k < HIST_CLUSTER; ++k)
          v += (i64)(((unsigned long long)rlo[k][e + HIST_LO_WORDS] << 32) | rlo[k][e]);
        if (b < nslots && v) atomic_add_i64(dst + 2 * b + ch, v);
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
// Which leaf a block of the histogram launch works for, and as which block of how many (false: nothing to do).
// leaf_arg >= 0: that leaf (step-wise API).  leaf_arg < 0: training -- round 0 builds the root, every later round the
// smaller child of each leaf being expanded; all of a round's leaves share the launch.  `unit` = blocks that stay
// together (a cluster).
__device__ __forceinline__ bool hist_pick_work(const DevView& V, int leaf_arg, int unit, DevLeaf* Lp, bool* is_root_p, int* bid_p, int* nblocks_p) {
  const DevCtl* c = V.ctl;
  DevLeaf L; bool is_root = false; int bid = blockIdx.x, nblocks = gridDim.x;
  if (leaf_arg >= 0) {
    L = V.leaves[leaf_arg]; is_root = leaf_arg == 0;
  } else {
    if (c->done) return false;
    if (c->round == 0) {
      if (c->skip_find) return false;
      L = V.leaves[0]; is_root = true;
    } else {
      // The clusters of the grid are dealt to the round's leaves in proportion to their rows (at least one each), so
      // that the leaves are histogrammed side by side; a cluster works for exactly one leaf.
      const int ni = c->num_items;
      const int nclusters = (int)gridDim.x / unit;
      const int my_cluster = (int)blockIdx.x / unit, rank = (int)blockIdx.x % unit;
      int rows[LGB_KMAX];
      long long total = 0;
      int live = 0;
#pragma unroll
      for (int e = 0; e < LGB_KMAX; ++e) {
        rows[e] = e < ni ? c->hist_rows[e] : 0;
        total += rows[e]; live += rows[e] != 0;
      }
      if (live == 0) return false;
      const int spare = nclusters - live;
      // floor(rows * spare / total) without a 64-bit division: scaled down a hair so that float rounding can never
      // hand out more clusters than there are (every block computes the same numbers)
      const float share = (float)spare / (float)total * 0.99999f;
      int off = 0, mine = -1;
#pragma unroll
      for (int e = 0; e < LGB_KMAX; ++e) {
        if (rows[e] == 0) continue;
        const int nc = 1 + (int)((float)rows[e] * share);
        if (mine < 0 && my_cluster >= off && my_cluster < off + nc) {
          mine = e;
          bid = (my_cluster - off) * unit + rank; nblocks = nc * unit;
        }
        off += nc;
      }
      if (mine < 0) return false;
      L = c->hist_leaf[mine];
    }
  }
  *Lp = L; *is_root_p = is_root; *bid_p = bid; *nblocks_p = nblocks;
  return true;
}

__global__ void __cluster_dims__(HIST_CLUSTER, 1, 1) __launch_bounds__(HIST_THREADS, HIST_BLOCKS_PER_SM) k_hist(DevView V, int leaf_arg, int mode, int min_rows, int small_rows) {
  extern __shared__ __align__(16) unsigned char hist_smem[];
  const DevCtl* c = V.ctl;
#ifdef LGBCU_STAGE_TIMERS
  if (blockIdx.x == 0 && threadIdx.x == 0) g_hist_t0 = clock64();
#endif
  STAMP(V.ctl, 0);
  DevLeaf L; bool is_root; int bid, nblocks;
  if (!hist_pick_work(V, leaf_arg, HIST_CLUSTER, &L, &is_root, &bid, &nblocks)) return;
  hist_phase(V, L, is_root, mode, min_rows, small_rows, bid, nblocks, hist_smem);
}


// The same launch for data with a wide feature.  k_hist's scheme -- a block-shared sub-histogram of 256 bins x 32
// features, lane = feature -- does not stretch to thousands of bins, so here a block takes ONE feature and a slice of
// the leaf's rows: the feature's whole histogram (16 bytes per slot, `smem_slots` of them fit) lives in shared memory as
// (low word, high word) pairs per sum, the rows' bins come from the column-major copy (both bytes of a wide feature,
// coalesced for the full-data root pass), a row's value is added with a 32-bit shared atomic per word and the carry
// handed to the high word -- the same exact 64-bit integer sums as everywhere -- and the block flushes its non-zero
// slots to the leaf's pool slot with global atomics.  Leaves too small to amortise zeroing and flushing, and features with
// more slots than fit, take the global-atomic path (hist_global_wide_path).  Measured at 2M x 28, max_bin = 1023:
// 2.5 ms of histogram time per iteration against 10.2 ms with global atomics alone (the 8-bit kernel: 0.74 ms at 255 bins).
#define WIDE_HIST_MIN_ELEMS 32768   /* rows x features below which a leaf goes straight to global atomics */
__device__ __forceinline__ void wide_add64(unsigned* lo_word, unsigned* hi_word, i64 v) {
  const unsigned lo = (unsigned)v;
  const unsigned old = atomicAdd(lo_word, lo);
  const unsigned carry = (old + lo) < old ? 1u : 0u;
  const unsigned hi = (unsigned)((unsigned long long)v >> 32) + carry;
  if (hi) atomicAdd(hi_word, hi);
}

__global__ void __launch_bounds__(256) k_hist_wide(DevView V, int leaf_arg, int smem_slots) {
  // four planes of one word per slot -- g low, g high, h low, h high -- so that the lanes of one atomic instruction, which
  // hit random slots, spread over all 32 banks (slot-major [slot][4] words put every low word on one of 8 banks)
  extern __shared__ __align__(16) unsigned wide_smem[];
  STAMP(V.ctl, 0);
  DevLeaf L; bool is_root; int bid, nblocks;
  if (!hist_pick_work(V, leaf_arg, 1, &L, &is_root, &bid, &nblocks)) return;
  const bool identity = V.ctl->root_identity && is_root && L.count == V.num_data;
  const int F = V.num_features, tid = threadIdx.x;
  if ((long long)L.count * F < WIDE_HIST_MIN_ELEMS) {
    hist_global_wide_path(V, L, identity, bid, nblocks);
  } else {
    const int* idx = V.indices[L.buf] + L.begin;
    const double scale_g = V.ctl->scale_g, scale_h = V.ctl->scale_h;
    int slices = nblocks / F; if (slices < 1) slices = 1;
    const int units = F * slices;
    const int rows_per_slice = (L.count + slices - 1) / slices;
    for (int u = bid; u < units; u += nblocks) {
      const int f = u % F, sl = u / F;   // neighbouring blocks: the same rows, different features (the gh records stay in L2)
      const DevFeat Ft = V.feats[f];
      const int r0 = sl * rows_per_slice, r1 = min(L.count, r0 + rows_per_slice);
      const uint8_t* col = V.bins_col + (size_t)Ft.col * V.col_stride;
      HistEntry* out = V.hist_pool + (size_t)L.slot * V.total_bin + Ft.hist_base;
      if (Ft.nslots > smem_slots) {      // does not fit: global atomics for this (feature, slice)
        for (int p = r0 + tid; p < r1; p += blockDim.x) {
          const int row = identity ? p : idx[p];
          int b = col[row]; if (Ft.wide) b |= (int)col[row + V.col_stride] << 8;
          const float2 gh = V.gh[row];
          atomic_add_i64(&out[b].g, to_fixed(gh.x, scale_g));
          atomic_add_i64(&out[b].h, to_fixed(gh.y, scale_h));
        }
        continue;
      }
      const int ns = Ft.nslots;
      unsigned* g_lo = wide_smem; unsigned* g_hi = wide_smem + ns; unsigned* h_lo = wide_smem + 2 * ns; unsigned* h_hi = wide_smem + 3 * ns;
      for (int i = tid; i < ns * 4; i += blockDim.x) wide_smem[i] = 0u;
      __syncthreads();
      for (int p = r0 + tid; p < r1; p += blockDim.x) {
        const int row = identity ? p : idx[p];
        int b = col[row]; if (Ft.wide) b |= (int)col[row + V.col_stride] << 8;
        const float2 gh = V.gh[row];
        wide_add64(&g_lo[b], &g_hi[b], to_fixed(gh.x, scale_g));
        wide_add64(&h_lo[b], &h_hi[b], to_fixed(gh.y, scale_h));
      }
      __syncthreads();
      for (int b = tid; b < ns; b += blockDim.x) {
        const i64 g = (i64)(((unsigned long long)g_hi[b] << 32) | g_lo[b]);
        const i64 h = (i64)(((unsigned long long)h_hi[b] << 32) | h_lo[b]);
        if (g) atomic_add_i64(&out[b].g, g);
        if (h) atomic_add_i64(&out[b].h, h);
      }
      __syncthreads();
    }
  }
  if (bid == 0 && threadIdx.x == 0) {
    atomicAdd(&V.ctl->prof_hist_calls, 1ull);
    atomicAdd(identity ? &V.ctl->prof_hist_rows : &V.ctl->prof_hist_rows_indexed, (unsigned long long)L.count);
  }
}
