// Data partition (replaces DataPartition::Split, src/treelearner/data_partition.hpp:101-120;
// ParallelPartitionRunner::Run, include/LightGBM/utils/threading.h:105-172; DenseBin::SplitInner,
// src/io/dense_bin.hpp:173-253 for single-feature groups) and the host bookkeeping that follows it in
// SerialTreeLearner::SplitInner (serial_tree_learner.cpp:541-659) and BeforeFindBestSplit (:274-314).
//
// Stable two-way compaction in three launches, all sized for the worst case and driven by DevCtl:
//   k_part_flags   : per row, gather the split feature's stored bin, decide left/right, warp-ballot the
//                    decisions into a bit mask and count lefts per 1024-row tile;
//   k_part_scan    : one block; exclusive scan of the tile counts, then the control step (leaf table, split
//                    record, smaller/larger choice, depth / min_data gates);
//   k_part_scatter : per row, rank = tile offset + popc prefix of the ballot words; write the row index into
//                    the other ping-pong buffer, lefts first -- order inside each side is preserved.
#pragma once
#include "dev_types.cuh"

#define PART_THREADS 256
#define PART_REPS 4
#define PART_TILE (PART_THREADS * PART_REPS) /* 1024 rows = 32 ballot words */

// DenseBin::SplitInner for a single-feature group (min_bin = 1, USE_MIN_BIN = false).
struct SplitRule {
  int th, t_zero, maxb;
  int default_left_side;   // side of stored value 0 (the most-frequent bin) when it is not the missing bin
  int missing_left;        // side of missing rows
  int miss_zero, miss_na, mfb_zero, mfb_na;
  int wide;                // min_bin < max_bin
  int maxbin_left;         // narrow case: side of the single non-default slot
};

__device__ __forceinline__ SplitRule make_rule(const DevFeat& F, int threshold, int default_left) {
  SplitRule r;
  r.maxb = F.nslots - 1;
  r.th = threshold + 1;
  r.t_zero = 1 + F.default_bin;
  if (F.mfb == 0) { --r.th; --r.t_zero; }
  r.default_left_side = F.mfb <= threshold;
  r.miss_zero = F.missing_type == LGBCU_MISSING_ZERO;
  r.miss_na = F.missing_type == LGBCU_MISSING_NAN;
  r.mfb_zero = r.miss_zero && F.default_bin == F.mfb;
  r.mfb_na = r.miss_na && (r.maxb == F.mfb + 1) && F.mfb > 0;
  r.missing_left = (r.miss_zero || r.miss_na) ? (default_left != 0) : 0;
  r.wide = 1 < r.maxb;
  r.maxbin_left = r.maxb <= r.th;
  return r;
}

__device__ __forceinline__ bool goes_left(const SplitRule& r, int bin) {
  if (r.wide) {
    if ((r.miss_zero && !r.mfb_zero && bin == r.t_zero) || (r.miss_na && !r.mfb_na && bin == r.maxb)) {
      return r.missing_left;
    } else if (bin == 0) {
      if ((r.miss_na && r.mfb_na) || (r.miss_zero && r.mfb_zero)) return r.missing_left;
      return r.default_left_side;
    }
    return !(bin > r.th);
  }
  if (r.miss_zero && !r.mfb_zero && bin == r.t_zero) return r.missing_left;
  if (bin != r.maxb) {
    if ((r.miss_na && r.mfb_na) || (r.miss_zero && r.mfb_zero)) return r.missing_left;
    return r.default_left_side;
  }
  if (r.miss_na && !r.mfb_na) return r.missing_left;
  return r.maxbin_left;
}

__device__ void part_flags_phase(const DevView& V, int bid, int nblocks) {
  const DevCtl* c = V.ctl;
  const DevLeaf L = V.leaves[c->split_leaf];
  const DevFeat F = V.feats[c->split_feature];
  const SplitRule rule = make_rule(F, c->split_threshold, c->split_default_left);
  const int* src = V.indices[L.buf] + L.begin;
  const uint8_t* col = V.bins + c->split_feature;
  const int ntiles = (L.count + PART_TILE - 1) / PART_TILE;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ int s_cnt[PART_THREADS / 32];
  for (int tile = bid; tile < ntiles; tile += nblocks) {
    int lefts = 0;
#pragma unroll
    for (int r = 0; r < PART_REPS; ++r) {
      const int pos = tile * PART_TILE + r * PART_THREADS + threadIdx.x;
      bool left = false;
      if (pos < L.count) {
        const int row = src[pos];
        left = goes_left(rule, col[(size_t)row * V.row_stride]);
      }
      const unsigned word = __ballot_sync(0xffffffffu, left);
      if (lane == 0) {
        V.part_bits[(size_t)tile * (PART_TILE / 32) + r * (PART_THREADS / 32) + warp] = word;
        lefts += __popc(word);
      }
    }
    if (lane == 0) s_cnt[warp] = lefts;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < PART_THREADS / 32; ++w) t += s_cnt[w];
      V.tile_left[tile] = t;
    }
    __syncthreads();
  }
}

__device__ void part_control(const DevView& V, int mode, int leaf, const DevLeaf& L, int left_cnt);

// mode 0: training (control step runs); mode 1: step-wise API (only counts/ranges are updated, sums untouched).
// One block (any multiple of 32 threads up to 1024).
__device__ void part_scan_phase(const DevView& V, int mode) {
  DevCtl* c = V.ctl;
  const int leaf = c->split_leaf;
  const DevLeaf L = V.leaves[leaf];
  const int ntiles = (L.count + PART_TILE - 1) / PART_TILE;
  if (mode == 0) {  // snapshot FeatureHistogram::is_splittable_ of the leaf being split for k_find_split
    const int8_t* src = V.splittable + (size_t)L.slot * V.num_features;
    for (int f = threadIdx.x; f < V.num_features; f += blockDim.x) V.parent_split[f] = src[f];
  }
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const int nthreads = blockDim.x;
  for (int base = 0; base < ntiles; base += nthreads) {
    const int i = base + threadIdx.x;
    const int v = i < ntiles ? V.tile_left[i] : 0;
    int inc = v;
    for (int o = 1; o < 32; o <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      int w = lane < nthreads / 32 ? s_warp[lane] : 0;
      int winc = w;
      for (int o = 1; o < 32; o <<= 1) {
        const int x = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += x;
      }
      s_warp[lane] = winc - w;  // exclusive prefix of warp sums
    }
    __syncthreads();
    const int excl = s_carry + s_warp[warp] + inc - v;
    if (i < ntiles) V.tile_left[i] = excl;
    __syncthreads();
    if (threadIdx.x == nthreads - 1) s_carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x != 0) return;
  part_control(V, mode, leaf, L, s_carry);
}

// The bookkeeping that follows a partition (one thread): leaf table, split record, smaller/larger, gates.
__device__ void part_control(const DevView& V, int mode, int leaf, const DevLeaf& L, int left_cnt) {
  DevCtl* c = V.ctl;
  const int right_cnt = L.count - left_cnt;
  c->split_left_count = left_cnt;
  const int R = c->split_right_leaf;
  DevLeaf A = L, B = L;  // A = left child (keeps the id), B = right child (new id R)
  A.count = left_cnt; A.buf = L.buf ^ 1; A.depth = L.depth + 1; A.gcount = left_cnt;
  B.begin = L.begin + left_cnt; B.count = right_cnt; B.buf = L.buf ^ 1; B.depth = L.depth + 1; B.gcount = right_cnt;
  A.tot_g = A.tot_h = B.tot_g = B.tot_h = 0;
  if (mode == 1) {
    B.slot = R;
    V.leaves[leaf] = A;
    V.leaves[R] = B;
    if (c->num_leaves < R + 1) c->num_leaves = R + 1;
    return;
  }
  const DevSplit s = V.best[leaf];
  // SplitInner: on one GPU (update_cnt = true) the counts come from the partition; in data-parallel mode
  // (update_cnt = false, data_parallel_tree_learner.cpp:250-256) they stay the global, hessian-derived counts of
  // the SplitInfo, which every rank computed identically.  Everything else always comes from the SplitInfo.
  const int gl = V.data_parallel ? s.left_count : left_cnt;
  const int gr = V.data_parallel ? s.right_count : right_cnt;
  A.gcount = gl; B.gcount = gr;
  LgbCuSplitRecord rec;
  rec.leaf = leaf; rec.right_leaf = R; rec.feature = s.feature; rec.real_feature = s.real_feature;
  rec.threshold = s.threshold; rec.default_left = s.default_left;
  rec.left_count = gl; rec.right_count = gr;
  rec.gain = (float)(s.gain + V.p.min_gain_to_split); rec.reserved = 0;
  rec.left_output = s.left_output; rec.right_output = s.right_output;
  rec.left_sum_hessian = s.left_sum_h; rec.right_sum_hessian = s.right_sum_h;
  rec.left_sum_gradient = s.left_sum_g; rec.right_sum_gradient = s.right_sum_g;
  V.records[c->num_splits] = rec;
  c->num_splits += 1;
  A.sum_g = s.left_sum_g; A.sum_h = s.left_sum_h; A.output = s.left_output;
  B.sum_g = s.right_sum_g; B.sum_h = s.right_sum_h; B.output = s.right_output;
  // smaller / larger (serial_tree_learner.cpp:629-649) and the histogram-pool hand-over (:298-312): the
  // larger child inherits the parent's slot (subtraction happens in place), the smaller gets a fresh one.
  int smaller, larger;
  if (gl < gr) { smaller = leaf; larger = R; A.slot = R; B.slot = L.slot; }
  else { smaller = R; larger = leaf; B.slot = R; A.slot = L.slot; }
  V.leaves[leaf] = A;
  V.leaves[R] = B;
  c->smaller = smaller; c->larger = larger;
  c->num_leaves = R + 1;
  // BeforeFindBestSplit gates (serial_tree_learner.cpp:274-297)
  int skip = 0;
  if (V.p.max_depth > 0 && A.depth >= V.p.max_depth) skip = 1;
  if (!skip && gr < V.p.min_data_in_leaf * 2 && gl < V.p.min_data_in_leaf * 2) skip = 1;
  if (skip) {
    V.best[leaf].gain = LGB_MIN_SCORE;
    V.leaf_key[leaf].gain = LGB_MIN_SCORE;
    DevSplit none; none.gain = LGB_MIN_SCORE; none.feature = -1; none.real_feature = -1;
    V.best[R] = none;
    { SplitKey k; k.gain = LGB_MIN_SCORE; k.real_feature = -1; k.pad = 0; V.leaf_key[R] = k; }
  }
  c->skip_find = skip;
  if (c->num_leaves >= V.p.num_leaves) c->done = 2;  // tree is full; nothing more to enqueue
}

__device__ void part_scatter_phase(const DevView& V, int bid, int nblocks) {
  const DevCtl* c = V.ctl;
  const int leaf = c->split_leaf;
  const DevLeaf A = V.leaves[leaf];  // already the left child: begin unchanged, buf flipped
  const int left_total = c->split_left_count;
  const int count = left_total + V.leaves[c->split_right_leaf].count;
  const int* src = V.indices[A.buf ^ 1] + A.begin;
  int* dst = V.indices[A.buf] + A.begin;
  const int ntiles = (count + PART_TILE - 1) / PART_TILE;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int tile = bid; tile < ntiles; tile += nblocks) {
    // every warp redundantly scans the tile's 32 ballot words (no block sync needed)
    const unsigned myword = V.part_bits[(size_t)tile * (PART_TILE / 32) + lane];
    const int pc = __popc(myword);
    int inc = pc;
    for (int o = 1; o < 32; o <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    const int excl = inc - pc;
    const int tile_off = V.tile_left[tile];
#pragma unroll
    for (int r = 0; r < PART_REPS; ++r) {
      const int w = r * (PART_THREADS / 32) + warp;
      const unsigned word = __shfl_sync(0xffffffffu, myword, w);
      const int wpre = __shfl_sync(0xffffffffu, excl, w);
      const int pos = tile * PART_TILE + w * 32 + lane;
      if (pos < count) {
        const unsigned lt = (1u << lane) - 1u;
        const int row = src[pos];
        if ((word >> lane) & 1u) {
          dst[tile_off + wpre + __popc(word & lt)] = row;
        } else {
          const int rights_before = (tile * PART_TILE - tile_off) + (w * 32 - wpre) + __popc(~word & lt);
          dst[left_total + rights_before] = row;
        }
      }
    }
  }
}

// stand-alone launches (step-wise API and the non-persistent training loop)
__global__ void __launch_bounds__(PART_THREADS) k_part_flags(DevView V) {
  if (V.ctl->done) return;
  part_flags_phase(V, blockIdx.x, gridDim.x);
}
__global__ void __launch_bounds__(1024) k_part_scan(DevView V, int mode) {
  if (V.ctl->done) return;
  part_scan_phase(V, mode);
}
__global__ void __launch_bounds__(PART_THREADS) k_part_scatter(DevView V, int unused) {
  // done == 2 is set by the k_part_scan that just ran for the LAST split: its scatter must still happen
  if (V.ctl->done == 1) return;
  (void)unused;
  part_scatter_phase(V, blockIdx.x, gridDim.x);
}

// ---------------------------------------------------------------------------------------------------------------
// Fused partition: flags -> (inter-block barrier) -> scatter + control in ONE launch.  The three-launch version
// above spends most of its time in launch ramps and dependent control loads for the typical 10^4..10^5-row leaf;
// here the only inter-block communication is one left-count per block and one arrive counter.
//   * blocks [0, P) own contiguous chunks of 1024-row tiles; block P (one extra) is the control block;
//   * phase 1: decisions -> ballot words (shared memory for a block's first PART_SMEM_TILES tiles, V.part_bits
//     beyond), block left-count -> V.tile_left[block]; arrive;
//   * barrier: spin until all P blocks have arrived (the launch is cooperative, so co-residency is guaranteed);
//   * phase 2: every block sums the counts of the blocks before it (its write offset) and of all blocks (where the
//     right side starts), then scatters its rows; the control block runs part_control meanwhile.
// The arrive/depart counters reset themselves: the last block to leave the barrier zeroes both.
#define PART_SMEM_TILES 8
#ifdef LGBCU_STAGE_TIMERS
#define PSTAGE(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) { long long t_ = clock64(); g_stage_cycles[8 + (i)] += (unsigned long long)(t_ - pt_); pt_ = t_; } } while (0)
#else
#define PSTAGE(i) do {} while (0)
#endif
__global__ void __launch_bounds__(PART_THREADS) k_partition(DevView V, int mode) {
  DevCtl* c = V.ctl;
#ifdef LGBCU_STAGE_TIMERS
  long long pt_ = clock64();
#endif
  if (c->done) return;
  STAMP(c, 2);
  const int leaf = c->split_leaf;
  const DevLeaf L = V.leaves[leaf];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ntiles = (L.count + PART_TILE - 1) / PART_TILE;
  const int maxp = (int)gridDim.x - 1;
  const int tpb = (ntiles + maxp - 1) / maxp;          // tiles per block
  const int P = (ntiles + tpb - 1) / tpb;              // blocks that own tiles
  const int bid = blockIdx.x;
  if (bid > P) return;
  const bool control = bid == P;
  __shared__ unsigned s_bits[PART_SMEM_TILES][PART_TILE / 32];
  __shared__ int s_cnt[PART_THREADS / 32];
  __shared__ int s_red[2][PART_THREADS / 32];
  const int t0 = bid * tpb;
  int t1 = t0 + tpb; if (t1 > ntiles) t1 = ntiles;
  const int* src = V.indices[L.buf] + L.begin;
  int* dst = V.indices[L.buf ^ 1] + L.begin;
  int row0[PART_REPS];  // rows of the block's first tile stay in registers for the scatter
  PSTAGE(0);
  if (!control) {
    const DevFeat F = V.feats[c->split_feature];
    const SplitRule rule = make_rule(F, c->split_threshold, c->split_default_left);
    const uint8_t* col = V.bins + c->split_feature;
    int lefts = 0;
    for (int tile = t0; tile < t1; ++tile) {
      const int lt = tile - t0;
#pragma unroll
      for (int r = 0; r < PART_REPS; ++r) {
        const int pos = tile * PART_TILE + r * PART_THREADS + tid;
        bool left = false;
        int row = 0;
        if (pos < L.count) {
          row = src[pos];
          left = goes_left(rule, col[(size_t)row * V.row_stride]);
        }
        if (lt == 0) row0[r] = row;
        const unsigned word = __ballot_sync(0xffffffffu, left);
        if (lane == 0) {
          const int wi = r * (PART_THREADS / 32) + warp;
          if (lt < PART_SMEM_TILES) s_bits[lt][wi] = word;
          else V.part_bits[(size_t)tile * (PART_TILE / 32) + wi] = word;
          lefts += __popc(word);
        }
      }
    }
    if (lane == 0) s_cnt[warp] = lefts;
    __syncthreads();
    PSTAGE(1);
    if (tid == 0) {
      int t = 0;
      for (int w = 0; w < PART_THREADS / 32; ++w) t += s_cnt[w];
      __stcg(&V.tile_left[bid], t);
      __threadfence();
      atomicAdd(&c->bar_arrive, 1u);
    }
  }
  if (tid == 0) {
    while (*reinterpret_cast<volatile unsigned*>(&c->bar_arrive) < (unsigned)P) { }
    __threadfence();
  }
  __syncthreads();
  PSTAGE(2);
  // offsets: lefts of the blocks before this one, and of all blocks
  int before = 0, total = 0;
  for (int b = tid; b < P; b += PART_THREADS) {
    const int v = __ldcg(&V.tile_left[b]);
    total += v;
    if (b < bid) before += v;
  }
  for (int o = 16; o > 0; o >>= 1) {
    before += __shfl_xor_sync(0xffffffffu, before, o);
    total += __shfl_xor_sync(0xffffffffu, total, o);
  }
  if (lane == 0) { s_red[0][warp] = before; s_red[1][warp] = total; }
  __syncthreads();
  before = 0; total = 0;
#pragma unroll
  for (int w = 0; w < PART_THREADS / 32; ++w) { before += s_red[0][w]; total += s_red[1][w]; }
  PSTAGE(3);
  if (tid == 0) {  // leave the barrier; the last one out re-arms it
    const unsigned d = atomicAdd(&c->bar_depart, 1u);
    if (d == (unsigned)P) { c->bar_arrive = 0; c->bar_depart = 0; }
  }
#ifdef LGBCU_STAGE_TIMERS
  const int split_ = V.ctl->num_splits;  // (racy by one against the control block: good enough for a probe)
#endif
  if (control) {
    if (mode == 0) {  // snapshot FeatureHistogram::is_splittable_ of the leaf being split for k_find_split
      const int8_t* sp = V.splittable + (size_t)L.slot * V.num_features;
      for (int f = tid; f < V.num_features; f += PART_THREADS) V.parent_split[f] = sp[f];
    }
    if (tid == 0) part_control(V, mode, leaf, L, total);
    return;
  }
  const int left_total = total;
  int tile_off = before;
  for (int tile = t0; tile < t1; ++tile) {
    const int lt = tile - t0;
    const unsigned myword = lt < PART_SMEM_TILES ? s_bits[lt][lane] : V.part_bits[(size_t)tile * (PART_TILE / 32) + lane];
    const int pc = __popc(myword);
    int inc = pc;
    for (int o = 1; o < 32; o <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    const int excl = inc - pc;
    const int tile_total = __shfl_sync(0xffffffffu, inc, 31);
#pragma unroll
    for (int r = 0; r < PART_REPS; ++r) {
      const int w = r * (PART_THREADS / 32) + warp;
      const unsigned word = __shfl_sync(0xffffffffu, myword, w);
      const int wpre = __shfl_sync(0xffffffffu, excl, w);
      const int pos = tile * PART_TILE + w * 32 + lane;
      if (pos < L.count) {
        const unsigned ltm = (1u << lane) - 1u;
        const int row = lt == 0 ? row0[r] : src[pos];
        if ((word >> lane) & 1u) {
          dst[tile_off + wpre + __popc(word & ltm)] = row;
        } else {
          const int rights_before = (tile * PART_TILE - tile_off) + (w * 32 - wpre) + __popc(~word & ltm);
          dst[left_total + rights_before] = row;
        }
      }
    }
    tile_off += tile_total;
  }
  PSTAGE(4);
  STAMP_END(split_, 2);
}
