// Data partition (replaces DataPartition::Split, src/treelearner/data_partition.hpp:101-120;
// ParallelPartitionRunner::Run, include/LightGBM/utils/threading.h:105-172; DenseBin::SplitInner,
// src/io/dense_bin.hpp:173-253 for single-feature groups) and the host bookkeeping that follows it in
// SerialTreeLearner::SplitInner (serial_tree_learner.cpp:541-659) and BeforeFindBestSplit (:274-314).
//
// Stable two-way compaction in ONE launch (k_partition, below), driven by DevCtl: per row, gather the split
// feature's stored bin, decide left/right, warp-ballot the decisions into bit masks, exchange one left-count per
// block, then write the row index into the other ping-pong buffer, lefts first -- order inside each side is
// preserved; one extra block does the control step (leaf table, split record, smaller/larger, gates).
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

// The bookkeeping that follows a partition (one thread).
// Step-wise API (lgbcu_partition): the children replace the leaf at once; sums are left to the caller.
__device__ void part_control_stepwise(const DevView& V, int leaf, const DevLeaf& L, int left_cnt) {
  DevCtl* c = V.ctl;
  const int right_cnt = L.count - left_cnt;
  c->split_left_count = left_cnt;
  const int R = c->split_right_leaf;
  DevLeaf A = L, B = L;  // A = left child (keeps the id), B = right child (new id R)
  A.count = left_cnt; A.buf = L.buf ^ 1; A.depth = L.depth + 1; A.gcount = left_cnt;
  B.begin = L.begin + left_cnt; B.count = right_cnt; B.buf = L.buf ^ 1; B.depth = L.depth + 1; B.gcount = right_cnt;
  B.slot = R;
  V.leaves[leaf] = A;
  V.leaves[R] = B;
  if (c->num_leaves < R + 1) c->num_leaves = R + 1;
}

// Training: item e of the round expands leaf `leaf`.  The children (ranges, sums and outputs from the leaf's
// SplitInfo as in SerialTreeLearner::SplitInner, serial_tree_learner.cpp:541-659; smaller / larger and the
// histogram-pool hand-over, :298-312, 629-649; BeforeFindBestSplit gates, :274-297) go to spec[leaf]; the leaf
// table is only touched when the expansion is committed (find_control).
__device__ void part_expand(const DevView& V, int e, int leaf, const DevLeaf& L, int left_cnt) {
  DevCtl* c = V.ctl;
  DevSpec* S = &V.spec[leaf];
  const int right_cnt = L.count - left_cnt;
  DevLeaf A = L, B = L;  // A = left child, B = right child
  A.count = left_cnt; A.buf = L.buf ^ 1; A.depth = L.depth + 1;
  B.begin = L.begin + left_cnt; B.count = right_cnt; B.buf = L.buf ^ 1; B.depth = L.depth + 1;
  const DevSplit s = V.best[leaf];
  // SplitInner: on one GPU (update_cnt = true) the counts come from the partition; in data-parallel mode
  // (update_cnt = false, data_parallel_tree_learner.cpp:250-256) they stay the global, hessian-derived counts of
  // the SplitInfo, which every rank computed identically.  Everything else always comes from the SplitInfo.
  const int gl = V.data_parallel ? s.left_count : left_cnt;
  const int gr = V.data_parallel ? s.right_count : right_cnt;
  A.gcount = gl; B.gcount = gr;
  A.sum_g = s.left_sum_g; A.sum_h = s.left_sum_h; A.output = s.left_output;
  B.sum_g = s.right_sum_g; B.sum_h = s.right_sum_h; B.output = s.right_output;
  if (V.use_mc) {
    // BasicLeafConstraints::Update (monotone_constraints.hpp:100-117): both children start from the leaf's interval (A
    // and B are copies of L); a numerical split on a constrained feature puts the mean of the two outputs between them
    const int type = V.monotone[s.feature];
    if (type != 0 && !V.feats[s.feature].cat) {
      const double mid = (s.left_output + s.right_output) / 2.0f;
      if (type < 0) { A.mc_min = fmax(mid, A.mc_min); B.mc_max = fmin(mid, B.mc_max); }
      else { A.mc_max = fmin(mid, A.mc_max); B.mc_min = fmax(mid, B.mc_min); }
    }
  }
  // the larger child inherits the parent's slot (subtraction happens in place); the smaller one gets a fresh slot
  // whose index depends on (round, item) only, so the host can address it blind (data-parallel all-reduce)
  const int fresh = 1 + (c->round - 1) * c->lookahead + e;
  int smaller;
  if (gl < gr) { smaller = 0; A.slot = fresh; B.slot = L.slot; }
  else { smaller = 1; B.slot = fresh; A.slot = L.slot; }
  int skip = 0;
  if (V.p.max_depth > 0 && A.depth >= V.p.max_depth) skip = 1;
  if (!skip && gr < V.p.min_data_in_leaf * 2 && gl < V.p.min_data_in_leaf * 2) skip = 1;
  S->child[0] = A; S->child[1] = B;
  S->left_count = left_cnt; S->smaller = smaller; S->skip_find = skip;
  const DevLeaf& Sm = smaller == 0 ? A : B;
  c->hist_leaf[e] = Sm;
  c->hist_rows[e] = skip ? 0 : (Sm.count > 0 ? Sm.count : 1);
}

// ---------------------------------------------------------------------------------------------------------------
// Fused partition: flags -> (inter-block exchange) -> scatter + control in ONE launch.
//   * blocks [0, P) own contiguous chunks of 1024-row tiles; block P (one extra) is the control block;
//   * phase 1: decisions -> ballot words (shared memory for a block's first PART_SMEM_TILES tiles, V.part_bits
//     beyond); the block's left-count is published as ONE 64-bit word (epoch << 32 | count) in V.part_sync[block];
//   * exchange: a block bumps a (never reset) arrive counter after publishing; one thread per block polls that
//     counter until all P blocks are in (polling the words themselves from every thread measured slower: the
//     readers starve the writers), then every block reads the P words once.  Count and "ready" tag travel in the
//     same word, so no fence is needed: a reader that still finds an old tag spins on that word.  The sums over the
//     blocks before it / over all blocks give a block its write offset and the start of the right side.  (The
//     launch is cooperative, so all blocks are resident and the spins cannot deadlock.)
//   * phase 2: scatter; the control block runs part_control meanwhile and bumps the epoch.
#define PART_SMEM_TILES 8
#define PART_BLOCKS_PER_SM 4   /* tile-owning blocks per SM (all must be resident: they meet at a barrier) */
#ifdef LGBCU_STAGE_TIMERS
#define PSTAGE(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) { long long t_ = clock64(); g_stage_cycles[8 + (i)] += (unsigned long long)(t_ - pt_); pt_ = t_; } } while (0)
#else
#define PSTAGE(i) do {} while (0)
#endif
// WIDE: the data holds a feature with more than 256 stored values (two byte columns); a separate instantiation, so that
// the 8-bit kernel keeps its registers
template <bool WIDE>
__global__ void __launch_bounds__(PART_THREADS, PART_BLOCKS_PER_SM + 1) k_partition(DevView V, int mode) {
  DevCtl* c = V.ctl;
#ifdef LGBCU_STAGE_TIMERS
  long long pt_ = clock64();
#endif
  if (c->done) return;
  STAMP(c, 2);
  // which item (leaf being expanded) this block works for, and its index among that item's blocks
  int e = 0, bid = blockIdx.x, maxp = (int)gridDim.x - 1;
  int leaf = c->split_leaf, feature = c->split_feature, threshold = c->split_threshold, default_left = c->split_default_left;
  if (mode == 0) {
    const int ni = c->num_items;
    if (ni <= 0) return;
    // the fresh histogram slots of this round's expansions (see part_expand) must be zero before k_hist adds to them:
    // every block of the grid clears a slice of each, whether or not it owns tiles
    for (int i = 0; i < ni; ++i) {
      uint4* z = reinterpret_cast<uint4*>(V.hist_pool + (size_t)(1 + (c->round - 1) * c->lookahead + i) * V.total_bin);
      const uint4 zero = make_uint4(0, 0, 0, 0);
      for (int k = blockIdx.x * PART_THREADS + threadIdx.x; k < V.total_bin; k += gridDim.x * PART_THREADS) z[k] = zero;
    }
    // the grid's blocks are dealt to the items in proportion to their 1024-row tiles; one control block per item
    int tiles[LGB_KMAX];
    int T = 0;
#pragma unroll
    for (int i = 0; i < LGB_KMAX; ++i) {
      tiles[i] = i < ni ? (c->item_rec[i].count + PART_TILE - 1) / PART_TILE : 0;
      T += tiles[i];
    }
    // floor(tiles * share / T) in float, scaled down a hair so that rounding never hands out more blocks than there are
    const float share = T > 0 ? (float)((int)gridDim.x - 2 * ni) / (float)T * 0.99999f : 0.f;
    int off = 0;
    bool found = false;
#pragma unroll
    for (int i = 0; i < LGB_KMAX; ++i) {
      if (i < ni && !found) {
        int nb = (int)((float)tiles[i] * share);
        if (nb < 1) nb = 1;
        if (nb > tiles[i] && tiles[i] > 0) nb = tiles[i];
        if ((int)blockIdx.x < off + nb + 1) { e = i; bid = (int)blockIdx.x - off; maxp = nb; found = true; }
        off += nb + 1;
      }
    }
    if (!found) return;
    leaf = c->item_leaf[e]; feature = c->item_feature[e]; threshold = c->item_threshold[e]; default_left = c->item_default_left[e];
  }
  const unsigned epoch = c->part_epoch[e] + 1u;  // every block of the item reads these before it publishes; the item's
  const unsigned bar_base = c->bar_base[e];      // control block updates them after all of them have arrived
  const DevLeaf L = mode == 0 ? c->item_rec[e] : V.leaves[leaf];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ntiles = (L.count + PART_TILE - 1) / PART_TILE;
  // an empty leaf (possible on one rank of a data-parallel run: none of a globally splittable leaf's rows live in this
  // shard) owns no tiles: P = 0, block 0 is the control block and records left_count = 0
  const int tpb = max(1, (ntiles + maxp - 1) / maxp);  // tiles per block
  const int P = (ntiles + tpb - 1) / tpb;              // blocks that own tiles
  if (bid > P) return;
  const bool control = bid == P;
  unsigned long long* psync = V.part_sync + (size_t)e * V.part_sync_stride;
  unsigned* pbits = V.part_bits + (size_t)e * V.bits_stride;
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
    const DevFeat F = mode == 0 ? c->item_feat[e] : V.feats[feature];
    const SplitRule rule = make_rule(F, threshold, default_left);
    // The split feature is read from the column-major copy, whatever the size of the leaf: a 32-byte sector there holds
    // this feature of 32 neighbouring rows (several of them in the leaf when it is large), a sector of the row-major
    // matrix one row's 32 features of which one byte is wanted.  (Round 1 used the row-major matrix below 1/16 of the
    // rows "to leave the row in L2 for the histogram pass": measured, the histogram pass does not notice and the
    // partition is 30 % faster without -- 1.15 vs 1.64 ms per iteration at 10.5M rows, 8.3 vs 12.0 ms at 100M.)
    const uint8_t* col = V.bins_col + (size_t)F.col * V.col_stride;
    const uint8_t* col_hi = col + V.col_stride;   // the high byte of a wide feature (max_bin > 255)
    // a categorical split (DenseBin::SplitCategoricalInner, dense_bin.hpp:309-341): left = the row's BIN is in the set;
    // the set is the pool entry the leaf's best split names (stored value 0 = the most frequent bin, which is in the set
    // only if it is not bin 0)
    __shared__ unsigned s_cat[LGB_CAT_WORDS];
    const bool is_cat = WIDE && F.cat != 0;
    if (is_cat) {
      if (tid < LGB_CAT_WORDS) s_cat[tid] = V.cat_pool[(size_t)V.best[leaf].threshold * LGB_CAT_WORDS + tid];
      __syncthreads();
    }
    int lefts = 0;
    // the row indices of tile k+1 are in flight while the bins of tile k are gathered
    int nrow[PART_REPS];
#pragma unroll
    for (int r = 0; r < PART_REPS; ++r) {
      const int pos = t0 * PART_TILE + r * PART_THREADS + tid;
      nrow[r] = pos < L.count ? src[pos] : -1;
    }
    for (int tile = t0; tile < t1; ++tile) {
      const int lt = tile - t0;
      int row[PART_REPS];
#pragma unroll
      for (int r = 0; r < PART_REPS; ++r) row[r] = nrow[r];
      if (tile + 1 < t1) {
#pragma unroll
        for (int r = 0; r < PART_REPS; ++r) {
          const int pos = (tile + 1) * PART_TILE + r * PART_THREADS + tid;
          nrow[r] = pos < L.count ? src[pos] : -1;
        }
      }
      int bin[PART_REPS];
#pragma unroll
      for (int r = 0; r < PART_REPS; ++r) bin[r] = row[r] >= 0 ? col[row[r]] : 0;
      if (WIDE && F.wide) {
#pragma unroll
        for (int r = 0; r < PART_REPS; ++r) if (row[r] >= 0) bin[r] |= (int)col_hi[row[r]] << 8;
      }
#pragma unroll
      for (int r = 0; r < PART_REPS; ++r) {
        bool left;
        if (is_cat) {
          const int b = bin[r] == 0 ? F.mfb : bin[r] - 1 + F.offset;
          left = row[r] >= 0 && ((s_cat[b >> 5] >> (b & 31)) & 1u) != 0;
        } else {
          left = row[r] >= 0 && goes_left(rule, bin[r]);
        }
        if (lt == 0) row0[r] = row[r] >= 0 ? row[r] : 0;
        const unsigned word = __ballot_sync(0xffffffffu, left);
        if (lane == 0) {
          const int wi = r * (PART_THREADS / 32) + warp;
          if (lt < PART_SMEM_TILES) s_bits[lt][wi] = word;
          else pbits[(size_t)tile * (PART_TILE / 32) + wi] = word;
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
      *reinterpret_cast<volatile unsigned long long*>(&psync[bid]) = ((unsigned long long)epoch << 32) | (unsigned)t;
      atomicAdd(&c->bar_arrive[e], 1u);
    }
  }
  // one thread per block waits on the arrive counter (light polling); the words are then read once -- should one
  // not be visible yet (no fence orders it before the counter), its reader spins on the epoch tag
  if (tid == 0) {
    while (*reinterpret_cast<volatile unsigned*>(&c->bar_arrive[e]) - bar_base < (unsigned)P) { }
  }
  __syncthreads();
  // offsets: lefts of the blocks before this one, and of all blocks
  int before = 0, total = 0;
  for (int b = tid; b < P; b += PART_THREADS) {
    unsigned long long w;
    do { w = *reinterpret_cast<volatile unsigned long long*>(&psync[b]); } while ((unsigned)(w >> 32) != epoch);
    const int v = (int)(unsigned)w;
    total += v;
    if (b < bid) before += v;
  }
  PSTAGE(2);
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
#ifdef LGBCU_STAGE_TIMERS
  const int split_ = V.ctl->num_splits;  // (racy by one against the control block: good enough for a probe)
#endif
  if (control) {
    if (mode == 0) {  // snapshot FeatureHistogram::is_splittable_ of the leaf being split for k_find_split
      const int8_t* sp = V.splittable + (size_t)L.slot * V.num_features;
      for (int f = tid; f < V.num_features; f += PART_THREADS) V.parent_split[(size_t)e * V.num_features + f] = sp[f];
    }
    if (tid == 0) {
      if (mode == 1) part_control_stepwise(V, leaf, L, total);
      else part_expand(V, e, leaf, L, total);
      c->part_epoch[e] = epoch; c->bar_base[e] = bar_base + (unsigned)P;
    }
    return;
  }
  const int left_total = total;
  int tile_off = before;
  for (int tile = t0; tile < t1; ++tile) {
    const int lt = tile - t0;
    const unsigned myword = lt < PART_SMEM_TILES ? s_bits[lt][lane] : pbits[(size_t)tile * (PART_TILE / 32) + lane];
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
