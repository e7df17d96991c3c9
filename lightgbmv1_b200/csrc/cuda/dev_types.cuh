// Device-side state of the B200 tree learner.  Everything a tree needs while it grows lives in HBM; the
// host only enqueues a fixed kernel sequence per split and reads the split records back once per tree.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "lgbcu.h"

// reference constants (include/LightGBM/meta.h:52-56): float literals widened to double
#define LGB_EPS        ((double)1e-15f)
#define LGB_ZERO_THR   ((double)1e-35f)
#define LGB_MIN_SCORE  (-__longlong_as_double(0x7ff0000000000000LL))  /* -inf */

typedef long long i64;

// One histogram entry: (sum_gradient, sum_hessian) in 64-bit fixed point.  Position in a leaf's buffer is
// identical to the reference's hist_t pair (include/LightGBM/bin.h:32-42).
struct __align__(16) HistEntry { i64 g; i64 h; };

struct DevFeat {
  int num_bin;       // BinMapper::num_bin
  int offset;        // FeatureMetainfo::offset = [most_freq_bin == 0]
  int mfb;           // most_freq_bin
  int default_bin;
  int missing_type;  // LGBCU_MISSING_*
  int hist_base;     // first slot of the feature's (single-feature) group = group_bin_boundaries_[g]
  int nslots;        // FeatureGroup::num_total_bin_ = 1 + num_bin - offset
  int real_index;
  // Storage: a feature with at most 256 stored values (nslots <= 256) is one byte column of the matrix; a WIDE feature
  // (max_bin > 255: the reference's 16-bit DenseBin, src/io/bin.cpp:635-645) is two adjacent byte columns, low byte
  // first.  `col` = the (first) byte column.  The matrices stay byte matrices, so the kernels of the 8-bit hot path are
  // untouched; data with a wide feature takes the general histogram and split-search paths.
  int col, wide;
  // cat = 1: a categorical feature (BinType::CategoricalBin): bin 0 = NaN / negative / rare categories, a split sends a
  // SET of bins left (a 256-bit set: categorical features are limited to 256 bins here)
  int cat, pad_;
  double penalty;
};
#define LGB_CAT_WORDS 8   /* words of a categorical split's bin set */

// stored value of feature F in a row-major row / in the column-major copy
__device__ __forceinline__ int stored_bin_row(const uint8_t* __restrict__ row, const DevFeat& F) {
  return F.wide ? (int)row[F.col] | ((int)row[F.col + 1] << 8) : (int)row[F.col];
}

struct DevSplit {    // SplitInfo, POD
  double gain;
  int feature;       // inner
  int real_feature;
  int threshold;
  int default_left;
  int left_count;
  int right_count;
  double left_output, right_output;
  double left_sum_g, left_sum_h, right_sum_g, right_sum_h;
};

struct __align__(16) SplitKey {  // what SplitInfo::operator> compares: compact twin of a DevSplit for arg-max passes
  double gain;
  int real_feature;
  int pad;
};

struct DevLeaf {     // DataPartition range + LeafSplits + pool slot of one leaf
  int begin, count;  // range in indices[buf]
  int slot;          // histogram pool slot
  int depth;
  int buf;           // which ping-pong index buffer holds this leaf's rows
  int gcount;        // rows of this leaf over ALL ranks (== count on one GPU); what the split search calls num_data
  double sum_g, sum_h;  // LeafSplits::sum_gradients/sum_hessians (from the parent's SplitInfo)
  double output;        // LeafSplits::weight (leaf output before shrinkage)
  // monotone constraints, basic method (BasicLeafConstraints, monotone_constraints.hpp:88-126): the interval the
  // leaf's output must stay in; inherited by the children, tightened by a split on a constrained feature
  double mc_min, mc_max;
};
#define LGB_MC_INF 1.7976931348623157e308   /* std::numeric_limits<double>::max() */

// Look-ahead growth.  The reference grows a tree best-first: split the open leaf with the largest gain, learn the
// gains of its two children, repeat (serial_tree_learner.cpp:181-199).  Which leaves get split, and in which order, is
// a function of the gains alone -- so the children of a leaf can be computed ("expanded": partition + histogram +
// split search) BEFORE its turn comes without changing the result.  Every round expands the up to K open leaves with
// the largest gains side by side (the kernels of a round are latency bound, so K leaves cost about as much as one),
// then commits, in the reference's order, every leaf whose turn has come and whose children are ready.  An expansion
// is only wasted if the leaf budget runs out before the leaf's turn.
#define LGB_KMAX 16
#define LGB_MAX_PEERS 8   /* ranks of a one-shot exchange (one NVLink box) */

struct DevSpec {     // an expanded leaf that is not committed yet: what Split + two FindBestThreshold runs produced
  DevLeaf child[2];  // [0] = left (keeps the parent's leaf id when committed), [1] = right (gets the next free id)
  DevSplit best[2];  // best split of each child
  SplitKey key[2];
  int left_count;    // rows that went left on this rank
  int smaller;       // which child got the freshly built histogram (the other one = parent - smaller)
  int skip_find;     // BeforeFindBestSplit said no: the children cannot be split
  int pad;
};

struct DevCtl {
  int num_leaves;    // leaves so far in the tree being grown
  int done;          // 1 once no leaf has positive gain / num_leaves reached
  int smaller, larger;  // step-wise API only: LeafSplits pair of the last lgbcu_partition
  int skip_find;     // BeforeFindBestSplit returned false (root round / step-wise API)
  int lookahead;     // K: leaves expanded per round (1..LGB_KMAX)
  int round;         // rounds run so far in this tree; round 0 = the root (histogram of leaf 0, search into best[0])
  int num_items;     // leaves being expanded in the current round
  int item_leaf[LGB_KMAX], item_feature[LGB_KMAX], item_threshold[LGB_KMAX], item_default_left[LGB_KMAX];
  int split_leaf, split_feature, split_threshold, split_default_left, split_right_leaf;
  int split_left_count;
  int num_splits;
  unsigned max_g_bits, max_h_bits;  // max |g|, |h| as float bits
  int root_identity; // root rows are 0..N-1 in order (no bagging): skip the index gather
  int num_data;      // rows in the root
  int kg, kh;        // fixed-point exponents
  double scale_g, scale_h, inv_g, inv_h;
  unsigned feat_done;  // "last block done" counter of k_find_split
  int cat_count;       // entries of the categorical bin-set pool taken by the tree being grown
  int cat_overflow;    // 1: the pool ran out (reported by lgbcu_train_tree)
  unsigned part_epoch[LGB_KMAX];  // per item: launches of k_partition so far (tags the words of V.part_sync); never reset
  unsigned bar_arrive[LGB_KMAX], bar_base[LGB_KMAX];  // k_partition, per item: blocks arrived so far (monotonic) / value when the launch began
  long long global_root;  // root rows over all ranks   } adjacent: one 2 x int64 sum all-reduce
  long long global_rows;  // dataset rows over all ranks }
  unsigned long long prof_hist_rows, prof_hist_rows_indexed, prof_hist_calls, prof_hist_ns;
  // what k_hist builds in the current round, written by k_partition (part_expand): per item the rows of the smaller
  // child (0 = nothing to build) and the child itself -- one dependent load for k_hist's preamble instead of the
  // item_leaf -> spec -> child chain
  unsigned long long xchg_epoch;  // one-shot exchange: rounds this rank has published (monotonic over the booster's life)
  int xchg_error;                 // a peer's flag did not arrive in time: the tree is void (checked by the host)
  unsigned xchg_done;             // "last block" counter of k_xchg_pull
  int hist_rows[LGB_KMAX];
  DevLeaf hist_leaf[LGB_KMAX];
  // what k_partition splits in the current round, written by the control step of k_find_split after its commits:
  // the items' leaves and the descriptors of their split features (again one dependent load instead of a chain)
  DevLeaf item_rec[LGB_KMAX];
  DevFeat item_feat[LGB_KMAX];
};

// Feature ownership (data-parallel, wide data): what a rank sends to the all-gather per (leaf being expanded, child):
// the best split over ITS features, with the key that SplitInfo::operator> compares.
struct __align__(16) CandRec { DevSplit split; SplitKey key; };

struct DevCatParams {   // Config: cat_smooth, cat_l2, max_cat_threshold, min_data_per_group, max_cat_to_onehot
  double cat_smooth, cat_l2;
  int max_cat_threshold, min_data_per_group, max_cat_to_onehot, pad;
};

struct DevParams {   // LgbCuTreeParams + derived switches
  int num_leaves, max_depth, min_data_in_leaf;
  int use_l1, use_max_output, use_smoothing;
  double min_sum_hessian, lambda_l1, lambda_l2, min_gain_to_split, max_delta_step, path_smooth;
};

// Everything the kernels need, passed by value (fits in the 4 KB parameter space with room to spare).
struct DevView {
  int num_data, num_features, row_stride, total_bin, num_tiles16;
  int num_storage, any_wide;   // byte columns of the matrices (= num_features unless a feature is wide)
  int data_parallel;         // 1: histograms are all-reduced across ranks, leaf counts are the global estimates
  int f_begin, f_end;        // features this rank scans in k_find_split: all of them, or its block in ownership mode
  int comm_size;             // ranks (ownership mode: rows of cand_all per (leaf, child))
  // One-shot exchange (data-parallel, comm.cuh): k_hist adds into pool_local; k_xchg_pull publishes the round, waits
  // for the peers and sums the ranks' local slots over NVLink into `pool`; no NCCL collective on the histogram.
  int xchg, comm_rank;
  HistEntry* pool_local;                      // this rank's local sums, slot for slot like `pool`
  HistEntry* hist_pool;                       // where k_hist adds and k_partition zeroes: pool_local in that mode, else pool
                                              // (one pointer, chosen on the host: a select in k_hist cost it a register spill)
  const HistEntry* xpool[LGB_MAX_PEERS];      // rank r's pool_local as mapped here (own entry = pool_local)
  unsigned long long* xflags;                 // [comm_size] epoch rank r has published up to (written by rank r)
  CandRec* cand_send;        // [2 * LGB_KMAX]              ownership mode: this rank's winners of the round
  CandRec* cand_all;         // [comm_size][2 * LGB_KMAX]   every rank's, after the all-gather
  const uint8_t* bins;
  const uint8_t* bins_col;   // column-major copy [F][col_stride]: the partition reads ONE feature of a leaf's rows -- a
  long long col_stride;      // sector of this copy holds that feature of 32 neighbouring rows, a sector of `bins` of one
  const DevFeat* feats;
  const int8_t* feat_used;   // ColSampler by-tree mask
  int8_t* splittable;        // [num_leaves slots][num_features]  FeatureHistogram::is_splittable_
  int8_t* parent_split;      // [LGB_KMAX][num_features] the expanded leaves' flags, snapshotted by k_partition for their children
  DevSpec* spec;             // [num_leaves] expansion results, indexed by the expanded leaf's id
  int* expanded;             // [num_leaves] 1 = spec[leaf] holds a finished expansion that is not committed yet
  int pool_slots;            // histogram slots in the pool
  long long bits_stride;     // words of part_bits per item
  double2* mfb_fix;          // [num_leaves slots][num_features] FixHistogram'd most-frequent-bin entry, in double
  const float2* gh;          // [num_data] interleaved (gradient, hessian) of the current iteration, written by k_absmax: 84 MB at
                             // 10.5M rows, so it stays L2-resident while a tree grows; converted to fixed point when a row is staged
  int* indices[2];
  HistEntry* pool;           // [num_leaves][total_bin]
  DevLeaf* leaves;
  DevSplit* best;            // best_split_per_leaf_
  DevSplit* feat_best;       // [2 * LGB_KMAX][num_features] per-feature candidates of (smaller, larger) of every item
  SplitKey* feat_key;        // [2 * LGB_KMAX][num_features] (gain, real feature) of feat_best, 16 B each: coalesced arg-max
  SplitKey* leaf_key;        // [num_leaves]      (gain, real feature) of best
  DevCtl* ctl;
  LgbCuSplitRecord* records;
  unsigned long long* part_sync;  // partition scratch: [LGB_KMAX][part_sync_stride] (epoch << 32 | left count) per block
  int part_sync_stride;
  unsigned* part_bits;       // partition scratch: 1 bit per row (1 = goes left)
  DevParams p;
  // categorical splits: an append-only pool of bin sets per tree.  A categorical candidate takes the next entry
  // (ctl->cat_count) and carries its index in DevSplit::threshold; partition, records and the host read the set there.
  int any_cat, cat_cap;
  int use_mc;                     // Config::monotone_constraints is not empty (USE_MC of feature_histogram.hpp)
  const signed char* monotone;    // [num_features] FeatureMetainfo::monotone_type
  double mc_penalty;              // Config::monotone_penalty
  unsigned* cat_pool;        // [cat_cap][LGB_CAT_WORDS]
  DevCatParams cat;
};

#ifdef LGBCU_STAGE_TIMERS
__device__ unsigned long long g_stamps[3 * 4096];
__device__ __forceinline__ void stamp(const DevCtl* c, int k) {
  // indexed by the round of tree growth (the control step of k_find_split has already bumped it when k_partition runs)
  const int r = c->round - (k == 2 ? 1 : 0);
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && r >= 0 && r < 4096 && !c->done) {
    unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_stamps[3 * r + k] = t;
  }
}
__device__ unsigned long long g_ends[3 * 4096];
__device__ __forceinline__ void stamp_end(int split, int k) {  // every block: latest block exit
  if (threadIdx.x == 0 && split < 4096) {
    unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    atomicMax(&g_ends[3 * split + k], t);
  }
}
#define STAMP(c, k) stamp(c, k)
#define STAMP_END(split, k) stamp_end(split, k)
#else
#define STAMP(c, k) do {} while (0)
#define STAMP_END(split, k) do {} while (0)
#endif

__device__ __forceinline__ int round_int(double x) { return (int)(x + 0.5f); }  // Common::RoundInt
