// Split finding (replaces FeatureHistogram::FindBestThreshold / FindBestThresholdSequentially,
// src/treelearner/feature_histogram.hpp:85-95,156-217,853-1049; Dataset::FixHistogram, src/io/dataset.cpp:1410-1428;
// FeatureHistogram::Subtract, feature_histogram.hpp:79-83; the per-leaf argmax of
// SerialTreeLearner::FindBestSplitsFromHistograms, serial_tree_learner.cpp:358-425, and the leaf choice of
// SerialTreeLearner::Train, serial_tree_learner.cpp:181-199).
//
// One block of two warps per feature: warp 0 scans the smaller leaf, warp 1 forms larger = parent - smaller
// and scans it.  A warp holds the <=256 bins of a feature 8 per lane; the reference's sequential
// accumulation becomes an exact integer warp prefix/suffix scan (the histogram is 64-bit fixed point), each
// lane evaluates its 8 thresholds, and an argmax reduction applies the reference's tie-break (reverse scan
// keeps the highest threshold among equal gains, forward scan the lowest; reverse wins ties with forward).
// The last block to finish reduces over features (SplitInfo::operator>, split_info.hpp:126-153), then picks
// the leaf to split next and writes the split descriptor the partition kernels read -- no host involved.
#pragma once
#include "dev_types.cuh"

#define FULL_MASK 0xffffffffu
#define BINS_PER_LANE 8

__device__ __forceinline__ int sign_of(double x) { return (x > 0.0) - (x < 0.0); }  // Common::Sign

// FeatureHistogram::ThresholdL1 (feature_histogram.hpp:734-737)
__device__ __forceinline__ double threshold_l1(double s, double l1) {
  const double reg_s = fmax(0.0, fabs(s) - l1);
  return sign_of(s) * reg_s;
}

// CalculateSplittedLeafOutput<USE_L1,USE_MAX_OUTPUT,USE_SMOOTHING> (feature_histogram.hpp:739-761)
__device__ __forceinline__ double leaf_output(double g, double h, const DevParams& P, int n, double parent_output) {
  double ret = P.use_l1 ? -threshold_l1(g, P.lambda_l1) / (h + P.lambda_l2) : -g / (h + P.lambda_l2);
  if (P.use_max_output) {
    if (P.max_delta_step > 0 && fabs(ret) > P.max_delta_step) ret = sign_of(ret) * P.max_delta_step;
  }
  if (P.use_smoothing) {
    ret = ret * (n / P.path_smooth) / (n / P.path_smooth + 1) + parent_output / (n / P.path_smooth + 1);
  }
  return ret;
}

// GetLeafGain / GetLeafGainGivenOutput (feature_histogram.hpp:822-851)
__device__ __forceinline__ double leaf_gain(double g, double h, const DevParams& P, int n, double parent_output) {
  if (!P.use_max_output && !P.use_smoothing) {
    if (P.use_l1) {
      const double sg = threshold_l1(g, P.lambda_l1);
      return (sg * sg) / (h + P.lambda_l2);
    }
    return (g * g) / (h + P.lambda_l2);
  }
  const double out = leaf_output(g, h, P, n, parent_output);
  const double sg = P.use_l1 ? threshold_l1(g, P.lambda_l1) : g;
  return -(2.0 * sg * out + (h + P.lambda_l2) * out * out);
}

struct ScanBest {
  double gain, left_g, left_h;
  int t;           // scan position of the winner
  int left_count;
};

// argmax over lanes; prefer_high: among equal gains keep the larger t (reverse scan) else the smaller t
__device__ __forceinline__ ScanBest warp_best(ScanBest b, bool prefer_high) {
  for (int o = 16; o > 0; o >>= 1) {
    ScanBest x;
    x.gain = __shfl_down_sync(FULL_MASK, b.gain, o);
    x.left_g = __shfl_down_sync(FULL_MASK, b.left_g, o);
    x.left_h = __shfl_down_sync(FULL_MASK, b.left_h, o);
    x.t = __shfl_down_sync(FULL_MASK, b.t, o);
    x.left_count = __shfl_down_sync(FULL_MASK, b.left_count, o);
    const bool take = (x.gain > b.gain) ||
                      (x.gain == b.gain && x.gain != LGB_MIN_SCORE && (prefer_high ? x.t > b.t : x.t < b.t));
    if (take) b = x;
  }
  b.gain = __shfl_sync(FULL_MASK, b.gain, 0);
  b.left_g = __shfl_sync(FULL_MASK, b.left_g, 0);
  b.left_h = __shfl_sync(FULL_MASK, b.left_h, 0);
  b.t = __shfl_sync(FULL_MASK, b.t, 0);
  b.left_count = __shfl_sync(FULL_MASK, b.left_count, 0);
  return b;
}

// exclusive prefix (sum over lower lanes) of a 64-bit value
__device__ __forceinline__ i64 warp_excl_prefix(i64 v, int lane) {
  i64 inc = v;
  for (int o = 1; o < 32; o <<= 1) {
    const i64 x = __shfl_up_sync(FULL_MASK, inc, o);
    if (lane >= o) inc += x;
  }
  return inc - v;
}
__device__ __forceinline__ i64 warp_sum(i64 v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
  return v;
}

// One warp: FindBestThreshold for one numerical feature whose nt = num_bin - offset histogram entries
// data_[t] sit in registers (gi/hi, 8 consecutive t per lane).  Writes *out (all lanes hold the same value).
__device__ void scan_feature_warp(const i64 (&gi)[BINS_PER_LANE], const i64 (&hi)[BINS_PER_LANE], int nt,
                                  const DevFeat& F, double sum_g, double leaf_sum_h, int num_data,
                                  double parent_output, const DevParams& P, double inv_g, double inv_h,
                                  DevSplit* out, int* splittable_out) {
  const int lane = threadIdx.x & 31;
  const double sum_h = leaf_sum_h + (double)(2 * 1e-15f);  // FindBestThreshold: sum_hessian + 2 * kEpsilon
  const double cnt_factor = num_data / sum_h;
  const double min_gain_shift = leaf_gain(sum_g, sum_h, P, num_data, parent_output) + P.min_gain_to_split;
  const int offset = F.offset;
  const bool two_way = F.num_bin > 2 && F.missing_type != LGBCU_MISSING_NONE;
  const bool skip_default = two_way && F.missing_type == LGBCU_MISSING_ZERO;
  const bool na_missing = two_way && F.missing_type == LGBCU_MISSING_NAN;

  int cnt[BINS_PER_LANE];
#pragma unroll
  for (int k = 0; k < BINS_PER_LANE; ++k) {
    const int t = lane * BINS_PER_LANE + k;
    cnt[k] = (t < nt) ? round_int(((double)hi[k] * inv_h) * cnt_factor) : 0;
  }
  int splittable = 0;

  // ---------------- REVERSE: t from nt-1-NA down to 1-offset, right side accumulates ----------------
  ScanBest rb; rb.gain = LGB_MIN_SCORE; rb.t = -1000; rb.left_g = 0; rb.left_h = 0; rb.left_count = 0;
  {
    const int t_hi = nt - 1 - (na_missing ? 1 : 0), t_lo = 1 - offset;
    i64 lg = 0, lh = 0, lc = 0;
    bool inc[BINS_PER_LANE];
#pragma unroll
    for (int k = 0; k < BINS_PER_LANE; ++k) {
      const int t = lane * BINS_PER_LANE + k;
      inc[k] = t >= t_lo && t <= t_hi && !(skip_default && (t + offset) == F.default_bin);
      if (inc[k]) { lg += gi[k]; lh += hi[k]; lc += cnt[k]; }
    }
    // sums over higher lanes = total - inclusive prefix
    const i64 tg = warp_sum(lg), th = warp_sum(lh), tc = warp_sum(lc);
    i64 ag = tg - (warp_excl_prefix(lg, lane) + lg);
    i64 ah = th - (warp_excl_prefix(lh, lane) + lh);
    i64 ac = tc - (warp_excl_prefix(lc, lane) + lc);
#pragma unroll
    for (int k = BINS_PER_LANE - 1; k >= 0; --k) {
      if (!inc[k]) continue;
      ag += gi[k]; ah += hi[k]; ac += cnt[k];
      const double right_g = (double)ag * inv_g;
      const double right_h = LGB_EPS + (double)ah * inv_h;
      const int right_cnt = (int)ac;
      if (right_cnt < P.min_data_in_leaf || right_h < P.min_sum_hessian) continue;
      const int left_cnt = num_data - right_cnt;
      if (left_cnt < P.min_data_in_leaf) continue;  // the reference breaks: monotone, same set
      const double left_h = sum_h - right_h;
      if (left_h < P.min_sum_hessian) continue;
      const double left_g = sum_g - right_g;
      const double gain = leaf_gain(left_g, left_h, P, left_cnt, parent_output) +
                          leaf_gain(right_g, right_h, P, right_cnt, parent_output);
      if (gain <= min_gain_shift) continue;
      splittable = 1;
      if (gain > rb.gain) {
        rb.gain = gain; rb.left_g = left_g; rb.left_h = left_h; rb.left_count = left_cnt;
        rb.t = lane * BINS_PER_LANE + k;
      }
    }
    rb = warp_best(rb, true);
  }
  splittable = __any_sync(FULL_MASK, splittable);

  DevSplit o;
  o.gain = LGB_MIN_SCORE; o.feature = -1; o.real_feature = -1; o.threshold = 0; o.default_left = 1;
  o.left_count = 0; o.right_count = 0; o.left_output = 0; o.right_output = 0;
  o.left_sum_g = 0; o.left_sum_h = 0; o.right_sum_g = 0; o.right_sum_h = 0;
  double best_gain_sel = LGB_MIN_SCORE, sel_left_g = 0, sel_left_h = 0;
  int sel_left_cnt = 0;
  bool have = false;
  if (splittable && rb.gain > o.gain + min_gain_shift) {
    have = true;
    o.threshold = rb.t - 1 + offset;
    o.default_left = 1;
    best_gain_sel = rb.gain; sel_left_g = rb.left_g; sel_left_h = rb.left_h; sel_left_cnt = rb.left_count;
    o.gain = rb.gain - min_gain_shift;
  }

  // ---------------- FORWARD (only with a missing bin and > 2 bins): left side accumulates ----------------
  if (two_way) {
    ScanBest fb; fb.gain = LGB_MIN_SCORE; fb.t = 100000; fb.left_g = 0; fb.left_h = 0; fb.left_count = 0;
    const int t_end = nt - 2;
    double base_g = 0.0, base_h = LGB_EPS;
    int base_c = 0;
    const bool implicit_first = na_missing && offset == 1;
    if (implicit_first) {
      i64 sg_all = 0, sh_all = 0, sc_all = 0;
#pragma unroll
      for (int k = 0; k < BINS_PER_LANE; ++k) {
        if (lane * BINS_PER_LANE + k < nt) { sg_all += gi[k]; sh_all += hi[k]; sc_all += cnt[k]; }
      }
      sg_all = warp_sum(sg_all); sh_all = warp_sum(sh_all); sc_all = warp_sum(sc_all);
      base_g = sum_g - (double)sg_all * inv_g;
      base_h = (sum_h - LGB_EPS) - (double)sh_all * inv_h;
      base_c = num_data - (int)sc_all;
    }
    i64 lg = 0, lh = 0, lc = 0;
    bool inc[BINS_PER_LANE];
#pragma unroll
    for (int k = 0; k < BINS_PER_LANE; ++k) {
      const int t = lane * BINS_PER_LANE + k;
      inc[k] = t <= t_end && !(skip_default && (t + offset) == F.default_bin);
      if (inc[k]) { lg += gi[k]; lh += hi[k]; lc += cnt[k]; }
    }
    i64 ag = warp_excl_prefix(lg, lane), ah = warp_excl_prefix(lh, lane), ac = warp_excl_prefix(lc, lane);
    int fsplittable = 0;
    // k = -1 is the implicit-first-bin candidate (t = -1), evaluated by lane 0 only
#pragma unroll
    for (int k = -1; k < BINS_PER_LANE; ++k) {
      if (k < 0) {
        if (!(implicit_first && lane == 0 && t_end >= -1)) continue;
      } else {
        if (!inc[k]) continue;
        ag += gi[k]; ah += hi[k]; ac += cnt[k];
      }
      const double left_g = base_g + (double)ag * inv_g;
      const double left_h = base_h + (double)ah * inv_h;
      const int left_cnt = base_c + (int)ac;
      if (left_cnt < P.min_data_in_leaf || left_h < P.min_sum_hessian) continue;
      const int right_cnt = num_data - left_cnt;
      if (right_cnt < P.min_data_in_leaf) continue;
      const double right_h = sum_h - left_h;
      if (right_h < P.min_sum_hessian) continue;
      const double right_g = sum_g - left_g;
      const double gain = leaf_gain(left_g, left_h, P, left_cnt, parent_output) +
                          leaf_gain(right_g, right_h, P, right_cnt, parent_output);
      if (gain <= min_gain_shift) continue;
      fsplittable = 1;
      if (gain > fb.gain) {
        fb.gain = gain; fb.left_g = left_g; fb.left_h = left_h; fb.left_count = left_cnt;
        fb.t = (k < 0) ? -1 : lane * BINS_PER_LANE + k;
      }
    }
    fb = warp_best(fb, false);
    if (__any_sync(FULL_MASK, fsplittable)) splittable = 1;
    if (splittable && fb.gain > o.gain + min_gain_shift) {
      have = true;
      o.threshold = fb.t + offset;
      o.default_left = 0;
      best_gain_sel = fb.gain; sel_left_g = fb.left_g; sel_left_h = fb.left_h; sel_left_cnt = fb.left_count;
      o.gain = fb.gain - min_gain_shift;
    }
  }
  if (have) {
    o.left_output = leaf_output(sel_left_g, sel_left_h, P, sel_left_cnt, parent_output);
    o.left_count = sel_left_cnt;
    o.left_sum_g = sel_left_g;
    o.left_sum_h = sel_left_h - LGB_EPS;
    o.right_output = leaf_output(sum_g - sel_left_g, sum_h - sel_left_h, P, num_data - sel_left_cnt, parent_output);
    o.right_count = num_data - sel_left_cnt;
    o.right_sum_g = sum_g - sel_left_g;
    o.right_sum_h = sum_h - sel_left_h - LGB_EPS;
  }
  (void)best_gain_sel;
  // a 2-bin (or <=2 bin) feature whose mapper carries a NaN bin: FuncForNumricalL3's last branch
  if (!two_way && F.missing_type == LGBCU_MISSING_NAN) o.default_left = 0;
  o.gain *= F.penalty;  // FindBestThreshold: output->gain *= meta_->penalty
  *out = o;
  *splittable_out = splittable;
}

// ComputeBestSplitForFeature (serial_tree_learner.cpp:709-718): the reference tests `leaf_index() == 0`, which holds
// for the root and for every later leaf that still carries id 0 (a left child keeps its parent's id): such a leaf's
// "parent output" is its own unsmoothed output; every other leaf uses the output its parent's split assigned.
__device__ __forceinline__ double parent_output_of(const DevLeaf& L, int leaf, const DevParams& P) {
  if (leaf != 0) return L.output;
  DevParams P0 = P;
  P0.use_smoothing = 0;
  return leaf_output(L.sum_g, L.sum_h, P0, L.gcount, 0.0);
}

// SplitInfo::operator> : larger gain, ties -> smaller real feature index (-1 counts as INT_MAX)
__device__ __forceinline__ bool split_better(double ga, int fa, double gb, int fb) {
  if (fa == -1) fa = 0x7fffffff;
  if (fb == -1) fb = 0x7fffffff;
  if (ga != gb) return ga > gb;
  return fa < fb;
}

// mode 0: training (smaller/larger from DevCtl, control step at the end);
// mode 1: explicit single leaf, no control (step-wise API / tests)
__global__ void __launch_bounds__(64) k_find_split(DevView V, int mode, int leaf_arg) {
  __shared__ int s_last;
  DevCtl* c = V.ctl;
  if (mode == 0 && c->done) return;
  const int f = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool active = (mode == 1) || !c->skip_find;
  const int smaller = (mode == 1) ? leaf_arg : c->smaller;
  const int larger = (mode == 1) ? -1 : c->larger;

  if (active && f < V.num_features) {
    const DevFeat F = V.feats[f];
    const DevLeaf Ls = V.leaves[smaller];
    int used = V.feat_used[f] != 0;
    if (larger >= 0) used = used && V.splittable[(size_t)V.leaves[larger].slot * V.num_features + f] != 0;
    const int nt = F.num_bin - F.offset;
    const double inv_g = c->inv_g, inv_h = c->inv_h;
    HistEntry* hs = V.pool + (size_t)Ls.slot * V.total_bin + F.hist_base;  // slot 0 of the group
    i64 gi[BINS_PER_LANE], hi[BINS_PER_LANE];
    if (warp == 0) {
      DevSplit o; int sp = 0;
      if (!used) {
        o.gain = LGB_MIN_SCORE; o.feature = -1; o.real_feature = -1;
      } else {
        // load data_[t] = slot 1+t ; FixHistogram: the most-frequent bin's entry is what stored-value 0 collected
        const HistEntry e0 = hs[0];
#pragma unroll
        for (int k = 0; k < BINS_PER_LANE; ++k) {
          const int t = lane * BINS_PER_LANE + k;
          HistEntry e; e.g = 0; e.h = 0;
          if (t < nt) {
            if (F.mfb > 0 && t == F.mfb) { e = e0; hs[1 + t] = e0; }
            else e = hs[1 + t];
          }
          gi[k] = e.g; hi[k] = e.h;
        }
        const double parent_output = parent_output_of(Ls, smaller, V.p);
        scan_feature_warp(gi, hi, nt, F, Ls.sum_g, Ls.sum_h, Ls.gcount, parent_output, V.p, inv_g, inv_h, &o, &sp);
        o.feature = f; o.real_feature = F.real_index;
      }
      if (lane == 0) {
        V.feat_best[f] = o;
        V.splittable[(size_t)Ls.slot * V.num_features + f] = (int8_t)sp;
      }
    }
    __syncthreads();
    if (warp == 1 && larger >= 0) {
      const DevLeaf Ll = V.leaves[larger];
      DevSplit o; int sp = 0;
      if (!used) {
        o.gain = LGB_MIN_SCORE; o.feature = -1; o.real_feature = -1;
      } else {
        HistEntry* hl = V.pool + (size_t)Ll.slot * V.total_bin + F.hist_base;  // holds the parent
#pragma unroll
        for (int k = 0; k < BINS_PER_LANE; ++k) {
          const int t = lane * BINS_PER_LANE + k;
          HistEntry e; e.g = 0; e.h = 0;
          if (t < nt) {
            const HistEntry a = hl[1 + t], b = hs[1 + t];
            e.g = a.g - b.g; e.h = a.h - b.h;
            hl[1 + t] = e;
          }
          gi[k] = e.g; hi[k] = e.h;
        }
        if (lane == 0) {  // the group's slot 0 (rows in the most-frequent bin)
          const HistEntry a = hl[0], b = hs[0];
          HistEntry e; e.g = a.g - b.g; e.h = a.h - b.h;
          hl[0] = e;
        }
        scan_feature_warp(gi, hi, nt, F, Ll.sum_g, Ll.sum_h, Ll.gcount, parent_output_of(Ll, larger, V.p), V.p, inv_g, inv_h, &o, &sp);
        o.feature = f; o.real_feature = F.real_index;
      }
      if (lane == 0) {
        V.feat_best[V.num_features + f] = o;
        if (used) V.splittable[(size_t)Ll.slot * V.num_features + f] = (int8_t)sp;
      }
    }
  }

  // ---- last block: reduce over features, then choose the next leaf to split ----
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned ticket = atomicAdd(&c->feat_done, 1u);
    s_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (warp != 0) return;
  if (active) {
    for (int sel = 0; sel < 2; ++sel) {
      const int leaf = sel == 0 ? smaller : larger;
      if (leaf < 0) continue;
      double bg = LGB_MIN_SCORE; int bf = -1, bi = -1;
      for (int j = lane; j < V.num_features; j += 32) {
        const DevSplit* s = &V.feat_best[sel * V.num_features + j];
        const double g = s->gain; const int rf = s->real_feature;
        if (bi < 0 || split_better(g, rf, bg, bf)) { bg = g; bf = rf; bi = j; }
      }
      for (int o = 16; o > 0; o >>= 1) {
        const double g2 = __shfl_down_sync(FULL_MASK, bg, o);
        const int f2 = __shfl_down_sync(FULL_MASK, bf, o);
        const int i2 = __shfl_down_sync(FULL_MASK, bi, o);
        if (i2 >= 0 && (bi < 0 || split_better(g2, f2, bg, bf))) { bg = g2; bf = f2; bi = i2; }
      }
      if (lane == 0) {
        DevSplit r;
        if (bi >= 0) r = V.feat_best[sel * V.num_features + bi];
        else { r.gain = LGB_MIN_SCORE; r.feature = -1; r.real_feature = -1; }
        V.best[leaf] = r;
      }
    }
  }
  __syncwarp();
  if (lane == 0) c->feat_done = 0;
  if (mode != 0) return;
  __threadfence_block();
  // ArrayArgs<SplitInfo>::ArgMax(best_split_per_leaf_): first leaf with the best (gain, feature)
  {
    const int nl = c->num_leaves;
    double bg = LGB_MIN_SCORE; int bf = -1, bl = -1;
    for (int l = lane; l < nl; l += 32) {
      const double g = V.best[l].gain; const int rf = V.best[l].real_feature;
      if (bl < 0 || split_better(g, rf, bg, bf)) { bg = g; bf = rf; bl = l; }
    }
    for (int o = 16; o > 0; o >>= 1) {
      const double g2 = __shfl_down_sync(FULL_MASK, bg, o);
      const int f2 = __shfl_down_sync(FULL_MASK, bf, o);
      const int l2 = __shfl_down_sync(FULL_MASK, bl, o);
      if (l2 >= 0) {
        if (bl < 0 || split_better(g2, f2, bg, bf) || (g2 == bg && f2 == bf && l2 < bl)) { bg = g2; bf = f2; bl = l2; }
      }
    }
    if (lane == 0) {
      if (bl < 0 || !(bg > 0.0)) {
        c->done = 1;  // "No further splits with positive gain"
      } else {
        const DevSplit s = V.best[bl];
        c->split_leaf = bl;
        c->split_feature = s.feature;
        c->split_threshold = s.threshold;
        c->split_default_left = s.default_left;
        c->split_right_leaf = nl;
      }
    }
  }
}

// Root LeafSplits::Init (leaf_splits.hpp:66-79; all-reduced over ranks in data_parallel_tree_learner.cpp:126-151):
// sum over every stored value of feature 0 of the (already rank-reduced) root histogram -- exact in fixed point.
__global__ void k_root_sums(DevView V) {
  const DevCtl* c = V.ctl;
  DevLeaf* L = &V.leaves[0];
  const DevFeat F = V.feats[0];
  const HistEntry* h = V.pool + (size_t)L->slot * V.total_bin + F.hist_base;
  i64 g = 0, hh = 0;
  for (int s = threadIdx.x; s < F.nslots; s += 32) { g += h[s].g; hh += h[s].h; }
  g = warp_sum(g); hh = warp_sum(hh);
  if (threadIdx.x == 0) {
    L->sum_g = (double)g * c->inv_g;
    L->sum_h = (double)hh * c->inv_h;
  }
}
