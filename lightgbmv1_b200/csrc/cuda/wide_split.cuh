// The GENERAL split search: wide features (more than 256 stored values, i.e. max_bin > 255: the reference's 16-bit
// DenseBin, src/io/bin.cpp:635-645), categorical features, and every feature of data under monotone constraints.
// Included by split_kernels.cuh in front of find_item, which hands such features over (only in the <true>
// instantiation of k_find_split: the 8-bit kernel carries none of this).
//
// The 8-bit search (scan_feature_block) is built around one thread per bin of a <= 256-bin feature and the closed-form
// gain.  Here the block forms larger = parent - smaller and the FixHistogram residue cooperatively, then scans the bins
// as FeatureHistogram::FindBestThresholdSequentially does (feature_histogram.hpp:853-1049) in chunks of 256 bins --
// sequential running sums, parallel gains (wide_scan_block) -- reading the 64-bit fixed-point entries from the pool and
// converting each to double (exact: an entry is an integer multiple of a power of two); a categorical feature is walked
// by one thread (cat_find_best_threshold).  Same outputs as find_item: feat_best / feat_key rows, splittable flags,
// mfb_fix.
#pragma once

struct WideData {   // data_[t] of FeatureHistogram for one (feature, leaf): entries of the pool + the fixed mfb entry
  const HistEntry* h;   // slot entries of the feature's group: h[1 + t] = data_[t]
  int mfb; double fix_g, fix_h, inv_g, inv_h;
  __device__ __forceinline__ double g(int t) const { return t == mfb ? fix_g : (double)h[1 + t].g * inv_g; }
  __device__ __forceinline__ double hh(int t) const { return t == mfb ? fix_h : (double)h[1 + t].h * inv_h; }
};


// Monotone constraints (USE_MC of feature_histogram.hpp:763-824).  `MC` = {use, min, max, type}: with use = 0 these are
// the plain forms.
struct McInfo { int use; double lo, hi; int type; };
__device__ __forceinline__ double mc_output(double g, double h, const DevParams& P, int n, double parent_output, const McInfo& M) {
  double ret = leaf_output(g, h, P, n, parent_output);
  if (M.use) { if (ret < M.lo) ret = M.lo; else if (ret > M.hi) ret = M.hi; }   // CalculateSplittedLeafOutput<USE_MC>
  return ret;
}
__device__ __forceinline__ double gain_given_output(double g, double h, const DevParams& P, double out) {   // GetLeafGainGivenOutput
  const double sg = P.use_l1 ? threshold_l1(g, P.lambda_l1) : g;
  return -(2.0 * sg * out + (h + P.lambda_l2) * out * out);
}
// GetSplitGains (feature_histogram.hpp:781-824)
__device__ __forceinline__ double mc_split_gains(double lg, double lh, double rg, double rh, const DevParams& P, int lc, int rc,
                                                 double parent_output, const McInfo& M) {
  if (!M.use) return leaf_gain(lg, lh, P, lc, parent_output) + leaf_gain(rg, rh, P, rc, parent_output);
  const double lo = mc_output(lg, lh, P, lc, parent_output, M);
  const double ro = mc_output(rg, rh, P, rc, parent_output, M);
  if ((M.type > 0 && lo > ro) || (M.type < 0 && lo < ro)) return 0;
  return gain_given_output(lg, lh, P, lo) + gain_given_output(rg, rh, P, ro);
}

struct WideBest { double gain, left_g, left_h; int left_count, threshold; bool found; };

// ---- FindBestThresholdSequentially (feature_histogram.hpp:853-1049) by the whole block ------------------------------
// One direction of the reference's scan -- reverse (high to low bins, the accumulated side is the right child), or
// forward where the feature has a missing bin -- has three parts per bin: the running sums (a dependent chain of additions -- their order
// fixes the rounding, so they stay sequential), the gain of the threshold (independent of the other thresholds), and the
// "first best wins" selection.  Per chunk of FS_THREADS bins in scan order:
//   1. thread j loads and converts bin j (g, h, RoundInt(h * cnt_factor));
//   2. three threads run the three chains (gradient, hessian, count) over the chunk from shared memory, each carrying
//      its running value from chunk to chunk in a register;
//   3. thread j evaluates threshold j from the prefix values -- the reference's `continue` / `break` tests included: an
//      iteration that breaks ends the scan, so a block-wide minimum of the breaking positions bounds the candidates;
//   4. a block arg-max (gain, then the earlier position: the reference keeps the first of equal gains) meets the best
//      of the chunks before it under the same strict ">".
// Same doubles in the same order as a one-thread walk of the reference's loop (which this replaced: 14.5 -> 2.9 ms of
// split search per iteration at 1023 bins x 28 features, 5.9 -> 1.3 ms for 255-bin features under monotone constraints).
struct WideScratch {
  double vg[FS_THREADS], vh[FS_THREADS];
  int vc[FS_THREADS];
  signed char skip[FS_THREADS];
  double wgain[FS_WARPS]; int wpos[FS_WARPS];
  int first_break, any_split;
};

template <bool REVERSE>
__device__ WideBest wide_scan_block(WideScratch& W, const WideData& D, const DevFeat& F, const DevParams& P, bool skip_default,
                                    bool na_missing, double sum_g, double sum_h, int num_data, double min_gain_shift,
                                    double parent_output, const McInfo& M, int* splittable) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int offset = F.offset;
  WideBest B; B.gain = LGB_MIN_SCORE; B.left_g = 0; B.left_h = 0; B.left_count = 0; B.threshold = F.num_bin; B.found = false;
  const double cnt_factor = num_data / sum_h;
  // scan positions p = 0 .. n-1 <-> bins t(p); forward scans with an implicit first bin start at t = -1
  int t0, n;
  bool implicit_first = false;
  if (REVERSE) { t0 = F.num_bin - 1 - offset - (na_missing ? 1 : 0); n = t0 - (1 - offset) + 1; }
  else { implicit_first = na_missing && offset == 1; t0 = implicit_first ? -1 : 0; n = (F.num_bin - 2 - offset) - t0 + 1; }
  // the chains' running values: thread 0 gradient, thread 32 hessian, thread 64 count
  double run = 0.0; int run_c = 0;
  if (tid == 32) run = LGB_EPS;
  if (!REVERSE && implicit_first) {
    // everything minus the stored bins (feature_histogram.hpp:951-964): three sequential chains over all bins
    const int nb = F.num_bin - offset;
    if (tid == 0) { run = sum_g; for (int i = 0; i < nb; ++i) run -= D.g(i); }
    if (tid == 32) { run = sum_h - LGB_EPS; for (int i = 0; i < nb; ++i) run -= D.hh(i); }
    if (tid == 64) { run_c = num_data; for (int i = 0; i < nb; ++i) run_c -= round_int(D.hh(i) * cnt_factor); }
  }
  if (tid == 0) { W.first_break = 0x7fffffff; W.any_split = 0; }
  __syncthreads();
  for (int base = 0; base < n; base += FS_THREADS) {
    const int m = min(FS_THREADS, n - base);
    // 1. load
    if (tid < m) {
      const int t = REVERSE ? t0 - (base + tid) : t0 + (base + tid);
      const bool sk = skip_default && (t + offset) == F.default_bin;
      W.skip[tid] = sk ? 1 : 0;
      double g = 0.0, h = 0.0; int c = 0;
      if (!sk && t >= 0) { g = D.g(t); h = D.hh(t); c = round_int(h * cnt_factor); }
      W.skip[tid] |= (t < 0) ? 2 : 0;   // the implicit first bin adds nothing
      W.vg[tid] = g; W.vh[tid] = h; W.vc[tid] = c;
    }
    __syncthreads();
    // 2. the three chains (a skipped bin leaves the running value as it is)
    if (tid == 0) { for (int j = 0; j < m; ++j) { if (!W.skip[j]) run += W.vg[j]; W.vg[j] = run; } }
    if (tid == 32) { for (int j = 0; j < m; ++j) { if (!W.skip[j]) run += W.vh[j]; W.vh[j] = run; } }
    if (tid == 64) { for (int j = 0; j < m; ++j) { if (!W.skip[j]) run_c += W.vc[j]; W.vc[j] = run_c; } }
    __syncthreads();
    // 3. evaluate
    double gain = LGB_MIN_SCORE; bool cand = false, brk = false;
    if (tid < m && !(W.skip[tid] & 1)) {
      const double ag = W.vg[tid], ah = W.vh[tid]; const int ac = W.vc[tid];   // accumulated side
      if (!(ac < P.min_data_in_leaf || ah < P.min_sum_hessian)) {
        const int oc = num_data - ac; const double oh = sum_h - ah;
        if (oc < P.min_data_in_leaf || oh < P.min_sum_hessian) brk = true;
        else {
          const double og = sum_g - ag;
          // GetSplitGains takes (left, right): the accumulated side is the right one in the reverse scan
          gain = REVERSE ? mc_split_gains(og, oh, ag, ah, P, oc, ac, parent_output, M)
                         : mc_split_gains(ag, ah, og, oh, P, ac, oc, parent_output, M);
          cand = gain > min_gain_shift;
        }
      }
    }
    if (brk) atomicMin(&W.first_break, base + tid);
    __syncthreads();
    const int fb = W.first_break;
    cand = cand && (base + tid) < fb;
    if (cand) W.any_split = 1;
    // 4. block arg-max: larger gain, then the earlier position
    const int w = warp_argmax(gain, (unsigned)tid, 0u, cand);
    if (lane == 0) { W.wpos[warp] = -1; W.wgain[warp] = LGB_MIN_SCORE; }
    __syncwarp();
    if (w >= 0 && lane == w) { W.wpos[warp] = tid; W.wgain[warp] = gain; }
    __syncthreads();
    int bp = -1; double bg = LGB_MIN_SCORE;
#pragma unroll
    for (int k = 0; k < FS_WARPS; ++k) if (W.wpos[k] >= 0 && (bp < 0 || W.wgain[k] > bg)) { bp = W.wpos[k]; bg = W.wgain[k]; }
    if (bp >= 0 && bg > B.gain) {
      const double ag = W.vg[bp], ah = W.vh[bp]; const int ac = W.vc[bp];
      const int t = REVERSE ? t0 - (base + bp) : t0 + (base + bp);
      if (REVERSE) { B.left_count = num_data - ac; B.left_g = sum_g - ag; B.left_h = sum_h - ah; B.threshold = t - 1 + offset; }
      else { B.left_count = ac; B.left_g = ag; B.left_h = ah; B.threshold = t + offset; }
      B.gain = bg; B.found = true;
    }
    __syncthreads();   // the scratch is rewritten by the next chunk
    if (fb != 0x7fffffff) break;
  }
  if (W.any_split) *splittable = 1;
  __syncthreads();
  return B;
}

// FindBestThreshold (feature_histogram.hpp:85-95, FuncForNumricalL3 :156-217) by the whole block: every thread returns
// the same result
__device__ void wide_find_best_threshold_block(WideScratch& W, const WideData& D, const DevFeat& F, const DevParams& P, double sum_g,
                                               double leaf_sum_h, int num_data, double parent_output, const McInfo& M, DevSplit* out,
                                               int* splittable_out) {
  DevSplit o;
  o.gain = LGB_MIN_SCORE; o.feature = -1; o.real_feature = -1; o.threshold = 0; o.default_left = 1;
  o.left_count = 0; o.right_count = 0; o.left_output = 0; o.right_output = 0;
  o.left_sum_g = 0; o.left_sum_h = 0; o.right_sum_g = 0; o.right_sum_h = 0;
  const double sum_h = leaf_sum_h + (double)(2 * 1e-15f);
  const double min_gain_shift = leaf_gain(sum_g, sum_h, P, num_data, parent_output) + P.min_gain_to_split;
  int splittable = 0;
  auto apply = [&](const WideBest& B, int default_left) {   // the tail of FindBestThresholdSequentially
    if (!(splittable && B.found && B.gain > o.gain + min_gain_shift)) return;
    o.threshold = B.threshold;
    o.left_output = mc_output(B.left_g, B.left_h, P, B.left_count, parent_output, M);
    o.left_count = B.left_count;
    o.left_sum_g = B.left_g;
    o.left_sum_h = B.left_h - LGB_EPS;
    o.right_output = mc_output(sum_g - B.left_g, sum_h - B.left_h, P, num_data - B.left_count, parent_output, M);
    o.right_count = num_data - B.left_count;
    o.right_sum_g = sum_g - B.left_g;
    o.right_sum_h = sum_h - B.left_h - LGB_EPS;
    o.gain = B.gain - min_gain_shift;
    o.default_left = default_left;
  };
  if (F.num_bin > 2 && F.missing_type != LGBCU_MISSING_NONE) {
    const bool zero = F.missing_type == LGBCU_MISSING_ZERO;
    apply(wide_scan_block<true>(W, D, F, P, zero, !zero, sum_g, sum_h, num_data, min_gain_shift, parent_output, M, &splittable), 1);
    apply(wide_scan_block<false>(W, D, F, P, zero, !zero, sum_g, sum_h, num_data, min_gain_shift, parent_output, M, &splittable), 0);
  } else {
    apply(wide_scan_block<true>(W, D, F, P, false, false, sum_g, sum_h, num_data, min_gain_shift, parent_output, M, &splittable), 1);
    if (F.missing_type == LGBCU_MISSING_NAN) o.default_left = 0;
  }
  o.gain *= F.penalty;
  *out = o;
  *splittable_out = splittable;
}

// ---- categorical features -------------------------------------------------------------------------------------------
// FeatureHistogram::FindBestThresholdCategoricalInner (feature_histogram.hpp:277-513; no extra_trees) by the block.
// G / H = data_[t] of the feature as doubles, staged in shared memory.  One-hot (num_bin <= max_cat_to_onehot): every
// bin against the rest, one thread per bin.  Otherwise: the bins with enough data are stably sorted by
// sum_gradient / (sum_hessian + cat_smooth) -- a rank sort, one thread per bin -- and accumulated from the low end and
// from the high end, up to max_cat_threshold bins a side: the accumulation with its `continue` / `break` tests and the
// min_data_per_group counter is a short sequential state machine (one thread per direction, additions only) that
// records the candidates it would score; the gains -- the expensive part -- are then computed side by side, and a block
// arg-max keeps the first best in the reference's order (direction +1 before -1, then position).  `bits` receives the
// set of BINS going left (thread 0's copy is the one used).
struct CatScratch {
  double ctr[256];
  double cand_g[256], cand_h[256];   // [direction * 128 + position]: left sums of a candidate
  int cand_c[256];
  unsigned char sorted[256], tmp[256], cand_ok[256];
  double wgain[FS_WARPS]; int wpos[FS_WARPS];
  int used_bin;
};

__device__ void cat_find_best_threshold_block(CatScratch& K, const double* __restrict__ G, const double* __restrict__ H, const DevFeat& F,
                                              const DevParams& P0, const DevCatParams& C, double sum_g, double leaf_sum_h, int num_data,
                                              double parent_output, const McInfo& M, unsigned* bits, DevSplit* out, int* splittable_out) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  DevSplit o;
  o.gain = LGB_MIN_SCORE; o.feature = -1; o.real_feature = -1; o.threshold = 0; o.default_left = 0;
  o.left_count = 0; o.right_count = 0; o.left_output = 0; o.right_output = 0;
  o.left_sum_g = 0; o.left_sum_h = 0; o.right_sum_g = 0; o.right_sum_h = 0;
  const double sum_h = leaf_sum_h + (double)(2 * 1e-15f);
  double gain_shift;
  if (P0.use_smoothing) {
    const double sg = P0.use_l1 ? threshold_l1(sum_g, P0.lambda_l1) : sum_g;   // GetLeafGainGivenOutput
    gain_shift = -(2.0 * sg * parent_output + (sum_h + P0.lambda_l2) * parent_output * parent_output);
  } else {
    gain_shift = leaf_gain(sum_g, sum_h, P0, num_data, 0.0);
  }
  const double min_gain_shift = gain_shift + P0.min_gain_to_split;
  const int offset = F.offset;
  const int bin_start = 1 - offset, bin_end = F.num_bin - offset;
  DevParams P = P0;   // l2 grows by cat_l2 in the many-vs-many search
  const bool use_onehot = F.num_bin <= C.max_cat_to_onehot;
  const double cnt_factor = num_data / sum_h;
  // this thread's candidate: (gain, left sums); `ok` = it beat min_gain_shift
  double gain = LGB_MIN_SCORE, c_lg = 0, c_lh = 0; int c_lc = 0; bool ok = false;
  int used_bin = 0;
  if (use_onehot) {
    const int t = bin_start + tid;
    if (t < bin_end) {
      const double grad = G[t], hess = H[t];
      const int cnt = round_int(hess * cnt_factor);
      const int other_count = num_data - cnt;
      const double sum_other_h = sum_h - hess - LGB_EPS;
      if (!(cnt < P.min_data_in_leaf || hess < P.min_sum_hessian) && !(other_count < P.min_data_in_leaf) &&
          !(sum_other_h < P.min_sum_hessian)) {
        gain = mc_split_gains(sum_g - grad, sum_other_h, grad, hess + LGB_EPS, P, other_count, cnt, parent_output, M);
        ok = gain > min_gain_shift;
        c_lg = grad; c_lh = hess + LGB_EPS; c_lc = cnt;
      }
    }
  } else {
    P.lambda_l2 += C.cat_l2;
    // the bins with enough data, in bin order (thread 0 compacts; the flags and ratios are computed side by side)
    const int t = bin_start + tid;
    bool use = false;
    if (t < bin_end) { use = round_int(H[t] * cnt_factor) >= C.cat_smooth; K.ctr[tid] = use ? G[t] / (H[t] + C.cat_smooth) : 0.0; K.tmp[tid] = use ? 1 : 0; }
    __syncthreads();
    if (tid == 0) {
      int n = 0;
      for (int i = 0; i < bin_end - bin_start; ++i) if (K.tmp[i]) K.sorted[n++] = (unsigned char)i;   // position in the bin range
      K.used_bin = n;
    }
    __syncthreads();
    used_bin = K.used_bin;
    // stable rank sort by ctr ascending (ties: the earlier bin first, as std::stable_sort leaves them)
    int my_bin = -1, rank = 0;
    if (tid < used_bin) {
      my_bin = K.sorted[tid];
      const double mine = K.ctr[my_bin];
      for (int j = 0; j < used_bin; ++j) {
        const int bj = K.sorted[j];
        const double cj = K.ctr[bj];
        rank += (cj < mine || (cj == mine && j < tid)) ? 1 : 0;
      }
    }
    __syncthreads();
    if (tid < used_bin) K.tmp[rank] = (unsigned char)(bin_start + my_bin);   // data index t of the rank-th bin
    __syncthreads();
    if (tid < used_bin) K.sorted[tid] = K.tmp[tid];
    K.cand_ok[tid] = 0;
    __syncthreads();
    const int max_num_cat = min(C.max_cat_threshold, (used_bin + 1) / 2);   // <= 128
    if (tid == 0 || tid == 32) {   // one thread per direction: the accumulation and its tests, no gains
      const int pass = tid == 0 ? 0 : 1;
      const int dir = pass == 0 ? 1 : -1;
      int start_pos = pass == 0 ? 0 : used_bin - 1;
      int cnt_cur_group = 0, left_count = 0;
      double lg = 0.0f, lh = LGB_EPS;
      for (int i = 0; i < used_bin && i < max_num_cat; ++i) {
        const int tt = K.sorted[start_pos];
        start_pos += dir;
        const double grad = G[tt], hess = H[tt];
        const int cnt = round_int(hess * cnt_factor);
        lg += grad; lh += hess; left_count += cnt; cnt_cur_group += cnt;
        if (left_count < P.min_data_in_leaf || lh < P.min_sum_hessian) continue;
        const int right_count = num_data - left_count;
        if (right_count < P.min_data_in_leaf || right_count < C.min_data_per_group) break;
        const double rh = sum_h - lh;
        if (rh < P.min_sum_hessian) break;
        if (cnt_cur_group < C.min_data_per_group) continue;
        cnt_cur_group = 0;
        const int k = pass * 128 + i;
        K.cand_g[k] = lg; K.cand_h[k] = lh; K.cand_c[k] = left_count; K.cand_ok[k] = 1;
      }
    }
    __syncthreads();
    if (K.cand_ok[tid]) {   // thread = direction * 128 + position: the reference's order of evaluation
      c_lg = K.cand_g[tid]; c_lh = K.cand_h[tid]; c_lc = K.cand_c[tid];
      gain = mc_split_gains(c_lg, c_lh, sum_g - c_lg, sum_h - c_lh, P, c_lc, num_data - c_lc, parent_output, M);
      ok = gain > min_gain_shift;
    }
  }
  // block arg-max over the candidates that beat min_gain_shift: larger gain, then the smaller thread id
  const int w = warp_argmax(gain, (unsigned)tid, 0u, ok);
  if (lane == 0) { K.wpos[warp] = -1; K.wgain[warp] = LGB_MIN_SCORE; }
  __syncwarp();
  if (w >= 0 && lane == w) { K.wpos[warp] = tid; K.wgain[warp] = gain; }
  __syncthreads();
  int bp = -1; double bg = LGB_MIN_SCORE;
#pragma unroll
  for (int k = 0; k < FS_WARPS; ++k) if (K.wpos[k] >= 0 && (bp < 0 || K.wgain[k] > bg)) { bp = K.wpos[k]; bg = K.wgain[k]; }
  // the winner publishes its sums (cand arrays double as the mailbox)
  if (tid == bp) { K.cand_g[0] = c_lg; K.cand_h[0] = c_lh; K.cand_c[0] = c_lc; }
  __syncthreads();
  for (int wd = 0; wd < LGB_CAT_WORDS; ++wd) bits[wd] = 0u;
  const int splittable = bp >= 0 ? 1 : 0;
  if (splittable) {
    const double best_lg = K.cand_g[0], best_lh = K.cand_h[0]; const int best_lc = K.cand_c[0];
    o.left_output = mc_output(best_lg, best_lh, P, best_lc, parent_output, M);
    o.left_count = best_lc;
    o.left_sum_g = best_lg;
    o.left_sum_h = best_lh - LGB_EPS;
    o.right_output = mc_output(sum_g - best_lg, sum_h - best_lh, P, num_data - best_lc, parent_output, M);
    o.right_count = num_data - best_lc;
    o.right_sum_g = sum_g - best_lg;
    o.right_sum_h = sum_h - best_lh - LGB_EPS;
    o.gain = bg - min_gain_shift;
    if (use_onehot) {
      const int b = bin_start + bp + offset;
      bits[b >> 5] |= 1u << (b & 31);
    } else {
      const int best_dir = bp < 128 ? 1 : -1, best_threshold = bp & 127;
      for (int i = 0; i <= best_threshold; ++i) {
        const int b = (best_dir == 1 ? K.sorted[i] : K.sorted[used_bin - 1 - i]) + offset;
        bits[b >> 5] |= 1u << (b & 31);
      }
    }
  }
  o.gain *= F.penalty;
  *out = o;
  *splittable_out = splittable;
  __syncthreads();
}

// The general counterpart of find_item: same inputs, same outputs, any number of bins, numerical or categorical.  All
// FS_THREADS threads call.
__device__ void find_item_wide(const DevView& V, ScanSmem& S, int f, int sel, const FindTarget& T, const DevFeat& F, bool used) {
  DevCtl* c = V.ctl;
  const int tid = threadIdx.x;
  const int nt = F.num_bin - F.offset;
  const double inv_g = c->inv_g, inv_h = c->inv_h;
  const DevLeaf Ls = T.Ls;
  const DevLeaf Lt = sel == 0 ? T.Ls : T.Ll;          // the leaf this block scans
  HistEntry* hs = V.pool + (size_t)Ls.slot * V.total_bin + F.hist_base;   // smaller leaf
  HistEntry* ht = V.pool + (size_t)Lt.slot * V.total_bin + F.hist_base;   // scanned leaf (sel 1: holds the parent)
  __shared__ double s_fix[2];
  // the block scan's scratch lies over the 8-bit search's (unused on this path once the block sums above are done)
  static_assert(sizeof(WideScratch) + sizeof(CatScratch) + 16 <= sizeof(ScanSmem), "the general search's scratch must fit ScanSmem");
  WideScratch& s_scan = *reinterpret_cast<WideScratch*>(&S);
  CatScratch& s_cat = *reinterpret_cast<CatScratch*>(reinterpret_cast<char*>(&S) + ((sizeof(WideScratch) + 15) & ~(size_t)15));
  DevSplit o; int sp = 0;
  o.gain = LGB_MIN_SCORE; o.feature = -1; o.real_feature = -1;
  if (used) {
    if (sel == 1) {
      // larger = parent - smaller (FeatureHistogram::Subtract), exact, in place.  The smaller leaf's most-frequent-bin
      // entry is being rewritten by the sibling block: its exact twin (stored value 0) is read instead.
      for (int i = tid; i < nt; i += FS_THREADS) {
        const HistEntry a = ht[1 + i];
        const HistEntry b = (F.mfb > 0 && i == F.mfb) ? hs[0] : hs[1 + i];
        HistEntry e; e.g = a.g - b.g; e.h = a.h - b.h;
        ht[1 + i] = e;
      }
      if (tid == 0) {
        const HistEntry a0 = ht[0], b0 = hs[0];
        HistEntry e0; e0.g = a0.g - b0.g; e0.h = a0.h - b0.h;
        ht[0] = e0;
      }
    }
    double fix_g = 0.0, fix_h = 0.0;
    if (F.mfb > 0) {
      // Dataset::FixHistogram (dataset.cpp:1410-1428): data[mfb] = leaf sum - sum of the other bins (of the SMALLER leaf;
      // the larger leaf's is the parent's fixed entry minus it, as in find_item)
      i64 og = 0, oh = 0;
      for (int i = tid; i < nt; i += FS_THREADS) {
        if (i == F.mfb) continue;
        const HistEntry e = hs[1 + i];
        og += e.g; oh += e.h;
      }
      og = block_sum_i64(og, S.wtot[0]);
      oh = block_sum_i64(oh, S.wtot[1]);
      const double sg_fix = Ls.sum_g - (double)og * inv_g, sh_fix = Ls.sum_h - (double)oh * inv_h;
      if (sel == 0) {
        fix_g = sg_fix; fix_h = sh_fix;
        if (tid == 0) {
          hs[1 + F.mfb] = hs[0];   // the exact integer twin of the most-frequent bin
          V.mfb_fix[(size_t)Ls.slot * V.num_features + f] = make_double2(fix_g, fix_h);
        }
      } else {
        double2* fix_l = V.mfb_fix + (size_t)Lt.slot * V.num_features + f;
        const double2 pf = *fix_l;
        fix_g = pf.x - sg_fix; fix_h = pf.y - sh_fix;
        __syncthreads();   // every thread has read *fix_l
        if (tid == 0) *fix_l = make_double2(fix_g, fix_h);
      }
    }
    if (tid == 0) { s_fix[0] = fix_g; s_fix[1] = fix_h; }
    __threadfence_block();
    __syncthreads();   // the subtracted entries are in place for the walking thread
    {
      // every thread builds the same views; the categorical search is one thread's walk, the numerical one the block's
      WideData D; D.h = ht; D.mfb = F.mfb > 0 ? F.mfb : -1000; D.fix_g = s_fix[0]; D.fix_h = s_fix[1]; D.inv_g = inv_g; D.inv_h = inv_h;
      const bool id0 = sel == 0 ? T.smaller_id0 : T.larger_id0;
      double parent_output = parent_output_of(Lt, id0 ? 0 : 1, V.p);
      // the leaf with id 0 takes its own unsmoothed output as "parent" output -- clamped to its interval
      // (CalculateSplittedLeafOutput<true, ...> with constraints_->Get(0), serial_tree_learner.cpp:709-715)
      if (id0 && V.use_mc) parent_output = fmin(fmax(parent_output, Lt.mc_min), Lt.mc_max);
      McInfo M; M.use = V.use_mc; M.lo = Lt.mc_min; M.hi = Lt.mc_max; M.type = 0;
      if (V.use_mc && !F.cat) M.type = V.monotone[f];
      if (F.cat) {
        for (int t = tid; t < nt; t += FS_THREADS) { s_scan.vg[t] = D.g(t); s_scan.vh[t] = D.hh(t); }   // nt <= 255 for a categorical feature
        __syncthreads();
        unsigned bits[LGB_CAT_WORDS];
        cat_find_best_threshold_block(s_cat, s_scan.vg, s_scan.vh, F, V.p, V.cat, Lt.sum_g, Lt.sum_h, Lt.gcount, parent_output, M, bits, &o, &sp);
        if (tid == 0 && sp) {   // the set of bins going left takes the next pool entry; the split carries its index as "threshold"
          const int e = atomicAdd(&c->cat_count, 1);
          if (e < V.cat_cap) {
            for (int w = 0; w < LGB_CAT_WORDS; ++w) V.cat_pool[(size_t)e * LGB_CAT_WORDS + w] = bits[w];
            o.threshold = e;
          } else {
            c->cat_overflow = 1; o.gain = LGB_MIN_SCORE; sp = 0;
          }
        }
      } else {
        wide_find_best_threshold_block(s_scan, D, F, V.p, Lt.sum_g, Lt.sum_h, Lt.gcount, parent_output, M, &o, &sp);
        if (M.type != 0) {
          // SerialTreeLearner::ComputeBestSplitForFeature (serial_tree_learner.cpp:728-732) with
          // LeafConstraintsBase::ComputeMonotoneSplitGainPenalty (monotone_constraints.hpp:66-75)
          const double pen = V.mc_penalty; const int depth = Lt.depth;
          double factor;
          if (pen >= depth + 1.) factor = LGB_EPS;
          else if (pen <= 1.) factor = 1. - pen / pow(2., (double)depth) + LGB_EPS;
          else factor = 1. - pow(2., pen - 1. - depth) + LGB_EPS;
          o.gain *= factor;
        }
      }
      o.feature = f; o.real_feature = F.real_index;
    }
  }
  if (tid == 0) {
    const size_t row = (size_t)(T.slot + sel) * V.num_features + f;
    V.feat_best[row] = o;
    SplitKey k; k.gain = o.gain; k.real_feature = o.real_feature;
    k.pad = o.feature >= 0 ? key_pack<true>(o.feature, F.cat ? 0 : o.threshold, o.default_left) : 0;
    V.feat_key[row] = k;
    if (sel == 0 || used) V.splittable[(size_t)Lt.slot * V.num_features + f] = (int8_t)sp;
  }
  __syncthreads();
}
