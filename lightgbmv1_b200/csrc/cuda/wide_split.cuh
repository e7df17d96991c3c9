// Split search for a WIDE feature (more than 256 stored values, i.e. max_bin > 255: the reference's 16-bit DenseBin,
// src/io/bin.cpp:635-645).  Included by split_kernels.cuh in front of find_item, which hands such features over.
//
// The 8-bit search (scan_feature_block) is built around one thread per bin of a <= 256-bin feature.  This is the general
// form: the block forms larger = parent - smaller and the FixHistogram residue cooperatively, then ONE thread walks the
// bins exactly as FeatureHistogram::FindBestThresholdSequentially does (feature_histogram.hpp:853-1049) -- reverse
// scan, then forward scan where the feature has a missing bin -- reading the 64-bit fixed-point entries from the pool
// and converting each to double (exact: an entry is an integer multiple of a power of two).  Slow by design (a
// 1024-bin feature costs ~50 us per (feature, child) block against ~7 us for the 8-bit search); correct for any
// number of bins.  Same outputs as find_item: feat_best / feat_key rows, splittable flags, mfb_fix.
#pragma once

struct WideData {   // data_[t] of FeatureHistogram for one (feature, leaf): entries of the pool + the fixed mfb entry
  const HistEntry* h;   // slot entries of the feature's group: h[1 + t] = data_[t]
  int mfb; double fix_g, fix_h, inv_g, inv_h;
  __device__ __forceinline__ double g(int t) const { return t == mfb ? fix_g : (double)h[1 + t].g * inv_g; }
  __device__ __forceinline__ double hh(int t) const { return t == mfb ? fix_h : (double)h[1 + t].h * inv_h; }
};

struct WideBest { double gain, left_g, left_h; int left_count, threshold; bool found; };

// One direction of FindBestThresholdSequentially.  Returns the best candidate of the direction (found = false: none
// beat min_gain_shift); *splittable is set when any threshold did.
template <bool REVERSE>
__device__ WideBest wide_scan(const WideData& D, const DevFeat& F, const DevParams& P, bool skip_default, bool na_missing,
                              double sum_g, double sum_h, int num_data, double min_gain_shift, double parent_output,
                              int* splittable) {
  const int offset = F.offset;
  WideBest B; B.gain = LGB_MIN_SCORE; B.left_g = 0; B.left_h = 0; B.left_count = 0; B.threshold = F.num_bin; B.found = false;
  const double cnt_factor = num_data / sum_h;
  if (REVERSE) {
    double rg = 0.0f, rh = LGB_EPS;
    int rc = 0;
    const int t_end = 1 - offset;
    for (int t = F.num_bin - 1 - offset - (na_missing ? 1 : 0); t >= t_end; --t) {
      if (skip_default && (t + offset) == F.default_bin) continue;
      const double grad = D.g(t), hess = D.hh(t);
      rg += grad; rh += hess; rc += round_int(hess * cnt_factor);
      if (rc < P.min_data_in_leaf || rh < P.min_sum_hessian) continue;
      const int lc = num_data - rc;
      if (lc < P.min_data_in_leaf) break;
      const double lh = sum_h - rh;
      if (lh < P.min_sum_hessian) break;
      const double lg = sum_g - rg;
      const double gain = leaf_gain(lg, lh, P, lc, parent_output) + leaf_gain(rg, rh, P, rc, parent_output);
      if (gain <= min_gain_shift) continue;
      *splittable = 1;
      if (gain > B.gain) { B.left_count = lc; B.left_g = lg; B.left_h = lh; B.threshold = t - 1 + offset; B.gain = gain; B.found = true; }
    }
  } else {
    double lg = 0.0f, lh = LGB_EPS;
    int lc = 0, t = 0;
    const int t_end = F.num_bin - 2 - offset;
    if (na_missing && offset == 1) {   // the implicit first bin: everything minus the stored bins
      lg = sum_g; lh = sum_h - LGB_EPS; lc = num_data;
      for (int i = 0; i < F.num_bin - offset; ++i) {
        const double grad = D.g(i), hess = D.hh(i);
        lg -= grad; lh -= hess; lc -= round_int(hess * cnt_factor);
      }
      t = -1;
    }
    for (; t <= t_end; ++t) {
      if (skip_default && (t + offset) == F.default_bin) continue;
      if (t >= 0) {
        const double hess = D.hh(t);
        lg += D.g(t); lh += hess; lc += round_int(hess * cnt_factor);
      }
      if (lc < P.min_data_in_leaf || lh < P.min_sum_hessian) continue;
      const int rc = num_data - lc;
      if (rc < P.min_data_in_leaf) break;
      const double rh = sum_h - lh;
      if (rh < P.min_sum_hessian) break;
      const double rg = sum_g - lg;
      const double gain = leaf_gain(lg, lh, P, lc, parent_output) + leaf_gain(rg, rh, P, rc, parent_output);
      if (gain <= min_gain_shift) continue;
      *splittable = 1;
      if (gain > B.gain) { B.left_count = lc; B.left_g = lg; B.left_h = lh; B.threshold = t + offset; B.gain = gain; B.found = true; }
    }
  }
  return B;
}

// FindBestThreshold (feature_histogram.hpp:85-95, FuncForNumricalL3 :156-217) of one (feature, leaf) by ONE thread
__device__ void wide_find_best_threshold(const WideData& D, const DevFeat& F, const DevParams& P, double sum_g, double leaf_sum_h,
                                         int num_data, double parent_output, DevSplit* out, int* splittable_out) {
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
    o.left_output = leaf_output(B.left_g, B.left_h, P, B.left_count, parent_output);
    o.left_count = B.left_count;
    o.left_sum_g = B.left_g;
    o.left_sum_h = B.left_h - LGB_EPS;
    o.right_output = leaf_output(sum_g - B.left_g, sum_h - B.left_h, P, num_data - B.left_count, parent_output);
    o.right_count = num_data - B.left_count;
    o.right_sum_g = sum_g - B.left_g;
    o.right_sum_h = sum_h - B.left_h - LGB_EPS;
    o.gain = B.gain - min_gain_shift;
    o.default_left = default_left;
  };
  if (F.num_bin > 2 && F.missing_type != LGBCU_MISSING_NONE) {
    const bool zero = F.missing_type == LGBCU_MISSING_ZERO;
    apply(wide_scan<true>(D, F, P, zero, !zero, sum_g, sum_h, num_data, min_gain_shift, parent_output, &splittable), 1);
    apply(wide_scan<false>(D, F, P, zero, !zero, sum_g, sum_h, num_data, min_gain_shift, parent_output, &splittable), 0);
  } else {
    apply(wide_scan<true>(D, F, P, false, false, sum_g, sum_h, num_data, min_gain_shift, parent_output, &splittable), 1);
    if (F.missing_type == LGBCU_MISSING_NAN) o.default_left = 0;
  }
  o.gain *= F.penalty;
  *out = o;
  *splittable_out = splittable;
}

// The wide counterpart of find_item: same inputs, same outputs, any number of bins.  All FS_THREADS threads call.
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
    if (tid == 0) {
      WideData D; D.h = ht; D.mfb = F.mfb > 0 ? F.mfb : -1000; D.fix_g = s_fix[0]; D.fix_h = s_fix[1]; D.inv_g = inv_g; D.inv_h = inv_h;
      const bool id0 = sel == 0 ? T.smaller_id0 : T.larger_id0;
      wide_find_best_threshold(D, F, V.p, Lt.sum_g, Lt.sum_h, Lt.gcount, parent_output_of(Lt, id0 ? 0 : 1, V.p), &o, &sp);
      o.feature = f; o.real_feature = F.real_index;
    }
  }
  if (tid == 0) {
    const size_t row = (size_t)(T.slot + sel) * V.num_features + f;
    V.feat_best[row] = o;
    SplitKey k; k.gain = o.gain; k.real_feature = o.real_feature;
    k.pad = o.feature >= 0 ? key_pack<true>(o.feature, o.threshold, o.default_left) : 0;
    V.feat_key[row] = k;
    if (sel == 0 || used) V.splittable[(size_t)Lt.slot * V.num_features + f] = (int8_t)sp;
  }
  __syncthreads();
}
