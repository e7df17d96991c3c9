// Split finding (replaces FeatureHistogram::FindBestThreshold / FindBestThresholdSequentially,
// src/treelearner/feature_histogram.hpp:85-95,156-217,853-1049; Dataset::FixHistogram, src/io/dataset.cpp:1410-1428;
// FeatureHistogram::Subtract, feature_histogram.hpp:79-83; the per-leaf argmax of
// SerialTreeLearner::FindBestSplitsFromHistograms, serial_tree_learner.cpp:358-425, and the leaf choice of
// SerialTreeLearner::Train, serial_tree_learner.cpp:181-199).
//
// One 256-thread block per (feature, child of a leaf being expanded).  Block "sel 0" scans the smaller child, block
// "sel 1" forms larger = parent - smaller in place and scans it.  Gradient sums and the reconstructed counts are exact
// integer prefix / suffix scans (the histogram is 64-bit fixed point) done by half of the block, while the other half
// replays the two kEpsilon-seeded hessian chains in the reference's order; each thread then evaluates one threshold,
// and REDUX arg-max reductions apply the reference's tie-breaks (reverse scan keeps the highest threshold among equal
// gains, forward scan the lowest; reverse wins ties with forward).  The last block to finish runs the control step:
// per-leaf reduction over features (SplitInfo::operator>, split_info.hpp:126-153), commit of the expansions whose
// turn has come in the reference's order, choice of the next round's leaves -- no host involved.
#pragma once
#include "dev_types.cuh"

#define FULL_MASK 0xffffffffu
#ifdef LGBCU_STAGE_TIMERS
__device__ unsigned long long g_stage_cycles[32];
__device__ long long g_stage_prev;
#define STAGE(i) do { if (blockIdx.x == 0 && blockIdx.y == 1 && threadIdx.x == 0) { long long t_ = clock64(); \
    g_stage_cycles[i] += (unsigned long long)(t_ - g_stage_prev); g_stage_prev = t_; } } while (0)
#define STAGE_BEGIN() do { if (blockIdx.x == 0 && blockIdx.y == 1 && threadIdx.x == 0) g_stage_prev = clock64(); } while (0)
#else
#define STAGE(i) do {} while (0)
#define STAGE_BEGIN() do {} while (0)
#endif
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

struct Cand {          // one threshold candidate
  double gain, left_g, left_h;
  int t, left_count;
};

__device__ __forceinline__ Cand cand_better(const Cand& a, const Cand& b, bool prefer_high) {
  const bool take_b = (b.gain > a.gain) ||
                      (b.gain == a.gain && b.gain != LGB_MIN_SCORE && (prefer_high ? b.t > a.t : b.t < a.t));
  return take_b ? b : a;
}

// Warp arg-max with the hardware reductions (REDUX on 32-bit integers): a double is mapped to an order-preserving
// 64-bit key, the maximum is found on its high word, then on the low word among the lanes still tied, then the
// smallest `tie` (and `tie2`) among those.  Four or five warp collectives instead of five shuffle rounds of compare
// logic.  Returns the winning lane, or -1 when no lane is valid.  NaN gains must be passed as invalid.
__device__ __forceinline__ void gain_key(double g, unsigned& hi, unsigned& lo) {
  unsigned long long u = (unsigned long long)__double_as_longlong(g);
  u = (u >> 63) ? ~u : (u | 0x8000000000000000ull);
  hi = (unsigned)(u >> 32); lo = (unsigned)u;
}
__device__ __forceinline__ int warp_argmax(double gain, unsigned tie, unsigned tie2, bool valid) {
  unsigned hi, lo;
  gain_key(gain, hi, lo);
  bool a = valid;
  const unsigned m1 = __reduce_max_sync(FULL_MASK, a ? hi : 0u);
  a = a && hi == m1;
  unsigned b = __ballot_sync(FULL_MASK, a);
  if ((b & (b - 1)) == 0) return b ? __ffs(b) - 1 : -1;  // the high word alone decides (the usual case)
  const unsigned m2 = __reduce_max_sync(FULL_MASK, a ? lo : 0u);
  a = a && lo == m2;
  b = __ballot_sync(FULL_MASK, a);
  if ((b & (b - 1)) == 0) return __ffs(b) - 1;
  const unsigned m3 = __reduce_min_sync(FULL_MASK, a ? tie : 0xffffffffu);
  a = a && tie == m3;
  const unsigned m4 = __reduce_min_sync(FULL_MASK, a ? tie2 : 0xffffffffu);
  a = a && tie2 == m4;
  b = __ballot_sync(FULL_MASK, a);
  return b ? __ffs(b) - 1 : -1;
}

// Best threshold candidate of a warp: larger gain, ties -> higher t (reverse scan) / lower t (forward scan).
// Only (gain, t) are reduced; the winner's sums are looked up again afterwards.
__device__ __forceinline__ Cand warp_best_cand(Cand c, bool prefer_high) {
  const bool valid = c.gain > LGB_MIN_SCORE;  // false for NaN as well: a NaN gain never wins in the reference either
  const unsigned tie = prefer_high ? (unsigned)(0x40000000 - c.t) : (unsigned)(c.t + 0x40000000);
  const int w = warp_argmax(c.gain, tie, 0u, valid);
  Cand r = c;
  if (w < 0) { r.gain = LGB_MIN_SCORE; r.t = prefer_high ? -1000 : 100000; return r; }
  r.gain = __shfl_sync(FULL_MASK, c.gain, w);
  r.t = __shfl_sync(FULL_MASK, c.t, w);
  return r;
}

__device__ __forceinline__ i64 warp_sum(i64 v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
  return v;
}

// One thread walks a shared-memory array in index order accumulating an fp64 chain exactly as a sequential CPU loop
// would; operands are fetched 8 at a time so that only the DADD latency is on the critical path.  `sign` = +1 adds,
// -1 subtracts (x - v is computed as x + (-v): identical in IEEE).  Entries equal to `skip_t` contribute +0.0, which
// leaves a running sum bit-unchanged.  (Used for the implicit-first-bin pre-pass; the two main chains walk a
// pre-laid-out input, chain_lin below.)
template <int DIR>
__device__ __forceinline__ void chain1(const double* __restrict__ h, double* oh, int t_first, int t_last, int skip_t,
                                       double sign, double* ah_io) {
  double ah = *ah_io;
  const int count = DIR > 0 ? t_last - t_first + 1 : t_first - t_last + 1;
  for (int base = 0; base < count; base += 8) {
    double vh[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int t = t_first + DIR * (base + k);
      vh[k] = ((base + k) < count && t != skip_t) ? sign * h[t] : 0.0;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      ah += vh[k];
      if (oh && (base + k) < count) oh[t_first + DIR * (base + k)] = ah;
    }
  }
  *ah_io = ah;
}
// The same walk over an input that was laid out in chain order beforehand (zero for skipped entries, zero padded by
// CHAIN_PAD): no bounds or skip tests in the loop, 16-byte loads, and the next eight operands are fetched before the
// current eight are added, so that the dependent DADDs are all the thread waits for.  Step i writes out0[DIR * i].
template <int DIR>
__device__ __forceinline__ void chain_lin(const double* __restrict__ in, double* out0, int count, double* ah_io) {
  double ah = *ah_io;
  const double2* in2 = reinterpret_cast<const double2*>(in);
  double2 a = in2[0], b = in2[1], c = in2[2], d = in2[3];
  int i = 0;
  for (; i + 8 <= count; i += 8) {
    const double2 na = in2[(i >> 1) + 4], nb = in2[(i >> 1) + 5], nc = in2[(i >> 1) + 6], nd = in2[(i >> 1) + 7];
    ah += a.x; out0[DIR * (i + 0)] = ah;
    ah += a.y; out0[DIR * (i + 1)] = ah;
    ah += b.x; out0[DIR * (i + 2)] = ah;
    ah += b.y; out0[DIR * (i + 3)] = ah;
    ah += c.x; out0[DIR * (i + 4)] = ah;
    ah += c.y; out0[DIR * (i + 5)] = ah;
    ah += d.x; out0[DIR * (i + 6)] = ah;
    ah += d.y; out0[DIR * (i + 7)] = ah;
    a = na; b = nb; c = nc; d = nd;
  }
  const double v[8] = {a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y};
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (i + k < count) { ah += v[k]; out0[DIR * (i + k)] = ah; }
  *ah_io = ah;
}
// SplitKey::pad = inner feature | threshold << 20 | default_left << 29 (bins are 8-bit, so a threshold fits 9 bits)
#define KEY_FEAT_BITS 20        /* 8-bit data: feature < 2^20, threshold in 10 bits */
#define KEY_FEAT_BITS_WIDE 14   /* data with a wide feature: feature < 2^14, threshold in 16 bits */
// pad = feature | threshold << fb | default_left << 30, fb by the data's kind (WIDE = the data has a wide feature;
// a template argument so that the 8-bit kernel shifts by constants)
template <bool WIDE>
__device__ __forceinline__ int key_pack(int feature, int threshold, int default_left) {
  constexpr int fb = WIDE ? KEY_FEAT_BITS_WIDE : KEY_FEAT_BITS;
  return feature | (threshold << fb) | (default_left << 30);
}
template <bool WIDE>
__device__ __forceinline__ void key_unpack(int pd, int* feature, int* threshold, int* default_left) {
  constexpr int fb = WIDE ? KEY_FEAT_BITS_WIDE : KEY_FEAT_BITS;
  *feature = pd & ((1 << fb) - 1);
  *threshold = (pd >> fb) & ((1 << (30 - fb)) - 1);
  *default_left = (pd >> 30) & 1;
}

#define FS_THREADS 256
#define FS_MODE_OWNED 2     /* k_find_split modes beside 0 (training) and 1 (one explicit leaf): owned scan -> send buffer */
#define FS_MODE_CONTROL 3   /* ... and the control step alone, from the gathered winners */
#define FS_WARPS (FS_THREADS / 32)

// Shared-memory work area of one (feature, leaf) block.
#define CHAIN_PAD 16
struct __align__(16) ScanSmem {
  double cr[256 + CHAIN_PAD], cf[256 + CHAIN_PAD];  // inputs of the reverse / forward hessian chain, in chain order, zero padded
  double hd[256];               // hessian data_[t] as doubles (after FixHistogram)
  i64 gi[256];                  // gradient entries in fixed point (exact prefix / suffix sums)
  i64 wtot[6][FS_WARPS];        // per-warp totals of the integer scans / reductions (one row per scan)
  double RG[256], RH[256];      // reverse scan: sum_right_gradient / sum_right_hessian after adding bin t
  double LG[256], LH[256];      // forward scan: sum_left_* after adding bin t
  int RC[256], LC[256];
  double base_g, base_h; int base_c;   // forward scan start (implicit first bin of NA features with offset 1)
  Cand wbest[2][FS_WARPS];
  int wsplit[FS_WARPS];
};

// block-wide sum of one int64 per thread (all FS_THREADS threads call; result broadcast)
__device__ __forceinline__ i64 block_sum_i64(i64 v, i64* wtot) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) wtot[warp] = v;
  __syncthreads();
  i64 t = 0;
#pragma unroll
  for (int w = 0; w < FS_WARPS; ++w) t += wtot[w];
  __syncthreads();
  return t;
}
// inclusive prefix (DIR=+1: over lower thread ids) or suffix (DIR=-1: over higher thread ids) sum across the block
template <int DIR>
__device__ __forceinline__ i64 block_scan_i64(i64 v, i64* wtot) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  i64 inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const i64 x = DIR > 0 ? __shfl_up_sync(FULL_MASK, inc, o) : __shfl_down_sync(FULL_MASK, inc, o);
    if (DIR > 0 ? lane >= o : lane + o < 32) inc += x;
  }
  if (DIR > 0 ? lane == 31 : lane == 0) wtot[warp] = inc;
  __syncthreads();
  i64 off = 0;
#pragma unroll
  for (int w = 0; w < FS_WARPS; ++w) if (DIR > 0 ? w < warp : w > warp) off += wtot[w];
  __syncthreads();
  return inc + off;
}

// 128 threads (the upper half of the block) scan 256 elements, two adjacent ones per thread (j = thread index
// within the group); inclusive prefix (DIR = +1) or suffix (DIR = -1).  One named barrier per call: every call gets
// its own wtot row.
#define GROUP_BAR() asm volatile("bar.sync 1, 128;" ::: "memory")
template <int DIR>
__device__ __forceinline__ void group_scan_i64(i64 v0, i64 v1, i64* o0, i64* o1, i64* wtot, int j) {
  const int lane = j & 31, w = j >> 5;
  const i64 pair = v0 + v1;
  i64 inc = pair;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const i64 x = DIR > 0 ? __shfl_up_sync(FULL_MASK, inc, o) : __shfl_down_sync(FULL_MASK, inc, o);
    if (DIR > 0 ? lane >= o : lane + o < 32) inc += x;
  }
  if (DIR > 0 ? lane == 31 : lane == 0) wtot[w] = inc;
  GROUP_BAR();
  i64 off = 0;
#pragma unroll
  for (int k = 0; k < FS_WARPS / 2; ++k) if (DIR > 0 ? k < w : k > w) off += wtot[k];
  const i64 excl = off + inc - pair;  // everything strictly before (DIR > 0) / after (DIR < 0) this thread's pair
  if (DIR > 0) { *o0 = excl + v0; *o1 = excl + pair; }
  else { *o1 = excl + v1; *o0 = excl + pair; }
}
__device__ __forceinline__ i64 group_sum_i64(i64 v, i64* wtot, int j) {
  v = warp_sum(v);
  if ((j & 31) == 0) wtot[j >> 5] = v;
  GROUP_BAR();
  i64 t = 0;
#pragma unroll
  for (int k = 0; k < FS_WARPS / 2; ++k) t += wtot[k];
  return t;
}

// FindBestThreshold for one numerical feature, one block.
// Arithmetic fidelity.  The reference accumulates bins sequentially in double.  Where its running sums are exactly
// representable (always for the small / medium leaves where gains tie) a sequential sum equals the exact sum, so
// gradient sums, FixHistogram's residual and the implicit first bin are formed from exact 64-bit integer
// prefix/suffix sums in parallel and rounded once.  The hessian scans are different: they start from kEpsilon
// (1e-15f), whose low bits are rounded away step by step as the sum grows, so their value depends on the ORDER of
// additions; those two chains (<= 256 dependent DADDs each) are replayed in the reference's order by one thread
// each, with operands prefetched so that only the DADD latency is on the critical path.  Gains (two fp64 divisions
// per threshold) are then evaluated one threshold per thread and reduced with the reference's tie rules.
__device__ void scan_feature_block(ScanSmem& S, int nt, const DevFeat& F, double sum_g, double leaf_sum_h,
                                   int num_data, double parent_output, const DevParams& P, double inv_g_in,
                                   DevSplit* out, int* splittable_out) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double sum_h = leaf_sum_h + (double)(2 * 1e-15f);  // FindBestThreshold: sum_hessian + 2 * kEpsilon
  const double cnt_factor = num_data / sum_h;
  const double min_gain_shift = leaf_gain(sum_g, sum_h, P, num_data, parent_output) + P.min_gain_to_split;
  const int offset = F.offset;
  const bool two_way = F.num_bin > 2 && F.missing_type != LGBCU_MISSING_NONE;
  const bool skip_default = two_way && F.missing_type == LGBCU_MISSING_ZERO;
  const bool na_missing = two_way && F.missing_type == LGBCU_MISSING_NAN;
  const bool implicit_first = na_missing && offset == 1;
  const int t_hi = nt - 1 - (na_missing ? 1 : 0), t_lo = 1 - offset;  // reverse range
  const int t_end = nt - 2;                                            // forward range [0, t_end]

  const int skip_t = skip_default ? (int)F.default_bin - offset : -1000;  // data_ index of the skipped default bin
  const double inv_g = inv_g_in;
  // Two things run side by side.  Warps 0 and 1: the two hessian chains (they start from kEpsilon, so the order of
  // the additions matters) are replayed sequentially by one thread each.  Warps 4..7: everything that is exact in
  // integers -- gradient sums (64-bit fixed point, rounded once) and the reconstructed counts -- as 128-thread
  // prefix / suffix scans over the <= 256 bins, two adjacent bins per thread, synchronised on named barrier 1.
  const int cnt_r = t_hi - t_lo + 1;
  for (int i = tid; i < 256 + CHAIN_PAD; i += FS_THREADS) {
    const int t = t_hi - i;
    S.cr[i] = (i < cnt_r && t != skip_t) ? S.hd[t] : 0.0;
    S.cf[i] = (two_way && i <= t_end && i != skip_t) ? S.hd[i] : 0.0;
  }
  __syncthreads();
  if (tid == 0) {  // reverse hessian chain
    double ah = LGB_EPS;
    chain_lin<-1>(S.cr, &S.RH[t_hi], cnt_r, &ah);

  } else if (tid == 32 && two_way) {  // forward hessian chain
    double lh = LGB_EPS;
    if (implicit_first) {
      lh = sum_h - LGB_EPS;
      chain1<1>(S.hd, nullptr, 0, nt - 1, -1000, -1.0, &lh);
    }
    S.base_h = lh;
    chain_lin<1>(S.cf, &S.LH[0], t_end + 1, &lh);
  } else if (warp >= FS_WARPS / 2) {
    {
    const int j = tid - FS_THREADS / 2;
    const int e0 = 2 * j, e1 = 2 * j + 1;
    const i64 g0 = S.gi[e0], g1 = S.gi[e1];
    const i64 c0 = e0 < nt ? (i64)round_int(S.hd[e0] * cnt_factor) : 0;
    const i64 c1 = e1 < nt ? (i64)round_int(S.hd[e1] * cnt_factor) : 0;
    const bool r0 = e0 >= t_lo && e0 <= t_hi && e0 != skip_t, r1 = e1 >= t_lo && e1 <= t_hi && e1 != skip_t;
    i64 a0, a1;
    group_scan_i64<-1>(r0 ? g0 : 0, r1 ? g1 : 0, &a0, &a1, S.wtot[0], j);
    S.RG[e0] = (double)a0 * inv_g; S.RG[e1] = (double)a1 * inv_g;
    group_scan_i64<-1>(r0 ? c0 : 0, r1 ? c1 : 0, &a0, &a1, S.wtot[1], j);
    S.RC[e0] = (int)a0; S.RC[e1] = (int)a1;
    if (two_way) {
      const bool f0 = e0 <= t_end && e0 != skip_t, f1 = e1 <= t_end && e1 != skip_t;
      double base_g = 0.0;
      i64 base_c = 0;
      if (implicit_first) {
        const i64 all_g = group_sum_i64((e0 < nt ? g0 : 0) + (e1 < nt ? g1 : 0), S.wtot[2], j);
        const i64 all_c = group_sum_i64(c0 + c1, S.wtot[3], j);
        base_g = sum_g - (double)all_g * inv_g;
        base_c = (i64)num_data - all_c;
      }
      group_scan_i64<1>(f0 ? g0 : 0, f1 ? g1 : 0, &a0, &a1, S.wtot[4], j);
      S.LG[e0] = base_g + (double)a0 * inv_g; S.LG[e1] = base_g + (double)a1 * inv_g;
      group_scan_i64<1>(f0 ? c0 : 0, f1 ? c1 : 0, &a0, &a1, S.wtot[5], j);
      S.LC[e0] = (int)(base_c + a0); S.LC[e1] = (int)(base_c + a1);
      if (j == 0) { S.base_g = base_g; S.base_c = (int)base_c; }
    }
    }
  }
  __syncthreads();
  STAGE(3);
  // ---- one threshold per thread
  int splittable = 0;
  Cand rb; rb.gain = LGB_MIN_SCORE; rb.t = -1000; rb.left_g = 0; rb.left_h = 0; rb.left_count = 0;
  Cand fb; fb.gain = LGB_MIN_SCORE; fb.t = 100000; fb.left_g = 0; fb.left_h = 0; fb.left_count = 0;
  {
    const int t = tid;
    if (t >= t_lo && t <= t_hi && !(skip_default && (t + offset) == F.default_bin)) {
      const double right_g = S.RG[t], right_h = S.RH[t];
      const int right_cnt = S.RC[t];
      if (!(right_cnt < P.min_data_in_leaf || right_h < P.min_sum_hessian)) {
        const int left_cnt = num_data - right_cnt;
        const double left_h = sum_h - right_h;
        if (!(left_cnt < P.min_data_in_leaf) && !(left_h < P.min_sum_hessian)) {  // the reference breaks here: monotone
          const double left_g = sum_g - right_g;
          const double gain = leaf_gain(left_g, left_h, P, left_cnt, parent_output) +
                              leaf_gain(right_g, right_h, P, right_cnt, parent_output);
          if (!(gain <= min_gain_shift)) {
            splittable = 1;
            rb.gain = gain; rb.left_g = left_g; rb.left_h = left_h; rb.left_count = left_cnt; rb.t = t;
          }
        }
      }
    }
  }
  if (two_way) {
    // thread t evaluates forward threshold t; thread FS_THREADS-1 also covers the implicit t = -1 candidate
    for (int pass = 0; pass < 2; ++pass) {
      int t; double left_g, left_h; int left_cnt;
      if (pass == 0) {
        t = tid;
        if (!(t <= t_end && !(skip_default && (t + offset) == F.default_bin))) continue;
        left_g = S.LG[t]; left_h = S.LH[t]; left_cnt = S.LC[t];
      } else {
        if (!(implicit_first && tid == FS_THREADS - 1)) continue;
        t = -1; left_g = S.base_g; left_h = S.base_h; left_cnt = S.base_c;
      }
      if (left_cnt < P.min_data_in_leaf || left_h < P.min_sum_hessian) continue;
      const int right_cnt = num_data - left_cnt;
      if (right_cnt < P.min_data_in_leaf) continue;
      const double right_h = sum_h - left_h;
      if (right_h < P.min_sum_hessian) continue;
      const double right_g = sum_g - left_g;
      const double gain = leaf_gain(left_g, left_h, P, left_cnt, parent_output) +
                          leaf_gain(right_g, right_h, P, right_cnt, parent_output);
      if (gain <= min_gain_shift) continue;
      splittable = 1;
      Cand c; c.gain = gain; c.left_g = left_g; c.left_h = left_h; c.left_count = left_cnt; c.t = t;
      fb = cand_better(fb, c, false);
    }
  }
  STAGE(24);
  rb = warp_best_cand(rb, true);
  fb = warp_best_cand(fb, false);
  splittable = __any_sync(FULL_MASK, splittable);
  if (lane == 0) { S.wbest[0][warp] = rb; S.wbest[1][warp] = fb; S.wsplit[warp] = splittable; }
  __syncthreads();
  if (warp != 0) return;
  {
    Cand x; x.gain = LGB_MIN_SCORE; x.t = -1000; x.left_g = 0; x.left_h = 0; x.left_count = 0;
    Cand y = x; y.t = 100000;
    int sp = 0;
    if (lane < FS_WARPS) { x = S.wbest[0][lane]; y = S.wbest[1][lane]; sp = S.wsplit[lane]; }
    rb = warp_best_cand(x, true);
    fb = warp_best_cand(y, false);
    splittable = __any_sync(FULL_MASK, sp);
  }
  STAGE(25);
  if (tid != 0) return;
  const int any_split = splittable;
  // the winners' sums, formed exactly as the evaluating thread formed them
  if (rb.gain > LGB_MIN_SCORE) {
    rb.left_count = num_data - S.RC[rb.t]; rb.left_h = sum_h - S.RH[rb.t]; rb.left_g = sum_g - S.RG[rb.t];
  }
  if (two_way && fb.gain > LGB_MIN_SCORE) {
    if (fb.t >= 0) { fb.left_g = S.LG[fb.t]; fb.left_h = S.LH[fb.t]; fb.left_count = S.LC[fb.t]; }
    else { fb.left_g = S.base_g; fb.left_h = S.base_h; fb.left_count = S.base_c; }
  }
  DevSplit o;
  o.gain = LGB_MIN_SCORE; o.feature = -1; o.real_feature = -1; o.threshold = 0; o.default_left = 1;
  o.left_count = 0; o.right_count = 0; o.left_output = 0; o.right_output = 0;
  o.left_sum_g = 0; o.left_sum_h = 0; o.right_sum_g = 0; o.right_sum_h = 0;
  double sel_left_g = 0, sel_left_h = 0;
  int sel_left_cnt = 0;
  bool have = false;
  // the reference runs the reverse scan first, then the forward scan; is_splittable_ set by either stays set
  const bool rev_split = rb.gain > LGB_MIN_SCORE;
  if (rev_split && rb.gain > o.gain + min_gain_shift) {
    have = true;
    o.threshold = rb.t - 1 + offset; o.default_left = 1;
    sel_left_g = rb.left_g; sel_left_h = rb.left_h; sel_left_cnt = rb.left_count;
    o.gain = rb.gain - min_gain_shift;
  }
  if (two_way && any_split && fb.gain > o.gain + min_gain_shift) {
    have = true;
    o.threshold = fb.t + offset; o.default_left = 0;
    sel_left_g = fb.left_g; sel_left_h = fb.left_h; sel_left_cnt = fb.left_count;
    o.gain = fb.gain - min_gain_shift;
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
  // a <=2-bin feature whose mapper carries a NaN bin: FuncForNumricalL3's last branch
  if (!two_way && F.missing_type == LGBCU_MISSING_NAN) o.default_left = 0;
  o.gain *= F.penalty;  // FindBestThreshold: output->gain *= meta_->penalty
  *out = o;
  *splittable_out = any_split;
  STAGE(26);
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

// FixHistogram / Subtract / FindBestThreshold of ONE (feature, leaf) pair by one block of FS_THREADS threads.
// sel = 0: the smaller leaf; sel = 1: the larger leaf (= parent - smaller, formed here, in the parent's slot).
// What one (feature, leaf) block of k_find_split works on.  The leaves are passed by value: in training they are the
// not-yet-committed children of an expanded leaf (DevSpec), in the step-wise API an entry of the leaf table.
struct FindTarget {
  DevLeaf Ls, Ll;            // smaller leaf, larger leaf
  int has_larger;            // 0: single leaf (root / step-wise)
  int smaller_id0, larger_id0;  // the leaf carries id 0 (ComputeBestSplitForFeature's leaf_index() == 0 quirk)
  int slot;                  // which (2 * item + sel) rows of feat_best / feat_key receive the results: slot + sel
  const int8_t* parent_split;   // the expanded leaf's is_splittable flags (nullptr: none)
};

// ---- one-shot exchange (data-parallel; comm.cuh) -----------------------------------------------------------------
// k_hist has added this rank's rows into pool_local.  One launch then does what the NCCL all-reduce did, sized on the
// device (only the slots of the round's live items, which the host cannot know):
//   * publish: block 0 writes this rank's next epoch into every peer's flag array (the launch follows this rank's
//     k_hist in stream order; a system-scope fence orders the histogram adds before the flag);
//   * wait: every block polls its OWN flag array until every rank has reached that epoch -- all ranks run the same
//     live rounds, so the epochs agree.  The wait is bounded: a peer that never arrives voids the tree (xchg_error)
//     instead of hanging the GPU;
//   * pull + reduce: entry by entry, pool[slot] = sum over the ranks' pool_local[slot], the peers' read over NVLink
//     with system-scope loads.  64-bit integers: exact, the same on every rank.
// The last block to finish records the epoch.  The split search afterwards reads `pool` exactly as it does on one GPU.
#define XCHG_TIMEOUT_CYCLES 40000000000ll   /* ~20 s: a peer may be held up on its host (first-iteration set-up, a page-fault storm); once a wait has
                                               timed out the error is sticky and later launches return at once */
#define XCHG_THREADS 256
struct XchgPeers { unsigned long long* flags[LGB_MAX_PEERS]; };   // rank r's flag array as mapped here
// publish this rank's epoch `need` (block 0) and wait until every rank has reached it (every block)
__device__ __forceinline__ void xchg_publish_and_wait(const DevView& V, const XchgPeers& P, unsigned long long need) {
  DevCtl* c = V.ctl;
  if (threadIdx.x == 0) {
    if (blockIdx.x == 0) {
      __threadfence_system();
      for (int r = 0; r < V.comm_size; ++r)
        if (r != V.comm_rank) *reinterpret_cast<volatile unsigned long long*>(P.flags[r] + V.comm_rank) = need;
    }
    const long long t0 = clock64();
    for (int r = 0; r < V.comm_size; ++r) {
      if (r == V.comm_rank) continue;
      const volatile unsigned long long* fl = V.xflags + r;
      while (*fl < need) {
        if (clock64() - t0 > XCHG_TIMEOUT_CYCLES) { c->xchg_error = 1; break; }
      }
    }
    __threadfence_system();
  }
  __syncthreads();
}
__device__ __forceinline__ void xchg_finish(const DevView& V, unsigned long long need) {
  DevCtl* c = V.ctl;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();   // (step 1 of the two-step form: the reduced slice must be out before the next flag)
    const unsigned ticket = atomicAdd(&c->xchg_done, 1u);
    if (ticket == gridDim.x - 1) { c->xchg_done = 0; c->xchg_epoch = need; }
  }
}
// step: 0 = the one-step form (two ranks): every entry of the live slots = sum over the ranks' local slots;
//       1 = two-step form, first half: this rank reduces ITS 1/N slice of every live slot over all ranks, in place in
//           its pool_local (nobody else reads that slice of this rank's pool_local) and into pool;
//       2 = second half: the other slices are copied from their owners' pool_local.  (N - 1) / N of a slot is read per
//           half instead of N - 1 slots in the one-step form.
__global__ void __launch_bounds__(XCHG_THREADS) k_xchg_pull(DevView V, XchgPeers P, int step) {
  DevCtl* c = V.ctl;
  if (c->done || c->xchg_error) return;
  const bool root_round = c->round == 0;
  if (root_round && c->skip_find) return;
  const unsigned long long need = c->xchg_epoch + 1;
  xchg_publish_and_wait(V, P, need);
  const int ni = root_round ? 1 : c->num_items;
  const long long per = V.total_bin;
  const int N = V.comm_size, me = V.comm_rank;
  const long long sl = (per + N - 1) / N;   // entries per slice
  const long long total = per * ni;
  for (long long e = (long long)blockIdx.x * XCHG_THREADS + threadIdx.x; e < total; e += (long long)gridDim.x * XCHG_THREADS) {
    const int i = (int)(e / per);
    const long long off = e - (long long)i * per;
    const int owner = (int)(off / sl);
    if (step == 1 && owner != me) continue;
    if (step == 2 && owner == me) continue;
    const size_t slot = root_round ? 0 : (size_t)(1 + (c->round - 1) * c->lookahead + i);
    const size_t idx = slot * (size_t)V.total_bin + (size_t)off;
    i64 g = 0, h = 0;
    if (step == 2) {
      // (volatile: the peers' memory must not be served from a stale L1 line)
      const volatile i64* p = reinterpret_cast<const volatile i64*>(V.xpool[owner] + idx);
      g = p[0]; h = p[1];
    } else {
      for (int r = 0; r < N; ++r) {
        const volatile i64* p = reinterpret_cast<const volatile i64*>(V.xpool[r] + idx);
        g += p[0]; h += p[1];
      }
    }
    HistEntry o; o.g = g; o.h = h;
    V.pool[idx] = o;
    if (step == 1) V.pool_local[idx] = o;
  }
  xchg_finish(V, need);
}

#include "wide_split.cuh"

template <bool WIDE>
__device__ void find_item(const DevView& V, ScanSmem& S, int f, int sel, const FindTarget& T) {
  DevCtl* c = V.ctl;
  const int tid = threadIdx.x;
  if ((sel == 1 && !T.has_larger) || f >= V.num_features) return;
  {

    const DevFeat F = V.feats[f];
    const DevLeaf Ls = T.Ls;
    int used = V.feat_used[f] != 0;
    // a feature with no valid threshold in the parent is skipped in both children (SerialTreeLearner::FindBestSplits,
    // serial_tree_learner.cpp:319-326); k_partition snapshotted the parent's flags, so sibling blocks cannot race on
    // them.  The data-parallel learner overrides FindBestSplits without that mask (data_parallel_tree_learner.cpp:
    // 155-173: every by-tree feature is histogrammed and scanned for every leaf), so there the flags are not applied.
    if (T.has_larger && T.parent_split && !V.data_parallel) used = used && T.parent_split[f] != 0;
    if (WIDE && (F.wide || F.cat || V.use_mc)) { find_item_wide(V, S, f, sel, T, F, used != 0); return; }   // more than 256 stored values, or categorical: the general search
    const int nt = F.num_bin - F.offset;
    const double inv_g = c->inv_g, inv_h = c->inv_h;
    HistEntry* hs = V.pool + (size_t)Ls.slot * V.total_bin + F.hist_base;  // smaller leaf, slot 0 of the group

    double2* fix_s = V.mfb_fix + (size_t)Ls.slot * V.num_features + f;
    DevSplit o; int sp = 0;
    o.gain = LGB_MIN_SCORE; o.feature = -1; o.real_feature = -1;
    if (sel == 0) {
      if (used) {
        HistEntry e; e.g = 0; e.h = 0;
        if (tid < nt) e = hs[1 + tid];
        if (F.mfb > 0) {
          // Dataset::FixHistogram (dataset.cpp:1410-1428): data[mfb] = leaf sum - sum of the other bins.  Every
          // step of the reference's sequential subtraction is exact while the bins are (the running value only
          // shrinks), so one subtraction of the exact integer sum gives the same double, residue included.
          const bool other = tid < nt && tid != F.mfb;
          const i64 og = block_sum_i64(other ? e.g : 0, S.wtot[0]);
          const i64 oh = block_sum_i64(other ? e.h : 0, S.wtot[1]);
          const double vg = Ls.sum_g - (double)og * inv_g, vh = Ls.sum_h - (double)oh * inv_h;
          if (tid == F.mfb) {
            e = hs[0];  // exact integer twin of the most-frequent bin (rows stored as 0)
            hs[1 + F.mfb] = e;
            *fix_s = make_double2(vg, vh);
          }
          S.hd[tid] = tid == F.mfb ? vh : (double)e.h * inv_h;
        } else {
          S.hd[tid] = (double)e.h * inv_h;
        }
        S.gi[tid] = e.g;
        __syncthreads();
        scan_feature_block(S, nt, F, Ls.sum_g, Ls.sum_h, Ls.gcount, parent_output_of(Ls, T.smaller_id0 ? 0 : 1, V.p), V.p, inv_g, &o, &sp);
        o.feature = f; o.real_feature = F.real_index;
      }
      if (tid == 0) {
        V.feat_best[(size_t)T.slot * V.num_features + f] = o;
        { SplitKey k; k.gain = o.gain; k.real_feature = o.real_feature; k.pad = o.feature >= 0 ? key_pack<WIDE>(o.feature, o.threshold, o.default_left) : 0; V.feat_key[(size_t)T.slot * V.num_features + f] = k; }
        V.splittable[(size_t)Ls.slot * V.num_features + f] = (int8_t)sp;
      }
    } else {
      const DevLeaf Ll = T.Ll;
      if (used) {
        HistEntry* hl = V.pool + (size_t)Ll.slot * V.total_bin + F.hist_base;  // holds the parent
        double2* fix_l = V.mfb_fix + (size_t)Ll.slot * V.num_features + f;
        HistEntry e; e.g = 0; e.h = 0; HistEntry b; b.g = 0; b.h = 0;
        if (tid < nt) {
          const HistEntry a = hl[1 + tid];
          // the smaller leaf's most-frequent-bin entry is being written by the sibling block: take its exact twin
          // (stored value 0) instead of racing on it
          b = (F.mfb > 0 && tid == F.mfb) ? hs[0] : hs[1 + tid];
          e.g = a.g - b.g; e.h = a.h - b.h;  // FeatureHistogram::Subtract, exact
          hl[1 + tid] = e;
        }
        if (tid == FS_THREADS - 1) {  // the group's slot 0 (rows in the most-frequent bin)
          const HistEntry a0 = hl[0], b0 = hs[0];
          HistEntry e0; e0.g = a0.g - b0.g; e0.h = a0.h - b0.h;
          hl[0] = e0;
        }
        if (F.mfb > 0) {
          // larger.data[mfb] = parent.data[mfb] (fixed when the parent was scanned) - smaller.data[mfb] (fixed now; the
          // smaller leaf's FixHistogram is redone here so that this block does not wait for its sibling block)
          const bool other = tid < nt && tid != F.mfb;
          const i64 og = block_sum_i64(other ? b.g : 0, S.wtot[0]);
          const i64 oh = block_sum_i64(other ? b.h : 0, S.wtot[1]);
          const double sg_fix = Ls.sum_g - (double)og * inv_g, sh_fix = Ls.sum_h - (double)oh * inv_h;
          const double2 pf = *fix_l;
          const double lg = pf.x - sg_fix, lh = pf.y - sh_fix;
          __syncthreads();  // every thread has read *fix_l
          if (tid == F.mfb) *fix_l = make_double2(lg, lh);
          S.hd[tid] = tid == F.mfb ? lh : (double)e.h * inv_h;
        } else {
          S.hd[tid] = (double)e.h * inv_h;
        }
        S.gi[tid] = e.g;
        __syncthreads();
        STAGE(1);
        scan_feature_block(S, nt, F, Ll.sum_g, Ll.sum_h, Ll.gcount, parent_output_of(Ll, T.larger_id0 ? 0 : 1, V.p), V.p, inv_g, &o, &sp);
        o.feature = f; o.real_feature = F.real_index;
      }
      if (tid == 0) {
        V.feat_best[(size_t)(T.slot + 1) * V.num_features + f] = o;
        { SplitKey k; k.gain = o.gain; k.real_feature = o.real_feature; k.pad = o.feature >= 0 ? key_pack<WIDE>(o.feature, o.threshold, o.default_left) : 0; V.feat_key[(size_t)(T.slot + 1) * V.num_features + f] = k; }
        if (used) V.splittable[(size_t)Ll.slot * V.num_features + f] = (int8_t)sp;
      }
    }
    }
  __syncthreads();  // S is reused by the caller's next item
}

// ---- control step (last block of k_find_split) ---------------------------------------------------------------------
// 1. reduce the per-feature candidates of every scanned leaf (SplitInfo::operator>) and store each leaf's best split
//    (training: into the DevSpec of the leaf being expanded, which thereby becomes "expanded");
// 2. commit, in the reference's order (ArgMax over best_split_per_leaf_, serial_tree_learner.cpp:181-199), every leaf
//    whose turn has come and whose expansion is ready: its children enter the leaf table, the split record is written;
// 3. choose the up to K open, not yet expanded leaves with the largest positive gains as the next round's items.
// A key carries (inner feature, threshold, default_left) packed in `pad`, so an item's split descriptor is known
// without fetching its 80-byte record.
#define DEVSPLIT_WORDS ((int)(sizeof(DevSplit) / 4))
#define DEVLEAF_WORDS ((int)(sizeof(DevLeaf) / 4))

// L2 (cache-global) load of a SplitKey another block of this launch has just written: the last-block pattern orders
// the writers (store, fence, ticket) -- the reader only has to bypass its own, non-coherent L1.
__device__ __forceinline__ SplitKey load_key_cg(const SplitKey* p) {
  const int4 v = __ldcg(reinterpret_cast<const int4*>(p));
  SplitKey k;
  k.gain = __hiloint2double(v.y, v.x); k.real_feature = v.z; k.pad = v.w;
  return k;
}
__device__ __forceinline__ void store_key(SplitKey* p, const SplitKey& k) {
  int4 v;
  v.x = __double2loint(k.gain); v.y = __double2hiint(k.gain); v.z = k.real_feature; v.w = k.pad;
  *reinterpret_cast<int4*>(p) = v;
}
// one warp copies `nwords` 32-bit words (both ends 4-byte aligned, nwords <= 32)
__device__ __forceinline__ void warp_copy_words(void* dst, const void* src, int nwords, int lane) {
  if (lane < nwords) reinterpret_cast<int*>(dst)[lane] = __ldcg(reinterpret_cast<const int*>(src) + lane);
}
__device__ __forceinline__ SplitKey none_key() {
  SplitKey k; k.gain = LGB_MIN_SCORE; k.real_feature = -1; k.pad = 0;
  return k;
}
__device__ __forceinline__ void store_none_split(DevSplit* d, int lane) {
  if (lane == 0) { d->gain = LGB_MIN_SCORE; d->feature = -1; d->real_feature = -1; }
}

struct PairBest { double gain; int real_feature, idx, pad; };

// one warp: best (gain, feature) of row `p` of feat_key over all features; result in every lane (idx < 0: none)
__device__ PairBest warp_reduce_feature_keys(const DevView& V, int p, int lane, int f_begin = 0, int f_end = -1) {
  const int F = V.num_features;
  if (f_end < 0) f_end = F;
  double bg = LGB_MIN_SCORE; int bf = -1, bi = -1, bp = 0;
  for (int j0 = f_begin + lane; j0 < f_end; j0 += 4 * 32) {
    SplitKey k[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = j0 + u * 32;
      k[u] = none_key();
      if (j < f_end) k[u] = load_key_cg(&V.feat_key[(size_t)p * F + j]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = j0 + u * 32;
      if (j < f_end && (bi < 0 || split_better(k[u].gain, k[u].real_feature, bg, bf))) {
        bg = k[u].gain; bf = k[u].real_feature; bp = k[u].pad; bi = j;
      }
    }
  }
  const unsigned tie = bf == -1 ? 0x7fffffffu : (unsigned)bf;
  int w = warp_argmax(bg, tie, 0u, bi >= 0 && bg == bg);
  if (w < 0) w = __ffs(__ballot_sync(FULL_MASK, bi >= 0)) - 1;  // only NaN gains: the first feature
  PairBest r; r.gain = LGB_MIN_SCORE; r.real_feature = -1; r.idx = -1; r.pad = 0;
  if (w < 0) return r;
  r.gain = __shfl_sync(FULL_MASK, bg, w); r.real_feature = __shfl_sync(FULL_MASK, bf, w);
  r.idx = __shfl_sync(FULL_MASK, bi, w); r.pad = __shfl_sync(FULL_MASK, bp, w);
  return r;
}

// one warp: store the winner of feat_best row p as the best split of one leaf
__device__ void store_leaf_best(const DevView& V, int p, const PairBest& b, DevSplit* dst_split, SplitKey* dst_key, int lane) {
  if (b.idx >= 0) warp_copy_words(dst_split, V.feat_best + ((size_t)p * V.num_features + b.idx), DEVSPLIT_WORDS, lane);
  else store_none_split(dst_split, lane);
  if (lane == 0) {
    SplitKey k; k.gain = b.idx >= 0 ? b.gain : LGB_MIN_SCORE; k.real_feature = b.idx >= 0 ? b.real_feature : -1; k.pad = b.idx >= 0 ? b.pad : 0;
    store_key(dst_key, k);
  }
}

// warp 0: arg-max over the open leaves.  want_expanded < 0: all leaves; otherwise only leaves whose expanded flag
// equals want_expanded and, if `taken` is given, that are not listed in it.  Returns (leaf, key) in every lane.
__device__ int best_open_leaf(const DevView& V, int nl, int want_expanded, const int* taken, int ntaken, SplitKey* out, int lane) {
  double bg = LGB_MIN_SCORE; int bf = -1, bl = -1, bp = 0;
  for (int l = lane; l < nl; l += 32) {
    if (want_expanded >= 0 && __ldcg(&V.expanded[l]) != want_expanded) continue;
    bool skip = false;
    for (int t = 0; t < ntaken; ++t) skip = skip || taken[t] == l;
    if (skip) continue;
    const SplitKey k = load_key_cg(&V.leaf_key[l]);
    if (bl < 0 || split_better(k.gain, k.real_feature, bg, bf)) { bg = k.gain; bf = k.real_feature; bl = l; bp = k.pad; }
  }
  const unsigned tie = bf == -1 ? 0x7fffffffu : (unsigned)bf;
  const int w = warp_argmax(bg, tie, (unsigned)bl, bl >= 0 && bg == bg);  // first leaf with the best (gain, feature)
  if (w < 0) return -1;
  out->gain = __shfl_sync(FULL_MASK, bg, w); out->real_feature = __shfl_sync(FULL_MASK, bf, w);
  out->pad = __shfl_sync(FULL_MASK, bp, w);
  return __shfl_sync(FULL_MASK, bl, w);
}

// warp 0: the expansion of `leaf` becomes part of the tree (SerialTreeLearner::SplitInner + Tree::Split bookkeeping)
// one warp: expansion of `leaf` -> leaf table, with the right child's id R and the record index given
__device__ void commit_leaf_as(const DevView& V, int leaf, int R, int record_index, int lane) {
  DevSpec* S = &V.spec[leaf];
  if (lane == 0) {
    const DevSplit s = V.best[leaf];  // the split that was applied
    LgbCuSplitRecord rec;
    rec.leaf = leaf; rec.right_leaf = R; rec.feature = s.feature; rec.real_feature = s.real_feature;
    rec.threshold = s.threshold; rec.default_left = s.default_left;
    rec.left_count = S->child[0].gcount; rec.right_count = S->child[1].gcount;
    rec.gain = (float)(s.gain + V.p.min_gain_to_split); rec.reserved = 0;
    rec.left_output = s.left_output; rec.right_output = s.right_output;
    rec.left_sum_hessian = s.left_sum_h; rec.right_sum_hessian = s.right_sum_h;
    rec.left_sum_gradient = s.left_sum_g; rec.right_sum_gradient = s.right_sum_g;
    V.records[record_index] = rec;
  }
  __syncwarp();
  warp_copy_words(&V.leaves[leaf], &S->child[0], DEVLEAF_WORDS, lane);
  warp_copy_words(&V.leaves[R], &S->child[1], DEVLEAF_WORDS, lane);
  warp_copy_words(&V.best[leaf], &S->best[0], DEVSPLIT_WORDS, lane);
  warp_copy_words(&V.best[R], &S->best[1], DEVSPLIT_WORDS, lane);
  warp_copy_words(&V.leaf_key[leaf], &S->key[0], 4, lane);
  warp_copy_words(&V.leaf_key[R], &S->key[1], 4, lane);
  if (lane == 0) { V.expanded[leaf] = 0; V.expanded[R] = 0; }
}

__device__ void commit_leaf(const DevView& V, int leaf, int lane) {
  DevCtl* c = V.ctl;
  const int R = c->num_leaves;
  DevSpec* S = &V.spec[leaf];
  if (lane == 0) {
    const DevSplit s = V.best[leaf];  // the split that was applied
    LgbCuSplitRecord rec;
    rec.leaf = leaf; rec.right_leaf = R; rec.feature = s.feature; rec.real_feature = s.real_feature;
    rec.threshold = s.threshold; rec.default_left = s.default_left;
    rec.left_count = S->child[0].gcount; rec.right_count = S->child[1].gcount;
    rec.gain = (float)(s.gain + V.p.min_gain_to_split); rec.reserved = 0;
    rec.left_output = s.left_output; rec.right_output = s.right_output;
    rec.left_sum_hessian = s.left_sum_h; rec.right_sum_hessian = s.right_sum_h;
    rec.left_sum_gradient = s.left_sum_g; rec.right_sum_gradient = s.right_sum_g;
    V.records[c->num_splits] = rec;
  }
  __syncwarp();
  warp_copy_words(&V.leaves[leaf], &S->child[0], DEVLEAF_WORDS, lane);
  warp_copy_words(&V.leaves[R], &S->child[1], DEVLEAF_WORDS, lane);
  warp_copy_words(&V.best[leaf], &S->best[0], DEVSPLIT_WORDS, lane);
  warp_copy_words(&V.best[R], &S->best[1], DEVSPLIT_WORDS, lane);
  warp_copy_words(&V.leaf_key[leaf], &S->key[0], 4, lane);
  warp_copy_words(&V.leaf_key[R], &S->key[1], 4, lane);
  if (lane == 0) {
    V.expanded[leaf] = 0; V.expanded[R] = 0;
    c->num_splits += 1;
    c->num_leaves = R + 1;
  }
  __threadfence_block();
  __syncwarp();
}

// warp 0, training: steps 2 and 3 of the control step -- general version, leaf keys re-read from memory every pass
template <bool WIDE>
__device__ void commit_and_select_global(const DevView& V, int lane) {
  DevCtl* c = V.ctl;
  int nl = c->num_leaves;
  const int max_leaves = V.p.num_leaves;
  bool done = false;
  SplitKey k;
  // 2. commit
  while (true) {
    const int bl = best_open_leaf(V, nl, -1, nullptr, 0, &k, lane);
    if (bl < 0 || !(k.gain > 0.0)) { done = true; if (lane == 0) c->done = 1; break; }  // "No further splits with positive gain"
    if (__ldcg(&V.expanded[bl]) == 0) break;                                              // its children are not known yet
    commit_leaf(V, bl, lane);
    nl += 1;
    if (nl >= max_leaves) { done = true; if (lane == 0) c->done = 2; break; }             // the tree is full
  }
  // 3. select the next round's items
  int taken[LGB_KMAX];
  int ni = 0;
  if (!done) {
    int K = c->lookahead;
    if (K > max_leaves - nl) K = max_leaves - nl;
    for (int e = 0; e < K; ++e) {
      const int bl = best_open_leaf(V, nl, 0, taken, ni, &k, lane);
      if (bl < 0 || !(k.gain > 0.0)) break;
      taken[ni] = bl;
      if (lane == 0) {
        c->item_leaf[ni] = bl;
        key_unpack<WIDE>(k.pad, &c->item_feature[ni], &c->item_threshold[ni], &c->item_default_left[ni]);
      }
      ++ni;
    }
  }
  if (lane == 0) { c->num_items = ni; c->round += 1; }
}

// Control step, trees of up to 256 leaves.  Leaf keys are handled as integers: `gk` = order-preserving 64-bit image
// of the gain (larger = better), `tie` = real feature << 16 | leaf (smaller = better; -1 counts as INT_MAX as in
// SplitInfo::operator>), so "a beats b" is two integer compares.
#define CTRL_SLOTS 8
#define CTRL_MAX_LEAVES (32 * CTRL_SLOTS)
struct CtrlShared {
  SplitKey ck[CTRL_MAX_LEAVES][2];          // children's keys of every expanded leaf (staged by the whole block)
  unsigned long long gk[CTRL_MAX_LEAVES];    // after the commit loop: final key of every open leaf ...
  unsigned long long tie[CTRL_MAX_LEAVES];
  int pad[CTRL_MAX_LEAVES];                  // ... its packed split descriptor
  int eligible[CTRL_MAX_LEAVES];             // ... and whether it may be expanded next (open, not expanded, gain > 0)
  int ex[CTRL_MAX_LEAVES], pos[CTRL_MAX_LEAVES];  // expanded flag, gain > 0
  int commit[CTRL_MAX_LEAVES];               // leaves to commit, in order
  int nc, done, nl;
};
__device__ __forceinline__ unsigned long long gain_key64(double g) {
  unsigned long long u = (unsigned long long)__double_as_longlong(g);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ unsigned long long tie_key64(int real_feature, int leaf) {
  return ((unsigned long long)(real_feature == -1 ? 0x7fffffffu : (unsigned)real_feature) << 16) | (unsigned)leaf;
}

// all threads: thread l loads the key of leaf l into the shared arrays
__device__ void ctrl_load_keys(const DevView& V, CtrlShared& C, int nl) {
  const int l = threadIdx.x;  // FS_THREADS == CTRL_MAX_LEAVES
  C.gk[l] = 0; C.tie[l] = ~0ull; C.pad[l] = 0; C.ex[l] = 0; C.pos[l] = 0;
  if (l < nl) {
    const SplitKey key = load_key_cg(&V.leaf_key[l]);
    C.gk[l] = gain_key64(key.gain); C.tie[l] = tie_key64(key.real_feature, l); C.pad[l] = key.pad; C.pos[l] = key.gain > 0.0;
    const int ex = __ldcg(&V.expanded[l]);
    C.ex[l] = ex;
    if (ex) { C.ck[l][0] = load_key_cg(&V.spec[l].key[0]); C.ck[l][1] = load_key_cg(&V.spec[l].key[1]); }
  }
}

// warp 0: step 2 -- which leaves are committed, in the reference's order (ArgMax over best_split_per_leaf_).  The
// keys live in shared memory (leaf = lane + 32 * slot: conflict-free); a pass is 8 local integer compares and one warp
// arg-max; a commit overwrites the two entries it changes with the children's keys.
__device__ void plan_commits(const DevView& V, int lane, CtrlShared& C) {
  DevCtl* c = V.ctl;
  const int max_leaves = V.p.num_leaves;
  int nl = c->num_leaves;
  int done = 0, nc = 0;
  while (true) {
    unsigned long long bg = 0, bt = ~0ull; int bl = -1;
#pragma unroll
    for (int k = 0; k < CTRL_SLOTS; ++k) {
      const int l = lane + 32 * k;
      if (l < nl) {
        const unsigned long long g = C.gk[l], t = C.tie[l];
        if (bl < 0 || g > bg || (g == bg && t < bt)) { bg = g; bt = t; bl = l; }
      }
    }
    const unsigned ghi = (unsigned)(bg >> 32), glo = (unsigned)bg;
    bool a = bl >= 0;
    const unsigned m1 = __reduce_max_sync(FULL_MASK, a ? ghi : 0u);
    a = a && ghi == m1;
    unsigned bal = __ballot_sync(FULL_MASK, a);
    if (bal & (bal - 1)) {
      const unsigned m2 = __reduce_max_sync(FULL_MASK, a ? glo : 0u);
      a = a && glo == m2;
      bal = __ballot_sync(FULL_MASK, a);
      if (bal & (bal - 1)) {
        const unsigned m3 = __reduce_min_sync(FULL_MASK, a ? (unsigned)(bt >> 32) : 0xffffffffu);
        a = a && (unsigned)(bt >> 32) == m3;
        const unsigned m4 = __reduce_min_sync(FULL_MASK, a ? (unsigned)bt : 0xffffffffu);
        a = a && (unsigned)bt == m4;
        bal = __ballot_sync(FULL_MASK, a);
      }
    }
    if (bal == 0) { done = 1; break; }
    bl = __shfl_sync(FULL_MASK, bl, __ffs(bal) - 1);
    if (!C.pos[bl]) { done = 1; break; }   // "No further splits with positive gain"
    if (!C.ex[bl]) break;                  // its children are not known yet
    const int R = nl;
    if (lane == 0) {
      C.commit[nc] = bl;
      const SplitKey k0 = C.ck[bl][0], k1 = C.ck[bl][1];
      C.gk[bl] = gain_key64(k0.gain); C.tie[bl] = tie_key64(k0.real_feature, bl); C.pad[bl] = k0.pad; C.pos[bl] = k0.gain > 0.0; C.ex[bl] = 0;
      C.gk[R] = gain_key64(k1.gain); C.tie[R] = tie_key64(k1.real_feature, R); C.pad[R] = k1.pad; C.pos[R] = k1.gain > 0.0; C.ex[R] = 0;
    }
    __syncwarp();
    ++nc;
    nl += 1;
    if (nl >= max_leaves) { done = 2; break; }  // the tree is full
  }
#pragma unroll
  for (int k = 0; k < CTRL_SLOTS; ++k) {
    const int l = lane + 32 * k;
    C.eligible[l] = (l < nl && !done && C.pos[l] && C.ex[l] == 0) ? 1 : 0;
  }
  if (lane == 0) { C.nc = nc; C.done = done; C.nl = nl; }
}

// Last act of the control step (all threads of the block, after the commits have been written): the leaf records and
// split-feature descriptors of the next round's items go next to the item list, for k_partition's preamble.
__device__ void publish_items(const DevView& V) {
  DevCtl* c = V.ctl;
  __threadfence_block();
  __syncthreads();
  if (c->done) return;
  const int ni = c->num_items;
  const int i = threadIdx.x >> 4, w = threadIdx.x & 15;   // 16 threads per item, one 4-byte word each per pass
  if (i < ni && i < LGB_KMAX) {
    const int leaf = c->item_leaf[i];
    static_assert(sizeof(DevLeaf) / 4 <= 16 && sizeof(DevFeat) / 4 <= 16, "16 threads copy one item's records");
    warp_copy_words(&c->item_rec[i], &V.leaves[leaf], (int)(sizeof(DevLeaf) / 4), w);
    warp_copy_words(&c->item_feat[i], &V.feats[c->item_feature[i]], (int)(sizeof(DevFeat) / 4), w);
  }
}

// mode 0: training.  grid = (num_features, 2 * LGB_KMAX): blockIdx.y = 2 * item + sel; sel 0 scans the smaller child of
//         the item's leaf, sel 1 forms larger = parent - smaller and scans it.  Round 0: only (item 0, sel 0) = root.
// mode 1: one explicit leaf of the leaf table (step-wise API / tests), grid = (num_features, 1); result -> best[leaf].
// The last block to finish runs the control step.
template <bool WIDE>
__global__ void __launch_bounds__(FS_THREADS, 4) k_find_split(DevView V, int mode, int leaf_arg) {
  __shared__ ScanSmem S;
  __shared__ int s_last;
  DevCtl* c = V.ctl;
  if (mode != 1 && c->done) return;
  STAGE_BEGIN();
  STAMP(c, 1);
  // training launches scan this rank's features only: all of them, or its block in ownership mode (comm.cuh)
  const int f = (mode == 1 ? 0 : V.f_begin) + blockIdx.x, f_end = mode == 1 ? V.num_features : V.f_end;
  const int y = blockIdx.y, e = y >> 1, sel = y & 1;
  const bool root_round = mode != 1 && c->round == 0;
  const int ni = mode != 1 ? (root_round ? 1 : c->num_items) : 1;
  FindTarget T;
  bool active = false;
  if (mode == 1 || root_round) {
    if (y == 0 && (mode == 1 || !c->skip_find)) {
      const int leaf = mode == 1 ? leaf_arg : 0;
      T.Ls = V.leaves[leaf]; T.Ll = T.Ls; T.has_larger = 0; T.smaller_id0 = leaf == 0; T.larger_id0 = 0; T.slot = 0;
      T.parent_split = nullptr;
      active = true;
    }
  } else if (e < ni) {
    const int parent = c->item_leaf[e];
    const DevSpec* sp = &V.spec[parent];
    if (!sp->skip_find) {
      const int sm = sp->smaller;
      T.Ls = sp->child[sm]; T.Ll = sp->child[1 - sm]; T.has_larger = 1;
      // the left child keeps the parent's id, the right child gets a new id >= 1
      T.smaller_id0 = parent == 0 && sm == 0; T.larger_id0 = parent == 0 && sm == 1;
      T.slot = 2 * e; T.parent_split = V.parent_split + (size_t)e * V.num_features;
      active = true;
    }
  }
  STAGE(0);
  if (mode != FS_MODE_CONTROL) {
    if (active && f < f_end) find_item<WIDE>(V, S, f, sel, T);
    STAGE(4);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned ticket = atomicAdd(&c->feat_done, 1u);
      s_last = (ticket == gridDim.x * gridDim.y - 1);
    }
    __syncthreads();
    STAGE(5);
    if (!s_last) return;
  }
  // ---- control step, last block only.  Step 1: every warp reduces feat_key rows (one leaf each) and stores that
  // leaf's best split where it belongs.
  __shared__ CtrlShared C;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#ifdef LGBCU_STAGE_TIMERS
  long long tc0 = clock64();
  long long tq_ = tc0;
#define QSTAGE(i) do { if (threadIdx.x == 0) { long long t_ = clock64(); g_stage_cycles[16 + (i)] += (unsigned long long)(t_ - tq_); tq_ = t_; } } while (0)
#else
#define QSTAGE(i) do {} while (0)
#endif
  if (threadIdx.x == 0 && mode != FS_MODE_CONTROL) c->feat_done = 0;
  if (mode == FS_MODE_OWNED) {
    // Ownership mode, first half: this rank's winners over ITS features go to the send buffer of the all-gather; the
    // control step runs in a second launch (FS_MODE_CONTROL) once every rank's winners have arrived.
    const int rows = root_round ? 1 : 2 * ni;
    for (int p = warp; p < rows; p += FS_WARPS) {
      bool live;
      if (root_round) live = !c->skip_find;
      else live = !V.spec[c->item_leaf[p >> 1]].skip_find;
      CandRec* dst = &V.cand_send[p];
      if (live) {
        const PairBest b = warp_reduce_feature_keys(V, p, lane, V.f_begin, V.f_end);
        store_leaf_best(V, p, b, &dst->split, &dst->key, lane);
      } else {
        store_none_split(&dst->split, lane);
        if (lane == 0) store_key(&dst->key, none_key());
      }
    }
    return;
  }
  if (mode == FS_MODE_CONTROL) {
    // Ownership mode, second half (one block): per (leaf, child) row the best of the ranks' winners -- the same
    // SplitInfo::operator> order, and the ranks' features are disjoint, so the winner is the one a scan of all
    // features would have found (SyncUpGlobalBestSplit, parallel_tree_learner.h:190-213).
    const int rows = root_round ? 1 : 2 * ni;
    for (int p = warp; p < rows; p += FS_WARPS) {
      DevSplit* dsplit; SplitKey* dkey;
      if (root_round) {
        if (c->skip_find) continue;
        dsplit = &V.best[0]; dkey = &V.leaf_key[0];
      } else {
        DevSpec* sp = &V.spec[c->item_leaf[p >> 1]];
        const int child = (p & 1) == 0 ? sp->smaller : 1 - sp->smaller;
        dsplit = &sp->best[child]; dkey = &sp->key[child];
      }
      SplitKey k = none_key();
      if (lane < V.comm_size) k = load_key_cg(&V.cand_all[(size_t)lane * (2 * LGB_KMAX) + p].key);
      const unsigned tie = k.real_feature == -1 ? 0x7fffffffu : (unsigned)k.real_feature;
      int w = warp_argmax(k.gain, tie, (unsigned)lane, lane < V.comm_size && k.gain == k.gain);
      if (w < 0) w = 0;  // only NaN gains: rank 0's record
      const CandRec* src = &V.cand_all[(size_t)w * (2 * LGB_KMAX) + p];
      warp_copy_words(dsplit, &src->split, DEVSPLIT_WORDS, lane);
      if (lane == 0) store_key(dkey, load_key_cg(&src->key));
    }
  } else if (mode == 1 || root_round) {
    const int leaf = mode == 1 ? leaf_arg : 0;
    if (warp == 0 && (mode == 1 || !c->skip_find)) {
      const PairBest b = warp_reduce_feature_keys(V, 0, lane);
      store_leaf_best(V, 0, b, &V.best[leaf], &V.leaf_key[leaf], lane);
    }
  } else {
    // Wide data with few items (K is small there): several warps share a row of feat_key, a slice of the features
    // each, and the row's first warp combines their winners in slice order -- the same total order (gain, then the
    // smaller real feature, then the earlier feature), so the same winner.
    int wpr = 1;
    if (V.num_features > 256) { while (wpr * 2 * (2 * ni) <= FS_WARPS) wpr *= 2; }
    if (wpr > 1) {
      __shared__ PairBest s_part[FS_WARPS];
      const int p = warp / wpr, sl = warp % wpr;
      DevSpec* sp = p < 2 * ni ? &V.spec[c->item_leaf[p >> 1]] : nullptr;
      const bool live = sp && !sp->skip_find;
      if (live) {
        const int per = (V.num_features + wpr - 1) / wpr;
        const int f0 = sl * per, f1 = min(V.num_features, f0 + per);
        const PairBest b = warp_reduce_feature_keys(V, p, lane, f0, f1);
        if (lane == 0) s_part[warp] = b;
      }
      __syncthreads();
      if (sp && sl == 0) {
        const int child = (p & 1) == 0 ? sp->smaller : 1 - sp->smaller;
        if (!live) {
          store_none_split(&sp->best[child], lane);
          if (lane == 0) store_key(&sp->key[child], none_key());
        } else {
          PairBest b = s_part[warp];
          for (int k = 1; k < wpr; ++k) {
            const PairBest o = s_part[warp + k];
            if (o.idx >= 0 && o.gain == o.gain && (b.idx < 0 || b.gain != b.gain || split_better(o.gain, o.real_feature, b.gain, b.real_feature))) b = o;
          }
          store_leaf_best(V, p, b, &sp->best[child], &sp->key[child], lane);
        }
      }
    } else
    for (int p = warp; p < 2 * ni; p += FS_WARPS) {
      const int i = p >> 1, which = p & 1;  // which = 0: the smaller child's row, 1: the larger child's
      DevSpec* sp = &V.spec[c->item_leaf[i]];
      const int child = which == 0 ? sp->smaller : 1 - sp->smaller;
      if (sp->skip_find) {
        store_none_split(&sp->best[child], lane);
        if (lane == 0) store_key(&sp->key[child], none_key());
      } else {
        const PairBest b = warp_reduce_feature_keys(V, p, lane);
        store_leaf_best(V, p, b, &sp->best[child], &sp->key[child], lane);
      }
    }
  }
  __threadfence_block();
  __syncthreads();
  QSTAGE(0);
  if (mode == 1) return;
  if (!root_round && threadIdx.x < ni) V.expanded[c->item_leaf[threadIdx.x]] = 1;
  __threadfence_block();
  __syncthreads();
  const int nl0 = c->num_leaves, ns0 = c->num_splits;
  if (V.p.num_leaves > CTRL_MAX_LEAVES) {  // big trees: the general, sequential version
    if (warp == 0) commit_and_select_global<WIDE>(V, lane);
    publish_items(V);
    return;
  }
  // step 2: stage every open leaf's key (and the children's keys of the expanded ones), let warp 0 decide the commits
  ctrl_load_keys(V, C, nl0);
  __syncthreads();
  QSTAGE(1);
  if (warp == 0) plan_commits(V, lane, C);
  __syncthreads();
  QSTAGE(2);
  // step 3, all threads: the next round's items = the K eligible leaves with the best keys; a leaf's item index is the
  // number of eligible leaves that beat it
  {
    int K = c->lookahead;
    if (K > V.p.num_leaves - C.nl) K = V.p.num_leaves - C.nl;
    const int l = threadIdx.x;  // FS_THREADS == CTRL_MAX_LEAVES
    const int mine = C.eligible[l];
    int rank = 0;
    // the eligible leaves' keys are compacted first (late in a tree they are a handful of 255 leaves), so that the
    // ranking loop is a dense, pipelined sweep
    __shared__ unsigned s_elig[FS_WARPS];
    __shared__ unsigned long long s_gk[CTRL_MAX_LEAVES], s_tie[CTRL_MAX_LEAVES];
    const unsigned long long g = C.gk[l], t = C.tie[l];
    const unsigned eb = __ballot_sync(FULL_MASK, mine != 0);
    if (lane == 0) s_elig[warp] = eb;
    __syncthreads();
    int n_el = 0, before = 0;
#pragma unroll
    for (int w = 0; w < FS_WARPS; ++w) {
      const int pc = __popc(s_elig[w]);
      before += w < warp ? pc : 0;
      n_el += pc;
    }
    if (mine) {
      const int pos = before + __popc(eb & ((1u << lane) - 1u));
      s_gk[pos] = g; s_tie[pos] = t;
    }
    __syncthreads();
    if (mine) {
#pragma unroll 4
      for (int m = 0; m < n_el; ++m) {
        const unsigned long long gm = s_gk[m], tm = s_tie[m];
        rank += (gm > g || (gm == g && tm < t)) ? 1 : 0;
      }
      if (rank < K) {
        const int pd = C.pad[l];
        c->item_leaf[rank] = l;
        key_unpack<WIDE>(pd, &c->item_feature[rank], &c->item_threshold[rank], &c->item_default_left[rank]);
      }
    }
    const int n_elig = n_el;
    if (threadIdx.x == 0) {
      c->num_items = n_elig < K ? n_elig : (K > 0 ? K : 0);
      c->round += 1;
      if (C.done) c->done = C.done;
    }
  }
  QSTAGE(3);
  const int nc = C.nc;
  for (int i = warp; i < nc; i += FS_WARPS) commit_leaf_as(V, C.commit[i], nl0 + i, ns0 + i, lane);
  QSTAGE(4);
  if (threadIdx.x == 0) { c->num_leaves = nl0 + nc; c->num_splits = ns0 + nc; }
  publish_items(V);
  QSTAGE(5);
#ifdef LGBCU_STAGE_TIMERS
  if (threadIdx.x == 0) { g_stage_cycles[6] += (unsigned long long)(clock64() - tc0); g_stage_cycles[7] += 1; }
#endif
}

// Root LeafSplits::Init (leaf_splits.hpp:66-79; all-reduced over ranks in data_parallel_tree_learner.cpp:126-151):
// sum over every stored value of feature 0 of the (already rank-reduced) root histogram -- exact in fixed point.
__device__ void root_sums_warp(const DevView& V) {  // one warp
  const DevCtl* c = V.ctl;
  DevLeaf* L = &V.leaves[0];
  const DevFeat F = V.feats[0];
  const HistEntry* h = V.pool + (size_t)L->slot * V.total_bin + F.hist_base;
  const int lane = threadIdx.x & 31;
  i64 g = 0, hh = 0;
  for (int s = lane; s < F.nslots; s += 32) { g += h[s].g; hh += h[s].h; }
  g = warp_sum(g); hh = warp_sum(hh);
  if (lane == 0) {
    L->sum_g = (double)g * c->inv_g;
    L->sum_h = (double)hh * c->inv_h;
  }
}
__global__ void k_root_sums(DevView V) {
  if (threadIdx.x < 32) root_sums_warp(V);
}
