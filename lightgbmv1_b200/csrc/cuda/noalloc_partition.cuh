// lgbcu_partition_rows_noalloc -- a stable row partition that allocates nothing.
//
// Who this is for: code that is not allowed to obtain device memory at the moment it needs a partition.  Examples
// are a host thread that runs under a memory budget owned by another component (an inference server or a
// dataframe engine that hands this library a fixed arena and forbids cudaMalloc behind its back, because an
// allocation may fail late, fragment the arena's neighbours or implicitly synchronise the whole device), and code
// on a latency-critical path where cudaMalloc's device-wide synchronisation is itself the problem.  Such a caller
// already owns the input and output buffers; what it cannot supply is the kernel's working memory.
//
// The working memory (one decision bit per row plus one counter per 1024-row tile) therefore lives in a
// file-scope __device__ array inside this library.  Consequences a caller has to know:
//   * every GPU that loads the library reserves NOALLOC_SCRATCH_BYTES (about 0.53 MB) until the process exits;
//   * at most LGBCU_NOALLOC_MAX_ROWS rows (4,194,304) per call -- larger inputs are refused, not truncated;
//   * all calls share the one array: two calls whose kernels may overlap on the same GPU (different streams,
//     different host threads) race on it.  Serialise them, e.g. by using one stream;
//   * the array keeps whatever the previous call left there.  Nothing in it needs to start at zero: every word a
//     call reads (bit masks and tile counters of its own tiles, the running total) is written earlier in that same
//     call, so no clearing pass exists or is needed;
//   * the first call on a given device looks up the array's device address (cudaGetSymbolAddress, with that
//     device current).  Make that first call outside of any stream capture.
// The entry point issues kernel launches only: no cudaMalloc / cudaFree / cudaMallocAsync, no library call that
// allocates, no host allocation either (errors are formatted into a fixed thread-local buffer).
#pragma once

// LGBCU_NOALLOC_MAX_ROWS comes from lgbcu.h
#define NOALLOC_MAX_TILES (LGBCU_NOALLOC_MAX_ROWS / PART_TILE)
#define NOALLOC_BITS_WORDS (LGBCU_NOALLOC_MAX_ROWS / 32)
// layout of the scratch array, in 32-bit words: [0, 64) control, then tile counters, then decision bits
#define NOALLOC_CTL_WORDS 64
#define NOALLOC_SCRATCH_WORDS (NOALLOC_CTL_WORDS + NOALLOC_MAX_TILES + NOALLOC_BITS_WORDS)
#define NOALLOC_SCRATCH_BYTES (NOALLOC_SCRATCH_WORDS * 4)

// In host code this name is only a handle for cudaGetSymbolAddress -- never pass it to a kernel.
static __device__ __align__(256) unsigned int g_noalloc_scratch[NOALLOC_SCRATCH_WORDS];

namespace {

__global__ void __launch_bounds__(PART_THREADS) k_na_flags(const uint8_t* __restrict__ col, long long row_stride,
                                                           SplitRule rule, const int* __restrict__ src, int n,
                                                           unsigned* __restrict__ bits, int* __restrict__ tile_left) {
  const int ntiles = (n + PART_TILE - 1) / PART_TILE;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ int s_cnt[PART_THREADS / 32];
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int lefts = 0;
#pragma unroll
    for (int r = 0; r < PART_REPS; ++r) {
      const int pos = tile * PART_TILE + r * PART_THREADS + threadIdx.x;
      bool left = false;
      if (pos < n) left = goes_left(rule, col[(size_t)src[pos] * row_stride]);
      const unsigned word = __ballot_sync(0xffffffffu, left);
      if (lane == 0) {
        bits[(size_t)tile * (PART_TILE / 32) + r * (PART_THREADS / 32) + warp] = word;
        lefts += __popc(word);
      }
    }
    if (lane == 0) s_cnt[warp] = lefts;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < PART_THREADS / 32; ++w) t += s_cnt[w];
      tile_left[tile] = t;
    }
    __syncthreads();
  }
}

// one block: exclusive scan of the tile counters in place; total -> ctl[0] and the caller's left_count
__global__ void __launch_bounds__(1024) k_na_scan(int* __restrict__ tile_left, int n, unsigned* __restrict__ ctl,
                                                  int* __restrict__ left_count_out) {
  const int ntiles = (n + PART_TILE - 1) / PART_TILE;
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < ntiles; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < ntiles ? tile_left[i] : 0;
    int inc = v;
    for (int o = 1; o < 32; o <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      const int w = s_warp[lane];
      int winc = w;
      for (int o = 1; o < 32; o <<= 1) {
        const int x = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += x;
      }
      s_warp[lane] = winc - w;
    }
    __syncthreads();
    const int excl = s_carry + s_warp[warp] + inc - v;
    if (i < ntiles) tile_left[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    ctl[0] = (unsigned)s_carry;
    *left_count_out = s_carry;
  }
}

__global__ void __launch_bounds__(PART_THREADS) k_na_scatter(const int* __restrict__ src, int n,
                                                             const unsigned* __restrict__ bits,
                                                             const int* __restrict__ tile_off,
                                                             const unsigned* __restrict__ ctl, int* __restrict__ dst) {
  const int left_total = (int)ctl[0];
  const int ntiles = (n + PART_TILE - 1) / PART_TILE;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const unsigned myword = bits[(size_t)tile * (PART_TILE / 32) + lane];
    const int pc = __popc(myword);
    int inc = pc;
    for (int o = 1; o < 32; o <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += x;
    }
    const int excl = inc - pc;
    const int off = tile_off[tile];
#pragma unroll
    for (int r = 0; r < PART_REPS; ++r) {
      const int w = r * (PART_THREADS / 32) + warp;
      const unsigned word = __shfl_sync(0xffffffffu, myword, w);
      const int wpre = __shfl_sync(0xffffffffu, excl, w);
      const int pos = tile * PART_TILE + w * 32 + lane;
      if (pos < n) {
        const unsigned lt = (1u << lane) - 1u;
        const int row = src[pos];
        if ((word >> lane) & 1u) dst[off + wpre + __popc(word & lt)] = row;
        else dst[left_total + (tile * PART_TILE - off) + (w * 32 - wpre) + __popc(~word & lt)] = row;
      }
    }
  }
}

// device address of the scratch array, looked up once per device
// (an atomic pointer per device: concurrent first calls may both resolve it, and both store the same value; a
// cudaDeviceReset() invalidates the cached address -- this library does not support being used across one)
std::atomic<unsigned int*> g_na_addr[64];

int noalloc_scratch(unsigned int** out) {
  int dev = 0;
  CU(cudaGetDevice(&dev));
  NEED(dev >= 0 && dev < 64, "lgbcu_partition_rows_noalloc: device index %d out of range", dev);
  unsigned int* cached = g_na_addr[dev].load(std::memory_order_acquire);
  if (!cached) {
    void* p = nullptr;
    // with `dev` current this yields THAT device's copy of the array; the host-side symbol itself is not an address
    CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    NEED(p != nullptr, "lgbcu_partition_rows_noalloc: cudaGetSymbolAddress returned NULL");
    cached = static_cast<unsigned int*>(p);
    g_na_addr[dev].store(cached, std::memory_order_release);
  }
  *out = cached;
  return 0;
}

}  // namespace

extern "C" int lgbcu_partition_rows_noalloc(const uint8_t* bins_dev, int64_t row_stride, int32_t column,
                                            const LgbCuFeatureMeta* meta, int32_t threshold, int32_t default_left,
                                            const int32_t* indices_in_dev, int32_t n, int32_t* indices_out_dev,
                                            int32_t* left_count_dev, void* stream) {
  NEED(bins_dev && meta && indices_in_dev && indices_out_dev && left_count_dev, "lgbcu_partition_rows_noalloc: NULL argument");
  NEED(n >= 0, "lgbcu_partition_rows_noalloc: negative row count");
  NEED(n <= LGBCU_NOALLOC_MAX_ROWS, "lgbcu_partition_rows_noalloc: %d rows exceed the built-in scratch (at most %d rows "
       "per call); partition in pieces or use lgbcu_partition", n, LGBCU_NOALLOC_MAX_ROWS);
  NEED(row_stride > 0 && column >= 0 && column < row_stride, "lgbcu_partition_rows_noalloc: column %d outside the row stride", column);
  NEED(indices_in_dev + n <= indices_out_dev || indices_out_dev + n <= indices_in_dev,
       "lgbcu_partition_rows_noalloc: input and output index buffers must not overlap");
  NEED(meta->num_bin >= 1 && meta->most_freq_bin >= 0 && meta->most_freq_bin < meta->num_bin, "lgbcu_partition_rows_noalloc: bad feature meta");
  unsigned int* scratch = nullptr;
  if (noalloc_scratch(&scratch)) return -1;
  unsigned* ctl = scratch;
  int* tile_left = reinterpret_cast<int*>(scratch + NOALLOC_CTL_WORDS);
  unsigned* bits = scratch + NOALLOC_CTL_WORDS + NOALLOC_MAX_TILES;
  DevFeat F;
  F.num_bin = meta->num_bin; F.mfb = meta->most_freq_bin; F.offset = meta->most_freq_bin == 0 ? 1 : 0;
  F.default_bin = meta->default_bin; F.missing_type = meta->missing_type; F.hist_base = 0;
  F.nslots = 1 + F.num_bin - F.offset; F.real_index = meta->real_index; F.penalty = meta->penalty;
  SplitRule rule;
  {  // make_rule is a __device__ function; the same arithmetic on the host (dense_bin.hpp:173-205)
    rule.maxb = F.nslots - 1; rule.th = threshold + 1; rule.t_zero = 1 + F.default_bin;
    if (F.mfb == 0) { --rule.th; --rule.t_zero; }
    rule.default_left_side = F.mfb <= threshold;
    rule.miss_zero = F.missing_type == LGBCU_MISSING_ZERO; rule.miss_na = F.missing_type == LGBCU_MISSING_NAN;
    rule.mfb_zero = rule.miss_zero && F.default_bin == F.mfb;
    rule.mfb_na = rule.miss_na && (rule.maxb == F.mfb + 1) && F.mfb > 0;
    rule.missing_left = (rule.miss_zero || rule.miss_na) ? (default_left != 0) : 0;
    rule.wide = 1 < rule.maxb; rule.maxbin_left = rule.maxb <= rule.th;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int ntiles = (n + PART_TILE - 1) / PART_TILE;
  int grid = ntiles < 1 ? 1 : (ntiles > 148 * 8 ? 148 * 8 : ntiles);
  LAUNCH(k_na_flags, grid, PART_THREADS, 0, s, bins_dev + column, (long long)row_stride, rule, indices_in_dev, n, bits, tile_left);
  LAUNCH(k_na_scan, 1, 1024, 0, s, tile_left, n, ctl, left_count_dev);
  LAUNCH(k_na_scatter, grid, PART_THREADS, 0, s, indices_in_dev, n, bits, tile_left, ctl, indices_out_dev);
  CU(cudaGetLastError());
  return 0;
}
