// extern "C" layer of the B200 tree learner: owns the device buffers and enqueues the kernels.
// See include/lgbcu.h for the contract of each entry point and the reference code it stands in for.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstddef>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <dlfcn.h>
#include <omp.h>
#include <sys/mman.h>
#include <unistd.h>
#include <nccl.h>

#include "lgbcu.h"
#include "dev_types.cuh"
#include "hist_kernels.cuh"
#include "split_kernels.cuh"
#include "partition_kernels.cuh"
#include "misc_kernels.cuh"

namespace {

thread_local char g_err[1024] = "everything is fine";
std::atomic<long long> g_launches{0};

int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return -1;
}

#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess) return fail("CUDA error %s at %s:%d (%s)", cudaGetErrorString(e__), __FILE__, __LINE__, #call); \
  } while (0)

// Every entry point works on its context's device and leaves the caller's current device as it found it: the host
// application (torch, for one) may be working on another GPU in the same thread.
struct ScopedDevice {
  int prev = -1; bool switched = false; cudaError_t err = cudaSuccess;
  explicit ScopedDevice(int dev) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != dev) { err = cudaSetDevice(dev); switched = err == cudaSuccess; }
  }
  ~ScopedDevice() { if (switched) cudaSetDevice(prev); }
};
#define USE_DEVICE(dev)                  \
  ScopedDevice scoped_device__(dev);     \
  if (scoped_device__.err != cudaSuccess) return fail("CUDA error %s selecting device %d at %s:%d", cudaGetErrorString(scoped_device__.err), (int)(dev), __FILE__, __LINE__)

#define LAUNCH(kernel, grid, block, smem, stream, ...)                                             \
  do {                                                                                             \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                                    \
    g_launches.fetch_add(1, std::memory_order_relaxed);                                            \
  } while (0)

#define NEED(cond, ...)                                                                            \
  do {                                                                                             \
    if (!(cond)) return fail(__VA_ARGS__);                                                         \
  } while (0)

#include "comm.cuh"

// ---- allocation.  Boosters and datasets of the same shape are created and destroyed again and again in one process
// (warm-up + timed run, cross-validation folds, parameter sweeps), and raw cudaMalloc / cudaFree / cudaMallocHost calls
// are where their set-up time is lost -- and lost unpredictably: the same lgbcu_create took 6 ms or 165 ms on the same
// box.  Device memory therefore comes from the device's stream-ordered pool with its release threshold raised, so
// that freed blocks stay with the process and the next booster's allocations are pointer bumps; small page-locked
// host buffers are kept on a size-keyed free list.  LGBCU_ALLOC_POOL=0 goes back to the plain calls.
// Semantics are those of cudaMalloc / cudaFree: an allocation is usable on every stream when the call returns, a free
// waits for the device first (the callers rely on both).
static bool pool_enabled() {
  static const bool on = [] { const char* e = std::getenv("LGBCU_ALLOC_POOL"); return !(e && std::atoi(e) == 0); }();
  return on;
}
static bool pool_ready(int dev) {
  static std::mutex mu; static int state[64] = {0};   // 0 unknown, 1 pool, -1 no pool
  if (dev < 0 || dev >= 64) return false;
  std::lock_guard<std::mutex> lock(mu);
  if (state[dev] == 0) {
    int ok = 0; cudaMemPool_t pool = nullptr;
    if (cudaDeviceGetAttribute(&ok, cudaDevAttrMemoryPoolsSupported, dev) == cudaSuccess && ok &&
        cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
      unsigned long long keep = ~0ull;
      state[dev] = cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep) == cudaSuccess ? 1 : -1;
    } else {
      state[dev] = -1;
    }
    cudaGetLastError();
  }
  return state[dev] == 1;
}
static cudaError_t lgb_dev_malloc(void** p, size_t bytes) {
  int dev = 0;
  if (pool_enabled() && cudaGetDevice(&dev) == cudaSuccess && pool_ready(dev)) {
    const cudaError_t e = cudaMallocAsync(p, bytes, cudaStreamPerThread);
    if (e != cudaSuccess) return e;
    return cudaStreamSynchronize(cudaStreamPerThread);
  }
  return cudaMalloc(p, bytes);
}
template <typename T>
static cudaError_t lgb_dev_malloc(T** p, size_t bytes) { return lgb_dev_malloc(reinterpret_cast<void**>(p), bytes); }
static cudaError_t lgb_dev_free(void* p) {
  if (!p) return cudaSuccess;
  int dev = 0;
  if (pool_enabled() && cudaGetDevice(&dev) == cudaSuccess && pool_ready(dev)) {
    cudaDeviceSynchronize();
    return cudaFreeAsync(p, cudaStreamPerThread);
  }
  return cudaFree(p);
}
struct HostCache {
  std::mutex mu; std::multimap<size_t, void*> free_; std::map<void*, size_t> live; size_t cached = 0;
};
static HostCache g_host;
static cudaError_t lgb_host_malloc(void** p, size_t bytes) {
  if (pool_enabled()) {
    std::lock_guard<std::mutex> lock(g_host.mu);
    auto it = g_host.free_.find(bytes);
    if (it != g_host.free_.end()) {
      *p = it->second; g_host.free_.erase(it); g_host.cached -= bytes; g_host.live[*p] = bytes;
      return cudaSuccess;
    }
  }
  const cudaError_t e = cudaMallocHost(p, bytes);
  if (e == cudaSuccess && pool_enabled()) { std::lock_guard<std::mutex> lock(g_host.mu); g_host.live[*p] = bytes; }
  return e;
}
template <typename T>
static cudaError_t lgb_host_malloc(T** p, size_t bytes) { return lgb_host_malloc(reinterpret_cast<void**>(p), bytes); }
static cudaError_t lgb_host_free(void* p) {
  if (!p) return cudaSuccess;
  if (pool_enabled()) {
    std::lock_guard<std::mutex> lock(g_host.mu);
    auto it = g_host.live.find(p);
    if (it != g_host.live.end()) {
      const size_t bytes = it->second;
      g_host.live.erase(it);
      if (bytes <= ((size_t)64 << 20) && g_host.cached + bytes <= ((size_t)512 << 20)) {
        g_host.free_.emplace(bytes, p); g_host.cached += bytes;
        return cudaSuccess;
      }
    }
  }
  return cudaFreeHost(p);
}

template <typename T>
cudaError_t dev_alloc(T** p, size_t n) {
  if (*p) { lgb_dev_free(*p); *p = nullptr; }
  if (n == 0) n = 1;
  return lgb_dev_malloc(reinterpret_cast<void**>(p), n * sizeof(T));
}

}  // namespace

struct RankParams {
  double sigmoid = 1.0; int truncation_level = 20; int norm = 1;
};

// NDCG tables of one score set (NDCGMetric::Init prepares them on the host)
struct NdcgTables {
  int* d_eval_at = nullptr; int num_eval = 0, max_k = 0;
  double* d_inv_max_dcg = nullptr; float* d_query_weight = nullptr; double sum_query_weights = 0.0;
  double* d_label_gain = nullptr; double* d_discount = nullptr;
  bool ready = false;
};

// A validation set: bins from the training mappers + its own score vector and meta data
struct ValidSet {
  int n = 0;
  uint8_t* d_bins = nullptr;   // row-major, same stride as the training matrix
  double* d_score = nullptr;        // the current class's plane of d_score_base
  double* d_score_base = nullptr;   // [num_class][n]
  float* d_label = nullptr; float* d_weight = nullptr;
  int* d_query = nullptr; int num_queries = 0;
  NdcgTables ndcg;
};

struct LgbCuContext {
  int device = 0;
  std::vector<ValidSet> valid;
  NdcgTables train_ndcg;
  // evaluation scratch (grown on demand)
  double* d_eval_partial = nullptr; double* d_eval_out = nullptr; double* h_eval_out = nullptr;
  unsigned long long* d_auc_keys[2] = {nullptr, nullptr}; double2* d_auc_vals[2] = {nullptr, nullptr};
  double* d_auc_pos = nullptr; int* d_auc_groups = nullptr; void* d_cub_tmp = nullptr; size_t cub_tmp_bytes = 0; int auc_cap = 0;
  double* d_ndcg_per_query = nullptr; size_t ndcg_cap = 0;
  // data-parallel
  ncclComm_t comm = nullptr; int rank = 0, world = 1;
  // one-shot exchange (comm.cuh): the ranks read each other's local histograms over NVLink; LGBCU_DP_ONESHOT
  bool xchg = false; int xchg_cfg = -1; bool xchg_tried = false; bool xchg_two_step = false;   // LGBCU_DP_ONESHOT=2: the two-step form
  HistEntry* d_pool_local = nullptr; size_t pool_local_entries = 0;       // plain cudaMalloc: exported with CUDA IPC
  unsigned long long* d_xflags = nullptr;
  void* xpeer_pool[LGB_MAX_PEERS] = {nullptr}; void* xpeer_flags[LGB_MAX_PEERS] = {nullptr};   // peers' mappings
  // feature ownership (comm.cuh): on for wide data, LGBCU_DP_OWNERSHIP forces it
  bool dp_owner = false; int dp_owner_cfg = -1;
  std::vector<int> own_begin;   // [world + 1] first feature of every rank's block
  CandRec* d_cand_send = nullptr; CandRec* d_cand_all = nullptr;
  int num_sms = 148;
  cudaStream_t stream = nullptr;
  // dataset
  int N = 0, F = 0, stride = 0, total_bin = 0, ntiles16 = 0;
  int FS = 0; bool any_wide = false;   // byte columns of the matrices; some feature has more than 256 stored values
  // categorical features: `general` = the data takes the general split search / partition instantiations (wide or
  // categorical feature present); the bin-set pool of the tree being grown; the sets staged for the tree walks
  bool any_cat = false, general = false; int cat_features = 0;
  // monotone constraints (basic method): per-feature types on the device; every feature then takes the general search
  bool use_mc = false; signed char* d_monotone = nullptr; double mc_penalty = 0.0;
  // rank_xendcg: one LCG state per query (seeded objective_seed + query), two doubles of scratch per document
  unsigned* d_xe_rng = nullptr; double* d_xe_rho = nullptr; double* d_xe_par = nullptr; bool xe_ready = false;
  int wide_smem_slots = 0;   // slots of a feature's histogram that k_hist_wide keeps in shared memory
  unsigned* d_cat_pool = nullptr; int cat_cap = 0;
  unsigned* d_tree_cat = nullptr; int tree_cat_count = 0, tree_cat_cap = 0;
  DevCatParams cat_params{10.0, 10.0, 32, 100, 4, 0};
  uint8_t* d_bins = nullptr;
  DevFeat* d_feats = nullptr;
  std::vector<DevFeat> h_feats;
  int8_t* d_feat_used = nullptr;
  // per-row state
  float2* d_gh = nullptr;  // interleaved (g, h) of the current iteration
  float *d_grad = nullptr, *d_hess = nullptr, *d_label = nullptr, *d_weight = nullptr;
  // RenewTreeOutput scratch (renew_kernels.cuh)
  double* d_renew_key[2] = {nullptr, nullptr}; float* d_renew_w[2] = {nullptr, nullptr}; double* d_renew_cdf = nullptr;
  double* d_renew_wd = nullptr; int* d_renew_seg = nullptr; int renew_cap = 0;
  int* d_renew_begin = nullptr; int* d_renew_end = nullptr; double* d_renew_out = nullptr; int* d_renew_nonempty = nullptr; int renew_leaf_cap = 0;
  // multiclass: K score / gradient / hessian planes ([class][row], as the reference's ScoreUpdater and gradients_ lay
  // them out); d_score / d_grad / d_hess always point at the CURRENT class's plane (lgbcu_set_class), so that every
  // kernel below trains and scores "a" class without knowing there are several
  int num_class = 1, cur_class = 0;
  std::vector<double> ova_w;   // multiclassova: (label weight of the negative, of the positive) per class
  float *d_grad_base = nullptr, *d_hess_base = nullptr; double* d_score_base = nullptr;
  bool has_weight = false, has_label = false;
  double* d_score = nullptr;
  int* d_indices[2] = {nullptr, nullptr};
  double* h_stage[2] = {nullptr, nullptr}; cudaEvent_t stage_ev[2] = {nullptr, nullptr};  // D2H staging of lgbcu_copy_score
  unsigned* d_part_bits = nullptr;
  unsigned long long* d_part_sync = nullptr;
  uint8_t* d_bins_col = nullptr; long long col_stride = 0;
  // device-side bagging (bagging_kernels.cuh)
  unsigned* d_bag_state = nullptr; unsigned* d_bag_bits = nullptr; int* d_bag_tile = nullptr; unsigned* d_bag_ctl = nullptr;
  float* d_goss = nullptr; unsigned* d_goss_amp = nullptr; float* d_goss_mult = nullptr; int goss_mult_cap = 0;  // GOSS: |g*h|, amplified-row bits, chunk multipliers
  int* d_bag = nullptr; int bag_count = 0; bool bag_ready = false;
  // query structure (lambdarank)
  int* d_query = nullptr; int num_queries = 0; int max_query = 0;
  std::vector<int> h_query;
  double* d_label_gain = nullptr; int num_label_gain = 0;
  double* d_inv_max_dcg = nullptr; float* d_sigmoid_table = nullptr;
  RankParams rank_params;
  bool rank_ready = false, rank_attr_set = false;
  // tree state
  DevParams params{}; bool params_set = false;
  int alloc_leaves = 0;
  int lookahead = LGB_KMAX, lookahead_cfg = 0, alloc_lookahead = 0, pool_slots = 0, last_rounds = 0;  // cfg 0 = choose from the data shape  // K leaves expanded per round (LGBCU_LOOKAHEAD)
  DevSpec* d_spec = nullptr; int* d_expanded = nullptr;
  long long bits_stride = 0; int part_sync_stride = 0;
  HistEntry* d_pool = nullptr;
  int8_t* d_splittable = nullptr; int8_t* d_parent_split = nullptr; double2* d_mfb_fix = nullptr;
  DevLeaf* d_leaves = nullptr;
  DevSplit *d_best = nullptr, *d_feat_best = nullptr;
  SplitKey *d_feat_key = nullptr, *d_leaf_key = nullptr;
  DevCtl* d_ctl = nullptr;
  LgbCuSplitRecord* d_records = nullptr;
  double* d_leaf_value = nullptr;
  double* d_export = nullptr;
  // tree upload for add_score_by_tree
  int* d_tree_i32 = nullptr; signed char* d_tree_i8 = nullptr; double* d_tree_f64 = nullptr; int tree_cap = 0;
  int* d_rows = nullptr; int rows_cap = 0; long long rows_tag = 0; int rows_count = 0;
  double* d_scratch = nullptr;
  // host staging (pinned)
  LgbCuSplitRecord* h_records = nullptr;
  DevCtl* h_ctl = nullptr;
  DevLeaf* h_leaf = nullptr;
  DevSplit* h_split = nullptr;
  // config
  int hist_mode = LGBCU_HIST_AUTO;
  int hist_min_rows = HIST_MIN_ROWS_PER_WARP;
  int hist_small_rows = HIST_SMALL_LEAF_ROWS;
  int part_blocks_per_sm = PART_BLOCKS_PER_SM;  // lowered at create time if fewer k_partition blocks fit on an SM
  int hist_grid = 0;  // clusters that can be resident at once x HIST_CLUSTER
  bool force_dp_counts = false;  // LGBCU_FORCE_DP_COUNTS=1: single rank, but leaf counts as in data-parallel mode (tests)
  std::vector<cudaGraphExec_t> chunk_graphs;  // further rounds in chunks of TREE_CHUNK_ROUNDS, for trees that outgrow the first batch
  int last_used_rounds = 0;                   // rounds the previous tree needed
  cudaGraphExec_t tree_graph = nullptr; bool use_graph = true; bool graph_comm = true;  // the per-tree launch sequence, captured once
  int objective = LGBCU_OBJ_NONE; double obj_p[3] = {0, 0, 0};
  int root_count = 0; bool tree_begun = false;
  // profiling
  bool profile = false;
  std::vector<cudaEvent_t> ev;
  double prof_hist_ms = 0; long long prof_hist_calls = 0; double prof_hist_bytes = 0;
  double loop_phase_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0}; bool loop_phase_valid = false;

  DevView view() const {
    DevView V;
    V.num_data = N; V.num_features = F; V.row_stride = stride; V.total_bin = total_bin; V.num_tiles16 = ntiles16;
    V.num_storage = FS; V.any_wide = any_wide ? 1 : 0;
    V.any_cat = any_cat ? 1 : 0; V.cat_cap = cat_cap; V.cat_pool = d_cat_pool; V.cat = cat_params;
    V.use_mc = use_mc ? 1 : 0; V.monotone = d_monotone; V.mc_penalty = mc_penalty;
    V.bins = d_bins; V.feats = d_feats; V.feat_used = d_feat_used; V.splittable = d_splittable;
    V.parent_split = d_parent_split; V.mfb_fix = d_mfb_fix; V.spec = d_spec; V.expanded = d_expanded; V.pool_slots = pool_slots;
    V.bits_stride = bits_stride; V.part_sync_stride = part_sync_stride;
    V.gh = d_gh; V.indices[0] = d_indices[0]; V.indices[1] = d_indices[1];
    V.pool = d_pool; V.leaves = d_leaves; V.best = d_best; V.feat_best = d_feat_best; V.feat_key = d_feat_key; V.leaf_key = d_leaf_key; V.ctl = d_ctl;
    V.records = d_records; V.part_sync = d_part_sync; V.bins_col = d_bins_col; V.col_stride = col_stride; V.part_bits = d_part_bits; V.p = params;
    V.data_parallel = (comm != nullptr || force_dp_counts) ? 1 : 0;
    V.f_begin = 0; V.f_end = F; V.comm_size = world; V.cand_send = d_cand_send; V.cand_all = d_cand_all;
    if (dp_owner && (int)own_begin.size() == world + 1) { V.f_begin = own_begin[rank]; V.f_end = own_begin[rank + 1]; }
    V.xchg = (xchg && comm != nullptr) ? 1 : 0; V.comm_rank = rank; V.pool_local = d_pool_local; V.xflags = d_xflags;
    V.hist_pool = V.xchg ? d_pool_local : d_pool;
    for (int r = 0; r < LGB_MAX_PEERS; ++r) V.xpool[r] = r == rank ? d_pool_local : static_cast<const HistEntry*>(xpeer_pool[r]);
    return V;
  }
};

namespace {

int grid_for(const LgbCuContext* c, long long work, int block, int per_sm = 8) {
  long long g = (work + block - 1) / block;
  const long long cap = (long long)c->num_sms * per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// Threads for the library's own host-side copy loops.  A launcher may pin OMP_NUM_THREADS to 1 per rank (torchrun
// does); the D2H staging copies are memory-bound helpers, not user compute, so they take a fair share of the cores.
int host_copy_threads(const LgbCuContext* c) {
  static const int hw = [] { const long n = sysconf(_SC_NPROCESSORS_ONLN); return n > 0 ? (int)n : 1; }();
  if (const char* e = std::getenv("LGBM_B200_HOST_THREADS")) { const int v = std::atoi(e); if (v >= 1) return v; }
  const int share = std::max(1, hw / std::max(1, c->world));
  return std::max(omp_get_max_threads(), std::min(8, share));
}

void drop_graph(LgbCuContext* c) {
  if (c->tree_graph) { cudaGraphExecDestroy(c->tree_graph); c->tree_graph = nullptr; }
  for (cudaGraphExec_t g : c->chunk_graphs) if (g) cudaGraphExecDestroy(g);
  c->chunk_graphs.clear();
}

int ensure_tree_buffers(LgbCuContext* c) {
  NEED(c->F > 0 && c->N > 0, "lgbcu: dataset not set");
  NEED(c->params_set, "lgbcu: tree params not set");
  const int L = c->params.num_leaves;
  // K leaves expanded per round.  Every expanded leaf costs 2 x F blocks in the split search of its round whether or
  // not its turn comes soon, so very wide data gets a smaller K (measured in round 2, ms per iteration: 137 features
  // K = 4 / 8 / 16 -> 9.6 / 8.8 / 8.3; 2000 features K = 1 / 2 / 4 -> 29.3 / 27.9 / 28.1; 28 features best at 16).
  int K = c->lookahead_cfg;
  if (K <= 0) { K = c->num_sms * 16 / std::max(1, c->F); if (K < 2) K = 2; if (K > LGB_KMAX) K = LGB_KMAX; }
  if (c->alloc_leaves == L && c->alloc_lookahead == K && c->d_pool) return 0;
  c->lookahead = K;  // the memory guard below may lower it; alloc_lookahead remembers what was asked for
  drop_graph(c);
  USE_DEVICE(c->device);
  // histogram slots: the root + one fresh slot per (round, item); a round commits at least one leaf, so there are at
  // most L-1 rounds after the root.  Keep the pool under 1/8 of the device memory by lowering K if necessary.
  {
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    while (c->lookahead > 1 && (1 + (size_t)(L - 1) * c->lookahead) * c->total_bin * sizeof(HistEntry) > total_b / 8) c->lookahead -= 1;
  }
  const size_t slots = std::max<size_t>((size_t)L, 1 + (size_t)(L > 1 ? L - 1 : 0) * c->lookahead);
  c->pool_slots = (int)slots;
  CU(dev_alloc(&c->d_pool, slots * c->total_bin));
  CU(dev_alloc(&c->d_splittable, slots * c->F));
  if (c->any_cat) {   // every (categorical feature, child of an expansion) that finds a split takes one bin set
    c->cat_cap = (int)std::min<size_t>((size_t)1 << 30, (size_t)c->cat_features * 2 * slots + 1);
    CU(dev_alloc(&c->d_cat_pool, (size_t)c->cat_cap * LGB_CAT_WORDS));
  }
  CU(dev_alloc(&c->d_mfb_fix, slots * c->F));
  CU(cudaMemsetAsync(c->d_mfb_fix, 0, sizeof(double2) * slots * c->F, c->stream));
  CU(dev_alloc(&c->d_leaves, (size_t)L));
  CU(dev_alloc(&c->d_best, (size_t)L));
  CU(dev_alloc(&c->d_leaf_key, (size_t)L));
  CU(dev_alloc(&c->d_spec, (size_t)L));
  CU(dev_alloc(&c->d_expanded, (size_t)L));
  CU(cudaMemsetAsync(c->d_expanded, 0, sizeof(int) * (size_t)L, c->stream));
  CU(dev_alloc(&c->d_records, (size_t)L));
  CU(dev_alloc(&c->d_leaf_value, (size_t)L));
  if (c->h_records) lgb_host_free(c->h_records);
  CU(lgb_host_malloc(reinterpret_cast<void**>(&c->h_records), sizeof(LgbCuSplitRecord) * (size_t)L));
  CU(cudaMemsetAsync(c->d_splittable, 0, slots * c->F, c->stream));
  c->alloc_lookahead = K;
  c->alloc_leaves = L;
  return 0;
}

int launch_hist(LgbCuContext* c, int leaf_arg) {
  const DevView V = c->view();
  if (c->any_wide) LAUNCH(k_hist_wide, c->num_sms * 8, 256, (size_t)c->wide_smem_slots * 16, c->stream, V, leaf_arg, c->wide_smem_slots);
  else LAUNCH(k_hist, c->hist_grid, HIST_THREADS, HIST_SMEM_BYTES, c->stream, V, leaf_arg, c->hist_mode, c->hist_min_rows, c->hist_small_rows);
  CU(cudaGetLastError());
  return 0;
}

// The fused partition spins on an inter-block barrier, so it is launched cooperatively (the runtime then guarantees
// that the whole grid is resident).  grid = blocks that own tiles (at most 4 per SM) + 1 control block.
int launch_partition(LgbCuContext* c, long long rows, int mode) {
  const DevView V = c->view();
  cudaLaunchConfig_t cfg = {};
  // training: the blocks are dealt to the round's items on the device; every item needs a tile block and a control block
  unsigned grid = (unsigned)grid_for(c, rows, PART_TILE, c->part_blocks_per_sm) + 1u;
  if (mode == 0) {
    const unsigned cap = (unsigned)(c->num_sms * c->part_blocks_per_sm);
    grid = std::min(cap, grid + 2u * (unsigned)c->lookahead);
    grid = std::max(grid, 2u * (unsigned)c->lookahead);
  }
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(PART_THREADS);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = c->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;
  at[0].val.cooperative = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  if (c->general) CU(cudaLaunchKernelEx(&cfg, k_partition<true>, V, mode));
  else CU(cudaLaunchKernelEx(&cfg, k_partition<false>, V, mode));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

// max|g|, max|h| must be the same on every rank: the fixed-point scale is part of the wire format.
int comm_allreduce_tree_setup(LgbCuContext* c) {
  if (!c->comm) return 0;
  char* base = reinterpret_cast<char*>(c->d_ctl);
  void* mx = base + offsetof(DevCtl, max_g_bits);
  void* rows = base + offsetof(DevCtl, global_root);
  NC(g_nccl.AllReduce(mx, mx, 2, ncclUint32, ncclMax, c->comm, c->stream));
  NC(g_nccl.AllReduce(rows, rows, 2, ncclInt64, ncclSum, c->comm, c->stream));
  return 0;
}

// Split s builds the smaller leaf's histogram in pool slot s (slot 0 = root, then the fresh slot of split s
// is s): known on the host without reading device state, so the collective can be enqueued blind.
int comm_reduce_hist(LgbCuContext* c, int first_slot, int num_slots, bool root_round) {
  if (!c->comm) return 0;
  HistEntry* slot = c->d_pool + (size_t)first_slot * c->total_bin;
  if (!c->dp_owner) {
    NC(g_nccl.AllReduce(slot, slot, (size_t)c->total_bin * 2 * num_slots, ncclInt64, ncclSum, c->comm, c->stream));
    return 0;
  }
  // Ownership mode (comm.cuh): every owner receives the global sums of its feature block, for each fresh slot.  The
  // root's feature 0 goes to every rank first: LeafSplits::Init's root sums are read off it (k_root_sums).
  int skip0 = 0;
  if (root_round) {
    const DevFeat& f0 = c->h_feats[0];
    HistEntry* b0 = slot + f0.hist_base;
    NC(g_nccl.AllReduce(b0, b0, (size_t)f0.nslots * 2, ncclInt64, ncclSum, c->comm, c->stream));
    skip0 = 1;
  }
  NC(g_nccl.GroupStart());
  for (int k = 0; k < num_slots; ++k) {
    for (int r = 0; r < c->world; ++r) {
      int fb = c->own_begin[r], fe = c->own_begin[r + 1];
      if (skip0 && fb == 0) fb = 1;   // already global everywhere
      if (fb >= fe) continue;
      const size_t e0 = (size_t)c->h_feats[fb].hist_base;
      const size_t e1 = fe < c->F ? (size_t)c->h_feats[fe].hist_base : (size_t)c->total_bin;
      HistEntry* blk = slot + (size_t)k * c->total_bin + e0;
      NC(g_nccl.Reduce(blk, blk, (e1 - e0) * 2, ncclInt64, ncclSum, r, c->comm, c->stream));
    }
  }
  NC(g_nccl.GroupEnd());
  return 0;
}

// ownership mode: every rank's per-(leaf, child) winners to every rank
int comm_gather_winners(LgbCuContext* c) {
  NC(g_nccl.AllGather(c->d_cand_send, c->d_cand_all, sizeof(CandRec) * 2 * LGB_KMAX, ncclChar, c->comm, c->stream));
  return 0;
}

// ---- one-shot exchange set-up: a second pool for this rank's LOCAL sums and a flag array, both plain cudaMalloc
// memory exported with CUDA IPC; the handles travel over the communicator; every peer's pair is mapped here.  Any
// failure (no peer access, IPC refused) leaves the NCCL all-reduce path in place.
void xchg_teardown(LgbCuContext* c) {
  for (int r = 0; r < LGB_MAX_PEERS; ++r) {
    if (c->xpeer_pool[r]) { cudaIpcCloseMemHandle(c->xpeer_pool[r]); c->xpeer_pool[r] = nullptr; }
    if (c->xpeer_flags[r]) { cudaIpcCloseMemHandle(c->xpeer_flags[r]); c->xpeer_flags[r] = nullptr; }
  }
  if (c->d_pool_local) { ::cudaFree(c->d_pool_local); c->d_pool_local = nullptr; }
  if (c->d_xflags) { ::cudaFree(c->d_xflags); c->d_xflags = nullptr; }
  c->pool_local_entries = 0; c->xchg = false;
  cudaGetLastError();
}

int xchg_setup(LgbCuContext* c) {
  const size_t entries = (size_t)c->pool_slots * c->total_bin;
  // measured (bench workload, ms per iteration, NCCL all-reduce -> one-shot): 2 ranks 3.34 -> 3.18; 8 ranks 2.87 -> 3.12
  // (every rank pulls 7 x the round's slots; NCCL's reduce-scatter + all-gather moves a quarter of that): on by
  // default for two ranks only, LGBCU_DP_ONESHOT=0/1 forces it
  bool want = c->comm != nullptr && c->world > 1 && c->world <= LGB_MAX_PEERS && !c->dp_owner &&
              (c->xchg_cfg == 1 || (c->xchg_cfg < 0 && c->world == 2));
  if (!want) { if (c->xchg) { drop_graph(c); xchg_teardown(c); } return 0; }
  if (c->xchg && c->pool_local_entries == entries) return 0;
  if (c->xchg_tried && !c->xchg && c->pool_local_entries == entries) return 0;   // failed before with this shape
  drop_graph(c);
  xchg_teardown(c);
  c->xchg_tried = true; c->pool_local_entries = entries;
  struct Handles { cudaIpcMemHandle_t pool, flags; };
  static_assert(sizeof(Handles) == 128, "two 64-byte IPC handles");
  bool ok = ::cudaMalloc(reinterpret_cast<void**>(&c->d_pool_local), entries * sizeof(HistEntry)) == cudaSuccess &&
            ::cudaMalloc(reinterpret_cast<void**>(&c->d_xflags), sizeof(unsigned long long) * LGB_MAX_PEERS) == cudaSuccess;
  Handles mine;
  std::memset(&mine, 0, sizeof(mine));
  if (ok) {
    CU(cudaMemsetAsync(c->d_pool_local, 0, entries * sizeof(HistEntry), c->stream));
    CU(cudaMemsetAsync(c->d_xflags, 0, sizeof(unsigned long long) * LGB_MAX_PEERS, c->stream));
    // the epoch restarts with the flags: every rank runs this set-up at the same point of its life
    CU(cudaMemsetAsync(reinterpret_cast<char*>(c->d_ctl) + offsetof(DevCtl, xchg_epoch), 0, sizeof(unsigned long long), c->stream));
    ok = cudaIpcGetMemHandle(&mine.pool, c->d_pool_local) == cudaSuccess && cudaIpcGetMemHandle(&mine.flags, c->d_xflags) == cudaSuccess;
  }
  // every rank takes part in the all-gather whether or not its own export worked: a zeroed pair means "no"
  if (!ok) std::memset(&mine, 0, sizeof(mine));
  char* d_h = nullptr;
  CU(dev_alloc(&d_h, sizeof(Handles) * (size_t)(c->world + 1)));
  struct Free { char* p; ~Free() { if (p) lgb_dev_free(p); } } fr{d_h};
  CU(cudaMemcpyAsync(d_h, &mine, sizeof(Handles), cudaMemcpyHostToDevice, c->stream));
  NC(g_nccl.AllGather(d_h, d_h + sizeof(Handles), sizeof(Handles), ncclChar, c->comm, c->stream));
  std::vector<Handles> all((size_t)c->world);
  CU(cudaMemcpyAsync(all.data(), d_h + sizeof(Handles), sizeof(Handles) * (size_t)c->world, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  const Handles zero = [] { Handles z; std::memset(&z, 0, sizeof(z)); return z; }();
  for (int r = 0; r < c->world && ok; ++r) if (std::memcmp(&all[r], &zero, sizeof(Handles)) == 0) ok = false;   // somebody could not export
  for (int r = 0; r < c->world && ok; ++r) {
    if (r == c->rank) continue;
    ok = cudaIpcOpenMemHandle(&c->xpeer_pool[r], all[r].pool, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess &&
         cudaIpcOpenMemHandle(&c->xpeer_flags[r], all[r].flags, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
  }
  // all ranks must agree (a rank that could not map a peer would otherwise wait on NCCL alone): sum of the verdicts
  {
    double v = ok ? 1.0 : 0.0;
    if (!c->d_scratch) CU(dev_alloc(&c->d_scratch, (size_t)64));
    CU(cudaMemcpyAsync(c->d_scratch, &v, sizeof(double), cudaMemcpyHostToDevice, c->stream));
    NC(g_nccl.AllReduce(c->d_scratch, c->d_scratch, 1, ncclDouble, ncclSum, c->comm, c->stream));
    CU(cudaMemcpyAsync(&v, c->d_scratch, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    ok = v > c->world - 0.5;
  }
  cudaGetLastError();
  if (std::getenv("LGBCU_VERBOSE")) std::fprintf(stderr, "[lgbcu] rank %d: one-shot histogram exchange %s\n", c->rank, ok ? "on (peers mapped with CUDA IPC)" : "unavailable, using the NCCL all-reduce");
  if (!ok) { xchg_teardown(c); c->pool_local_entries = entries; return 0; }
  c->xchg = true;
  return 0;
}

// Decided once the communicator and the data shape are known: ownership pays when the replicated split search is what
// a round waits for (wide data); for narrow data the search costs no wall time and the extra collective would.
int setup_ownership(LgbCuContext* c) {
  bool want = c->comm != nullptr && c->world > 1 && c->F >= 256;
  if (c->dp_owner_cfg == 0) want = false;
  if (c->dp_owner_cfg == 1) want = c->comm != nullptr && c->world > 1;
  if (want && !g_nccl.has_ownership_calls()) want = false;
  if (want && c->F < c->world) want = false;
  // a categorical winner names its bin set by an index into the RANK's own pool: such records cannot be gathered from
  // other ranks, so categorical data keeps the replicated search on the reduced histograms
  if (want && c->any_cat) want = false;
  if (want == c->dp_owner && (!want || (int)c->own_begin.size() == c->world + 1)) return 0;
  drop_graph(c);
  c->dp_owner = want;
  if (!want) return 0;
  // contiguous feature blocks balanced by histogram entries; identical on every rank (same mappers)
  c->own_begin.assign((size_t)c->world + 1, c->F);
  c->own_begin[0] = 0;
  int f = 0;
  for (int r = 0; r < c->world; ++r) {
    c->own_begin[r] = f;
    const long long target = (long long)c->total_bin * (r + 1) / c->world;
    const int must_leave = c->world - 1 - r;   // at least one feature for every later rank
    while (f < c->F - must_leave && (f == c->own_begin[r] || (long long)c->h_feats[f].hist_base + c->h_feats[f].nslots <= target)) ++f;
  }
  c->own_begin[c->world] = c->F;
  if (!c->d_cand_send) CU(dev_alloc(&c->d_cand_send, (size_t)2 * LGB_KMAX));
  if (c->d_cand_all) { lgb_dev_free(c->d_cand_all); c->d_cand_all = nullptr; }
  CU(dev_alloc(&c->d_cand_all, (size_t)c->world * 2 * LGB_KMAX));
  return 0;
}

// host tree arrays -> device (shared by the training-set and validation-set tree walks)
int upload_tree(LgbCuContext* c, int32_t num_leaves, const int32_t* split_feature_inner, const int32_t* threshold_bin,
                const int8_t* decision_type, const int32_t* left_child, const int32_t* right_child,
                const double* leaf_value, TreeDev* T) {
  NEED(split_feature_inner && threshold_bin && decision_type && left_child && right_child, "add_score_by_tree: NULL tree array");
  const int ni = num_leaves - 1;
  if (c->tree_cap < num_leaves) {
    CU(dev_alloc(&c->d_tree_i32, (size_t)4 * num_leaves));
    CU(dev_alloc(&c->d_tree_i8, (size_t)num_leaves));
    CU(dev_alloc(&c->d_tree_f64, (size_t)num_leaves));
    c->tree_cap = num_leaves;
  }
  for (int i = 0; i < ni; ++i) {
    NEED(split_feature_inner[i] >= 0 && split_feature_inner[i] < c->F, "add_score_by_tree: split feature out of range");
    if (decision_type[i] & 1) {
      NEED(threshold_bin[i] >= 0 && threshold_bin[i] < c->tree_cat_count, "add_score_by_tree: categorical node %d names set %d, %d were staged (lgbcu_set_tree_cat_sets)", i, threshold_bin[i], c->tree_cat_count);
      NEED(c->h_feats[split_feature_inner[i]].cat, "add_score_by_tree: categorical node on the numerical feature %d", split_feature_inner[i]);
    }
  }
  int* d_feat = c->d_tree_i32, *d_thr = d_feat + num_leaves, *d_lc = d_thr + num_leaves, *d_rc = d_lc + num_leaves;
  CU(cudaMemcpyAsync(d_feat, split_feature_inner, sizeof(int) * ni, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d_thr, threshold_bin, sizeof(int) * ni, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d_lc, left_child, sizeof(int) * ni, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d_rc, right_child, sizeof(int) * ni, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(c->d_tree_i8, decision_type, ni, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(c->d_tree_f64, leaf_value, sizeof(double) * num_leaves, cudaMemcpyHostToDevice, c->stream));
  T->split_feature = d_feat; T->threshold_bin = d_thr; T->decision_type = c->d_tree_i8; T->left_child = d_lc; T->right_child = d_rc;
  T->leaf_value = c->d_tree_f64; T->num_leaves = num_leaves; T->cat_sets = c->d_tree_cat;
  return 0;
}

void free_valid_sets(LgbCuContext* c);
int rank_prepare(LgbCuContext* c);
int rank_launch(LgbCuContext* c);

}  // namespace

#include "rank_kernels.cuh"
#include "eval_kernels.cuh"

extern "C" {

const char* lgbcu_last_error(void) { return g_err; }

int lgbcu_device_count(int* out) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { *out = 0; cudaGetLastError(); return 0; }
  *out = n;
  return 0;
}

int lgbcu_pin_host_buffer(void* ptr, size_t bytes) {
  if (!ptr || bytes == 0) return -1;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return -1; }
  if (cudaHostRegister(ptr, bytes, cudaHostRegisterPortable) != cudaSuccess) { cudaGetLastError(); return -1; }
  return 0;
}

int lgbcu_unpin_host_buffer(void* ptr) {
  if (!ptr) return -1;
  if (cudaHostUnregister(ptr) != cudaSuccess) { cudaGetLastError(); return -1; }
  return 0;
}

int64_t lgbcu_launch_count(void) { return g_launches.load(); }

int lgbcu_create(int device_id, LgbCuContext** out) {
  NEED(out != nullptr, "lgbcu_create: out is NULL");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail("lgbcu_create: no CUDA device available (%s); this library has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  }
  NEED(device_id >= 0 && device_id < n, "lgbcu_create: device %d out of range (have %d)", device_id, n);
  USE_DEVICE(device_id);
  LgbCuContext* c = new LgbCuContext();
  struct Guard { LgbCuContext* c; ~Guard() { if (c) lgbcu_destroy(c); } } guard{c};   // a failure below frees the half-built context
  c->device = device_id;
  int sm_count = 0;
  CU(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device_id));  // (cudaGetDeviceProperties costs milliseconds)
  c->num_sms = sm_count;
  if (const char* e = std::getenv("LGBCU_HIST_MIN_ROWS")) { const int v = std::atoi(e); if (v >= 1) c->hist_min_rows = v; }
  if (const char* e = std::getenv("LGBCU_DP_OWNERSHIP")) c->dp_owner_cfg = std::atoi(e) != 0 ? 1 : 0;
  if (const char* e = std::getenv("LGBCU_DP_ONESHOT")) { const int v = std::atoi(e); c->xchg_cfg = v != 0 ? 1 : 0; c->xchg_two_step = v == 2; }
  if (const char* e = std::getenv("LGBCU_HIST_SMALL_ROWS")) { const int v = std::atoi(e); if (v >= 0) c->hist_small_rows = v; }
  CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CU(dev_alloc(&c->d_ctl, 1));
  CU(cudaMemset(c->d_ctl, 0, sizeof(DevCtl)));
  CU(lgb_host_malloc(reinterpret_cast<void**>(&c->h_ctl), sizeof(DevCtl)));
  CU(lgb_host_malloc(reinterpret_cast<void**>(&c->h_leaf), sizeof(DevLeaf)));
  CU(lgb_host_malloc(reinterpret_cast<void**>(&c->h_split), sizeof(DevSplit)));
  CU(cudaFuncSetAttribute(k_hist, cudaFuncAttributeMaxDynamicSharedMemorySize, HIST_SMEM_BYTES));
  {
    // the histogram kernel is persistent over its work units: as many clusters as the GPU can hold at once
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(c->num_sms * HIST_BLOCKS_PER_SM / HIST_CLUSTER * HIST_CLUSTER));
    cfg.blockDim = dim3(HIST_THREADS);
    cfg.dynamicSmemBytes = HIST_SMEM_BYTES;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = HIST_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int nclusters = 0;
    CU(cudaOccupancyMaxActiveClusters(&nclusters, k_hist, &cfg));
    NEED(nclusters >= 1, "lgbcu_create: device %d cannot hold one %d-block cluster of the histogram kernel", device_id, HIST_CLUSTER);
    const int want = c->num_sms * HIST_BLOCKS_PER_SM / HIST_CLUSTER;
    if (nclusters > want) nclusters = want;
    if (const char* e = std::getenv("LGBCU_HIST_CLUSTERS")) { const int v = std::atoi(e); if (v >= 1 && v < nclusters) nclusters = v; }
    c->hist_grid = nclusters * HIST_CLUSTER;
    if (std::getenv("LGBCU_VERBOSE")) std::fprintf(stderr, "[lgbcu] histogram kernel: %d clusters of %d blocks\n", nclusters, HIST_CLUSTER);
  }
  {
    int coop = 0;
    CU(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device_id));
    NEED(coop != 0, "lgbcu_create: device %d cannot do cooperative launches (needed by the fused partition)", device_id);
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_partition<false>, PART_THREADS, 0));
    { int per_sm_w = 0; CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_w, k_partition<true>, PART_THREADS, 0)); per_sm = std::min(per_sm, per_sm_w); }
    NEED(per_sm >= 2, "lgbcu_create: only %d k_partition block(s) fit on an SM", per_sm);
    // one slot per SM is left for the control block
    if (per_sm - 1 < c->part_blocks_per_sm) c->part_blocks_per_sm = per_sm - 1;
    if (const char* e = std::getenv("LGBCU_GRAPH")) c->use_graph = std::atoi(e) != 0;
    if (const char* e = std::getenv("LGBCU_LOOKAHEAD")) { const int v = std::atoi(e); if (v >= 1 && v <= LGB_KMAX) c->lookahead_cfg = v; }
    if (const char* e = std::getenv("LGBCU_GRAPH_COMM")) c->graph_comm = std::atoi(e) != 0;
    if (const char* e = std::getenv("LGBCU_FORCE_DP_COUNTS")) c->force_dp_counts = std::atoi(e) != 0;
  }
  guard.c = nullptr;
  *out = c;
  return 0;
}

int lgbcu_destroy(LgbCuContext* c) {
  if (!c) return 0;
  ScopedDevice dev_(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  drop_graph(c);
  xchg_teardown(c);
  lgbcu_comm_destroy(c);
  free_valid_sets(c);
  if (c->d_cat_pool) { lgb_dev_free(c->d_cat_pool); c->d_cat_pool = nullptr; }
  if (c->d_monotone) { lgb_dev_free(c->d_monotone); c->d_monotone = nullptr; }
  for (void* q : {(void*)c->d_xe_rng, (void*)c->d_xe_rho, (void*)c->d_xe_par}) if (q) lgb_dev_free(q);
  c->d_xe_rng = nullptr; c->d_xe_rho = nullptr; c->d_xe_par = nullptr;
  if (c->d_tree_cat) { lgb_dev_free(c->d_tree_cat); c->d_tree_cat = nullptr; }
  void* ptrs[] = {c->d_bins, c->d_feats, c->d_feat_used, c->d_grad_base, c->d_hess_base, c->d_gh, c->d_label, c->d_weight, c->d_score_base,
                  c->d_indices[0], c->d_indices[1], c->d_part_bits, c->d_part_sync, c->d_bins_col, c->d_query, c->d_label_gain,
                  c->d_inv_max_dcg, c->d_sigmoid_table, c->d_pool, c->d_splittable, c->d_leaves, c->d_best,
                  c->d_feat_best, c->d_ctl, c->d_records, c->d_leaf_value, c->d_export, c->d_tree_i32, c->d_tree_i8,
                  c->d_tree_f64, c->d_rows, c->d_scratch, c->d_parent_split, c->d_mfb_fix, c->d_feat_key, c->d_leaf_key, c->d_spec, c->d_expanded,
                  c->d_bag_state, c->d_bag_bits, c->d_bag_tile, c->d_bag_ctl, c->d_bag, c->d_goss, c->d_goss_amp, c->d_goss_mult,
                  c->d_cand_send, c->d_cand_all, c->d_renew_key[0], c->d_renew_key[1], c->d_renew_w[0], c->d_renew_w[1], c->d_renew_cdf,
                  c->d_renew_wd, c->d_renew_seg, c->d_renew_begin, c->d_renew_end, c->d_renew_out, c->d_renew_nonempty};
  for (void* p : ptrs) if (p) lgb_dev_free(p);
  if (c->h_records) lgb_host_free(c->h_records);
  if (c->h_ctl) lgb_host_free(c->h_ctl);
  if (c->h_leaf) lgb_host_free(c->h_leaf);
  if (c->h_split) lgb_host_free(c->h_split);
  for (int i = 0; i < 2; ++i) { if (c->h_stage[i]) lgb_host_free(c->h_stage[i]); if (c->stage_ev[i]) cudaEventDestroy(c->stage_ev[i]); }
  for (auto e : c->ev) cudaEventDestroy(e);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
  return 0;
}

int lgbcu_synchronize(LgbCuContext* c) {
  NEED(c, "lgbcu: NULL context");
  USE_DEVICE(c->device);
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

static int setup_dataset_common(LgbCuContext* c, int32_t num_data, int32_t num_features, const LgbCuFeatureMeta* meta) {
  NEED(c, "lgbcu: NULL context");
  NEED(num_data > 0 && num_features > 0 && meta, "lgbcu_set_dataset: empty dataset");
  USE_DEVICE(c->device);
  c->N = num_data; c->F = num_features;
  c->h_feats.resize(num_features);
  int base = 0, scol = 0;
  c->any_wide = false; c->any_cat = false; c->cat_features = 0;
  for (int f = 0; f < num_features; ++f) {
    DevFeat& d = c->h_feats[f];
    const LgbCuFeatureMeta& m = meta[f];
    NEED(m.num_bin >= 1 && m.most_freq_bin >= 0 && m.most_freq_bin < m.num_bin, "lgbcu_set_dataset: bad meta for feature %d", f);
    d.num_bin = m.num_bin; d.mfb = m.most_freq_bin; d.offset = m.most_freq_bin == 0 ? 1 : 0;
    d.default_bin = m.default_bin; d.missing_type = m.missing_type;
    d.hist_base = base; d.nslots = 1 + d.num_bin - d.offset;
    NEED(d.nslots <= 65536, "lgbcu_set_dataset: feature %d needs %d stored values; at most 16-bit bins are supported", f, d.nslots);
    d.wide = d.nslots > 256 ? 1 : 0; d.col = scol; scol += d.wide ? 2 : 1;
    c->any_wide = c->any_wide || d.wide;
    d.cat = m.bin_type == LGBCU_BIN_CATEGORICAL ? 1 : 0; d.pad_ = 0;
    NEED(m.bin_type == LGBCU_BIN_NUMERICAL || m.bin_type == LGBCU_BIN_CATEGORICAL, "lgbcu_set_dataset: bad bin_type of feature %d", f);
    NEED(!d.cat || d.num_bin <= 256, "lgbcu_set_dataset: the categorical feature %d has %d bins; at most 256 are supported", f, d.num_bin);
    if (d.cat) { c->any_cat = true; c->cat_features += 1; }
    d.real_index = m.real_index; d.penalty = m.penalty;
    base += d.nslots;
  }
  {
    c->use_mc = false;   // lgbcu_set_monotone follows the dataset
    c->general = c->any_wide || c->any_cat;
    if (c->any_wide) {
      // k_hist_wide: the largest feature's histogram in shared memory if it fits (16 bytes a slot, 192 KB at most, and
      // small enough for several blocks per SM when the features allow it)
      int max_slots = 0;
      for (int f = 0; f < num_features; ++f) max_slots = std::max(max_slots, c->h_feats[f].nslots);
      c->wide_smem_slots = std::min(max_slots, 12288);
      CU(cudaFuncSetAttribute(k_hist_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, c->wide_smem_slots * 16));
    }
    const int fb = c->general ? KEY_FEAT_BITS_WIDE : KEY_FEAT_BITS;
    NEED(num_features < (1 << fb), "lgbcu_set_dataset: %d features exceed the supported %d%s", num_features, (1 << fb) - 1,
         c->general ? " (for data with a categorical feature or more than 256 bins in a feature)" : "");
  }
  c->total_bin = base;
  c->FS = scol;
  c->stride = (c->FS + 15) & ~15;
  c->ntiles16 = c->stride / 16;
  CU(dev_alloc(&c->d_feats, (size_t)num_features));
  CU(cudaMemcpyAsync(c->d_feats, c->h_feats.data(), sizeof(DevFeat) * num_features, cudaMemcpyHostToDevice, c->stream));
  CU(dev_alloc(&c->d_bins, (size_t)num_data * c->stride));
  c->col_stride = ((long long)num_data + 255) / 256 * 256;
  CU(dev_alloc(&c->d_bins_col, (size_t)c->col_stride * c->FS));
  CU(dev_alloc(&c->d_feat_used, (size_t)num_features));
  CU(cudaMemsetAsync(c->d_feat_used, 1, num_features, c->stream));
  CU(dev_alloc(&c->d_feat_best, (size_t)2 * LGB_KMAX * num_features));
  CU(dev_alloc(&c->d_feat_key, (size_t)2 * LGB_KMAX * num_features));
  CU(dev_alloc(&c->d_parent_split, (size_t)LGB_KMAX * num_features));
  CU(cudaMemsetAsync(c->d_parent_split, 1, (size_t)LGB_KMAX * num_features, c->stream));
  CU(dev_alloc(&c->d_grad_base, (size_t)num_data * c->num_class));
  CU(dev_alloc(&c->d_hess_base, (size_t)num_data * c->num_class));
  CU(dev_alloc(&c->d_gh, (size_t)num_data));
  CU(dev_alloc(&c->d_score_base, (size_t)num_data * c->num_class));
  CU(cudaMemsetAsync(c->d_score_base, 0, sizeof(double) * num_data * c->num_class, c->stream));
  c->cur_class = 0; c->d_grad = c->d_grad_base; c->d_hess = c->d_hess_base; c->d_score = c->d_score_base;
  CU(dev_alloc(&c->d_indices[0], (size_t)num_data));
  CU(dev_alloc(&c->d_indices[1], (size_t)num_data));
  const size_t ntiles = ((size_t)num_data + PART_TILE - 1) / PART_TILE;
  c->part_sync_stride = c->num_sms * 4 + 2;
  CU(dev_alloc(&c->d_part_sync, (size_t)LGB_KMAX * c->part_sync_stride));
  CU(cudaMemset(c->d_part_sync, 0, sizeof(unsigned long long) * (size_t)LGB_KMAX * c->part_sync_stride));
  c->bits_stride = (long long)ntiles * (PART_TILE / 32);
  CU(dev_alloc(&c->d_part_bits, (size_t)LGB_KMAX * c->bits_stride));
  CU(dev_alloc(&c->d_export, (size_t)2 * c->total_bin));
  c->alloc_leaves = 0;  // pool depends on total_bin
  drop_graph(c);
  c->has_label = c->has_weight = false;
  c->tree_begun = false;
  return 0;
}

int lgbcu_set_dataset(LgbCuContext* c, int32_t num_data, int32_t num_features, const uint8_t* bins,
                      const LgbCuFeatureMeta* meta) {
  NEED(c && bins, "lgbcu_set_dataset: NULL argument");
  USE_DEVICE(c->device);
  if (setup_dataset_common(c, num_data, num_features, meta)) return -1;
  {
    // one contiguous H2D copy, then pad rows to the 16-byte stride on the device (a strided 2-D copy of 28-byte rows
    // crawls: one DMA descriptor per row)
    struct Tmp { uint8_t* p = nullptr; ~Tmp() { if (p) lgb_dev_free(p); } } packed;   // released on every way out
    // (the matrix has one byte column per feature, two -- low byte first -- per wide feature: c->FS in all)
    CU(dev_alloc(&packed.p, (size_t)num_data * c->FS));
    CU(cudaMemcpyAsync(packed.p, bins, (size_t)num_data * c->FS, cudaMemcpyHostToDevice, c->stream));
    LAUNCH(k_pad_rows, grid_for(c, (long long)num_data * (c->stride / 4), 256), 256, 0, c->stream, (long long)num_data,
           c->FS, c->stride, packed.p, c->d_bins);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(c->stream));
  }
  LAUNCH(k_transpose_bins, grid_for(c, num_data, 256), 256, 0, c->stream, num_data, c->FS, c->stride, c->d_bins,
         c->d_bins_col, c->col_stride);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_set_dataset_columns(LgbCuContext* c, int32_t num_data, int32_t num_features, const uint8_t* const* columns,
                              const LgbCuFeatureMeta* meta) {
  NEED(c && columns, "lgbcu_set_dataset_columns: NULL argument");
  USE_DEVICE(c->device);
  if (setup_dataset_common(c, num_data, num_features, meta)) return -1;
  CU(cudaMemsetAsync(c->d_bins, 0, (size_t)num_data * c->stride, c->stream));
  struct Tmp { uint8_t* p[2] = {nullptr, nullptr}; ~Tmp() { for (auto q : p) if (q) lgb_dev_free(q); } } T;   // released on every way out
  uint8_t** tmp = T.p;
  CU(dev_alloc(&tmp[0], (size_t)num_data));
  CU(dev_alloc(&tmp[1], (size_t)num_data));
  NEED(!c->any_wide, "lgbcu_set_dataset_columns: a feature has more than 256 stored values; pass the byte matrix to lgbcu_set_dataset");
  for (int f = 0; f < num_features; ++f) {
    uint8_t* t = tmp[f & 1];
    if (f >= 2) CU(cudaStreamSynchronize(c->stream));  // the staging buffer two columns back is free again
    CU(cudaMemcpyAsync(t, columns[f], num_data, cudaMemcpyHostToDevice, c->stream));
    LAUNCH(k_scatter_column, grid_for(c, num_data, 256), 256, 0, c->stream, num_data, t, c->d_bins, c->stride, f);
  }
  LAUNCH(k_transpose_bins, grid_for(c, num_data, 256), 256, 0, c->stream, num_data, c->FS, c->stride, c->d_bins,
         c->d_bins_col, c->col_stride);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_set_tree_params(LgbCuContext* c, const LgbCuTreeParams* p) {
  NEED(c && p, "lgbcu_set_tree_params: NULL argument");
  NEED(p->num_leaves >= 2, "lgbcu_set_tree_params: num_leaves must be >= 2");
  DevParams d;
  d.num_leaves = p->num_leaves; d.max_depth = p->max_depth; d.min_data_in_leaf = p->min_data_in_leaf;
  d.min_sum_hessian = p->min_sum_hessian_in_leaf; d.lambda_l1 = p->lambda_l1; d.lambda_l2 = p->lambda_l2;
  d.min_gain_to_split = p->min_gain_to_split; d.max_delta_step = p->max_delta_step; d.path_smooth = p->path_smooth;
  d.use_l1 = p->lambda_l1 > 0; d.use_max_output = p->max_delta_step > 0; d.use_smoothing = p->path_smooth > LGB_EPS;
  c->params = d; c->params_set = true;
  drop_graph(c);  // kernel arguments (DevParams) are baked into the captured graph
  return 0;
}

int lgbcu_set_hist_mode(LgbCuContext* c, int mode) {
  NEED(c, "lgbcu: NULL context");
  NEED(mode >= 0 && mode <= 2, "lgbcu_set_hist_mode: unknown mode %d", mode);
  c->hist_mode = mode;
  drop_graph(c);
  return 0;
}

int lgbcu_set_label(LgbCuContext* c, const float* label, int32_t n) {
  NEED(c && label, "lgbcu_set_label: NULL argument");
  NEED(n == c->N, "lgbcu_set_label: length %d != num_data %d", n, c->N);
  USE_DEVICE(c->device);
  CU(dev_alloc(&c->d_label, (size_t)n));
  CU(cudaMemcpyAsync(c->d_label, label, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  c->has_label = true; c->rank_ready = false;
  return 0;
}

int lgbcu_set_weight(LgbCuContext* c, const float* weight, int32_t n) {
  NEED(c, "lgbcu: NULL context");
  if (!weight) { c->has_weight = false; return 0; }
  NEED(n == c->N, "lgbcu_set_weight: length %d != num_data %d", n, c->N);
  USE_DEVICE(c->device);
  CU(dev_alloc(&c->d_weight, (size_t)n));
  CU(cudaMemcpyAsync(c->d_weight, weight, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  c->has_weight = true;
  return 0;
}

int lgbcu_set_query(LgbCuContext* c, const int32_t* boundaries, int32_t num_queries) {
  NEED(c && boundaries && num_queries > 0, "lgbcu_set_query: bad argument");
  NEED(boundaries[0] == 0 && boundaries[num_queries] == c->N, "lgbcu_set_query: boundaries must span [0, num_data]");
  USE_DEVICE(c->device);
  c->h_query.assign(boundaries, boundaries + num_queries + 1);
  c->num_queries = num_queries;
  c->max_query = 0;
  for (int q = 0; q < num_queries; ++q) {
    NEED(boundaries[q + 1] >= boundaries[q], "lgbcu_set_query: boundaries not monotone");
    c->max_query = std::max(c->max_query, boundaries[q + 1] - boundaries[q]);
  }
  CU(dev_alloc(&c->d_query, (size_t)num_queries + 1));
  c->xe_ready = false;
  CU(cudaMemcpyAsync(c->d_query, boundaries, sizeof(int) * (num_queries + 1), cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  c->rank_ready = false;
  return 0;
}

int lgbcu_set_objective(LgbCuContext* c, int kind, double p0, double p1, double p2) {
  NEED(c, "lgbcu: NULL context");
  NEED(kind >= LGBCU_OBJ_NONE && kind <= LGBCU_OBJ_XENDCG, "lgbcu_set_objective: unknown kind %d", kind);
  if (kind == LGBCU_OBJ_BINARY || kind == LGBCU_OBJ_LAMBDARANK) NEED(p0 > 0, "Sigmoid parameter %f should be greater than zero", p0);
  c->objective = kind; c->obj_p[0] = p0; c->obj_p[1] = p1; c->obj_p[2] = p2;
  c->rank_ready = false; c->xe_ready = false;
  return 0;
}

int lgbcu_set_label_gain(LgbCuContext* c, const double* label_gain, int32_t n) {
  NEED(c && label_gain && n > 0, "lgbcu_set_label_gain: bad argument");
  USE_DEVICE(c->device);
  CU(dev_alloc(&c->d_label_gain, (size_t)n));
  CU(cudaMemcpyAsync(c->d_label_gain, label_gain, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  c->num_label_gain = n; c->rank_ready = false;
  return 0;
}

// Multiclass: K planes of scores / gradients / hessians.  Call before the dataset is set (or the planes are rebuilt,
// zeroed).  lgbcu_set_class picks the plane every later call trains / scores / reads.
int lgbcu_set_num_class(LgbCuContext* c, int32_t num_class) {
  NEED(c && num_class >= 1 && num_class <= 4096, "lgbcu_set_num_class: bad argument");
  if (num_class == c->num_class) return 0;
  USE_DEVICE(c->device);
  c->num_class = num_class; c->cur_class = 0;
  if (c->N > 0) {
    CU(dev_alloc(&c->d_grad_base, (size_t)c->N * num_class));
    CU(dev_alloc(&c->d_hess_base, (size_t)c->N * num_class));
    CU(dev_alloc(&c->d_score_base, (size_t)c->N * num_class));
    CU(cudaMemsetAsync(c->d_score_base, 0, sizeof(double) * (size_t)c->N * num_class, c->stream));
    c->d_grad = c->d_grad_base; c->d_hess = c->d_hess_base; c->d_score = c->d_score_base;
  }
  NEED(c->valid.empty(), "lgbcu_set_num_class: validation sets were added before the number of classes was known");
  return 0;
}

int lgbcu_set_ova_label_weights(LgbCuContext* c, const double* neg_pos, int32_t num_class) {
  NEED(c && neg_pos && num_class >= 1, "lgbcu_set_ova_label_weights: bad argument");
  c->ova_w.assign(neg_pos, neg_pos + 2 * (size_t)num_class);
  return 0;
}

int lgbcu_set_class(LgbCuContext* c, int32_t k) {
  NEED(c && k >= 0 && k < c->num_class, "lgbcu_set_class: class %d out of range", (int)k);
  c->cur_class = k;
  if (c->N > 0) {
    c->d_grad = c->d_grad_base + (size_t)k * c->N; c->d_hess = c->d_hess_base + (size_t)k * c->N;
    c->d_score = c->d_score_base + (size_t)k * c->N;
  }
  for (ValidSet& v : c->valid) v.d_score = v.d_score_base + (size_t)k * v.n;
  return 0;
}

int lgbcu_get_gradients(LgbCuContext* c) {
  NEED(c && c->N > 0, "lgbcu_get_gradients: dataset not set");
  NEED(c->has_label, "lgbcu_get_gradients: label not set");
  USE_DEVICE(c->device);
  const float* w = c->has_weight ? c->d_weight : nullptr;
  const int g = grid_for(c, c->N, 256);
  if (c->objective == LGBCU_OBJ_BINARY) {
    LAUNCH(k_grad_binary, g, 256, 0, c->stream, c->N, c->d_score, c->d_label, w, c->obj_p[0], c->obj_p[1], c->obj_p[2], c->d_grad, c->d_hess);
  } else if (c->objective == LGBCU_OBJ_L2) {
    LAUNCH(k_grad_l2, g, 256, 0, c->stream, c->N, c->d_score, c->d_label, w, c->d_grad, c->d_hess, c->obj_p[0] != 0.0 ? 1 : 0);
  } else if (c->objective == LGBCU_OBJ_LAMBDARANK) {
    if (rank_prepare(c)) return -1;
    if (rank_launch(c)) return -1;
  } else if (c->objective == LGBCU_OBJ_XENDCG) {
    NEED(c->d_query && c->num_queries > 0, "Ranking tasks require query information");
    if (!c->xe_ready) {
      std::vector<unsigned> st((size_t)c->num_queries);
      for (int q = 0; q < c->num_queries; ++q) st[q] = (unsigned)((int)c->obj_p[0] + q);   // Random(seed_ + i)
      CU(dev_alloc(&c->d_xe_rng, (size_t)c->num_queries));
      CU(dev_alloc(&c->d_xe_rho, (size_t)c->N)); CU(dev_alloc(&c->d_xe_par, (size_t)c->N));
      CU(cudaMemcpyAsync(c->d_xe_rng, st.data(), sizeof(unsigned) * st.size(), cudaMemcpyHostToDevice, c->stream));
      CU(cudaStreamSynchronize(c->stream));
      c->xe_ready = true;
    }
    LAUNCH(k_grad_xendcg, grid_for(c, c->num_queries, 128), 128, 0, c->stream, c->num_queries, c->d_query, c->d_score, c->d_label, w,
           c->d_xe_rng, c->d_xe_rho, c->d_xe_par, c->d_grad, c->d_hess);
  } else if (c->objective == LGBCU_OBJ_MULTICLASSOVA) {
    NEED((int)c->obj_p[0] == c->num_class && (int)c->ova_w.size() == 2 * c->num_class, "lgbcu_get_gradients: multiclassova needs lgbcu_set_num_class and lgbcu_set_ova_label_weights");
    for (int k = 0; k < c->num_class; ++k) {   // one BinaryLogloss per class on its own planes
      const size_t off = (size_t)k * c->N;
      LAUNCH(k_grad_binary, g, 256, 0, c->stream, c->N, c->d_score_base + off, c->d_label, w, c->obj_p[1], c->ova_w[2 * k], c->ova_w[2 * k + 1],
             c->d_grad_base + off, c->d_hess_base + off, k);
    }
  } else if (c->objective == LGBCU_OBJ_HUBER || c->objective == LGBCU_OBJ_FAIR || c->objective == LGBCU_OBJ_POISSON ||
             c->objective == LGBCU_OBJ_GAMMA || c->objective == LGBCU_OBJ_TWEEDIE || c->objective == LGBCU_OBJ_XENTROPY ||
             c->objective == LGBCU_OBJ_XENTLAMBDA) {
    LAUNCH(k_grad_regression, g, 256, 0, c->stream, c->N, c->objective, c->obj_p[0], c->d_score, c->d_label, w, c->d_grad, c->d_hess);
  } else if (c->objective == LGBCU_OBJ_L1) {
    LAUNCH(k_grad_l1, g, 256, 0, c->stream, c->N, c->d_score, c->d_label, w, c->d_grad, c->d_hess, 0);
  } else if (c->objective == LGBCU_OBJ_MAPE) {
    // (with weights the reference declares the hessian constant although it is not, dataset.cpp's count-times-first-
    // hessian histogram then disagrees with the leaf sums: that case is refused instead of imitated)
    NEED(!w, "lgbcu_get_gradients: the mape objective with weights is outside this library's path");
    LAUNCH(k_grad_l1, g, 256, 0, c->stream, c->N, c->d_score, c->d_label, w, c->d_grad, c->d_hess, 1);
  } else if (c->objective == LGBCU_OBJ_QUANTILE) {
    LAUNCH(k_grad_quantile, g, 256, 0, c->stream, c->N, c->d_score, c->d_label, w, (float)c->obj_p[0], c->d_grad, c->d_hess);
  } else if (c->objective == LGBCU_OBJ_MULTICLASS) {
    NEED((int)c->obj_p[0] == c->num_class, "lgbcu_get_gradients: the objective has %d classes, the learner %d (lgbcu_set_num_class)", (int)c->obj_p[0], c->num_class);
    LAUNCH(k_grad_multiclass, g, 256, 0, c->stream, c->N, c->num_class, c->d_score_base, c->d_label, w, c->obj_p[1], c->d_grad_base, c->d_hess_base);
  } else {
    return fail("lgbcu_get_gradients: no objective set");
  }
  CU(cudaGetLastError());
  return 0;
}

int lgbcu_set_gradients(LgbCuContext* c, const float* grad, const float* hess, int32_t n) {
  NEED(c && grad && hess, "lgbcu_set_gradients: NULL argument");
  NEED(n == c->N, "lgbcu_set_gradients: length %d != num_data %d", n, c->N);
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(c->d_grad, grad, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(c->d_hess, hess, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_copy_gradients(LgbCuContext* c, float* grad, float* hess, int32_t n) {
  NEED(c && grad && hess && n == c->N, "lgbcu_copy_gradients: bad argument");
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(grad, c->d_grad, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(hess, c->d_hess, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_set_score(LgbCuContext* c, const double* score, int32_t n) {
  NEED(c && c->N > 0, "lgbcu_set_score: dataset not set");
  USE_DEVICE(c->device);
  if (!score) { CU(cudaMemsetAsync(c->d_score, 0, sizeof(double) * c->N, c->stream)); return 0; }
  NEED(n == c->N, "lgbcu_set_score: length %d != num_data %d", n, c->N);
  CU(cudaMemcpyAsync(c->d_score, score, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_add_score_const(LgbCuContext* c, double val) {
  NEED(c && c->N > 0, "lgbcu_add_score_const: dataset not set");
  USE_DEVICE(c->device);
  LAUNCH(k_add_const, grid_for(c, c->N, 256), 256, 0, c->stream, c->N, c->d_score, val);
  CU(cudaGetLastError());
  return 0;
}

// ScoreUpdater::MultiplyScore of the current class's training scores (boosting=rf keeps the scores as running means)
int lgbcu_multiply_score(LgbCuContext* c, double val) {
  NEED(c && c->N > 0, "lgbcu_multiply_score: dataset not set");
  USE_DEVICE(c->device);
  LAUNCH(k_mul_const, grid_for(c, c->N, 256), 256, 0, c->stream, c->N, c->d_score, val);
  CU(cudaGetLastError());
  return 0;
}

int lgbcu_copy_score(LgbCuContext* c, double* score, int32_t n) {
  NEED(c && score && n == c->N, "lgbcu_copy_score: bad argument");
  USE_DEVICE(c->device);
  // The caller's buffer is pageable (and usually untouched): stream the scores through two page-locked staging
  // buffers and let the host threads move each chunk on while the next one is in flight.
  const size_t CH = (size_t)1 << 19;  // doubles per chunk (4 MB)
  const int nthreads = host_copy_threads(c);
  if ((size_t)n <= CH) {
    CU(cudaMemcpyAsync(score, c->d_score, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return 0;
  }
  {
    // The destination is usually a freshly allocated (untouched) host buffer: populate its pages up front, in
    // parallel and without one trap per page (MADV_POPULATE_WRITE, Linux >= 5.14; ignored where unsupported), so that
    // the copy loop below runs at memory speed instead of page-fault speed.
#ifndef MADV_POPULATE_WRITE
#define MADV_POPULATE_WRITE 23
#endif
    const size_t page = (size_t)sysconf(_SC_PAGESIZE);
    const uintptr_t lo = ((uintptr_t)score + page - 1) / page * page, hi = ((uintptr_t)(score + n)) / page * page;
    if (hi > lo) {
      const size_t slice = (size_t)8 << 20;
      const long long pieces = (long long)((hi - lo + slice - 1) / slice);
#pragma omp parallel for schedule(static) num_threads(nthreads)
      for (long long p = 0; p < pieces; ++p) {
        const uintptr_t a = lo + (uintptr_t)p * slice, b = std::min<uintptr_t>(hi, a + slice);
        (void)madvise(reinterpret_cast<void*>(a), b - a, MADV_POPULATE_WRITE);
      }
    }
  }
  if (!c->h_stage[0]) {
    for (int i = 0; i < 2; ++i) {
      CU(lgb_host_malloc(reinterpret_cast<void**>(&c->h_stage[i]), CH * sizeof(double)));
      CU(cudaEventCreateWithFlags(&c->stage_ev[i], cudaEventDisableTiming));
    }
  }
  const size_t nchunk = ((size_t)n + CH - 1) / CH;
  auto chunk_len = [&](size_t k) { return std::min(CH, (size_t)n - k * CH); };
  CU(cudaMemcpyAsync(c->h_stage[0], c->d_score, sizeof(double) * chunk_len(0), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaEventRecord(c->stage_ev[0], c->stream));
  for (size_t k = 0; k < nchunk; ++k) {
    const int cur = (int)(k & 1), nxt = cur ^ 1;
    if (k + 1 < nchunk) {
      CU(cudaMemcpyAsync(c->h_stage[nxt], c->d_score + (k + 1) * CH, sizeof(double) * chunk_len(k + 1), cudaMemcpyDeviceToHost, c->stream));
      CU(cudaEventRecord(c->stage_ev[nxt], c->stream));
    }
    CU(cudaEventSynchronize(c->stage_ev[cur]));
    const size_t len = chunk_len(k);
    double* dst = score + k * CH;
    const double* src = c->h_stage[cur];
    const long long pieces = (long long)((len + 32767) / 32768);
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (long long p = 0; p < pieces; ++p) {
      const size_t a = (size_t)p * 32768, b = std::min(len, a + 32768);
      std::memcpy(dst + a, src + a, (b - a) * sizeof(double));
    }
  }
  return 0;
}

int lgbcu_add_score_by_partition(LgbCuContext* c, const double* leaf_value, int32_t num_leaves) {
  NEED(c && leaf_value, "lgbcu_add_score_by_partition: NULL argument");
  NEED(c->tree_begun, "lgbcu_add_score_by_partition: no tree has been grown");
  NEED(num_leaves >= 1 && num_leaves <= c->alloc_leaves, "lgbcu_add_score_by_partition: num_leaves %d out of range", num_leaves);
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(c->d_leaf_value, leaf_value, sizeof(double) * num_leaves, cudaMemcpyHostToDevice, c->stream));
  const DevView V = c->view();
  // one block per 4096-position chunk; a leaf's last chunk may be partial, hence the num_leaves spare blocks
  const int chunks = (c->root_count + ADD_SCORE_CHUNK - 1) / ADD_SCORE_CHUNK + num_leaves;
  LAUNCH(k_add_score_partition, chunks, ADD_SCORE_THREADS, 0, c->stream, V, num_leaves, c->d_leaf_value, c->d_score);
  CU(cudaGetLastError());
  // leaf_value is pageable host memory: the async copy above has been staged by the time it returns
  return 0;
}

// Categorical splits.  Config's cat_smooth / cat_l2 / max_cat_threshold / min_data_per_group / max_cat_to_onehot
// (feature_histogram.hpp:277-513 reads them through meta_->config)
int lgbcu_set_cat_params(LgbCuContext* c, double cat_smooth, double cat_l2, int32_t max_cat_threshold, int32_t min_data_per_group,
                         int32_t max_cat_to_onehot) {
  NEED(c, "lgbcu: NULL context");
  NEED(cat_smooth >= 0 && cat_l2 >= 0 && max_cat_threshold > 0 && min_data_per_group > 0 && max_cat_to_onehot > 0,
       "lgbcu_set_cat_params: parameter out of range");
  c->cat_params = DevCatParams{cat_smooth, cat_l2, max_cat_threshold, min_data_per_group, max_cat_to_onehot, 0};
  drop_graph(c);   // the captured launches hold the parameters by value
  return 0;
}

// Monotone constraints, basic method (Config::monotone_constraints per inner feature: -1 / 0 / +1, and monotone_penalty;
// LeafConstraintsBase / BasicLeafConstraints, src/treelearner/monotone_constraints.hpp).  types = NULL: none.
int lgbcu_set_monotone(LgbCuContext* c, const int8_t* types, int32_t num_features, double penalty) {
  NEED(c && c->F > 0, "lgbcu_set_monotone: dataset not set");
  USE_DEVICE(c->device);
  drop_graph(c);   // the captured launches hold the view by value
  if (!types) { c->use_mc = false; c->general = c->any_wide || c->any_cat; return 0; }
  NEED(num_features == c->F, "lgbcu_set_monotone: %d types for %d features", num_features, c->F);
  NEED(penalty >= 0.0, "lgbcu_set_monotone: monotone_penalty should be >= 0");
  for (int f = 0; f < num_features; ++f) {
    NEED(types[f] >= -1 && types[f] <= 1, "lgbcu_set_monotone: type of feature %d is %d", f, (int)types[f]);
    NEED(types[f] == 0 || !c->h_feats[f].cat, "The output cannot be monotone with respect to categorical features");
  }
  NEED(c->F < (1 << KEY_FEAT_BITS_WIDE), "lgbcu_set_monotone: %d features exceed the supported %d for constrained data", c->F, (1 << KEY_FEAT_BITS_WIDE) - 1);
  CU(dev_alloc(&c->d_monotone, (size_t)num_features));
  CU(cudaMemcpyAsync(c->d_monotone, types, (size_t)num_features, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  c->use_mc = true; c->mc_penalty = penalty; c->general = true;
  return 0;
}

// The bin sets of the tree lgbcu_train_tree just grew: a categorical record's `threshold` is an index into them
int lgbcu_tree_cat_sets(LgbCuContext* c, uint32_t* out, int32_t cap_sets, int32_t* num_sets) {
  NEED(c && num_sets, "lgbcu_tree_cat_sets: NULL argument");
  NEED(c->tree_begun && c->h_ctl, "lgbcu_tree_cat_sets: no tree has been grown");
  const int n = c->any_cat ? c->h_ctl->cat_count : 0;
  *num_sets = n;
  if (n == 0 || out == nullptr) return 0;   // out = NULL: the count alone
  NEED(cap_sets >= n, "lgbcu_tree_cat_sets: room for %d sets, the tree has %d", cap_sets, n);
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(out, c->d_cat_pool, sizeof(unsigned) * LGB_CAT_WORDS * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

// Stage the bin sets ([num_sets][8] words) that the categorical nodes of the trees passed to lgbcu_add_score_by_tree /
// lgbcu_valid_add_tree from now on name in threshold_bin (Tree::cat_threshold_inner_, tree.h)
int lgbcu_set_tree_cat_sets(LgbCuContext* c, const uint32_t* sets, int32_t num_sets) {
  NEED(c && num_sets >= 0 && (sets || num_sets == 0), "lgbcu_set_tree_cat_sets: bad argument");
  USE_DEVICE(c->device);
  if (c->tree_cat_cap < num_sets) { CU(dev_alloc(&c->d_tree_cat, (size_t)num_sets * LGB_CAT_WORDS)); c->tree_cat_cap = num_sets; }
  if (num_sets > 0) {
    CU(cudaMemcpyAsync(c->d_tree_cat, sets, sizeof(unsigned) * LGB_CAT_WORDS * (size_t)num_sets, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  c->tree_cat_count = num_sets;
  return 0;
}

int lgbcu_add_score_by_tree(LgbCuContext* c, int32_t num_leaves, const int32_t* split_feature_inner,
                            const int32_t* threshold_bin, const int8_t* decision_type, const int32_t* left_child,
                            const int32_t* right_child, const double* leaf_value, const int32_t* rows, int32_t num_rows,
                            int64_t rows_tag) {
  NEED(c && c->N > 0, "lgbcu_add_score_by_tree: dataset not set");
  NEED(num_leaves >= 1 && leaf_value, "lgbcu_add_score_by_tree: bad tree");
  USE_DEVICE(c->device);
  const bool device_oob = rows == LGBCU_DEVICE_OOB;   // the out-of-bag tail of the bag lgbcu_bagging drew
  if (device_oob) { NEED(c->bag_ready, "lgbcu_add_score_by_tree: no bag has been drawn"); rows = nullptr; num_rows = c->N - c->bag_count; }
  const int n = (rows || device_oob) ? num_rows : c->N;
  if (n <= 0) return 0;
  if (num_leaves == 1) {  // constant tree: score += leaf_value[0] (tree.cpp:110-118)
    if (leaf_value[0] == 0.0) return 0;
    NEED(rows == nullptr && !device_oob, "lgbcu_add_score_by_tree: constant tree over a row subset is not supported");
    return lgbcu_add_score_const(c, leaf_value[0]);
  }
  TreeDev T;
  if (upload_tree(c, num_leaves, split_feature_inner, threshold_bin, decision_type, left_child, right_child, leaf_value, &T)) return -1;
  const int* d_rows = device_oob ? c->d_bag + c->bag_count : nullptr;
  if (rows) {
    const bool reuse = rows_tag != 0 && rows_tag == c->rows_tag && num_rows == c->rows_count;
    if (!reuse) {
      if (c->rows_cap < num_rows) { CU(dev_alloc(&c->d_rows, (size_t)num_rows)); c->rows_cap = num_rows; }
      CU(cudaMemcpyAsync(c->d_rows, rows, sizeof(int) * num_rows, cudaMemcpyHostToDevice, c->stream));
      c->rows_tag = rows_tag; c->rows_count = num_rows;
    }
    d_rows = c->d_rows;
  }
  LAUNCH(k_add_score_tree, grid_for(c, n, 256), 256, 0, c->stream, c->view(), T, d_rows, n, c->d_score);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(c->stream));  // the caller's host arrays may go away
  return 0;
}

int lgbcu_begin_tree(LgbCuContext* c, const int32_t* bag_indices, int32_t bag_count, const int8_t* feature_used) {
  NEED(c, "lgbcu: NULL context");
  USE_DEVICE(c->device);
  if (ensure_tree_buffers(c)) return -1;
  if (setup_ownership(c)) return -1;
  if (xchg_setup(c)) return -1;
  int identity = 1;
  if (bag_indices == LGBCU_DEVICE_BAG) {   // the bag lgbcu_bagging drew, already on the device
    NEED(c->bag_ready, "lgbcu_begin_tree: no bag has been drawn (lgbcu_bagging)");
    CU(cudaMemcpyAsync(c->d_indices[0], c->d_bag, sizeof(int) * (size_t)c->bag_count, cudaMemcpyDeviceToDevice, c->stream));
    c->root_count = c->bag_count;
    identity = 0;
    bag_indices = nullptr;
  } else if (bag_indices) {
    NEED(bag_count > 0 && bag_count <= c->N, "lgbcu_begin_tree: bag_count %d out of range", bag_count);
    CU(cudaMemcpyAsync(c->d_indices[0], bag_indices, sizeof(int) * bag_count, cudaMemcpyHostToDevice, c->stream));
    c->root_count = bag_count;
    identity = 0;
  } else {
    LAUNCH(k_iota, grid_for(c, c->N, 256), 256, 0, c->stream, c->N, c->d_indices[0]);
    c->root_count = c->N;
  }
  if (feature_used) {
    CU(cudaMemcpyAsync(c->d_feat_used, feature_used, c->F, cudaMemcpyHostToDevice, c->stream));
  } else {
    CU(cudaMemsetAsync(c->d_feat_used, 1, c->F, c->stream));
  }
  // only the root's slot: every later slot is zeroed by the partition launch that hands it out
  CU(cudaMemsetAsync(c->xchg ? c->d_pool_local : c->d_pool, 0, sizeof(HistEntry) * (size_t)c->total_bin, c->stream));
  LAUNCH(k_reset_ctl_max, 1, 1, 0, c->stream, c->d_ctl, c->root_count, c->N);
  LAUNCH(k_absmax, grid_for(c, c->N, 256), 256, 0, c->stream, c->N, c->d_grad, c->d_hess, c->d_gh, c->d_ctl);
  if (comm_allreduce_tree_setup(c)) return -1;
  LAUNCH(k_prepare_tree, 1, 256, 0, c->stream, c->view(), c->root_count, identity, c->lookahead);
  CU(cudaGetLastError());
  if (bag_indices || feature_used) CU(cudaStreamSynchronize(c->stream));  // caller may reuse its buffers
  c->tree_begun = true;
  return 0;
}

int lgbcu_num_total_bin(LgbCuContext* c, int32_t* out) {
  NEED(c && out, "lgbcu_num_total_bin: NULL argument");
  *out = c->total_bin;
  return 0;
}

int lgbcu_hist_build(LgbCuContext* c, int32_t leaf) {
  NEED(c && c->tree_begun, "lgbcu_hist_build: call lgbcu_begin_tree first");
  NEED(leaf >= 0 && leaf < c->params.num_leaves, "lgbcu_hist_build: leaf %d out of range", leaf);
  USE_DEVICE(c->device);
  const DevView V = c->view();
  LAUNCH(k_zero_slot, grid_for(c, c->total_bin, 256), 256, 0, c->stream, V, leaf);
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (c->profile) {
    if (c->ev.size() < 2) { c->ev.resize(2); CU(cudaEventCreate(&c->ev[0])); CU(cudaEventCreate(&c->ev[1])); }
    e0 = c->ev[0]; e1 = c->ev[1];
    CU(cudaEventRecord(e0, c->stream));
  }
  if (launch_hist(c, leaf)) return -1;
  if (c->profile) {
    CU(cudaEventRecord(e1, c->stream));
    CU(cudaEventSynchronize(e1));
    float ms = 0; CU(cudaEventElapsedTime(&ms, e0, e1));
    c->prof_hist_ms += ms;
  }
  if (leaf == 0) LAUNCH(k_root_sums, 1, 32, 0, c->stream, V);
  CU(cudaGetLastError());
  return 0;
}

// k_find_split comes in two instantiations so that data without wide features runs the kernel without the wide search
// in it (the 8-bit kernel sits at its register cap: the extra code cost 0.1 ms per iteration when it was one kernel)
static void launch_find_split(LgbCuContext* c, dim3 grid, const DevView& V, int mode, int leaf) {
  if (c->general) LAUNCH(k_find_split<true>, grid, FS_THREADS, 0, c->stream, V, mode, leaf);
  else LAUNCH(k_find_split<false>, grid, FS_THREADS, 0, c->stream, V, mode, leaf);
}

int lgbcu_hist_read(LgbCuContext* c, int32_t leaf, double* out, int64_t out_len) {
  NEED(c && out && c->tree_begun, "lgbcu_hist_read: bad argument / no tree begun");
  NEED(out_len == 2LL * c->total_bin, "lgbcu_hist_read: out_len %lld != 2*total_bin %d", (long long)out_len, 2 * c->total_bin);
  NEED(leaf >= 0 && leaf < c->params.num_leaves, "lgbcu_hist_read: leaf %d out of range", leaf);
  USE_DEVICE(c->device);
  LAUNCH(k_hist_export, c->F, 256, 0, c->stream, c->view(), leaf, c->d_export);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(out, c->d_export, sizeof(double) * 2 * c->total_bin, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_hist_subtract(LgbCuContext* c, int32_t parent_leaf, int32_t smaller_leaf) {
  NEED(c && c->tree_begun, "lgbcu_hist_subtract: no tree begun");
  USE_DEVICE(c->device);
  LAUNCH(k_hist_subtract, grid_for(c, c->total_bin, 256), 256, 0, c->stream, c->view(), parent_leaf, smaller_leaf);
  CU(cudaGetLastError());
  return 0;
}

int lgbcu_find_best_split(LgbCuContext* c, int32_t leaf, LgbCuSplitInfo* out) {
  NEED(c && out && c->tree_begun, "lgbcu_find_best_split: bad argument / no tree begun");
  NEED(leaf >= 0 && leaf < c->params.num_leaves, "lgbcu_find_best_split: leaf %d out of range", leaf);
  USE_DEVICE(c->device);
  launch_find_split(c, dim3(c->F, 1), c->view(), 1, leaf);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(c->h_split, c->d_best + leaf, sizeof(DevSplit), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  const DevSplit& s = *c->h_split;
  out->feature = s.feature; out->real_feature = s.real_feature; out->threshold = s.threshold;
  out->default_left = s.default_left; out->left_count = s.left_count; out->right_count = s.right_count;
  out->gain = s.gain; out->left_output = s.left_output; out->right_output = s.right_output;
  out->left_sum_gradient = s.left_sum_g; out->left_sum_hessian = s.left_sum_h;
  out->right_sum_gradient = s.right_sum_g; out->right_sum_hessian = s.right_sum_h;
  return 0;
}

int lgbcu_partition(LgbCuContext* c, int32_t leaf, int32_t inner_feature, int32_t threshold, int32_t default_left,
                    int32_t right_leaf, int32_t* left_count) {
  NEED(c && left_count && c->tree_begun, "lgbcu_partition: bad argument / no tree begun");
  NEED(leaf >= 0 && leaf < c->params.num_leaves && right_leaf > 0 && right_leaf < c->params.num_leaves && leaf != right_leaf,
       "lgbcu_partition: leaf ids out of range");
  NEED(inner_feature >= 0 && inner_feature < c->F, "lgbcu_partition: feature %d out of range", inner_feature);
  USE_DEVICE(c->device);
  const DevView V = c->view();
  LAUNCH(k_set_split, 1, 1, 0, c->stream, V, leaf, inner_feature, threshold, default_left, right_leaf);
  if (launch_partition(c, c->N, 1)) return -1;
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(c->h_ctl, c->d_ctl, sizeof(DevCtl), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  *left_count = c->h_ctl->split_left_count;
  return 0;
}

int lgbcu_leaf_indices(LgbCuContext* c, int32_t leaf, int32_t* out, int32_t capacity, int32_t* count) {
  NEED(c && count && c->tree_begun, "lgbcu_leaf_indices: bad argument / no tree begun");
  NEED(leaf >= 0 && leaf < c->params.num_leaves, "lgbcu_leaf_indices: leaf %d out of range", leaf);
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(c->h_leaf, c->d_leaves + leaf, sizeof(DevLeaf), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  const DevLeaf L = *c->h_leaf;
  *count = L.count;
  if (out) {
    NEED(capacity >= L.count, "lgbcu_leaf_indices: capacity %d < leaf count %d", capacity, L.count);
    if (L.count > 0) {
      CU(cudaMemcpyAsync(out, c->d_indices[L.buf] + L.begin, sizeof(int) * L.count, cudaMemcpyDeviceToHost, c->stream));
      CU(cudaStreamSynchronize(c->stream));
    }
  }
  return 0;
}

int lgbcu_set_leaf_sums(LgbCuContext* c, int32_t leaf, double sum_g, double sum_h, double output) {
  NEED(c && c->tree_begun, "lgbcu_set_leaf_sums: no tree begun");
  NEED(leaf >= 0 && leaf < c->params.num_leaves, "lgbcu_set_leaf_sums: leaf %d out of range", leaf);
  USE_DEVICE(c->device);
  LAUNCH(k_set_leaf_sums, 1, 1, 0, c->stream, c->view(), leaf, sum_g, sum_h, output);
  CU(cudaGetLastError());
  return 0;
}

int lgbcu_get_leaf_sums(LgbCuContext* c, int32_t leaf, double* sum_g, double* sum_h, int32_t* count) {
  NEED(c && c->tree_begun, "lgbcu_get_leaf_sums: no tree begun");
  NEED(leaf >= 0 && leaf < c->params.num_leaves, "lgbcu_get_leaf_sums: leaf %d out of range", leaf);
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(c->h_leaf, c->d_leaves + leaf, sizeof(DevLeaf), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (sum_g) *sum_g = c->h_leaf->sum_g;
  if (sum_h) *sum_h = c->h_leaf->sum_h;
  if (count) *count = c->h_leaf->count;
  return 0;
}

// One round of tree growth = {k_hist, [all-reduce], k_find_split (+ control), k_partition}.  Round 0 handles the root;
// every later round expands up to K leaves and commits at least one (see DevSpec in dev_types.cuh).
#define TREE_CHUNK_ROUNDS 16
static int enqueue_round(LgbCuContext* c, int round, cudaEvent_t* ev) {
  const DevView V = c->view();
  const int K = c->lookahead;
  if (ev) CU(cudaEventRecord(ev[0], c->stream));
  if (launch_hist(c, -1)) return -1;
  if (ev) CU(cudaEventRecord(ev[4], c->stream));  // end of the histogram kernel itself (the all-reduce follows)
  if (c->comm) {  // slots of a round are a function of (round, item): the collective is enqueued blind
    const int first = round == 0 ? 0 : 1 + (round - 1) * K, count = round == 0 ? 1 : K;
    NEED(first + count <= c->pool_slots, "lgbcu_train_tree: round %d needs histogram slots [%d, %d) but the pool has %d", round, first, first + count, c->pool_slots);
    if (c->xchg) {
      XchgPeers P;
      for (int r = 0; r < LGB_MAX_PEERS; ++r) P.flags[r] = r == c->rank ? c->d_xflags : static_cast<unsigned long long*>(c->xpeer_flags[r]);
      if (c->xchg_two_step) {
        LAUNCH(k_xchg_pull, c->num_sms * 2, XCHG_THREADS, 0, c->stream, V, P, 1);
        LAUNCH(k_xchg_pull, c->num_sms * 2, XCHG_THREADS, 0, c->stream, V, P, 2);
      } else {
        LAUNCH(k_xchg_pull, c->num_sms * 2, XCHG_THREADS, 0, c->stream, V, P, 0);
      }
    } else if (comm_reduce_hist(c, first, count, round == 0)) return -1;
  }
  if (round == 0) LAUNCH(k_root_sums, 1, 32, 0, c->stream, V);
  if (ev) CU(cudaEventRecord(ev[1], c->stream));
  if (c->comm && c->dp_owner) {
    // owned scan -> all-gather of the winners -> control step
    launch_find_split(c, dim3(V.f_end - V.f_begin, 2 * K), V, FS_MODE_OWNED, -1);
    if (comm_gather_winners(c)) return -1;
    launch_find_split(c, dim3(1, 1), V, FS_MODE_CONTROL, -1);
  } else {
    launch_find_split(c, dim3(c->F, 2 * K), V, 0, -1);
  }
  if (ev) CU(cudaEventRecord(ev[2], c->stream));
  if (launch_partition(c, c->root_count, 0)) return -1;
  if (ev) CU(cudaEventRecord(ev[3], c->stream));
  return 0;
}

int lgbcu_train_tree(LgbCuContext* c, LgbCuSplitRecord* records, int32_t* num_splits) {
  NEED(c && records && num_splits, "lgbcu_train_tree: NULL argument");
  NEED(c->tree_begun, "lgbcu_train_tree: call lgbcu_begin_tree first");
  USE_DEVICE(c->device);
  const int L = c->params.num_leaves;
  const int K = c->lookahead;
  const bool prof = c->profile;
  const int max_rounds = L;  // the root round + at most L-1 rounds that commit one leaf each
  // rounds enqueued before the host looks at `done` for the first time: what a tree needs when most expansions are
  // committed in the round after they were made (measured: > 95 % on the benchmark workloads), plus some slack
  int first_batch = 2 + ((L - 1) * 8 / 5 + K - 1) / K;
  if (first_batch > max_rounds) first_batch = max_rounds;
  int round = 0;
  const int EV = 5;  // events per round: before hist, after hist + all-reduce, after find, after partition; [4] = after hist alone
  if (prof && (int)c->ev.size() < EV * max_rounds) {
    const size_t old = c->ev.size();
    c->ev.resize(EV * max_rounds);
    for (size_t i = old; i < c->ev.size(); ++i) CU(cudaEventCreate(&c->ev[i]));
  }
  const bool graphs = c->use_graph && !prof && (!c->comm || c->graph_comm);
  // the launch sequence of a range of rounds is fixed (all sizes are read on the device): capture once, replay per tree
  auto capture_rounds = [&](int r0, int r1, cudaGraphExec_t* out) -> int {
    cudaGraph_t g = nullptr;
    CU(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    const long long before = g_launches.load(std::memory_order_relaxed);  // captured launches are not executions
    bool ok = true;
    for (int r = r0; r < r1 && ok; ++r) ok = enqueue_round(c, r, nullptr) == 0;
    cudaError_t ce = cudaStreamEndCapture(c->stream, &g);
    g_launches.fetch_add(before - g_launches.load(std::memory_order_relaxed), std::memory_order_relaxed);
    if (!ok || ce != cudaSuccess || !g) { if (g) cudaGraphDestroy(g); return fail("lgbcu_train_tree: graph capture failed (%s)", cudaGetErrorString(ce)); }
    ce = cudaGraphInstantiate(out, g, 0);
    cudaGraphDestroy(g);
    if (ce != cudaSuccess) { *out = nullptr; return fail("lgbcu_train_tree: graph instantiate failed (%s)", cudaGetErrorString(ce)); }
    return 0;
  };
  // rounds [first_batch + j * TREE_CHUNK_ROUNDS, ...) as one more graph launch
  auto launch_chunk = [&]() -> int {
    const int j = (round - first_batch) / TREE_CHUNK_ROUNDS;
    const int r1 = std::min(max_rounds, round + TREE_CHUNK_ROUNDS);
    if ((int)c->chunk_graphs.size() <= j) c->chunk_graphs.resize((size_t)j + 1, nullptr);
    if (!c->chunk_graphs[j] && capture_rounds(round, r1, &c->chunk_graphs[j])) return -1;
    CU(cudaGraphLaunch(c->chunk_graphs[j], c->stream));
    g_launches.fetch_add((long long)(r1 - round) * 3, std::memory_order_relaxed);
    round = r1;
    return 0;
  };
  if (graphs) {
    if (!c->tree_graph && capture_rounds(0, first_batch, &c->tree_graph)) return -1;
    CU(cudaGraphLaunch(c->tree_graph, c->stream));
    g_launches.fetch_add((long long)first_batch * 3 + 1, std::memory_order_relaxed);
    round = first_batch;
    // Late in a long run the trees are deep and need several times the first batch: enqueue what the previous tree
    // needed before looking at `done` at all (a round enqueued after the tree is finished costs three empty launches)
    while (round < max_rounds && round < c->last_used_rounds + 2)
      if (launch_chunk()) return -1;
  } else {
    for (; round < first_batch; ++round)
      if (enqueue_round(c, round, prof ? &c->ev[EV * round] : nullptr)) return -1;
  }
  // look at the control block; a tree whose expansions were often not committed at once needs more rounds
  while (true) {
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(c->h_ctl, c->d_ctl, sizeof(DevCtl), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    if (c->h_ctl->done || round >= max_rounds) break;
    if (graphs) {
      if (launch_chunk()) return -1;
    } else {
      for (int i = 0; i < 8 && round < max_rounds; ++i, ++round)
        if (enqueue_round(c, round, prof ? &c->ev[EV * round] : nullptr)) return -1;
    }
  }
  c->last_used_rounds = c->h_ctl->round;
  NEED(!c->h_ctl->xchg_error, "lgbcu_train_tree: a peer's histograms did not arrive within the exchange time-out (is every rank training?)");
  NEED(!c->h_ctl->cat_overflow, "lgbcu_train_tree: the pool of categorical bin sets (%d entries) ran out", c->cat_cap);
  NEED(c->h_ctl->done, "lgbcu_train_tree: tree not finished after %d rounds", round);
  const int n = c->h_ctl->num_splits;
  NEED(n >= 0 && n <= L - 1, "lgbcu_train_tree: corrupt split count %d", n);
  if (n > 0) {
    CU(cudaMemcpyAsync(c->h_records, c->d_records, sizeof(LgbCuSplitRecord) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    std::memcpy(records, c->h_records, sizeof(LgbCuSplitRecord) * (size_t)n);
  }
  *num_splits = n;
  c->last_rounds = round;
  if (prof) {
    for (int i = 0; i < 8; ++i) c->loop_phase_ms[i] = 0;
    const int slot[3] = {0, 2, 4};  // hist (+ all-reduce), find + control, partition
    for (int r = 0; r < round; ++r) {
      for (int k = 0; k < 3; ++k) {
        float ms = 0; CU(cudaEventElapsedTime(&ms, c->ev[EV * r + k], c->ev[EV * r + k + 1]));
        c->loop_phase_ms[slot[k]] += ms;
      }
      float hms = 0; CU(cudaEventElapsedTime(&hms, c->ev[EV * r], c->ev[EV * r + 4]));
      c->prof_hist_ms += hms;               // the kernel alone: what the roofline figure is about
      c->loop_phase_ms[1] += hms;           // [0] = histogram + all-reduce, [1] = histogram kernel alone
    }
    c->loop_phase_valid = true;
    if (const char* path = std::getenv("LGBCU_PROF_DUMP")) {  // per-round phase times (us), for offline analysis
      if (FILE* fp = std::fopen(path, "a")) {
        for (int r = 0; r < round; ++r) {
          float t[3];
          for (int k = 0; k < 3; ++k) cudaEventElapsedTime(&t[k], c->ev[EV * r + k], c->ev[EV * r + k + 1]);
          std::fprintf(fp, "%d,%.2f,%.2f,%.2f\n", r, t[0] * 1e3f, t[1] * 1e3f, t[2] * 1e3f);
        }
        std::fclose(fp);
      }
    }
  }
  return 0;
}

int lgbcu_profile_phases(LgbCuContext* c, double* ms8) {
  NEED(c && ms8, "lgbcu_profile_phases: NULL argument");
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(c->h_ctl, c->d_ctl, sizeof(DevCtl), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  for (int i = 0; i < 8; ++i) ms8[i] = c->loop_phase_valid ? c->loop_phase_ms[i] : 0.0;
  ms8[6] = (double)c->last_rounds;       // rounds enqueued for the last tree
  ms8[7] = (double)c->h_ctl->round;      // rounds the tree actually needed (device counter)
  return 0;
}

#ifdef LGBCU_STAGE_TIMERS
extern "C" int lgbcu_debug_stamps(unsigned long long* out, int n) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_stamps, sizeof(unsigned long long) * (size_t)n);
  cudaMemcpyFromSymbol(out + n, g_ends, sizeof(unsigned long long) * (size_t)n);
  static unsigned long long z[3 * 4096];
  cudaMemcpyToSymbol(g_ends, z, sizeof(z));
  cudaMemcpyToSymbol(g_stamps, z, sizeof(z));
  return 0;
}
extern "C" int lgbcu_debug_hist_stages(unsigned long long* out8, int reset) {
  unsigned long long z[8] = {0};
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out8, g_hist_cycles, sizeof(z));
  if (reset) cudaMemcpyToSymbol(g_hist_cycles, z, sizeof(z));
  return 0;
}
extern "C" int lgbcu_debug_stages(unsigned long long* out16, int reset) {
  unsigned long long z[32] = {0};
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out16, g_stage_cycles, sizeof(z));
  if (reset) cudaMemcpyToSymbol(g_stage_cycles, z, sizeof(z));
  return 0;
}
#endif

int lgbcu_profile_enable(LgbCuContext* c, int on) {
  NEED(c, "lgbcu: NULL context");
  c->profile = on != 0;
  return 0;
}

int lgbcu_profile_read(LgbCuContext* c, double* hist_ms, int64_t* hist_calls, double* hist_bytes, int reset) {
  NEED(c, "lgbcu: NULL context");
  USE_DEVICE(c->device);
  CU(cudaMemcpyAsync(c->h_ctl, c->d_ctl, sizeof(DevCtl), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  const unsigned long long rows = c->h_ctl->prof_hist_rows, rows_idx = c->h_ctl->prof_hist_rows_indexed;
  const unsigned long long calls = c->h_ctl->prof_hist_calls;
  // B_hist(n) = n*(G + 4 + 4) [+ 4n if indexed] + total_bin*16
  const double bytes = (double)rows * (c->F + 8) + (double)rows_idx * (c->F + 12) + (double)calls * c->total_bin * 16.0;
  if (hist_ms) *hist_ms = c->prof_hist_ms;
  if (hist_calls) *hist_calls = (int64_t)calls;
  if (hist_bytes) *hist_bytes = bytes;
  if (reset) {
    c->prof_hist_ms = 0;
    CU(cudaMemsetAsync(&c->d_ctl->prof_hist_rows, 0, 4 * sizeof(unsigned long long), c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  return 0;
}

int lgbcu_comm_unique_id(uint8_t* id128) {
  NEED(id128, "lgbcu_comm_unique_id: NULL argument");
  if (nccl_load()) return -1;
  ncclUniqueId id;
  NC(g_nccl.GetUniqueId(&id));
  static_assert(sizeof(id) == 128, "ncclUniqueId is expected to be 128 bytes");
  std::memcpy(id128, &id, 128);
  return 0;
}

// NCCL communicators live as long as the rendezvous they belong to, not as long as a booster (the reference's
// Network::Init runs once per process, src/network/network.cpp:45-58): the first context that names a (unique id,
// rank, world, device) creates the communicator, later contexts attach to the cached one.  lgbcu_comm_release_all()
// (LGBM_NetworkFree) destroys them; no context may still be using one then.
struct CachedComm { uint8_t id[128]; int rank, world, device; ncclComm_t comm; };
static std::vector<CachedComm> g_comms;
static std::mutex g_comms_mu;

int lgbcu_comm_init(LgbCuContext* c, const uint8_t* id128, int32_t rank, int32_t world_size) {
  NEED(c && id128, "lgbcu_comm_init: NULL argument");
  NEED(world_size >= 1 && rank >= 0 && rank < world_size, "lgbcu_comm_init: bad rank %d / world %d", rank, world_size);
  if (world_size == 1) { c->rank = 0; c->world = 1; return 0; }
  if (nccl_load()) return -1;
  USE_DEVICE(c->device);
  std::lock_guard<std::mutex> lock(g_comms_mu);
  for (const CachedComm& k : g_comms) {
    if (k.rank == rank && k.world == world_size && k.device == c->device && std::memcmp(k.id, id128, 128) == 0) {
      c->comm = k.comm; c->rank = rank; c->world = world_size;
      return 0;
    }
  }
  ncclUniqueId id;
  std::memcpy(&id, id128, 128);
  CachedComm k;
  NC(g_nccl.CommInitRank(&k.comm, world_size, id, rank));
  std::memcpy(k.id, id128, 128); k.rank = rank; k.world = world_size; k.device = c->device;
  g_comms.push_back(k);
  c->comm = k.comm; c->rank = rank; c->world = world_size;
  return 0;
}

int lgbcu_comm_release_all(void) {
  std::lock_guard<std::mutex> lock(g_comms_mu);
  for (const CachedComm& k : g_comms) { ScopedDevice dev_(k.device); g_nccl.CommDestroy(k.comm); }
  g_comms.clear();
  return 0;
}

int lgbcu_comm_allreduce_sum_f64(LgbCuContext* c, double* inout, int32_t n) {
  NEED(c && inout && n > 0 && n <= 64, "lgbcu_comm_allreduce_sum_f64: bad argument");
  if (!c->comm) return 0;
  USE_DEVICE(c->device);
  if (!c->d_scratch) CU(dev_alloc(&c->d_scratch, (size_t)64));
  CU(cudaMemcpyAsync(c->d_scratch, inout, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
  NC(g_nccl.AllReduce(c->d_scratch, c->d_scratch, (size_t)n, ncclDouble, ncclSum, c->comm, c->stream));
  CU(cudaMemcpyAsync(inout, c->d_scratch, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

int lgbcu_comm_destroy(LgbCuContext* c) {  // detach: the communicator itself stays cached (see lgbcu_comm_init)
  if (!c || !c->comm) return 0;
  c->comm = nullptr; c->world = 1; c->rank = 0;
  drop_graph(c);
  return 0;
}

}  // extern "C"

#include "valid_eval.cuh"

#include "noalloc_partition.cuh"

#include "bagging_kernels.cuh"

#include "binning_kernels.cuh"

#include <cub/device/device_segmented_sort.cuh>
#include "renew_kernels.cuh"
