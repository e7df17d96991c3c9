// Data-parallel plumbing (A11; replaces Network::ReduceScatter of histograms + SyncUpGlobalBestSplit,
// src/treelearner/data_parallel_tree_learner.cpp:125-173,176-247, and the root (count, sum_g, sum_h)
// Allreduce at :126-151) over NCCL on NVLink.
//
// One process per GPU.  Per round ONE collective: an int64 sum all-reduce of the round's fresh histogram slots, in
// place in the pool.  Because the histogram is 64-bit fixed point the sum is exact and
// independent of rank order, so every rank holds a bit-identical global histogram, runs the same split search
// and reaches the same decision: no best-split exchange is needed (the reference's reduce-scatter +
// SplitInfo all-reduce moves the same bytes in two steps).  NCCL is resolved with dlopen so that a single-GPU
// process needs no NCCL at all, and so that a host process that already loaded an NCCL (torch) shares it.
//
// Two ranks (one-shot exchange): no collective on the histograms at all.  k_hist adds into a second pool that is
// exported with CUDA IPC; one launch per round (k_xchg_pull, split_kernels.cuh) publishes the round with a flag in the
// peer's memory, waits for the peer's flag (bounded: a peer that never arrives voids the tree, it does not hang the
// GPU), and sums the ranks' local slots -- the peer's read over NVLink with plain system-scope loads -- into the pool
// the split search reads.  Sized on the device: only the live items' slots travel, which the host-enqueued NCCL call
// cannot know.  Measured: 3.18 vs 3.34 ms per iteration at 2 ranks, 3.12 vs 2.87 at 8 (every rank pulls 7 x the
// slots), hence two ranks only by default.
//
// Wide data (ownership mode, the reference's own scheme): each rank owns a contiguous block of features (balanced by
// bins).  Per round: a grouped ncclReduce per (fresh slot, owner) leaves every owner with the global histogram of its
// block (the reference's ReduceScatter, data_parallel_tree_learner.cpp:176-186), the split search scans only the
// owned features (:188-247: 1/N of the search blocks per rank), and one small ncclAllGather of the per-leaf winners
// followed by the same SplitInfo::operator> arg-max on every rank replaces SyncUpGlobalBestSplit
// (parallel_tree_learner.h:190-213).  For narrow data the replicated search costs no wall time and the extra
// collective would, so the mode is chosen by shape (LGBCU_DP_OWNERSHIP=0/1 forces it).
#pragma once
// <dlfcn.h> and <nccl.h> are included by lgbcu.cu at global scope (this file is included inside a namespace)

struct NcclApi {
  void* lib = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  // feature ownership (wide data): reduce of each owner's block, all-gather of the per-leaf winners
  decltype(&ncclReduce) Reduce = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  bool has_ownership_calls() const { return Reduce && AllGather && GroupStart && GroupEnd; }
};

static NcclApi g_nccl;

static int nccl_load() {
  if (g_nccl.lib) return 0;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
  if (!h) return fail("lgbcu_comm: cannot load libnccl.so.2 (%s)", dlerror());
  g_nccl.GetUniqueId = reinterpret_cast<decltype(&ncclGetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
  g_nccl.CommInitRank = reinterpret_cast<decltype(&ncclCommInitRank)>(dlsym(h, "ncclCommInitRank"));
  g_nccl.CommDestroy = reinterpret_cast<decltype(&ncclCommDestroy)>(dlsym(h, "ncclCommDestroy"));
  g_nccl.AllReduce = reinterpret_cast<decltype(&ncclAllReduce)>(dlsym(h, "ncclAllReduce"));
  g_nccl.GetErrorString = reinterpret_cast<decltype(&ncclGetErrorString)>(dlsym(h, "ncclGetErrorString"));
  g_nccl.Reduce = reinterpret_cast<decltype(&ncclReduce)>(dlsym(h, "ncclReduce"));
  g_nccl.AllGather = reinterpret_cast<decltype(&ncclAllGather)>(dlsym(h, "ncclAllGather"));
  g_nccl.GroupStart = reinterpret_cast<decltype(&ncclGroupStart)>(dlsym(h, "ncclGroupStart"));
  g_nccl.GroupEnd = reinterpret_cast<decltype(&ncclGroupEnd)>(dlsym(h, "ncclGroupEnd"));
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllReduce || !g_nccl.GetErrorString) {
    return fail("lgbcu_comm: libnccl is missing a required symbol");
  }
  g_nccl.lib = h;
  return 0;
}

#define NC(call)                                                                                         \
  do {                                                                                                   \
    ncclResult_t r__ = (call);                                                                           \
    if (r__ != ncclSuccess) return fail("NCCL error %s at %s:%d", g_nccl.GetErrorString(r__), __FILE__, __LINE__); \
  } while (0)
