// Data-parallel plumbing (A11; replaces Network::ReduceScatter of histograms + SyncUpGlobalBestSplit,
// src/treelearner/data_parallel_tree_learner.cpp:125-173,176-247, and the root (count, sum_g, sum_h)
// Allreduce at :126-151) over NCCL on NVLink.
//
// One process per GPU.  Per split ONE collective: an int64 sum all-reduce of the smaller leaf's local
// histogram, in place in its pool slot.  Because the histogram is 64-bit fixed point the sum is exact and
// independent of rank order, so every rank holds a bit-identical global histogram, runs the same split search
// and reaches the same decision: no best-split exchange is needed (the reference's reduce-scatter +
// SplitInfo all-reduce moves the same bytes in two steps).  NCCL is resolved with dlopen so that a single-GPU
// process needs no NCCL at all, and so that a host process that already loaded an NCCL (torch) shares it.
#pragma once
// <dlfcn.h> and <nccl.h> are included by lgbcu.cu at global scope (this file is included inside a namespace)

struct NcclApi {
  void* lib = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
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
