// k_grow_tree: SerialTreeLearner::Train (src/treelearner/serial_tree_learner.cpp:152-202) as ONE persistent
// cooperative kernel: one 256-thread CTA per SM stays resident for the whole tree and walks the
// {histogram, split search, control, partition flags, partition scan + bookkeeping, scatter} phases of every split,
// separated by grid-wide barriers instead of kernel launches.  A 255-leaf tree is 254 strictly sequential splits
// whose data-dependent sizes are only known on the device; the persistent form removes ~1500 launches per tree
// (and their dependency latency) and keeps the 192 KB of private sub-histograms allocated across splits.
// Every phase is the same __device__ function the stand-alone kernels wrap, so the step-wise API, the
// launch-per-phase loop (used with NCCL between histogram and split search) and this kernel cannot drift apart.
#pragma once
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#define GROW_THREADS 256
static_assert(GROW_THREADS == HIST_THREADS && GROW_THREADS == FS_THREADS && GROW_THREADS == PART_THREADS,
              "all phases share one block shape");
static_assert(sizeof(ScanSmem) <= HIST_SMEM_BYTES, "the split-search work area aliases the histogram area");

__device__ __forceinline__ unsigned long long grow_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(GROW_THREADS, 1) k_grow_tree(DevView V, int hist_mode, int min_rows, int profile) {
  extern __shared__ __align__(16) unsigned char grow_smem[];
  cg::grid_group grid = cg::this_grid();
  DevCtl* c = V.ctl;
  const int bid = blockIdx.x, nblocks = gridDim.x;
  ScanSmem& S = *reinterpret_cast<ScanSmem*>(grow_smem);
  const int max_splits = V.p.num_leaves - 1;
  for (int s = 0; s < max_splits; ++s) {
    if (c->done) break;                      // grid-uniform: written before the last barrier
    const bool active = !c->skip_find;
    const int smaller = c->smaller, larger = c->larger;
    // ---- histogram of the smaller leaf
    const bool timer = profile && bid == 0 && threadIdx.x == 0;
    unsigned long long t0 = 0, t1;
#define GROW_TICK(i) if (timer) { t1 = grow_now_ns(); c->prof_phase_ns[i] += t1 - t0; t0 = t1; }
    if (timer) t0 = grow_now_ns();
    if (active && smaller >= 0) hist_phase(V, smaller, hist_mode, min_rows, bid, nblocks, grow_smem);
    grid.sync();
    if (timer) { t1 = grow_now_ns(); c->prof_hist_ns += t1 - t0; c->prof_phase_ns[0] += t1 - t0; t0 = t1; }
    if (s == 0) {                            // root LeafSplits::Init from the finished histogram
      if (bid == 0 && threadIdx.x < 32) root_sums_warp(V);
      grid.sync();
      GROW_TICK(1)
    }
    // ---- FixHistogram / Subtract / FindBestThreshold: (feature, leaf) items over the CTAs
    if (active) {
      const int nitems = V.num_features * 2;
      for (int item = bid; item < nitems; item += nblocks) {
        const int f = item >> 1, sel = item & 1;
        find_item(V, S, f, sel, smaller, larger);
      }
    }
    grid.sync();
    GROW_TICK(2)
    // ---- per-leaf arg-max, choice of the next leaf (one warp)
    if (bid == 0 && threadIdx.x < 32) find_control(V, 0, active, smaller, larger);
    grid.sync();
    GROW_TICK(3)
    if (c->done) break;
    // ---- DataPartition::Split of the chosen leaf
    part_flags_phase(V, bid, nblocks);
    grid.sync();
    GROW_TICK(4)
    if (bid == 0) part_scan_phase(V, 0);
    grid.sync();
    GROW_TICK(5)
    part_scatter_phase(V, bid, nblocks);
    grid.sync();
    GROW_TICK(6)
    if (profile > 1) { grid.sync(); GROW_TICK(7) }
  }
}
