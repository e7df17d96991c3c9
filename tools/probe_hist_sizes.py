"""Scratch probe: k_hist time vs leaf size, for random (sorted) row subsets and for contiguous ones."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lightgbmv1_b200 import _native as N
n, F = 10_500_000, 28
only = int(sys.argv[1]) if len(sys.argv) > 1 else 0
rng = np.random.default_rng(0)
metas = [(3, 0, 0, 0) if f % 9 == 8 else ((130, 0, 0, 0) if f % 7 == 6 else (255, 127, 127, 0)) for f in range(F)]
bins = rng.integers(0, 3, size=(n, F), dtype=np.uint8)
for f, m in enumerate(metas):
    if m[0] > 3: bins[:, f] = rng.integers(0, m[0] - 1, size=n).astype(np.uint8)
g = rng.standard_normal(n).astype(np.float32); h = (rng.random(n).astype(np.float32) * 0.25 + 0.01).astype(np.float32)
L = N.CudaLearner(); L.set_dataset(bins, metas)
L.set_tree_params(N.make_tree_params(num_leaves=31, min_data_in_leaf=0, min_sum_hessian_in_leaf=1.0))
L.set_gradients(g, h); L.profile_enable(True)
sizes = [only] if only else [2048, 8192, 32768, 131072, 524288, 2097152, 8388608]
for k in sizes:
    for kind in ("random", "contiguous"):
        bag = np.sort(rng.choice(n, size=k, replace=False)).astype(np.int32) if kind == "random" else np.arange(k, dtype=np.int32)
        L.begin_tree(bag)
        ts = []
        for rep in range(4):
            L.profile_read(reset=True); L.hist_build(0); L.synchronize()
            ms, calls, nbytes = L.profile_read(); ts.append(ms * 1e3)
        extra = ""
        if hasattr(N.lib, "lgbcu_debug_hist_stages"):
            import ctypes
            o = (ctypes.c_ulonglong * 8)()
            N.lib.lgbcu_debug_hist_stages(o, 1); L.hist_build(0); L.synchronize(); N.lib.lgbcu_debug_hist_stages(o, 1)
            extra = "  block0 us [load+zero, accumulate, sync1, reduce+red, sync2]: " + str([round(o[i] / 1.965e3, 2) for i in range(5)])
        print(f"rows {k:8d} {kind:10s}: {min(ts):8.1f} us   ({nbytes / min(ts) / 1e3:7.1f} GB/s algorithmic){extra}")
