"""Scratch probe (debug build): k_partition block-0 stage times vs leaf size."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lightgbmv1_b200 import _native as N
n, F = 10_500_000, 28
rng = np.random.default_rng(0)
metas = [(255, 127, 127, 0)] * F
bins = rng.integers(0, 255, size=(n, F), dtype=np.uint8)
g = rng.standard_normal(n).astype(np.float32); h = np.full(n, 0.25, dtype=np.float32)
L = N.CudaLearner(); L.set_dataset(bins, metas)
L.set_tree_params(N.make_tree_params(num_leaves=31, min_data_in_leaf=0, min_sum_hessian_in_leaf=1.0))
L.set_gradients(g, h)
o = (ctypes.c_ulonglong * 32)()
for k in [2048, 8192, 32768, 131072, 524288, 2097152, 8388608]:
    bag = np.sort(rng.choice(n, size=k, replace=False)).astype(np.int32)
    res = []
    for rep in range(3):
        L.begin_tree(bag); L.hist_build(0); L.synchronize()
        N.lib.lgbcu_debug_stages(o, 1)
        left = L.partition(0, 3, 100, True, 1)
        N.lib.lgbcu_debug_stages(o, 1)
        res = [round(o[8 + i] / 1.965e3, 2) for i in range(5)]
    print(f"rows {k:8d}: left {left:8d}  block0 us [ctl+leaf, flags, barrier, prefix, scatter] = {res}  sum {sum(res):.1f}")
