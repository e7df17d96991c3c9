"""Scratch probe: histogram kernel + whole-tree timings on synthetic binned data (not a bench line)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lightgbmv1_b200 import _native as N

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_500_000
F = int(sys.argv[2]) if len(sys.argv) > 2 else 28
leaves = int(sys.argv[3]) if len(sys.argv) > 3 else 255
rng = np.random.default_rng(0)
metas = []
for f in range(F):
    if f % 9 == 8: metas.append((3, 0, 0, 0))
    elif f % 7 == 6: metas.append((130, 0, 0, 0))
    else: metas.append((255, 127, 127, 0))
bins = np.empty((n, F), dtype=np.uint8)
for f, m in enumerate(metas):
    b = rng.integers(0, m[0], size=n, dtype=np.int32)
    bins[:, f] = np.where(b == m[1], 0, b + 1 - (1 if m[1] == 0 else 0)).astype(np.uint8)
g = rng.standard_normal(n).astype(np.float32)
g += (0.5 * np.sign(bins[:, 0].astype(np.float32) - 100) + 0.3 * (bins[:, 1] > 50) - 0.2 * (bins[:, 2] > 200)).astype(np.float32)
h = (rng.random(n).astype(np.float32) * 0.25 + 0.01).astype(np.float32)
L = N.CudaLearner()
t = time.time(); L.set_dataset(bins, metas); print("upload s", time.time() - t)
L.set_tree_params(N.make_tree_params(num_leaves=leaves, min_data_in_leaf=0, min_sum_hessian_in_leaf=100.0))
L.set_gradients(g, h)
L.profile_enable(True)
for mode, name in ((N.HIST_GLOBAL_ATOMIC, "global"), (N.HIST_WARP_PRIVATE, "warp")):
    L.set_hist_mode(mode)
    L.begin_tree()
    for rep in range(3):
        L.profile_read(reset=True)
        L.hist_build(0)
        L.synchronize()
        ms, calls, nbytes = L.profile_read()
        print(f"root hist {name}: {ms:.3f} ms  {nbytes/ms/1e6:.1f} GB/s  ({nbytes/1e6:.1f} MB)")
for mode, name in ((N.HIST_AUTO, "auto"), (N.HIST_WARP_PRIVATE, "warp"), (N.HIST_GLOBAL_ATOMIC, "global")):
    L.set_hist_mode(mode)
    for rep in range(3):
        L.profile_read(reset=True)
        L.begin_tree(); L.synchronize()
        t = time.time()
        recs = L.train_tree()
        dt = time.time() - t
        ms, calls, nbytes = L.profile_read()
        print(f"tree {name}: {dt*1e3:.2f} ms wall, {len(recs)} splits, hist {ms:.2f} ms in {calls} calls, {nbytes/max(ms,1e-9)/1e6:.1f} GB/s, launches {N.launch_count()}")
    L.profile_enable(False)
    L.begin_tree(); L.synchronize()
    t = time.time(); recs = L.train_tree(); print(f"tree {name} no-profile: {(time.time()-t)*1e3:.2f} ms wall")
    L.profile_enable(True)
