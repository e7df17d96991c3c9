"""Scratch probe: histogram kernel time (root pass, whole tree) on synthetic binned data."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lightgbmv1_b200 import _native as N

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_500_000
F = int(sys.argv[2]) if len(sys.argv) > 2 else 28
leaves = int(sys.argv[3]) if len(sys.argv) > 3 else 255
bits_list = [0]
tree_reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
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
for bits in bits_list:
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=leaves, min_data_in_leaf=0, min_sum_hessian_in_leaf=100.0))
    L.set_gradients(g, h)
    L.profile_enable(True)
    L.set_hist_mode(N.HIST_AUTO)
    L.begin_tree()
    best = 1e9
    for rep in range(4):
        L.profile_read(reset=True)
        L.hist_build(0)
        L.synchronize()
        ms, calls, nbytes = L.profile_read()
        best = min(best, ms)
    print(f"bits {bits}: root hist {best:.3f} ms  {nbytes/best/1e6:.1f} GB/s", flush=True)
    for rep in range(tree_reps):
        L.profile_read(reset=True)
        L.begin_tree(); L.synchronize()
        t = time.time()
        recs = L.train_tree()
        dt = time.time() - t
        ms, calls, nbytes = L.profile_read()
        ph = L.profile_phases()
        print(f"bits {bits}: tree {dt*1e3:.2f} ms wall, {len(recs)} splits, hist {ms:.2f} ms in {calls} calls, {nbytes/max(ms,1e-9)/1e6:.1f} GB/s; "
              f"phases hist {ph[0]:.2f} find {ph[2]:.2f} part {ph[4]:.2f} rounds {ph[7]:.0f}", flush=True)
    L.profile_enable(False)
    if tree_reps > 0:
        L.begin_tree(); L.synchronize()
        t = time.time(); recs = L.train_tree(); print(f"bits {bits}: tree no-profile {(time.time()-t)*1e3:.2f} ms wall", flush=True)
    L.close()
