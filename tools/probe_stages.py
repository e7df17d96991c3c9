"""Scratch probe: per-stage cycle counts of k_find_split (debug build with -DLGBCU_STAGE_TIMERS)."""
import sys, os, ctypes, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lightgbmv1_b200 import _native as N
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_500_000
F = int(sys.argv[2]) if len(sys.argv) > 2 else 28
leaves = 255
rng = np.random.default_rng(0)
metas = [(3,0,0,0) if f % 9 == 8 else ((130,0,0,0) if f % 7 == 6 else (255,127,127,0)) for f in range(F)]
bins = np.empty((n, F), dtype=np.uint8)
for f, m in enumerate(metas):
    b = rng.integers(0, m[0], size=n, dtype=np.int32)
    bins[:, f] = np.where(b == m[1], 0, b + 1 - (1 if m[1] == 0 else 0)).astype(np.uint8)
g = rng.standard_normal(n).astype(np.float32)
g += (0.5 * np.sign(bins[:, 0].astype(np.float32) - 100) + 0.3 * (bins[:, 1] > 50) - 0.2 * (bins[:, 2] > 200)).astype(np.float32)
h = (rng.random(n).astype(np.float32) * 0.25 + 0.01).astype(np.float32)
L = N.CudaLearner(); L.set_dataset(bins, metas)
L.set_tree_params(N.make_tree_params(num_leaves=leaves, min_data_in_leaf=0, min_sum_hessian_in_leaf=100.0))
L.set_gradients(g, h)
lib = N.lib
out = (ctypes.c_ulonglong * 32)()
for rep in range(3):
    L.begin_tree(); L.synchronize(); lib.lgbcu_debug_stages(out, 1)
    t0 = time.time(); recs = L.train_tree(); dt = time.time() - t0
    lib.lgbcu_debug_stages(out, 0)
    print(f"tree {dt*1e3:.2f} ms; find stages (us per call @1.9GHz): ", [round(out[i] / 1.9e3 / max(1, len(recs)), 2) for i in range(8)],
          " part stages [ctl+leaf, flags, barrier, prefix, scatter]:", [round(out[8 + i] / 1.9e3 / max(1, len(recs)), 2) for i in range(3)],
          " control [prefetch, feature argmax+loads, leaf argmax, writes]:", [round(out[16 + i] / 1.9e3 / max(1, len(recs)), 2) for i in range(4)],
          " gains [eval, reduce, epilogue]:", [round(out[20 + i] / 1.9e3 / max(1, len(recs)), 2) for i in range(3)])
# per-split start stamps of the three kernels (graph path): deltas = kernel duration + gap
ns = len(recs)
st = (ctypes.c_ulonglong * (6 * ns))()
lib.lgbcu_debug_stamps(st, 3 * ns)
a = np.array(list(st)[:3 * ns], dtype=np.float64).reshape(ns, 3)
e = np.array(list(st)[3 * ns:], dtype=np.float64).reshape(ns, 3)
ex = e - a   # execution: first block start -> last block exit
print("exec-only means us: hist %.2f find(to ticket) %.2f part %.2f" % tuple(np.median(ex, axis=0) / 1e3))
hist = a[:, 1] - a[:, 0]; find = a[:, 2] - a[:, 1]; part = np.append(a[1:, 0] - a[:-1, 2], np.nan)
rows = np.array([n] + [min(r.left_count, r.right_count) for r in recs[:-1]])
prow = np.array([r.left_count + r.right_count for r in recs])
print("totals ms: hist %.2f find %.2f part %.2f" % (hist.sum() / 1e6, find.sum() / 1e6, np.nansum(part) / 1e6))
for lo, hi in [(0, 4096), (4096, 20000), (20000, 100000), (100000, 500000), (500000, 2e6), (2e6, 2e7)]:
    m = (rows >= lo) & (rows < hi)
    if m.sum(): print(f"hist rows [{lo:.0f},{hi:.0f}) n={m.sum():3d} mean rows {rows[m].mean():9.0f} hist {hist[m].mean()/1e3:7.2f}us (exec {ex[m,0].mean()/1e3:6.2f}) find {find[m].mean()/1e3:6.2f}us (exec {ex[m,1].mean()/1e3:6.2f})")
for lo, hi in [(0, 20000), (20000, 100000), (100000, 500000), (500000, 2e6), (2e6, 2e7)]:
    m = (prow >= lo) & (prow < hi)
    if m.sum(): print(f"part rows [{lo:.0f},{hi:.0f}) n={m.sum():3d} mean {prow[m].mean():9.0f} part {np.nanmean(part[m])/1e3:6.2f}us (exec {ex[m,2].mean()/1e3:6.2f})")
