"""Scratch probe: where the end-to-end (host buffers) time of bench.py goes."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench, capi
ours = capi.ours()
X, y = bench.higgs_shaped(bench.N_ROWS)
t = time.perf_counter(); ds = ours.dataset_from_mat(X, bench.PARAMS); ours.set_field(ds, "label", y); print("dataset %.1f ms" % ((time.perf_counter() - t) * 1e3))
from lightgbmv1_b200 import _native as N
N.CudaLearner(0).close()
for rep in range(2):
    t = time.perf_counter(); b = ours.booster_create(ds, bench.PARAMS); t1 = time.perf_counter()
    its = []
    for _ in range(6):
        a = time.perf_counter(); ours.update(b); its.append((time.perf_counter() - a) * 1e3)
    a = time.perf_counter(); s = ours.get_predict(b, bench.N_ROWS); t2 = time.perf_counter()
    print("create %.1f ms; iters %s; get_predict %.1f ms" % ((t1 - t) * 1e3, [round(x, 1) for x in its], (t2 - a) * 1e3))
    ours.booster_free(b)
# with the debug build (tools/_dbg/lib_dbg.so copied over the product library): per-kernel split of the last tree
if hasattr(ours.lib, "lgbcu_debug_stamps"):
    import ctypes
    b = ours.booster_create(ds, bench.PARAMS)
    for _ in range(5):
        ours.update(b)
    ns = 254
    st = (ctypes.c_ulonglong * (6 * ns))()
    ours.lib.lgbcu_debug_stamps(st, 3 * ns)
    a = np.array(list(st)[:3 * ns], dtype=np.float64).reshape(ns, 3)
    e = np.array(list(st)[3 * ns:], dtype=np.float64).reshape(ns, 3)
    hist = a[:, 1] - a[:, 0]; find = a[:, 2] - a[:, 1]; part = a[1:, 0] - a[:-1, 2]
    print("last tree, ms: hist %.2f find %.2f part %.2f | medians us: hist %.1f find %.1f part %.1f" % (
        hist.sum() / 1e6, find.sum() / 1e6, part.sum() / 1e6, np.median(hist) / 1e3, np.median(find) / 1e3, np.median(part) / 1e3))
    hs = np.sort(hist)[::-1] / 1e3
    print("hist us, largest 12:", [round(x) for x in hs[:12]], " sum of the rest: %.0f" % hs[12:].sum())
