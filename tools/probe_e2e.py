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
