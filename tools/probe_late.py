"""Scratch probe: per-round phase times of late boosting iterations of the bench workload (deep, unbalanced trees:
many rounds of small leaves).  usage: probe_late.py [iterations before profiling] [rows]"""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi, workloads as W
skip = int(sys.argv[1]) if len(sys.argv) > 1 else 470
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 10_500_000
dump = os.path.join(ROOT, "gpurun_out", "late_rounds.csv")
if os.path.exists(dump): os.remove(dump)
ours = capi.ours(); lib = ours.lib
X, y = W.higgs_shaped(rows)
ds = ours.dataset_from_mat(X, W.HIGGS_PARAMS + " num_threads=16"); ours.set_field(ds, "label", y)
b = ours.booster_create(ds, W.HIGGS_PARAMS)
import time
t0 = time.perf_counter()
for _ in range(skip): ours.update(b)
print("ms/iter over the first", skip, ":", 1e3 * (time.perf_counter() - t0) / max(skip, 1))
t0 = time.perf_counter()
for _ in range(10): ours.update(b)
print("ms/iter next 10:", 1e3 * (time.perf_counter() - t0) / 10)
ms, calls, nb = C.c_double(), C.c_int64(), C.c_double()
ours.check(lib.LGBM_B200_BoosterProfile(b, 1, 1, C.byref(ms), C.byref(calls), C.byref(nb)))
os.environ["LGBCU_PROF_DUMP"] = dump
ph = (C.c_double * 8)()
for _ in range(3):
    ours.update(b)
    ours.check(lib.LGBM_B200_BoosterPhases(b, ph))
    print("phases ms: hist+ar %.2f hist %.2f find+control %.2f partition %.2f; rounds enqueued %d used %d" % (ph[0], ph[1], ph[2], ph[4], ph[6], ph[7]))
a = np.loadtxt(dump, delimiter=",")
print("per-round us (hist, find, partition): median", np.median(a[:, 1:], axis=0), "mean", a[:, 1:].mean(axis=0), "p10", np.percentile(a[:, 1:], 10, axis=0))
