"""Scratch probe: LGBM_DatasetCreateFromMat on the bench matrix several times in a row (LGBM_B200_TIMING=1 prints the
stages), to see the spread of the from-the-matrix fixed cost."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi, workloads as W
X, y = W.higgs_shaped(10_500_000)
ours = capi.ours()
for i in range(6):
    t0 = time.perf_counter()
    ds = ours.dataset_from_mat(X, W.HIGGS_PARAMS + " num_threads=16")
    print(f"create {i}: {1e3 * (time.perf_counter() - t0):.1f} ms", flush=True)
    if i % 2 == 1: ours.dataset_free(ds)
