"""Scratch probe: per-phase time and round counts of one boosting iteration of bench.py's workload (profile mode)."""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench, capi
ours = capi.ours(); lib = ours.lib
X, y = bench.higgs_shaped(bench.N_ROWS)
ds = ours.dataset_from_mat(X, bench.PARAMS); ours.set_field(ds, "label", y)
b = ours.booster_create(ds, bench.PARAMS)
for _ in range(3): ours.update(b)
ms, calls, nb = C.c_double(), C.c_int64(), C.c_double()
ours.check(lib.LGBM_B200_BoosterProfile(b, 1, 1, C.byref(ms), C.byref(calls), C.byref(nb)))
ph = (C.c_double * 8)()
for _ in range(3):
    ours.update(b)
    ours.check(lib.LGBM_B200_BoosterPhases(b, ph))
    print("K=%s phases ms: hist %.2f find+control %.2f partition %.2f; rounds enqueued %d used %d" % (
        os.environ.get("LGBCU_LOOKAHEAD", "default"), ph[0], ph[2], ph[4], ph[6], ph[7]))
if hasattr(lib, "lgbcu_debug_stages"):   # debug build (tools/build_dbg.sh): block-0 stage times
    out = (C.c_ulonglong * 32)()
    lib.lgbcu_debug_stages(out, 1)
    ours.update(b)
    lib.lgbcu_debug_stages(out, 0)
    ours.check(lib.LGBM_B200_BoosterPhases(b, ph))
    r = max(1.0, ph[7])
    us = lambda i: round(out[i] / 1.965e3 / r, 2)
    print("per round, block 0 (us): find [ctl, load+fix, -, chains+scans, gains, ticket, control]:", [us(i) for i in range(7)],
          " partition [ctl+leaf, flags, barrier, prefix, scatter]:", [us(8 + i) for i in range(5)],
          " control [reduce+store, stage keys, plan, select, copies]:", [us(16 + i) for i in range(5)])
if hasattr(lib, "lgbcu_debug_hist_stages"):
    o8 = (C.c_ulonglong * 8)()
    lib.lgbcu_debug_hist_stages(o8, 1)
    ours.update(b)
    lib.lgbcu_debug_hist_stages(o8, 0)
    ours.check(lib.LGBM_B200_BoosterPhases(b, ph))
    r = max(1.0, ph[7])
    print("hist block 0 per round (us) [load+zero, accumulate, sync1, reduce+red, sync2]:", [round(o8[i] / 1.965e3 / r, 2) for i in range(5)])
