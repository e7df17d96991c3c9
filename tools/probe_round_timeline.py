"""Scratch probe (debug build, tools/build_dbg.sh; run with LGBM_B200_LIB=tools/_dbg/lib_dbg.so): the timeline of one
tree's rounds inside the captured graph, from %globaltimer stamps taken by the first block of each kernel -- no events,
no profile mode -- plus the stage cycle counters of block 0 of each kernel.
usage: probe_round_timeline.py [iterations before the probed tree] [rows]"""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi, workloads as W
skip = int(sys.argv[1]) if len(sys.argv) > 1 else 470
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 10_500_000
ours = capi.ours(); lib = ours.lib
shape = sys.argv[3] if len(sys.argv) > 3 else "higgs"
if shape == "epsilon":
    X, y = W.epsilon_shaped(rows if len(sys.argv) > 2 else 400_000, 2000); PARAMS = W.EPSILON_PARAMS
else:
    X, y = W.higgs_shaped(rows); PARAMS = W.HIGGS_PARAMS
ds = ours.dataset_from_mat(X, PARAMS + " num_threads=16"); ours.set_field(ds, "label", y)
b = ours.booster_create(ds, PARAMS)
for _ in range(skip): ours.update(b)
NR = 1024
st = (C.c_ulonglong * (6 * NR))()
fs = (C.c_ulonglong * 32)(); hs = (C.c_ulonglong * 8)()
lib.lgbcu_debug_stamps(st, 3 * NR); lib.lgbcu_debug_stages(fs, 1); lib.lgbcu_debug_hist_stages(hs, 1)
for rep in range(3):
    ours.update(b)
    lib.lgbcu_debug_stamps(st, 3 * NR); lib.lgbcu_debug_stages(fs, 1); lib.lgbcu_debug_hist_stages(hs, 1)
    a = np.array(list(st)[:3 * NR], dtype=np.float64).reshape(NR, 3)
    used = int((a[:, 0] > 0).sum())
    a = a[:used]
    hist = a[:, 1] - a[:, 0]; find = a[:, 2] - a[:, 1]; part = a[1:, 0] - a[:-1, 2]
    ok = (a[:, 2] > 0)
    if used > 2 and ok.sum() > 1: print(f"tree {skip + rep}: rounds {used}; span {(a[-1, 1] - a[0, 0]) / 1e6:.3f} ms; per round us  median hist {np.median(hist[ok]) / 1e3:.2f} "
          f"find {np.median(find[ok]) / 1e3:.2f} part {np.median(part[ok[:-1]]) / 1e3:.2f} | p10 {np.percentile(hist[ok], 10) / 1e3:.2f} "
          f"{np.percentile(find[ok], 10) / 1e3:.2f} {np.percentile(part[ok[:-1]], 10) / 1e3:.2f} | sum ms {hist[ok].sum() / 1e6:.2f} {find[ok].sum() / 1e6:.2f} {part[ok[:-1]].sum() / 1e6:.2f}")
    f = np.array(list(fs), dtype=np.float64) / 1.9e3 / used
    h = np.array(list(hs), dtype=np.float64) / 1.9e3 / used
    print("  find block(0,1) us/round: setup %.2f | find_item %.2f | ticket %.2f ; control total %.2f = reduce %.2f load %.2f plan %.2f items %.2f commit %.2f"
          % (f[0], f[1] + f[2] + f[3] + f[4], f[5], f[6], f[16], f[17], f[18], f[19], f[20]))
    print("  find_item stages us/round [1,2,3,4 | 24,25,26]:", np.round(f[1:5], 2), np.round(f[24:27], 2))
    print("  part block 0 us/round:", np.round(f[8:13], 2), "  hist block 0 us/round [zero+prefetch, rows, sync, flush, sync, preamble]:", np.round(h[:6], 2))
