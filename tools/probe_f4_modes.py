"""Scratch probe: ms per iteration of the F4 paths on a Higgs-shaped matrix -- the 8-bit baseline, max_bin=1023 (wide
features), categorical features, monotone constraints -- with the reference CPU learner (oracle/_ref) timed beside each.
usage: probe_f4_modes.py [rows] [iterations]"""
import os, sys, time, ctypes as C, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi, workloads as W
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
cores = os.cpu_count() or 4
X, y = W.higgs_shaped(rows)
Xc = X.copy()
rng = np.random.default_rng(5)
for c in (3, 7, 11, 15):                       # four columns become 50-level categories (skewed)
    Xc[:, c] = np.minimum(49, np.floor(np.abs(Xc[:, c]) * 12)).astype(np.float32)
base = "objective=binary num_leaves=255 learning_rate=0.1 min_data_in_leaf=0 min_sum_hessian_in_leaf=100 verbosity=-1"
modes = [("8-bit baseline (max_bin=255)", X, base + " max_bin=255"),
         ("wide features (max_bin=1023)", X, base + " max_bin=1023"),
         ("4 categorical features of 50 levels", Xc, base + " max_bin=255 categorical_feature=3,7,11,15"),
         ("monotone constraints on 2 features", X, base + " max_bin=255 monotone_constraints=1,-1" + ",0" * 26)]
out = {}
for name, M, params in modes:
    ours = capi.ours(); lib = ours.lib
    ds = ours.dataset_from_mat(M, params + f" num_threads={cores}"); ours.set_field(ds, "label", y)
    b = ours.booster_create(ds, params)
    for _ in range(3): ours.update(b)
    ours.get_predict(b, rows)
    t0 = time.perf_counter()
    for _ in range(iters): ours.update(b)
    ours.get_predict(b, rows)
    ms = 1e3 * (time.perf_counter() - t0) / iters
    ph = (C.c_double * 8)(); ms_, calls, nb = C.c_double(), C.c_int64(), C.c_double()
    ours.check(lib.LGBM_B200_BoosterProfile(b, 1, 1, C.byref(ms_), C.byref(calls), C.byref(nb)))
    acc = np.zeros(8)
    for _ in range(3):
        ours.update(b); ours.check(lib.LGBM_B200_BoosterPhases(b, ph)); acc += np.array(list(ph))
    acc /= 3
    ours.check(lib.LGBM_B200_BoosterProfile(b, 0, 0, C.byref(ms_), C.byref(calls), C.byref(nb)))
    ours.booster_free(b); ours.dataset_free(ds)
    ref = capi.reference()
    p = params + f" force_col_wise=true num_threads={cores}"
    ds = ref.dataset_from_mat(M, p); ref.set_field(ds, "label", y)
    b = ref.booster_create(ds, p)
    ref.update(b)
    t0 = time.perf_counter()
    for _ in range(3): ref.update(b)
    ref_ms = 1e3 * (time.perf_counter() - t0) / 3
    ref.booster_free(b); ref.dataset_free(ds)
    out[name] = {"rows": rows, "ms_per_iter": ms, "phases_ms": {"hist": acc[1], "find_control": acc[2], "partition": acc[4]},
                 "reference_ms_per_iter_col_wise": ref_ms, "cores": cores}
    print(name, json.dumps(out[name]), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "f4_modes.json"), "w"), indent=1)
