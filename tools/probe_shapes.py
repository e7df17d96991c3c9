"""Scratch probe: ms per iteration and per-phase times of the Epsilon-shaped (config #3) and MSLR-shaped (config #4)
workloads through the C API, optionally with the reference CPU learner (oracle/_ref) timed beside them.
usage: probe_shapes.py [epsilon|mslr|both] [iterations] [ref]"""
import os, sys, time, ctypes as C, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi, workloads as W
which = sys.argv[1] if len(sys.argv) > 1 else "both"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 8
with_ref = len(sys.argv) > 3 and sys.argv[3] == "ref"
cores = os.cpu_count() or 4
out = {}
def run(name, X, y, params, group=None):
    ours = capi.ours(); lib = ours.lib
    t0 = time.perf_counter()
    ds = ours.dataset_from_mat(X, params + f" num_threads={cores}"); ours.set_field(ds, "label", y)
    if group is not None: ours.set_field(ds, "group", group)
    t_ds = time.perf_counter() - t0
    b = ours.booster_create(ds, params)
    for _ in range(3): ours.update(b)
    ours.get_predict(b, X.shape[0])   # drains the stream
    t0 = time.perf_counter()
    for _ in range(iters): ours.update(b)
    ours.get_predict(b, X.shape[0])
    ms = 1e3 * (time.perf_counter() - t0) / iters
    ms_, calls, nb = C.c_double(), C.c_int64(), C.c_double()
    ours.check(lib.LGBM_B200_BoosterProfile(b, 1, 1, C.byref(ms_), C.byref(calls), C.byref(nb)))
    ph = (C.c_double * 8)()
    acc = np.zeros(8)
    for _ in range(3):
        ours.update(b); ours.check(lib.LGBM_B200_BoosterPhases(b, ph)); acc += np.array(list(ph))
    acc /= 3
    ours.check(lib.LGBM_B200_BoosterProfile(b, 0, 0, C.byref(ms_), C.byref(calls), C.byref(nb)))
    r = {"rows": int(X.shape[0]), "features": int(X.shape[1]), "ms_per_iter": ms, "dataset_s": t_ds,
         "phases_ms": {"hist": acc[1], "find_control": acc[2], "partition": acc[4], "rounds_enqueued": acc[6], "rounds_used": acc[7]},
         "hist_gb_per_s": (nb.value / 1e9) / (ms_.value / 1e3) if ms_.value > 0 else None}
    ours.booster_free(b); ours.dataset_free(ds)
    if with_ref:
        ref = capi.reference()
        p = params + f" force_col_wise=true num_threads={cores}"
        ds = ref.dataset_from_mat(X, p); ref.set_field(ds, "label", y)
        if group is not None: ref.set_field(ds, "group", group)
        b = ref.booster_create(ds, p)
        ref.update(b)
        t0 = time.perf_counter()
        for _ in range(3): ref.update(b)
        r["reference_ms_per_iter"] = 1e3 * (time.perf_counter() - t0) / 3
        r["cores"] = cores
        ref.booster_free(b); ref.dataset_free(ds)
    out[name] = r
    print(name, json.dumps(r), flush=True)
if which in ("epsilon", "both"):
    X, y = W.epsilon_shaped(400_000, 2000); run("config3_epsilon_shaped", X, y, W.EPSILON_PARAMS); del X
if which in ("mslr", "both"):
    X, y, sizes = W.mslr_shaped(2_270_000, 137); run("config4_mslr_shaped", X, y, W.MSLR_PARAMS, group=sizes)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "shapes.json"), "w"), indent=1)
