"""Record for north_star's target: Higgs-shaped 10.5M x 28, 255 leaves, 500 iterations -- the reference CPU learner
(oracle/_ref, tree_learner=serial force_col_wise, all host threads) against this library on the same matrix.
Writes profiles/r2_parity_higgs500.json: number of identical trees, training logloss / AUC of both, the largest score
difference, this library's ms per iteration at the start / middle / end of the run and rounds per tree.
    python tools/parity_higgs500.py [iterations] [rows]        (on a GPU box; the reference takes ~0.3 s / iteration)"""
import ctypes as C, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi
import workloads as W
from test_e2e_gpu import auc, logloss

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 500
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 10_500_000
cores = os.cpu_count() or 1
X, y = W.higgs_shaped(rows)
out = {"workload": f"Higgs-shaped {rows}x28, {W.HIGGS_PARAMS}", "iterations": iters, "host_cores": cores}

ours = capi.ours(); lib = ours.lib
ds = ours.dataset_from_mat(X, W.HIGGS_PARAMS + f" num_threads={cores}")
ours.set_field(ds, "label", y)
b = ours.booster_create(ds, W.HIGGS_PARAMS)
ms = np.zeros(iters); rounds = np.zeros(iters)
ph = (C.c_double * 8)()
for i in range(iters):
    t0 = time.perf_counter()
    ours.update(b)
    ms[i] = 1e3 * (time.perf_counter() - t0)
    ours.check(lib.LGBM_B200_BoosterPhases(b, ph)); rounds[i] = ph[7]
text_ours = ours.model_string(b)
pred_ours = ours.get_predict(b, rows)
ours.booster_free(b); ours.dataset_free(ds)
win = lambda a, lo, hi: float(np.mean(a[max(0, lo):min(iters, hi)])) if iters > lo else None
out["ours"] = {"ms_per_iter_1_25": win(ms, 0, 25), "ms_per_iter_240_260": win(ms, 240, 260), "ms_per_iter_480_500": win(ms, 480, 500),
               "ms_per_iter_mean": float(ms.mean()), "rounds_per_tree_mean": float(rounds.mean()), "rounds_per_tree_max": float(rounds.max()),
               "rounds_per_tree_1_25": win(rounds, 0, 25), "rounds_per_tree_480_500": win(rounds, 480, 500)}

ref = capi.reference()
rp = W.HIGGS_PARAMS + f" force_col_wise=true num_threads={cores}"
rds = ref.dataset_from_mat(X, rp)
ref.set_field(rds, "label", y)
rb = ref.booster_create(rds, rp)
t0 = time.perf_counter()
for i in range(iters):
    ref.update(rb)
out["reference"] = {"ms_per_iter_mean": 1e3 * (time.perf_counter() - t0) / iters}
text_ref = ref.model_string(rb)
pred_ref = ref.get_predict(rb, rows)

same, n, first = capi.compare_models(text_ref, text_ours, rtol_value=1e-9)
out["trees_compared"] = n; out["trees_identical"] = same; out["first_mismatch"] = first
out["max_abs_score_diff"] = float(np.max(np.abs(pred_ref - pred_ours)))
out["logloss"] = {"reference": logloss(pred_ref, y), "ours": logloss(pred_ours, y)}
sub = np.random.default_rng(0).choice(rows, size=min(rows, 2_000_000), replace=False)   # AUC on a 2M-row sample (python sort)
out["auc_2M_row_sample"] = {"reference": auc(pred_ref[sub], y[sub]), "ours": auc(pred_ours[sub], y[sub])}
out["abs_diff"] = {"logloss": abs(out["logloss"]["reference"] - out["logloss"]["ours"]),
                   "auc": abs(out["auc_2M_row_sample"]["reference"] - out["auc_2M_row_sample"]["ours"])}
out["speedup_vs_reference_device_resident"] = out["reference"]["ms_per_iter_mean"] / out["ours"]["ms_per_iter_mean"]
path = os.path.join(ROOT, "gpurun_out", f"r2_parity_higgs{iters}.json")
os.makedirs(os.path.dirname(path), exist_ok=True)
json.dump(out, open(path, "w"), indent=1)
print(json.dumps(out))
