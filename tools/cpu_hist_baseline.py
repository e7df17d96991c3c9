"""BASELINE.md 3, CPU side of the roofline: the reference's own -DTIMETAG build (oracle/_ref/lib_lightgbm_timetag.so,
`make -C oracle ref_timetag`) trains the bench workload for a few iterations; Common::Timer prints every stage's
accumulated seconds at process exit.  CPU histogram GB/s = the SAME algorithmic bytes the GPU roofline uses (rows
histogrammed x 36 bytes: 28 bin bytes + 8 bytes of (g, h); + the histogram written) / its
"SerialTreeLearner::ConstructHistograms" seconds.  Rows histogrammed are read off the trained model (root + the smaller
child of every split), so both arms are charged for the same work.
usage: cpu_hist_baseline.py [iterations] [col|row]   -> one JSON line (also written to gpurun_out/cpu_hist_baseline_<mode>.json)"""
import json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
mode = sys.argv[2] if len(sys.argv) > 2 else "col"
LIB = os.path.join(ROOT, "oracle", "_ref", "lib_lightgbm_timetag.so")
CHILD = r'''
import os, sys, time
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
os.environ["LGBM_REF_LIB"] = %(lib)r
import numpy as np, capi, workloads as W
ref = capi.LGBMLib(%(lib)r, True)
X, y = W.higgs_shaped(10_500_000)
cores = os.cpu_count() or 4
p = W.HIGGS_PARAMS.replace("verbosity=-1", "verbosity=1") + " force_%(mode)s_wise=true num_threads=%%d" %% cores
ds = ref.dataset_from_mat(X, p); ref.set_field(ds, "label", y)
b = ref.booster_create(ds, p)
t0 = time.perf_counter()
for _ in range(%(iters)d): ref.update(b)
print("TRAIN_SECONDS", time.perf_counter() - t0, "CORES", cores, flush=True)
open(%(model)r, "w").write(ref.model_string(b))
'''
model = os.path.join(ROOT, "gpurun_out", "cpu_hist_model.txt")
os.makedirs(os.path.dirname(model), exist_ok=True)
out = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT, "lib": LIB, "mode": mode, "iters": iters, "model": model}],
                     capture_output=True, text=True, timeout=3600)
log = out.stdout + out.stderr
if out.returncode != 0:
    print(log[-3000:]); sys.exit(1)
stages = {m.group(1): float(m.group(2)) for m in re.finditer(r"\] ([A-Za-z:_]+) costs:\s+([0-9.eE+-]+)", log)}
train_s = float(re.search(r"TRAIN_SECONDS ([0-9.]+)", log).group(1)); cores = int(re.search(r"CORES (\d+)", log).group(1))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import capi
trees = capi.parse_trees(open(model).read())
rows = 0
for t in trees:
    if not t.get("internal_count"): continue
    lc, rc = [int(v) for v in t["left_child"]], [int(v) for v in t["right_child"]]
    ic, lcnt = [int(v) for v in t["internal_count"]], [int(v) for v in t["leaf_count"]]
    rows += ic[0]
    cnt = lambda c: ic[c] if c >= 0 else lcnt[~c]
    for i in range(len(lc)): rows += min(cnt(lc[i]), cnt(rc[i]))
F, total_bin = 28, None
bytes_alg = rows * 36.0
hist_s = stages.get("SerialTreeLearner::ConstructHistograms")
res = {"workload": "Higgs-shaped 10500000x28, 255 leaves, binary", "mode": "force_%s_wise" % mode, "iterations": iters, "cores": cores,
       "train_seconds": train_s, "iters_per_s": iters / train_s, "rows_histogrammed": rows, "algorithmic_bytes": bytes_alg,
       "construct_histograms_seconds": hist_s, "cpu_hist_gb_per_s": (bytes_alg / 1e9 / hist_s) if hist_s else None,
       "stages_seconds": {k: v for k, v in sorted(stages.items()) if v >= 0.01}}
print(json.dumps(res))
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "cpu_hist_baseline_%s.json" % mode), "w"), indent=1)
