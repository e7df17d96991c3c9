"""Scratch probe: MSLR-shaped lambdarank -- time per boosting iteration through the C API."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi
rng = np.random.default_rng(3)
n_target = int(sys.argv[1]) if len(sys.argv) > 1 else 2_270_000
sizes = np.clip(rng.poisson(120, size=n_target // 118), 1, 1000).astype(np.int32)
n = int(sizes.sum()); F = 137
X = rng.standard_normal((n, F)).astype(np.float32)
X[:, 100:] = np.round(np.abs(X[:, 100:]) * 3)
latent = X[:, 0] + 0.5 * X[:, 101] + 0.5 * rng.standard_normal(n)
y = np.digitize(latent, np.quantile(latent, [0.52, 0.84, 0.97, 0.99])).astype(np.float32)
params = "objective=lambdarank num_leaves=255 learning_rate=0.1 min_data_in_leaf=20 verbosity=-1"
ours = capi.ours()
t = time.time(); ds = ours.dataset_from_mat(X, params); ours.set_field(ds, "label", y); ours.set_field(ds, "group", sizes)
print(f"rows {n}, queries {len(sizes)}; dataset {time.time()-t:.1f} s")
b = ours.booster_create(ds, params)
for i in range(6):
    t = time.time(); ours.update(b); print(f"iter {i}: {(time.time()-t)*1e3:.1f} ms")
t = time.time(); ev = ours.get_eval(b, 0); print("ndcg@1..5 on train:", np.round(ev, 4), f"({(time.time()-t)*1e3:.1f} ms)")
