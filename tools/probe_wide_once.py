"""Scratch probe: a few iterations on Higgs-shaped data with max_bin=1023 (wide features), for ncu captures of
k_hist_wide / k_find_split<true>.  usage: probe_wide_once.py [rows] [iterations]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import capi, workloads as W
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
X, y = W.higgs_shaped(rows)
ours = capi.ours()
p = "objective=binary num_leaves=255 learning_rate=0.1 min_data_in_leaf=0 min_sum_hessian_in_leaf=100 max_bin=1023 verbosity=-1"
ds = ours.dataset_from_mat(X, p); ours.set_field(ds, "label", y)
b = ours.booster_create(ds, p)
for _ in range(iters): ours.update(b)
ours.get_predict(b, rows)
print("done")
