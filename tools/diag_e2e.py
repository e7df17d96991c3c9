import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, capi
import test_e2e_gpu as T
X, y = T.higgs_like(7, 7000)
w = (np.random.default_rng(8).random(7000) + 0.5).astype(np.float32)
params = ("objective=binary num_leaves=63 learning_rate=0.1 feature_fraction=0.8 bagging_fraction=0.8 bagging_freq=5 "
          "min_data_in_leaf=50 min_sum_hessian_in_leaf=5.0 max_bin=255 verbosity=-1")
ref, ours, br, bo, nr, no = T.run_both(X, y, params, 30, weight=w)
ta, tb = capi.parse_trees(ref.model_string(br)), capi.parse_trees(ours.model_string(bo))
for i,(a,b) in enumerate(zip(ta,tb)):
    bad = [k for k in ("split_feature","threshold","decision_type","left_child","right_child","leaf_count") if a.get(k)!=b.get(k)]
    if bad:
        print("tree", i, bad)
        for j,(x,z) in enumerate(zip(a["threshold"], b["threshold"])):
            if x!=z or a["split_feature"][j]!=b["split_feature"][j]:
                print(" node", j, "feat", a["split_feature"][j], b["split_feature"][j], "thr", x, z, "gain", a["split_gain"][j], b["split_gain"][j], "cnt", a["internal_count"][j], b["internal_count"][j], "dt", a["decision_type"][j], b["decision_type"][j])
                break
        print(" ref leaf_count", a["leaf_count"][:12]); print(" our leaf_count", b["leaf_count"][:12])
        break
