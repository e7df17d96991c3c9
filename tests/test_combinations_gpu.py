"""The F4 paths together: categorical + wide + constrained + bundled features in one dataset, under the other boosting
modes, objectives and samplers.  Each case: the reference on the CPU and this library on the GPU, identical trees (a
mirrored one-vs-rest tie aside, see test_categorical_gpu.py) and equal training scores.
"""
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import run_both
from test_oracle_vs_reference import categorical_matrix, sparse_matrix

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


def combined_matrix(seed, n):
    """7 columns of categorical_matrix (0, 2, 4, 5 categorical) + 6 continuous columns + 12 sparse one-hot columns"""
    rng = np.random.default_rng(seed)
    A = categorical_matrix(seed, n)
    B = rng.standard_normal((n, 6))
    B[:, 1] = np.where(rng.random(n) < 0.05, np.nan, B[:, 1])
    cat = rng.integers(0, 12, size=n)
    Cm = (cat[:, None] == np.arange(12)[None, :]).astype(np.float64)
    X = np.concatenate([A, B, Cm], axis=1)
    eff = np.random.default_rng(5).standard_normal(400)
    z = (eff[np.nan_to_num(X[:, 2]).astype(int) % 400] + 0.8 * eff[np.nan_to_num(X[:, 4]).astype(int) % 400] + 1.2 * X[:, 7]
         - 0.9 * X[:, 9] + 0.5 * np.nan_to_num(X[:, 8]) * X[:, 10] + 0.3 * eff[cat] + 0.4 * rng.standard_normal(n))
    return X, z


BASE = "categorical_feature=2,4,5 num_leaves=31 min_data_in_leaf=20 learning_rate=0.1 verbosity=-1"
MC = " monotone_constraints=0,0,0,0,0,0,0,1,0,-1" + ",0" * 15


@pytest.mark.parametrize("name,params,objective", [
    ("wide+cat+mc+bundles", BASE + " max_bin=600" + MC, "regression"),
    ("goss", BASE + " boosting=goss top_rate=0.3 other_rate=0.2" + MC, "binary"),
    ("dart", BASE + " boosting=dart drop_rate=0.4 skip_drop=0 max_bin=300", "binary"),
    ("rf", BASE + " boosting=rf bagging_fraction=0.7 bagging_freq=1 feature_fraction=0.8" + MC, "regression"),
    ("multiclass", BASE + " objective=multiclass num_class=3 max_bin=400 bagging_fraction=0.8 bagging_freq=1", "multiclass"),
    ("l1+renew", BASE + " objective=regression_l1 max_bin=500" + MC, "regression_l1"),
])
def test_f4_paths_together(name, params, objective):
    n = 30_000
    X, z = combined_matrix(3, n)
    if objective == "binary":
        y = (z > np.median(z)).astype(np.float32); params += " objective=binary"
    elif objective == "multiclass":
        y = np.digitize(z, np.quantile(z, [0.33, 0.66])).astype(np.float32)
    elif objective == "regression":
        y = z.astype(np.float32); params += " objective=regression"
    else:
        y = z.astype(np.float32)
    ref, ours, br, bo, nr, no = run_both(X, y, params, 8)
    assert nr == no == 8
    tr, to = ref.model_string(br), ours.model_string(bo)
    same, n_trees, first = capi.compare_models(tr, to, stop_at_first=False)
    assert same >= n_trees - 2, f"{name}: {same}/{n_trees} trees identical; first mismatch: {first}"
    K = 3 if objective == "multiclass" else 1
    assert np.max(np.abs(ref.get_predict(br, n * K) - ours.get_predict(bo, n * K))) <= 1e-9
    assert any(int(t["num_cat"][0]) > 0 for t in capi.parse_trees(to)), "no categorical split in the case"
