"""Monotone constraints, basic method (monotone_constraints=..., monotone_penalty; src/treelearner/monotone_constraints.hpp:
19-126 and the USE_MC forms of feature_histogram.hpp:763-824): every leaf carries the interval its output must stay in,
candidate splits are scored from the clamped outputs, a split on a constrained feature whose outputs are in the wrong order
scores 0, the children of such a split divide the interval at the mean of the two outputs.  Identical trees with the
reference, and predictions that are monotone in the constrained features.
"""
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import assert_same_model, run_both

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


def make(seed, n, F=8):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, F)).astype(np.float32)
    X[:, 3] = np.where(rng.random(n) < 0.05, np.nan, X[:, 3])
    # increasing in x0, decreasing in x1 on average, but noisy and with interactions that tempt the other way
    z = (1.5 * X[:, 0] - 1.0 * X[:, 1] + 0.8 * np.sin(3 * X[:, 0]) + 0.5 * X[:, 1] * X[:, 2] + 0.5 * np.nan_to_num(X[:, 3]) ** 2
         + 0.7 * rng.standard_normal(n))
    return X, z


@pytest.mark.parametrize("seed,extra,objective", [
    (1, "monotone_constraints=1,-1,0,0,0,0,0,0", "regression"),
    (2, "monotone_constraints=1,-1,0,1,0,0,0,0 monotone_penalty=1.5 lambda_l2=1.0", "regression"),
    (3, "monotone_constraints=1,-1,0,0,0,0,0,0 lambda_l1=0.3 max_delta_step=0.6 path_smooth=3 bagging_fraction=0.8 bagging_freq=1", "binary"),
    (4, "monotone_constraints=0,0,0,0,0,0,0,0", "regression"),
])
def test_identical_trees_under_monotone_constraints(seed, extra, objective):
    n = 30_000
    X, z = make(seed, n)
    y = (z > np.median(z)).astype(np.float32) if objective == "binary" else z.astype(np.float32)
    params = f"objective={objective} num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 {extra} verbosity=-1"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10)
    assert nr == no == 10
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, n), ours.get_predict(bo, n)
    assert np.max(np.abs(pr - po)) <= 1e-9
    if "1,-1" in extra:
        # the model is monotone: sweep x0 up and x1 down on a few rows, raw scores never decrease
        base = X[:200].copy(); base[:, 3] = np.nan_to_num(base[:, 3])
        grid = np.linspace(-3, 3, 41, dtype=np.float32)
        for col, sign in ((0, 1.0), (1, -1.0)):
            prev = None
            for g in (grid if sign > 0 else grid[::-1]):
                Z = base.copy(); Z[:, col] = g
                s = ours.predict_mat(bo, Z, raw=True)
                if prev is not None:
                    assert np.all(s >= prev - 1e-12)
                prev = s


def test_monotone_constraints_are_checked():
    X, z = make(9, 2000)
    ours = capi.ours()
    with pytest.raises(Exception, match="monotone"):
        capi.train(ours, X, z.astype(np.float32), "objective=regression monotone_constraints=1,0 verbosity=-1", 1)
    with pytest.raises(Exception, match="monotone"):
        capi.train(ours, X, z.astype(np.float32), "objective=regression monotone_constraints=1,0,0,0,0,0,0,0 monotone_constraints_method=intermediate verbosity=-1", 1)
