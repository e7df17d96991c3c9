"""Percentile objectives against the unmodified reference (oracle/_ref): regression_l1 and quantile
(src/objective/regression_objective.hpp:204-283, 479-571).  Their gradients are signs / constants, and after every
tree SerialTreeLearner::RenewTreeOutput (serial_tree_learner.cpp:661-699) replaces each leaf's value by the (weighted)
percentile of label - score over the leaf's in-bag rows -- here a segmented sort + pick on the device
(lgbcu_renew_leaf_outputs).  Identical trees and leaf values are the bar, with and without weights and bagging."""
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import assert_same_model, run_both

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


def make(n, seed, F=8):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, F)).astype(np.float32)
    y = (2.0 * X[:, 0] + np.abs(X[:, 1]) * 1.5 - X[:, 2] * X[:, 3] + rng.standard_t(3, n) * 0.7).astype(np.float32)
    return X, y


CASES = [("regression_l1", "", False, False), ("regression_l1", "", True, False), ("regression_l1", "", False, True),
         ("quantile", " alpha=0.3", False, False), ("quantile", " alpha=0.9", True, False), ("quantile", " alpha=0.75", True, True)]


@pytest.mark.parametrize("objective,extra,weighted,bagged", CASES)
def test_percentile_objectives_identical_trees_and_leaf_values(objective, extra, weighted, bagged):
    X, y = make(40_000, 7)
    w = (np.random.default_rng(2).random(len(y)).astype(np.float32) * 2 + 0.25) if weighted else None
    params = (f"objective={objective}{extra} num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 verbosity=-1 num_threads=4"
              + (" bagging_fraction=0.7 bagging_freq=1 bagging_seed=5" if bagged else ""))
    ref, ours, br, bo, nr, no = run_both(X, y, params, 12, weight=w)
    assert nr == no == 12
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, len(y)), ours.get_predict(bo, len(y))
    np.testing.assert_allclose(po, pr, rtol=1e-9, atol=1e-9)
    for k in range(len(ref.eval_names(br))):
        np.testing.assert_allclose(ours.get_eval(bo, 0)[k], ref.get_eval(br, 0)[k], rtol=1e-9)
    assert ours.eval_names(bo) == ref.eval_names(br) == ["l1" if objective == "regression_l1" else "quantile"]
