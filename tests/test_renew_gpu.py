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


@pytest.mark.parametrize("objective,weighted", [("huber alpha=1.2", False), ("huber", True), ("fair fair_c=0.8", False),
                                                ("fair", True), ("poisson", False), ("poisson poisson_max_delta_step=0.5", True),
                                                ("gamma", False), ("gamma", True), ("tweedie", False),
                                                ("tweedie tweedie_variance_power=1.2", True)])
def test_pointwise_regression_objectives_identical_trees(objective, weighted):
    """RegressionHuberLoss / FairLoss / PoissonLoss / GammaLoss / TweedieLoss (regression_objective.hpp:290-477, 672-748): pointwise
    gradients on the device, the objective's own metric (through its ConvertOutput: exp for poisson) on the device."""
    X, y = make(30_000, 11)
    name = objective.split()[0]
    if name in ("poisson", "tweedie"):
        rng = np.random.default_rng(3)
        y = rng.poisson(np.exp(0.4 * X[:, 0] + 0.2 * X[:, 1] + 0.5)).astype(np.float32)
    if name == "gamma":   # strictly positive targets
        rng = np.random.default_rng(3)
        y = rng.gamma(2.0, np.exp(0.3 * X[:, 0] - 0.2 * X[:, 1]) / 2.0).astype(np.float32) + 1e-3
    w = (np.random.default_rng(4).random(len(y)).astype(np.float32) + 0.5) if weighted else None
    params = f"objective={objective} num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 verbosity=-1 num_threads=4"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10, weight=w)
    assert nr == no == 10
    assert_same_model(ref, ours, br, bo)
    np.testing.assert_allclose(ours.get_predict(bo, len(y)), ref.get_predict(br, len(y)), rtol=1e-9, atol=1e-9)
    assert ours.eval_names(bo) == ref.eval_names(br) == [name]
    np.testing.assert_allclose(ours.get_eval(bo, 0), ref.get_eval(br, 0), rtol=1e-9)


@pytest.mark.parametrize("bagged", [False, True])
def test_mape_identical_trees_and_leaf_values(bagged):
    """RegressionMAPELOSS (regression_objective.hpp:577-669): L1 gradients scaled by 1 / max(1, |label|), and the same
    weights in the leaf renewal's weighted median.  (With user weights it is refused: see lgbcu_get_gradients.)"""
    X, y = make(40_000, 21)
    y = (y * 3.0 + 4.0).astype(np.float32)     # labels on both sides of |label| = 1
    params = ("objective=mape num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 verbosity=-1 num_threads=4"
              + (" bagging_fraction=0.6 bagging_freq=1 bagging_seed=9" if bagged else ""))
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10)
    assert nr == no == 10
    assert_same_model(ref, ours, br, bo)
    np.testing.assert_allclose(ours.get_predict(bo, len(y)), ref.get_predict(br, len(y)), rtol=1e-9, atol=1e-9)
    assert ours.eval_names(bo) == ref.eval_names(br) == ["mape"]
    np.testing.assert_allclose(ours.get_eval(bo, 0), ref.get_eval(br, 0), rtol=1e-9)
    with pytest.raises(Exception, match="mape"):
        capi.train(ours, X, y, params, 1, weight=np.ones(len(y), dtype=np.float32))


@pytest.mark.parametrize("weighted,metric", [(False, ""), (True, " metric=cross_entropy,kullback_leibler,binary_logloss")])
def test_cross_entropy_identical_trees(weighted, metric):
    """CrossEntropy (xentropy_objective.hpp:44-147): labels are probabilities in [0, 1]; cross_entropy and
    kullback_leibler metrics (xentropy_metric.hpp) through the objective's sigmoid."""
    X, _ = make(40_000, 31)
    rng = np.random.default_rng(32)
    p = 1.0 / (1.0 + np.exp(-(1.5 * X[:, 0] - X[:, 1] * X[:, 2] + 0.3 * rng.standard_normal(len(X)))))
    y = np.where(rng.random(len(X)) < 0.3, (rng.random(len(X)) < p).astype(np.float64), p).astype(np.float32)   # hard and soft labels
    w = (rng.random(len(y)).astype(np.float32) + 0.5) if weighted else None
    params = f"objective=cross_entropy{metric} num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 verbosity=-1 num_threads=4"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10, weight=w)
    assert nr == no == 10
    assert_same_model(ref, ours, br, bo)
    np.testing.assert_allclose(ours.get_predict(bo, len(y)), ref.get_predict(br, len(y)), rtol=1e-9, atol=1e-12)
    assert ours.eval_names(bo) == ref.eval_names(br)
    assert "cross_entropy" in ours.eval_names(bo)
    np.testing.assert_allclose(ours.get_eval(bo, 0), ref.get_eval(br, 0), rtol=1e-9)
    # the reference loads the model and converts with the same sigmoid
    loaded = ref.load_model_from_string(ours.model_string(bo))
    Xt, _ = make(3000, 33)
    np.testing.assert_allclose(ours.predict_mat(bo, Xt), ref.predict_mat(loaded, Xt), rtol=0, atol=1e-12)


@pytest.mark.parametrize("weighted", [False, True])
def test_cross_entropy_lambda_identical_trees(weighted):
    """CrossEntropyLambda (xentropy_objective.hpp:150-283): the weights sit inside the link (a separate gradient form when
    weights are given); its metric keeps them inside the loss and averages over the rows."""
    X, _ = make(40_000, 41)
    rng = np.random.default_rng(42)
    p = 1.0 / (1.0 + np.exp(-(1.2 * X[:, 0] - X[:, 1] * X[:, 2] + 0.3 * rng.standard_normal(len(X)))))
    y = np.where(rng.random(len(X)) < 0.3, (rng.random(len(X)) < p).astype(np.float64), p).astype(np.float32)
    w = (rng.random(len(y)).astype(np.float32) * 1.5 + 0.5) if weighted else None
    params = "objective=cross_entropy_lambda num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 verbosity=-1 num_threads=4"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10, weight=w)
    assert nr == no == 10
    assert_same_model(ref, ours, br, bo)
    np.testing.assert_allclose(ours.get_predict(bo, len(y)), ref.get_predict(br, len(y)), rtol=1e-9, atol=1e-12)
    assert ours.eval_names(bo) == ref.eval_names(br) == ["cross_entropy_lambda"]
    np.testing.assert_allclose(ours.get_eval(bo, 0), ref.get_eval(br, 0), rtol=1e-9)
    loaded = ref.load_model_from_string(ours.model_string(bo))
    Xt, _ = make(3000, 43)
    np.testing.assert_allclose(ours.predict_mat(bo, Xt), ref.predict_mat(loaded, Xt), rtol=0, atol=1e-12)
    mine = ours.load_model_from_string(ref.model_string(br))
    np.testing.assert_allclose(ours.predict_mat(mine, Xt), ref.predict_mat(br, Xt), rtol=0, atol=1e-12)
