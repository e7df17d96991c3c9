"""Validation data and device metrics through the drop-in C API: LGBM_BoosterAddValidData,
LGBM_BoosterGetEvalCounts / GetEvalNames / GetEval and LGBM_BoosterGetPredict(data_idx > 0) against the same calls on the
reference's lib_lightgbm.so.  Bar: metric names identical; values within 1e-9 relative (north_star asks for 1e-5; the
trees are identical, what differs is the order of the double-precision sums)."""
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import REF_EXTRA, higgs_like

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


def _run(lib, Xt, yt, Xv, yv, params, iters, wt=None, wv=None, qt=None, qv=None, valid_late=False, extra=""):
    p = params + extra
    dt = lib.dataset_from_mat(Xt, p)
    lib.set_field(dt, "label", yt)
    dv = lib.dataset_from_mat(Xv, p, reference=dt)
    lib.set_field(dv, "label", yv)
    if wt is not None:
        lib.set_field(dt, "weight", wt); lib.set_field(dv, "weight", wv)
    if qt is not None:
        lib.set_field(dt, "group", qt); lib.set_field(dv, "group", qv)
    b = lib.booster_create(dt, p)
    if not valid_late:
        lib.add_valid(b, dv)
    names = lib.eval_names(b)
    hist = []
    for it in range(iters):
        lib.update(b)
        if valid_late and it == iters // 2:
            lib.add_valid(b, dv)   # trees grown so far are replayed onto the new set
        if not valid_late or it >= iters // 2:
            hist.append((lib.get_eval(b, 0), lib.get_eval(b, 1)))
    pred_v = lib.get_predict(b, len(yv), data_idx=1)
    return names, hist, pred_v


def _compare(a, b, rtol=1e-9):
    names_a, hist_a, pv_a = a
    names_b, hist_b, pv_b = b
    assert names_a == names_b
    assert len(hist_a) == len(hist_b)
    for (ta, va), (tb, vb) in zip(hist_a, hist_b):
        np.testing.assert_allclose(tb, ta, rtol=rtol, atol=1e-12)
        np.testing.assert_allclose(vb, va, rtol=rtol, atol=1e-12)
    np.testing.assert_allclose(pv_b, pv_a, rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("weighted", [False, True])
def test_binary_logloss_and_auc(weighted):
    X, y = higgs_like(11, 60_000)
    Xt, yt, Xv, yv = X[:40_000], y[:40_000], X[40_000:], y[40_000:]
    rng = np.random.default_rng(5)
    wt = (rng.random(len(yt)) + 0.5).astype(np.float32) if weighted else None
    wv = (rng.random(len(yv)) + 0.5).astype(np.float32) if weighted else None
    params = "objective=binary metric=binary_logloss,auc,binary_error num_leaves=31 min_data_in_leaf=20 learning_rate=0.1 verbosity=-1"
    ref = _run(capi.reference(), Xt, yt, Xv, yv, params, 8, wt, wv, extra=REF_EXTRA)
    ours = _run(capi.ours(), Xt, yt, Xv, yv, params, 8, wt, wv)
    assert ref[0] == ["binary_logloss", "auc", "binary_error"]
    _compare(ref, ours)


def test_auc_with_tied_scores_and_default_metric():
    # 3 coarse features -> few distinct leaves -> many tied scores: the tie groups of the AUC matter
    rng = np.random.default_rng(2)
    X = rng.integers(0, 3, size=(20_000, 3)).astype(np.float32)
    y = ((X[:, 0] + rng.random(20_000) * 2) > 2).astype(np.float32)
    params = "objective=binary metric=auc num_leaves=4 min_data_in_leaf=50 verbosity=-1"
    ref = _run(capi.reference(), X[:12_000], y[:12_000], X[12_000:], y[12_000:], params, 4, extra=REF_EXTRA)
    ours = _run(capi.ours(), X[:12_000], y[:12_000], X[12_000:], y[12_000:], params, 4)
    _compare(ref, ours)
    # no `metric` given: the objective names it
    params = "objective=binary num_leaves=4 min_data_in_leaf=50 verbosity=-1"
    ours = _run(capi.ours(), X[:12_000], y[:12_000], X[12_000:], y[12_000:], params, 2)
    assert ours[0] == ["binary_logloss"]


def test_regression_l2_rmse_and_late_valid_set():
    rng = np.random.default_rng(3)
    X = rng.standard_normal((30_000, 12)).astype(np.float32)
    y = (X[:, 0] * 2 + np.sin(X[:, 1]) + 0.1 * rng.standard_normal(30_000)).astype(np.float32)
    params = "objective=regression metric=l2,rmse num_leaves=31 min_data_in_leaf=20 learning_rate=0.1 verbosity=-1"
    args = (X[:20_000], y[:20_000], X[20_000:], y[20_000:], params, 8)
    ref = _run(capi.reference(), *args, valid_late=True, extra=REF_EXTRA)
    ours = _run(capi.ours(), *args, valid_late=True)
    assert ref[0] == ["l2", "rmse"]
    _compare(ref, ours)


@pytest.mark.parametrize("weighted", [False, True])
def test_lambdarank_ndcg(weighted):
    rng = np.random.default_rng(4)
    def make(nq):
        sizes = rng.integers(1, 40, size=nq).astype(np.int32)
        n = int(sizes.sum())
        X = rng.standard_normal((n, 10)).astype(np.float32)
        rel = X[:, 0] + 0.5 * X[:, 1] + 0.5 * rng.standard_normal(n)
        y = np.clip(np.floor(rel + 1.5), 0, 4).astype(np.float32)
        y[: sizes[0]] = 0          # first query: all documents irrelevant -> counts as NDCG 1
        return X, y, sizes
    Xt, yt, qt = make(600)
    Xv, yv, qv = make(300)
    wt = (rng.random(len(yt)) + 0.5).astype(np.float32) if weighted else None
    wv = (rng.random(len(yv)) + 0.5).astype(np.float32) if weighted else None
    params = ("objective=lambdarank metric=ndcg eval_at=1,3,5,10 num_leaves=15 min_data_in_leaf=10 "
              "learning_rate=0.1 verbosity=-1")
    ref = _run(capi.reference(), Xt, yt, Xv, yv, params, 6, wt, wv, qt, qv, extra=REF_EXTRA)
    ours = _run(capi.ours(), Xt, yt, Xv, yv, params, 6, wt, wv, qt, qv)
    assert ref[0] == ["ndcg@1", "ndcg@3", "ndcg@5", "ndcg@10"]
    _compare(ref, ours)


@pytest.mark.parametrize("weighted", [False, True])
def test_lambdarank_map_beside_ndcg(weighted):
    """MapMetric (map_metric.hpp): map@k on the device beside ndcg@k, binary relevance, a query without positives = 1"""
    rng = np.random.default_rng(14)
    def make(nq):
        sizes = rng.integers(1, 40, size=nq).astype(np.int32)
        n = int(sizes.sum())
        X = rng.standard_normal((n, 10)).astype(np.float32)
        rel = X[:, 0] + 0.5 * X[:, 1] + 0.5 * rng.standard_normal(n)
        y = (rel > 0.6).astype(np.float32)
        y[: sizes[0]] = 0
        return X, y, sizes
    Xt, yt, qt = make(500)
    Xv, yv, qv = make(250)
    wt = (rng.random(len(yt)) + 0.5).astype(np.float32) if weighted else None
    wv = (rng.random(len(yv)) + 0.5).astype(np.float32) if weighted else None
    params = ("objective=lambdarank metric=map,ndcg eval_at=1,2,5,20 num_leaves=15 min_data_in_leaf=10 "
              "learning_rate=0.1 verbosity=-1")
    ref = _run(capi.reference(), Xt, yt, Xv, yv, params, 6, wt, wv, qt, qv, extra=REF_EXTRA)
    ours = _run(capi.ours(), Xt, yt, Xv, yv, params, 6, wt, wv, qt, qv)
    assert ref[0] == ["map@1", "map@2", "map@5", "map@20", "ndcg@1", "ndcg@2", "ndcg@5", "ndcg@20"]
    _compare(ref, ours)


@pytest.mark.parametrize("weighted", [False, True])
def test_rank_xendcg_identical_run(weighted):
    """RankXENDCG (rank_objective.hpp:288-366): the randomised gains come from one LCG per query (objective_seed + query),
    advanced a draw per document per iteration -- reproduced on the device, so the whole run is the reference's"""
    rng = np.random.default_rng(24)
    def make(nq):
        sizes = rng.integers(1, 60, size=nq).astype(np.int32)
        n = int(sizes.sum())
        X = rng.standard_normal((n, 10)).astype(np.float32)
        rel = X[:, 0] + 0.5 * X[:, 1] + 0.5 * rng.standard_normal(n)
        y = np.clip(np.floor(rel + 1.5), 0, 4).astype(np.float32)
        return X, y, sizes
    Xt, yt, qt = make(700)
    Xv, yv, qv = make(300)
    wt = (rng.random(len(yt)) + 0.5).astype(np.float32) if weighted else None
    wv = (rng.random(len(yv)) + 0.5).astype(np.float32) if weighted else None
    params = ("objective=rank_xendcg metric=ndcg eval_at=1,5,10 objective_seed=17 num_leaves=15 min_data_in_leaf=10 "
              "learning_rate=0.1 verbosity=-1")
    ref = _run(capi.reference(), Xt, yt, Xv, yv, params, 8, wt, wv, qt, qv, extra=REF_EXTRA)
    ours = _run(capi.ours(), Xt, yt, Xv, yv, params, 8, wt, wv, qt, qv)
    _compare(ref, ours, rtol=1e-7)


@pytest.mark.parametrize("weighted", [False, True])
def test_reg_sqrt_objective_and_converted_metrics(weighted):
    """reg_sqrt (RegressionL2loss, regression_objective.hpp:96-160): the objective fits sign(y) sqrt(|y|), predictions and
    metrics see sign(s) s^2"""
    rng = np.random.default_rng(34)
    def make(n):
        X = rng.standard_normal((n, 8)).astype(np.float32)
        y = ((2 * X[:, 0] + X[:, 1] * X[:, 2]) ** 3 + rng.standard_normal(n)).astype(np.float32)
        return X, y
    Xt, yt = make(20_000); Xv, yv = make(6_000)
    wt = (rng.random(len(yt)) + 0.5).astype(np.float32) if weighted else None
    wv = (rng.random(len(yv)) + 0.5).astype(np.float32) if weighted else None
    params = "objective=regression reg_sqrt=true metric=l2,l1,rmse num_leaves=15 min_data_in_leaf=20 learning_rate=0.1 verbosity=-1"
    ref = _run(capi.reference(), Xt, yt, Xv, yv, params, 6, wt, wv, extra=REF_EXTRA)
    ours = _run(capi.ours(), Xt, yt, Xv, yv, params, 6, wt, wv)
    _compare(ref, ours)


def test_gamma_deviance_metric():
    rng = np.random.default_rng(15)
    def make(n):
        X = rng.standard_normal((n, 8)).astype(np.float32)
        y = (rng.gamma(2.0, np.exp(0.3 * X[:, 0] - 0.2 * X[:, 1]) / 2.0) + 1e-3).astype(np.float32)
        return X, y
    Xt, yt = make(20_000); Xv, yv = make(6_000)
    wt = (rng.random(len(yt)) + 0.5).astype(np.float32); wv = (rng.random(len(yv)) + 0.5).astype(np.float32)
    params = "objective=gamma metric=gamma,gamma_deviance num_leaves=15 min_data_in_leaf=20 learning_rate=0.1 verbosity=-1"
    ref = _run(capi.reference(), Xt, yt, Xv, yv, params, 6, wt, wv, extra=REF_EXTRA)
    ours = _run(capi.ours(), Xt, yt, Xv, yv, params, 6, wt, wv)
    assert ref[0] == ["gamma", "gamma_deviance"]
    _compare(ref, ours)


def test_misaligned_valid_set_and_unknown_metric_are_errors():
    ours = capi.ours()
    X, y = higgs_like(1, 5_000)
    dt = ours.dataset_from_mat(X, "verbosity=-1"); ours.set_field(dt, "label", y)
    dv = ours.dataset_from_mat(X[:1000] * 3 + 1, "verbosity=-1"); ours.set_field(dv, "label", y[:1000])  # its own bins
    b = ours.booster_create(dt, "objective=binary verbosity=-1")
    with pytest.raises(capi.LGBMError, match="different bin mappers"):
        ours.add_valid(b, dv)
    with pytest.raises(capi.LGBMError, match="outside this library's path"):
        ours.booster_create(dt, "objective=binary metric=auc_mu verbosity=-1")
