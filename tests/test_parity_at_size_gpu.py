"""Parity against the unmodified reference (oracle/_ref, CPU, tree_learner=serial force_col_wise) AT THE SIZES
BASELINE.json's configs name -- not properties, the same LGBM_* calls on the same matrix:
  config #1  the reference's own examples/binary_classification fixture + its train.conf (weights, column sampling,
             subset bagging, a validation set)
  config #2  Higgs-shaped 10.5M x 28, 255 leaves, binary
  config #3  Epsilon-shaped 400K x 2000, 255 leaves, binary (wide-feature path)
  config #4  MSLR-shaped 2.27M x 137, lambdarank, queries of ~120 documents
Bars (north_star): identical split feature / threshold per node except on gain near-ties; loss / AUC within 1e-5.
The iteration counts are what the reference's CPU learner finishes in about a minute on the box's host cores; the
500-iteration run of config #2 is a committed record (profiles/, tools/parity_higgs500.py), not a test."""
import os

import numpy as np
import pytest

import capi
import helpers as H
import workloads as W
from test_e2e_gpu import assert_same_model, auc, logloss

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]

CORES = os.cpu_count() or 4
REF_EXTRA = f" force_col_wise=true num_threads={CORES}"


def _both(X, y, params, iters, **kw):
    ref, ours = capi.reference(), capi.ours()
    dsr, br, nr = capi.train(ref, X, y, params + REF_EXTRA, iters, **kw)
    tr = ref.model_string(br)
    pr = ref.get_predict(br, len(y))
    ref.booster_free(br); ref.dataset_free(dsr)
    dso, bo, no = capi.train(ours, X, y, params + f" num_threads={CORES}", iters, **kw)
    to = ours.model_string(bo)
    po = ours.get_predict(bo, len(y))
    ours.booster_free(bo); ours.dataset_free(dso)
    assert nr == no == iters
    return tr, to, pr, po


def _assert_models(tr, to, iters):
    same, n, first = capi.compare_models(tr, to, rtol_value=H.VALUE_RTOL)
    assert n == iters
    assert same == n, f"{same}/{n} trees identical; first mismatch: {first}"


def test_config1_real_fixture_with_train_conf():
    fx = np.load(os.path.join(H.ROOT, "tests", "golden", "config1_binary_classification.npz"))
    params = str(fx["params"]) + " verbosity=-1"
    Xtr, ytr, wtr = fx["X_train"], fx["y_train"], fx["w_train"]
    Xte, yte, wte = fx["X_test"], fx["y_test"], fx["w_test"]
    assert Xtr.shape == (7000, 28) and Xte.shape == (500, 28)
    out = []
    for lib, extra in ((capi.reference(), " force_col_wise=true num_threads=4"), (capi.ours(), "")):
        ds = lib.dataset_from_mat(Xtr, params + extra)
        lib.set_field(ds, "label", ytr); lib.set_field(ds, "weight", wtr)
        dv = lib.dataset_from_mat(Xte, params + extra, reference=ds)
        lib.set_field(dv, "label", yte); lib.set_field(dv, "weight", wte)
        b = lib.booster_create(ds, params + extra)
        lib.add_valid(b, dv)
        evals = []
        for it in range(100):   # train.conf: num_trees = 100
            assert lib.update(b) == 0
            if it % 10 == 9:
                evals.append((lib.get_eval(b, 0), lib.get_eval(b, 1)))
        out.append((lib.model_string(b), lib.get_predict(b, 7000), evals, lib.eval_names(b)))
    (tr, pr, er, nr), (to, po, eo, no) = out
    assert nr == no == ["binary_logloss", "auc"]
    _assert_models(tr, to, 100)
    for (a0, a1), (b0, b1) in zip(er, eo):
        np.testing.assert_allclose(b0, a0, rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(b1, a1, rtol=1e-6, atol=1e-9)
    assert abs(logloss(pr, ytr) - logloss(po, ytr)) <= 1e-5
    assert abs(auc(pr, ytr) - auc(po, ytr)) <= 1e-5


def test_config2_higgs_shaped_full_size():
    X, y = W.higgs_shaped(10_500_000)
    iters = 10
    tr, to, pr, po = _both(X, y, W.HIGGS_PARAMS, iters)
    _assert_models(tr, to, iters)
    assert np.max(np.abs(pr - po)) <= 1e-6
    assert abs(logloss(pr, y) - logloss(po, y)) <= 1e-5


def test_config3_epsilon_shaped_full_size():
    X, y = W.epsilon_shaped(400_000, 2000)
    iters = 3
    tr, to, pr, po = _both(X, y, W.EPSILON_PARAMS, iters)
    _assert_models(tr, to, iters)
    assert np.max(np.abs(pr - po)) <= 1e-6
    assert abs(logloss(pr, y) - logloss(po, y)) <= 1e-5


def test_config4_mslr_shaped_full_size():
    X, y, sizes = W.mslr_shaped(2_270_000, 137)
    iters = 3
    tr, to, pr, po = _both(X, y, W.MSLR_PARAMS, iters, group=sizes)
    _assert_models(tr, to, iters)
    assert np.max(np.abs(pr - po)) <= 1e-6 * max(1.0, float(np.max(np.abs(pr))))
