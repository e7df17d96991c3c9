"""objective=multiclass (softmax) against the unmodified reference (oracle/_ref): MulticlassSoftmax::GetGradients
(src/objective/multiclass_objective.hpp:86-130) on the device over K score planes, one tree per class and iteration
(GBDT::TrainOneIter's cur_tree_id loop, src/boosting/gbdt.cpp:353-412), per-class boost-from-average, multi_logloss /
multi_error on the device, [class][row] training scores and [row][class] predictions as the reference lays them out."""
import ctypes as C
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import REF_EXTRA, assert_same_model

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


def make(n, K, seed, F=10):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, F)).astype(np.float32)
    X[rng.random((n, F)) < 0.02] = np.nan
    centers = rng.standard_normal((K, F)).astype(np.float32)
    logits = np.nan_to_num(X) @ centers.T + 0.8 * rng.standard_normal((n, K)).astype(np.float32)
    y = np.argmax(logits, axis=1).astype(np.float32)
    return X, y


@pytest.mark.parametrize("K,n,weighted,objective", [(3, 30_000, False, "multiclass"), (5, 20_000, True, "multiclass"),
                                                    (4, 12_000, False, "multiclass"),
                                                    # one-vs-all (multiclass_objective.hpp:186-276): a binary logloss per class, sigmoid outputs
                                                    (3, 20_000, False, "multiclassova"), (4, 15_000, True, "multiclassova is_unbalance=true sigmoid=1.5")])
def test_multiclass_identical_trees_scores_metrics_predictions(K, n, weighted, objective):
    X, y = make(n, K, 100 + K)
    w = (np.random.default_rng(1).random(n).astype(np.float32) + 0.5) if weighted else None
    Xv, yv = make(4000, K, 200 + K)
    params = (f"objective={objective} num_class={K} num_leaves=15 learning_rate=0.2 min_data_in_leaf=20 "
              f"metric=multi_logloss,multi_error verbosity=-1 num_threads=4")
    out = []
    for lib, extra in ((capi.reference(), REF_EXTRA), (capi.ours(), "")):
        ds = lib.dataset_from_mat(X, params + extra)
        lib.set_field(ds, "label", y)
        if w is not None:
            lib.set_field(ds, "weight", w)
        dv = lib.dataset_from_mat(Xv, params + extra, reference=ds)
        lib.set_field(dv, "label", yv)
        b = lib.booster_create(ds, params + extra)
        lib.add_valid(b, dv)
        evals = []
        for it in range(8):
            assert lib.update(b) == 0
            evals.append((lib.get_eval(b, 0), lib.get_eval(b, 1)))
        k = C.c_int(0)
        lib.check(lib.lib.LGBM_BoosterGetNumClasses(b, C.byref(k))); assert k.value == K
        lib.check(lib.lib.LGBM_BoosterNumModelPerIteration(b, C.byref(k))); assert k.value == K
        lib.check(lib.lib.LGBM_BoosterNumberOfTotalModel(b, C.byref(k))); assert k.value == 8 * K
        lib.check(lib.lib.LGBM_BoosterGetCurrentIteration(b, C.byref(k))); assert k.value == 8
        out.append((lib, b, lib.eval_names(b), evals, lib.get_predict(b, n * K), lib.get_predict(b, 4000 * K, data_idx=1),
                    lib.predict_mat(b, Xv, num_class=K), lib.predict_mat(b, Xv, raw=True, num_class=K)))
    (ref, br, nr, er, pr, pvr, mr, rr), (ours, bo, no, eo, po, pvo, mo, ro) = out
    assert nr == no == ["multi_logloss", "multi_error"]
    tr, to = assert_same_model(ref, ours, br, bo)
    assert "num_class=%d" % K in to and "num_tree_per_iteration=%d" % K in to
    for (a0, a1), (b0, b1) in zip(er, eo):
        np.testing.assert_allclose(b0, a0, rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(b1, a1, rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(po, pr, rtol=0, atol=1e-9)       # [class][row] training probabilities
    np.testing.assert_allclose(pvo, pvr, rtol=0, atol=1e-9)
    np.testing.assert_allclose(mo, mr, rtol=0, atol=1e-9)       # [row][class] from raw features
    np.testing.assert_allclose(ro, rr, rtol=0, atol=1e-9)
    if objective == "multiclass":
        assert abs(po.reshape(K, n).sum(axis=0) - 1.0).max() < 1e-12
    # rolling the last iteration back removes K trees and their scores in both libraries
    for lib, b in ((ref, br), (ours, bo)):
        lib.check(lib.lib.LGBM_BoosterRollbackOneIter(b))
    np.testing.assert_allclose(ours.get_predict(bo, n * K), ref.get_predict(br, n * K), rtol=0, atol=1e-9)


def test_multiclass_needs_num_class_and_matching_metric():
    X, y = make(2000, 3, 1)
    ours = capi.ours()
    with pytest.raises(Exception, match="Number of classes"):
        ds = ours.dataset_from_mat(X, "objective=multiclass verbosity=-1")
    ds = ours.dataset_from_mat(X, "objective=multiclass num_class=3 verbosity=-1")
    ours.set_field(ds, "label", y)
    with pytest.raises(Exception, match="don't match"):
        ours.booster_create(ds, "objective=multiclass num_class=3 metric=auc verbosity=-1")
    bad = y.copy(); bad[5] = 7
    ours.set_field(ds, "label", bad)
    with pytest.raises(Exception, match=r"Label must be in \[0, 3\)"):
        ours.booster_create(ds, "objective=multiclass num_class=3 verbosity=-1")
    ours.dataset_free(ds)


def test_multiclass_with_bagging_and_column_sampling():
    """one bag per iteration shared by the iteration's K trees (GBDT::Bagging runs once per TrainOneIter), one column
    sample per tree (ColSampler::ResetByTree in every BeforeTrain): 3 x 10 trees identical to the reference's"""
    K, n = 3, 25_000
    X, y = make(n, K, 77)
    params = (f"objective=multiclass num_class={K} num_leaves=15 learning_rate=0.1 min_data_in_leaf=10 verbosity=-1 num_threads=4 "
              "bagging_fraction=0.7 bagging_freq=2 bagging_seed=3 feature_fraction=0.6 feature_fraction_seed=4")
    from test_e2e_gpu import run_both
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10)
    assert nr == no == 10
    assert_same_model(ref, ours, br, bo)
    np.testing.assert_allclose(ours.get_predict(bo, n * K), ref.get_predict(br, n * K), rtol=0, atol=1e-9)
