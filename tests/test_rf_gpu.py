"""boosting=rf (src/boosting/rf.hpp): the gradients are taken once at the initial scores, every iteration grows a tree on
a fresh bag at shrinkage 1, the scores are kept as the running mean of the trees and predictions are averaged
(average_output in the model header).  Identical models, training scores, validation metrics and predictions.
"""
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import assert_same_model, higgs_like

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


@pytest.mark.parametrize("params", [
    "objective=binary metric=binary_logloss bagging_fraction=0.7 bagging_freq=1 feature_fraction=0.8",
    "objective=regression metric=l2 bagging_fraction=0.6 bagging_freq=2 bagging_seed=9",
    "objective=multiclass num_class=3 metric=multi_logloss bagging_fraction=0.75 bagging_freq=1",
])
def test_rf_identical_models_scores_and_predictions(params):
    X, y = higgs_like(41, 30_000, F=12)
    Xv, yv = higgs_like(42, 8_000, F=12)
    if "multiclass" in params:
        y = (np.digitize(X[:, 0] + 0.5 * X[:, 1], [-0.5, 0.5])).astype(np.float32)
        yv = (np.digitize(Xv[:, 0] + 0.5 * Xv[:, 1], [-0.5, 0.5])).astype(np.float32)
    elif "regression" in params:
        y = (X[:, 0] * 2 + X[:, 1] ** 2 + 0.1 * X[:, 2]).astype(np.float32)
        yv = (Xv[:, 0] * 2 + Xv[:, 1] ** 2 + 0.1 * Xv[:, 2]).astype(np.float32)
    params = params + " boosting=rf num_leaves=31 min_data_in_leaf=20 verbosity=-1"
    K = 3 if "multiclass" in params else 1
    out = {}
    for name, lib, more in (("ref", capi.reference(), " force_col_wise=true num_threads=4"), ("ours", capi.ours(), "")):
        ds, b, _ = capi.train(lib, X, y, params + more, 0)
        dv = lib.dataset_from_mat(Xv, params + more, reference=ds)
        lib.set_field(dv, "label", yv)
        lib.add_valid(b, dv)
        evals = []
        for _ in range(10):
            lib.update(b)
            evals.append(lib.get_eval(b, 1)[0])
        out[name] = (lib, b, np.array(evals))
    (ref, br, er), (ours, bo, eo) = out["ref"], out["ours"]
    tr, to = assert_same_model(ref, ours, br, bo)
    assert "average_output" in to.split("Tree=")[0]
    np.testing.assert_allclose(eo, er, rtol=1e-9)
    assert np.max(np.abs(ref.get_predict(br, len(y) * K) - ours.get_predict(bo, len(y) * K))) <= 1e-9
    for raw in (False, True):
        a = ref.predict_mat(br, Xv, raw=raw, num_class=K); b = ours.predict_mat(bo, Xv, raw=raw, num_class=K)
        assert np.max(np.abs(a - b)) <= 1e-9
    # the model text round-trips through the other library with the averaging intact
    assert np.max(np.abs(ref.predict_mat(ref.load_model_from_string(to), Xv, num_class=K) - ours.predict_mat(ours.load_model_from_string(tr), Xv, num_class=K))) <= 1e-12


def test_rf_needs_bagging():
    X, y = higgs_like(43, 2000, F=6)
    with pytest.raises(Exception, match="bagging"):
        capi.train(capi.ours(), X, y, "objective=binary boosting=rf verbosity=-1", 1)
