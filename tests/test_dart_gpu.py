"""boosting=dart (src/boosting/dart.hpp): before the gradients of an iteration are taken a random subset of the earlier
trees is dropped from the training scores; after the new tree is grown the dropped trees return scaled by k / (k + 1)
(and the new one enters at learning_rate / (k + 1)), in the validation scores too.  The drops come from the reference's
LCG with drop_seed, so the whole run is reproducible: identical models, training scores and validation metrics.
"""
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import assert_same_model, higgs_like

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


@pytest.mark.parametrize("extra", ["", " drop_rate=0.5 skip_drop=0.2 max_drop=3", " uniform_drop=true drop_rate=0.3 skip_drop=0",
                                   " xgboost_dart_mode=true drop_rate=0.4 skip_drop=0.1 drop_seed=11",
                                   " objective=multiclass num_class=3 drop_rate=0.4 skip_drop=0"])
def test_dart_identical_models_scores_and_validation_metric(extra):
    X, y = higgs_like(31, 30_000, F=12)
    Xv, yv = higgs_like(32, 8_000, F=12)
    if "multiclass" in extra:
        y = (np.digitize(X[:, 0] + 0.5 * X[:, 1], [-0.5, 0.5])).astype(np.float32)
        yv = (np.digitize(Xv[:, 0] + 0.5 * Xv[:, 1], [-0.5, 0.5])).astype(np.float32)
        params = "boosting=dart num_leaves=15 learning_rate=0.2 min_data_in_leaf=20 verbosity=-1" + extra
    else:
        params = "objective=binary metric=binary_logloss boosting=dart num_leaves=31 learning_rate=0.2 min_data_in_leaf=20 verbosity=-1" + extra
    out = {}
    for name, lib, more in (("ref", capi.reference(), " force_col_wise=true num_threads=4"), ("ours", capi.ours(), "")):
        ds, b, _ = capi.train(lib, X, y, params + more, 0)
        dv = lib.dataset_from_mat(Xv, params + more, reference=ds)
        lib.set_field(dv, "label", yv)
        lib.add_valid(b, dv)
        evals = []
        for _ in range(15):
            lib.update(b)
            evals.append(lib.get_eval(b, 1)[0])
        out[name] = (lib, b, np.array(evals))
    (ref, br, er), (ours, bo, eo) = out["ref"], out["ours"]
    assert_same_model(ref, ours, br, bo)
    np.testing.assert_allclose(eo, er, rtol=1e-9)
    K = 3 if "multiclass" in extra else 1
    pr = ref.get_predict(br, len(y) * K); po = ours.get_predict(bo, len(y) * K)
    assert np.max(np.abs(pr - po)) <= 1e-9
