"""The inner boundary of INTEGRATION.md section 2, run for real: oracle/b200_tree_learner.cpp is a subclass of the
REFERENCE's `TreeLearner` (compiled against the reference's headers, linked to the unmodified oracle/_ref/lib_lightgbm.so)
that forwards Train / SetBaggingData / AddPredictionToScore to this repository's lgbcu_* layer, and the reference's own
GBDT -- its Dataset, its host objective, its score updater, its model writer -- trains around it.  The model must be the
one the reference's stock serial learner produces on the same Dataset."""
import ctypes as C
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import REF_EXTRA, higgs_like

SHIM = os.path.join(H.ROOT, "oracle", "_ref", "libb200_learner_shim.so")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (os.path.exists(H.REF_SO) and os.path.exists(SHIM)),
                                                 reason="oracle/_ref (reference + learner shim) not built")]


def shim_train(ref_dataset, params, iters):
    lib = C.CDLL(SHIM)
    buf = C.create_string_buffer(8 << 20)
    ln = C.c_longlong(0)
    rc = lib.b200_shim_train(ref_dataset, C.c_char_p(params.encode()), C.c_int(iters), buf, C.c_longlong(len(buf)), C.byref(ln))
    assert rc == 0, buf.value.decode()
    assert 0 < ln.value <= len(buf)
    return buf.value.decode()


@pytest.mark.parametrize("params,iters", [
    ("objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 verbosity=-1", 8),
    ("objective=regression num_leaves=63 learning_rate=0.05 min_data_in_leaf=5 lambda_l2=1.5 verbosity=-1", 6),
    ("objective=binary num_leaves=31 bagging_fraction=0.8 bagging_freq=1 bagging_seed=9 verbosity=-1", 8),   # bag by indices
])
def test_reference_gbdt_around_the_b200_tree_learner_gives_the_stock_model(params, iters):
    X, y = higgs_like(31, 30_000)
    if "regression" in params:
        y = (X[:, 0] * 2 - X[:, 1] + 0.1 * np.random.default_rng(0).standard_normal(len(y))).astype(np.float32)
    ref = capi.reference()
    p = params + REF_EXTRA
    ds = ref.dataset_from_mat(X, p)
    ref.set_field(ds, "label", y)
    b = ref.booster_create(ds, p)
    for _ in range(iters):
        ref.update(b)
    stock = ref.model_string(b)
    swapped = shim_train(ds, p, iters)
    same, n, first = capi.compare_models(stock, swapped, rtol_value=H.VALUE_RTOL)
    assert n == iters and same == n, f"{same}/{n} trees identical; first mismatch: {first}"
    hs, hw = capi.header_lines(stock), capi.header_lines(swapped)
    for k in ("version", "num_class", "max_feature_idx", "objective", "feature_names", "feature_infos"):
        assert hs.get(k) == hw.get(k), k


@pytest.mark.parametrize("kind", ["categorical", "monotone"])
def test_categorical_and_monotone_through_the_reference_tree_learner_interface(kind):
    """The same swap for the F4 paths: categorical splits reach the reference's Tree::SplitCategorical from the device's
    bin sets, monotone constraints go to the device from the reference's Config"""
    from test_oracle_vs_reference import categorical_matrix
    from test_categorical_gpu import target
    n, iters = 25_000, 6
    ref = capi.reference()
    if kind == "categorical":
        X = categorical_matrix(21, n)
        z = target(X, np.random.default_rng(22))
        y = (z > np.median(z)).astype(np.float32)
        p = "objective=binary num_leaves=31 min_data_in_leaf=20 categorical_feature=2,4,5 enable_bundle=false verbosity=-1" + REF_EXTRA
    else:
        X, _ = higgs_like(23, n, F=8)
        y = (1.5 * X[:, 0] - X[:, 1] + 0.5 * X[:, 1] * X[:, 2] + 0.5 * np.random.default_rng(24).standard_normal(n)).astype(np.float32)
        p = "objective=regression num_leaves=31 min_data_in_leaf=20 monotone_constraints=1,-1,0,0,0,0,0,0 monotone_penalty=0.5 verbosity=-1" + REF_EXTRA
    ds = ref.dataset_from_mat(X, p)
    ref.set_field(ds, "label", y)
    b = ref.booster_create(ds, p)
    for _ in range(iters):
        ref.update(b)
    stock = ref.model_string(b)
    swapped = shim_train(ds, p, iters)
    same, n_trees, first = capi.compare_models(stock, swapped, rtol_value=H.VALUE_RTOL)
    assert n_trees == iters and same == n_trees, f"{same}/{n_trees} trees identical; first mismatch: {first}"
    if kind == "categorical":
        assert any(int(t["num_cat"][0]) > 0 for t in capi.parse_trees(swapped))
