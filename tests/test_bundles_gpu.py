"""Exclusive feature bundling (enable_bundle, on by default): the reference stores several sparse features in one column;
this library keeps a column per feature and reproduces what the shared column does to the data (rows where two members
are set keep only the last pushed value).  Bundled data must train to the reference's trees.
"""
import ctypes as C
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import assert_same_model, run_both
from test_oracle_vs_reference import sparse_matrix

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


def dump(lib, ds, n):
    nf = C.c_int(); lib.check(lib.lib.LGBM_B200_DatasetNumInnerFeatures(ds, C.byref(nf)))
    bins = np.empty((n, nf.value), dtype=np.int32); real = np.empty(nf.value, dtype=np.int32)
    lib.check(lib.lib.LGBM_B200_DatasetDumpBins(ds, bins.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
    return bins, real


@pytest.mark.parametrize("order", ["C", "F"])
def test_device_binning_applies_the_bundle_rule(order, monkeypatch):
    X = sparse_matrix(1, 30_000)
    ours = capi.ours()
    params = "max_bin=255 verbosity=-1"
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "1")
    bh, rh = dump(ours, ours.dataset_from_mat(X, params), len(X))
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "0")
    Xd = np.asfortranarray(X) if order == "F" else X
    bd, rd = dump(ours, ours.dataset_from_mat(Xd, params), len(X))
    assert np.array_equal(rh, rd)
    assert np.array_equal(bh, bd)
    # and the rule changed something (the host form is pinned against the reference in test_oracle_vs_reference.py)
    bn, rn = dump(ours, ours.dataset_from_mat(X, params + " enable_bundle=false"), len(X))
    by_real = {int(r): i for i, r in enumerate(rn)}
    assert sum(int(np.sum(bd[:, f] != bn[:, by_real[int(rd[f])]])) for f in range(len(rd))) > 0


def target(X, rng):
    Z = np.nan_to_num(X).astype(np.float64)
    w = np.random.default_rng(99).standard_normal(X.shape[1])
    return Z @ w + 0.5 * Z[:, 0] * Z[:, 3] + 0.3 * rng.standard_normal(len(X))


@pytest.mark.parametrize("seed,n,params", [
    (1, 30_000, "objective=binary num_leaves=31 min_data_in_leaf=20 learning_rate=0.1 verbosity=-1"),
    (3, 25_000, "objective=binary num_leaves=63 min_data_in_leaf=5 zero_as_missing=true verbosity=-1"),
    (2, 42_000, "objective=regression num_leaves=31 max_bin=63 min_data_in_bin=1 bagging_fraction=0.8 bagging_freq=1 verbosity=-1"),
])
def test_identical_trees_on_bundled_data(seed, n, params):
    X = sparse_matrix(seed, n)
    z = target(X, np.random.default_rng(seed + 100))
    y = (z > np.median(z)).astype(np.float32) if "binary" in params else z.astype(np.float32)
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10)
    assert nr == no == 10
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, n), ours.get_predict(bo, n)
    assert np.max(np.abs(pr - po)) <= 1e-9


def test_validation_set_beside_bundled_training_data():
    # an aligned dataset is NOT bundled in the reference (Dataset::CreateValid): its conflicting rows keep both values
    A = sparse_matrix(1, 39_000)
    X, Xv = np.ascontiguousarray(A[:30_000]), np.ascontiguousarray(A[30_000:])
    rng = np.random.default_rng(5)
    z = target(X, rng); zv = target(Xv, rng)
    y = (z > np.median(z)).astype(np.float32); yv = (zv > np.median(z)).astype(np.float32)
    params = "objective=binary metric=binary_logloss num_leaves=31 verbosity=-1"
    out = {}
    for name, lib, extra in (("ref", capi.reference(), " force_col_wise=true num_threads=4"), ("ours", capi.ours(), "")):
        ds, b, _ = capi.train(lib, X, y, params + extra, 0)
        dv = lib.dataset_from_mat(Xv, params + extra, reference=ds)
        lib.set_field(dv, "label", yv)
        lib.add_valid(b, dv)
        for _ in range(8):
            lib.update(b)
        out[name] = lib.get_eval(b, 1)[0]
    assert abs(out["ref"] - out["ours"]) <= 1e-9


def test_training_from_csr_and_csr_prediction():
    """LGBM_DatasetCreateFromCSR / LGBM_BoosterPredictForCSR (densified on the host, then the dense path): the reference
    trained from the same CSR matrix gives the same trees; predictions from CSR equal predictions from the dense matrix."""
    sp = pytest.importorskip("scipy.sparse")
    n = 30_000
    X = np.nan_to_num(sparse_matrix(6, n, dtype=np.float64))
    z = target(X, np.random.default_rng(61))
    y = (z > np.median(z)).astype(np.float32)
    S = sp.csr_matrix(X)
    params = "objective=binary num_leaves=31 min_data_in_leaf=20 verbosity=-1"
    out = {}
    for name, lib, extra in (("ref", capi.reference(), " force_col_wise=true num_threads=4"), ("ours", capi.ours(), "")):
        ds = lib.dataset_from_sparse(S, params + extra)
        lib.set_field(ds, "label", y)
        b = lib.booster_create(ds, params + extra)
        for _ in range(8):
            lib.update(b)
        out[name] = (lib, b)
    (ref, br), (ours, bo) = out["ref"], out["ours"]
    assert_same_model(ref, ours, br, bo)
    Sv = sp.csr_matrix(np.nan_to_num(sparse_matrix(7, 5000, dtype=np.float64)))
    a = ref.predict_csr(br, Sv); b = ours.predict_csr(bo, Sv); c = ours.predict_mat(bo, Sv.toarray())
    assert np.max(np.abs(a - b)) <= 1e-12
    assert np.array_equal(b, c)
