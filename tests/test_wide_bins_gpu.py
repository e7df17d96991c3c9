"""max_bin > 255: features with more than 256 stored values (the reference's 16-bit DenseBin, src/io/bin.cpp:635-645).

The binned matrix keeps such a feature in two adjacent byte columns (low byte first); histograms come from
k_hist_wide, the split search from find_item_wide, the partition from k_partition<true>.  Same bar as every other
case: the reference on the CPU and this library on the GPU give identical trees.
"""
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import assert_same_model, run_both

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]


def wide_matrix(seed, n, F=10, dtype=np.float32):
    """continuous columns (many distinct values -> max_bin bins), some with NaN, some sparse, some discrete"""
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, F)).astype(dtype)
    X[:, 1] = np.where(rng.random(n) < 0.08, np.nan, X[:, 1])          # NaN missing, wide
    X[:, 2] = np.where(rng.random(n) < 0.6, 0.0, X[:, 2])              # mostly zero: most_freq_bin != 0, wide
    X[:, 3] = rng.integers(0, 5, size=n)                               # narrow (5 bins) beside wide ones
    X[:, 4] = np.exp(X[:, 4])                                          # positive only
    X[:, 5] = rng.integers(0, 300, size=n)                             # just over one byte of bins
    w = np.random.default_rng(seed + 1).standard_normal(F)
    z = np.nan_to_num(X).astype(np.float64) @ w * 0.3 + np.sin(3 * np.nan_to_num(X[:, 0])) + 0.3 * rng.standard_normal(n)
    return X, z


def dump(lib, ds, n):
    import ctypes as C
    nf = C.c_int(); lib.check(lib.lib.LGBM_B200_DatasetNumInnerFeatures(ds, C.byref(nf)))
    bins = np.empty((n, nf.value), dtype=np.int32); real = np.empty(nf.value, dtype=np.int32)
    lib.check(lib.lib.LGBM_B200_DatasetDumpBins(ds, bins.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
    return bins, real


@pytest.mark.parametrize("max_bin", [1000, 4000])
def test_device_binning_of_wide_features_equals_host_binning(max_bin, monkeypatch):
    X, _ = wide_matrix(3, 120_001)
    ours = capi.ours()
    params = f"max_bin={max_bin} min_data_in_bin=1 verbosity=-1"
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "1")
    ds_h = ours.dataset_from_mat(X, params)
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "0")
    ds_d = ours.dataset_from_mat(X, params)
    bh, rh = dump(ours, ds_h, len(X))
    bd, rd = dump(ours, ds_d, len(X))
    assert np.array_equal(rh, rd)
    assert bh.max() > 255          # the case is what it says it is
    assert np.array_equal(bh, bd)


@pytest.mark.parametrize("max_bin,extra", [(1000, ""), (511, " zero_as_missing=true"), (3000, " use_missing=false min_data_in_bin=1")])
def test_binary_identical_trees_with_wide_bins(max_bin, extra):
    X, z = wide_matrix(11, 50_000)
    y = (z > np.median(z)).astype(np.float32)
    params = f"objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 max_bin={max_bin}{extra} verbosity=-1"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 8)
    assert nr == no == 8
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, len(y)), ours.get_predict(bo, len(y))
    assert np.max(np.abs(pr - po)) <= 1e-9


def test_regression_identical_trees_with_wide_bins_and_bagging():
    X, z = wide_matrix(12, 40_000, dtype=np.float64)
    y = z.astype(np.float32)
    params = ("objective=regression num_leaves=63 learning_rate=0.1 min_data_in_leaf=5 max_bin=2000 bagging_fraction=0.7 "
              "bagging_freq=1 lambda_l2=0.5 verbosity=-1")
    ref, ours, br, bo, nr, no = run_both(X, y, params, 6)
    assert_same_model(ref, ours, br, bo)


def test_valid_metric_and_prediction_with_wide_bins():
    X, z = wide_matrix(13, 30_000)
    y = (z > np.median(z)).astype(np.float32)
    Xv, zv = wide_matrix(14, 8_000)
    yv = (zv > np.median(z)).astype(np.float32)
    params = "objective=binary metric=binary_logloss num_leaves=31 max_bin=1500 verbosity=-1"
    ref, ours = capi.reference(), capi.ours()
    out = {}
    for name, lib, extra in (("ref", ref, " force_col_wise=true num_threads=4"), ("ours", ours, "")):
        ds, b, _ = capi.train(lib, X, y, params + extra, 0)
        dv = lib.dataset_from_mat(Xv, params + extra, reference=ds)
        lib.set_field(dv, "label", yv)
        lib.add_valid(b, dv)
        for _ in range(6):
            lib.update(b)
        out[name] = (lib.get_eval(b, 1), lib.predict_mat(b, Xv))
    assert abs(out["ref"][0][0] - out["ours"][0][0]) <= 1e-9
    assert np.max(np.abs(out["ref"][1] - out["ours"][1])) <= 1e-12
