"""On-device binning (lgbcu_bin_matrix, reached through LGBM_DatasetCreateFromMat) must be bit-identical to the host
map (BinMapper::ValueToBin per element, forced with LGBM_B200_HOST_BINNING=1) -- and through it, by the CPU tests of
tests/test_oracle_vs_reference.py, to the reference's own BinIterator::RawGet."""
import ctypes as C
import os

import numpy as np
import pytest

import capi

pytestmark = pytest.mark.gpu


def mixed_matrix(rng, n, ncol, dtype):
    X = rng.standard_normal((n, ncol)).astype(dtype)
    X[:, 1] = np.maximum(0, X[:, 1])                       # half zeros
    X[:, 2] = rng.integers(0, 3, size=n)                   # three levels
    X[rng.random(n) < 0.07, 3] = np.nan                    # NaN missing
    X[:, 4] = np.where(rng.random(n) < 0.9, 0.0, X[:, 4])  # sparse
    X[:, 5] = 1.5                                          # constant -> dropped
    X[:, 6] = np.round(X[:, 6] * 4) / 4                    # few distinct values, ties on the boundaries
    X[rng.random(n) < 0.01, 7] = np.inf
    return X


def dump(lib, ds, n):
    nf = C.c_int(); lib.check(lib.lib.LGBM_B200_DatasetNumInnerFeatures(ds, C.byref(nf)))
    bins = np.empty((n, nf.value), dtype=np.int32); real = np.empty(nf.value, dtype=np.int32)
    lib.check(lib.lib.LGBM_B200_DatasetDumpBins(ds, bins.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
    return bins, real


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("params", ["max_bin=255 verbosity=-1", "max_bin=63 zero_as_missing=true verbosity=-1",
                                    "max_bin=15 use_missing=false min_data_in_bin=1 verbosity=-1"])
def test_device_bins_equal_host_bins(dtype, params, monkeypatch):
    rng = np.random.default_rng(17)
    n, ncol = 150_003, 12
    X = mixed_matrix(rng, n, ncol, dtype)
    ours = capi.ours()
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "1")
    d_host = ours.dataset_from_mat(X, params)
    monkeypatch.delenv("LGBM_B200_HOST_BINNING")
    d_dev = ours.dataset_from_mat(X, params)
    bh, rh = dump(ours, d_host, n)
    bd, rd = dump(ours, d_dev, n)
    assert np.array_equal(rh, rd)
    assert np.array_equal(bh, bd)
    # a second matrix binned against the first one's mappers (the validation-set / row-shard path)
    X2 = mixed_matrix(np.random.default_rng(18), 20_001, ncol, dtype)
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "1")
    v_host = ours.dataset_from_mat(X2, params, reference=d_host)
    monkeypatch.delenv("LGBM_B200_HOST_BINNING")
    v_dev = ours.dataset_from_mat(X2, params, reference=d_dev)
    assert np.array_equal(dump(ours, v_host, 20_001)[0], dump(ours, v_dev, 20_001)[0])
    for d in (d_host, d_dev, v_host, v_dev):
        ours.dataset_free(d)


def test_column_major_input_and_training_on_the_device_matrix():
    rng = np.random.default_rng(19)
    n, ncol = 60_000, 10
    X = mixed_matrix(rng, n, ncol, np.float32)
    y = (np.nan_to_num(X[:, 0]) + np.nan_to_num(X[:, 3]) > 0).astype(np.float32)
    ours = capi.ours()
    params = "objective=binary num_leaves=31 verbosity=-1"
    models = []
    for col_major in (False, True):
        h = C.c_void_p()
        A = np.asfortranarray(X) if col_major else np.ascontiguousarray(X)
        ours.check(ours.lib.LGBM_DatasetCreateFromMat(A.ctypes.data_as(C.c_void_p), C.c_int(0), C.c_int32(n), C.c_int32(ncol),
                                                      C.c_int(0 if col_major else 1), C.c_char_p(params.encode()), None, C.byref(h)))
        ours.set_field(h, "label", y)
        b = ours.booster_create(h, params)
        for _ in range(5):
            ours.update(b)
        models.append(ours.model_string(b).split("parameters:")[0])
    assert models[0] == models[1]
