"""CPU-side checks: the C-ABI library loads, exports every symbol include/*.h declares, reports errors the
reference's way, and fails loudly (no fallback) when no CUDA device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import capi
import helpers as H
from lightgbmv1_b200 import _native as N


def _declared(header, prefix):
    text = open(os.path.join(H.ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # declarations only, not names mentioned in comments
    return sorted(set(re.findall(r"\b(" + prefix + r"[A-Za-z0-9_]+)\s*\(", text)))


def test_every_declared_lgbcu_symbol_is_exported():
    names = _declared("lgbcu.h", "lgbcu_")
    assert len(names) >= 35
    assert sorted(names) == sorted(N.LGBCU_SYMBOLS)
    for n in names:
        assert hasattr(N.lib, n), n


def test_every_declared_lgbm_symbol_is_exported():
    names = _declared("lgbm_b200.h", "LGBM_")
    assert "LGBM_DatasetCreateFromMat" in names and "LGBM_BoosterCreate" in names and "LGBM_BoosterUpdateOneIter" in names
    lib = C.CDLL(capi.OURS_SO)
    for n in names:
        assert hasattr(lib, n), n


def test_pod_layouts_match_the_header():
    assert C.sizeof(N.FeatureMeta) == 32
    assert C.sizeof(N.TreeParams) == 64
    assert C.sizeof(N.SplitInfo) == 80
    assert C.sizeof(N.SplitRecord) == 88


def test_dataset_side_needs_no_gpu_and_reports_errors():
    ours = capi.ours()
    X = np.random.default_rng(0).standard_normal((300, 5)).astype(np.float32)
    ds = ours.dataset_from_mat(X, "max_bin=31 verbosity=-1")
    n = C.c_int()
    ours.check(ours.lib.LGBM_DatasetGetNumData(ds, C.byref(n)))
    assert n.value == 300
    ours.check(ours.lib.LGBM_DatasetGetNumFeature(ds, C.byref(n)))
    assert n.value == 5
    with pytest.raises(capi.LGBMError, match="Length of label"):
        ours.set_field(ds, "label", np.zeros(7, np.float32))
    with pytest.raises(capi.LGBMError):
        ours.dataset_from_mat(X, "num_leaves=1")
    ours.dataset_free(ds)


@pytest.mark.skipif(N.device_count() > 0, reason="checks the no-GPU behaviour")
def test_no_gpu_means_loud_failure_not_fallback():
    with pytest.raises(N.LgbCuError, match="no CUDA device"):
        N.CudaLearner()
    ours = capi.ours()
    X = np.random.default_rng(0).standard_normal((300, 5)).astype(np.float32)
    ds = ours.dataset_from_mat(X, "verbosity=-1")
    ours.set_field(ds, "label", (X[:, 0] > 0).astype(np.float32))
    with pytest.raises(capi.LGBMError, match="no CUDA device"):
        ours.booster_create(ds, "objective=binary verbosity=-1")


def test_config_aliases_and_seed_fanout_match_reference():
    # same parameter string => same binning decisions: proves alias handling + seed fan-out (config.cpp:190-200)
    if not H.have_ref():
        pytest.skip("oracle/_ref not built")
    C.CDLL(H.REF_SO, mode=C.RTLD_GLOBAL)
    shim = C.CDLL(H.REF_SHIM_SO)
    ref, ours = capi.reference(), capi.ours()
    X = np.random.default_rng(1).standard_normal((5000, 6)).astype(np.float32)
    p = "seed=17 subsample_for_bin=1000 min_data_per_leaf=5 verbose=-1"
    dr, do = ref.dataset_from_mat(X, p), ours.dataset_from_mat(X, p)
    for f in range(6):
        mr = (C.c_int32 * 5)(); br = (C.c_double * 300)(); mo = (C.c_int32 * 5)(); bo = (C.c_double * 300)()
        shim.ref_feature_info(dr, C.c_int(f), mr, br, C.c_int(300))
        ours.check(ours.lib.LGBM_B200_DatasetFeatureInfo(do, C.c_int(f), mo, bo, C.c_int(300)))
        assert list(mr) == list(mo)
        assert np.array_equal(np.array(br[:mr[0]]), np.array(bo[:mo[0]]), equal_nan=True)
