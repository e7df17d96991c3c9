"""Boundary proof: the reference's OWN python package (python-package/lightgbm, unmodified, read
from /root/reference -- absent on the GPU box, where this test is skipped) bound to lib_lightgbm_b200.so instead of
lib_lightgbm.so.  lgb.train on a numpy matrix must run through every LGBM_* call basic.py / engine.py make and give the
model the reference library gives.  The CPU half checks that every symbol that path touches is exported."""
import ctypes as C
import importlib
import os
import re
import sys

import numpy as np
import pytest

import capi
import helpers as H

PKG = os.environ.get("LGBM_REF_PYTHON_PACKAGE", "/root/reference/python-package")   # (absent on the GPU box)


def _train_path_symbols():
    """LGBM_* names used by the functions lgb.train(params, lgb.Dataset(X, y)) executes (found by tracing the package
    on the reference library once; kept as a list so that the check needs no reference tree)."""
    return ["LGBM_GetLastError", "LGBM_RegisterLogCallback", "LGBM_DatasetCreateFromMat", "LGBM_DatasetCreateFromCSR", "LGBM_DatasetCreateFromCSC",
            "LGBM_DatasetGetSubset", "LGBM_BoosterPredictForCSR", "LGBM_DatasetSetField", "LGBM_DatasetGetField",
            "LGBM_DatasetSetFeatureNames", "LGBM_DatasetGetFeatureNames", "LGBM_DatasetUpdateParamChecking",
            "LGBM_DatasetGetNumData", "LGBM_DatasetGetNumFeature", "LGBM_DatasetFree", "LGBM_BoosterCreate",
            "LGBM_BoosterFree", "LGBM_BoosterAddValidData", "LGBM_BoosterGetNumClasses", "LGBM_BoosterGetEvalCounts",
            "LGBM_BoosterGetEvalNames", "LGBM_BoosterGetEval", "LGBM_BoosterUpdateOneIter", "LGBM_BoosterUpdateOneIterCustom",
            "LGBM_BoosterRollbackOneIter", "LGBM_BoosterGetCurrentIteration", "LGBM_BoosterNumModelPerIteration",
            "LGBM_BoosterNumberOfTotalModel", "LGBM_BoosterResetParameter", "LGBM_BoosterSaveModelToString",
            "LGBM_BoosterLoadModelFromString", "LGBM_BoosterGetNumFeature", "LGBM_BoosterGetFeatureNames",
            "LGBM_BoosterFeatureImportance", "LGBM_BoosterPredictForMat", "LGBM_BoosterCalcNumPredict",
            "LGBM_BoosterGetPredict", "LGBM_BoosterGetNumPredict", "LGBM_BoosterSaveModel", "LGBM_NetworkInit", "LGBM_NetworkFree"]


def test_every_symbol_of_the_python_train_path_is_exported():
    lib = C.CDLL(capi.OURS_SO)
    missing = [n for n in _train_path_symbols() if not hasattr(lib, n)]
    assert not missing, missing


def test_loaded_model_round_trip_and_prediction_cpu():
    """LGBM_BoosterLoadModelFromString on a model the REFERENCE trained: same predictions as the reference, same text back"""
    if not H.have_ref():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(3)
    X = rng.standard_normal((4000, 8)); X[::9, 2] = np.nan
    y = (np.nan_to_num(X[:, 0]) + X[:, 1] ** 2 > 1).astype(np.float32)
    ref, ours = capi.reference(), capi.ours()
    _, br, _ = capi.train(ref, X, y, "objective=binary num_leaves=15 verbosity=-1 num_threads=2", 6)
    text = ref.model_string(br)
    bo = ours.load_model_from_string(text)
    for raw in (True, False):
        np.testing.assert_allclose(ours.predict_mat(bo, X, raw=raw), ref.predict_mat(br, X, raw=raw), rtol=0, atol=1e-15)
    assert ours.model_string(bo) == text
    n = C.c_int()
    ours.check(ours.lib.LGBM_BoosterGetNumFeature(bo, C.byref(n))); assert n.value == 8
    imp_o = np.zeros(8); imp_r = np.zeros(8)
    ours.check(ours.lib.LGBM_BoosterFeatureImportance(bo, C.c_int(-1), C.c_int(1), imp_o.ctypes.data_as(C.POINTER(C.c_double))))
    ref.check(ref.lib.LGBM_BoosterFeatureImportance(br, C.c_int(-1), C.c_int(1), imp_r.ctypes.data_as(C.POINTER(C.c_double))))
    np.testing.assert_allclose(imp_o, imp_r, rtol=1e-6)   # the text keeps the gains with 6 significant digits
    with pytest.raises(capi.LGBMError, match="loaded from a model string"):
        ours.update(bo)
    with pytest.raises(capi.LGBMError, match="LGBM_NetworkInitNCCL"):
        ours.check(ours.lib.LGBM_NetworkInit(C.c_char_p(b"127.0.0.1:1,127.0.0.1:2"), C.c_int(1), C.c_int(1), C.c_int(2)))


def _package_bound_to(so_path, name):
    """import the reference's python package (its files, unmodified) with its library lookup pointed at `so_path`:
    lightgbm/libpath.py is loaded first and its find_lib_path replaced, then the package's own __init__ runs"""
    import importlib.util
    pkg_dir = os.path.join(PKG, "lightgbm")
    if not os.path.isdir(pkg_dir):
        pytest.skip("reference python package not present")
    for k in [k for k in sys.modules if k == "lightgbm" or k.startswith("lightgbm.")]:
        del sys.modules[k]
    spec = importlib.util.spec_from_file_location("lightgbm", os.path.join(pkg_dir, "__init__.py"),
                                                  submodule_search_locations=[pkg_dir])
    pkg = importlib.util.module_from_spec(spec)
    sys.modules["lightgbm"] = pkg
    lspec = importlib.util.spec_from_file_location("lightgbm.libpath", os.path.join(pkg_dir, "libpath.py"))
    libpath = importlib.util.module_from_spec(lspec)
    lspec.loader.exec_module(libpath)
    libpath.find_lib_path = lambda: [so_path]
    sys.modules["lightgbm.libpath"] = libpath
    spec.loader.exec_module(pkg)
    return pkg


@pytest.mark.gpu
def test_lgb_train_through_the_reference_python_package():
    if not H.have_ref():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(5)
    X = rng.standard_normal((30_000, 10)).astype(np.float32)
    y = (X[:, 0] * X[:, 1] + X[:, 2] > 0).astype(np.float32)
    Xv, yv = X[:5000], y[:5000]
    params = {"objective": "binary", "num_leaves": 31, "learning_rate": 0.1, "metric": ["binary_logloss", "auc"], "verbosity": -1}
    out = []
    for so, extra in ((capi.REF_SO, {"force_col_wise": True, "num_threads": 4}), (capi.OURS_SO, {})):
        lgb = _package_bound_to(so, so)
        dtrain = lgb.Dataset(X, label=y, feature_name=[f"f{i}" for i in range(10)])
        dvalid = lgb.Dataset(Xv, label=yv, reference=dtrain)
        evals = {}
        bst = lgb.train({**params, **extra}, dtrain, num_boost_round=12, valid_sets=[dvalid], valid_names=["v"],
                        evals_result=evals, verbose_eval=False)
        out.append((bst.model_to_string(), bst.predict(Xv), evals, bst.feature_importance("gain"), bst.feature_name()))
    (mr, pr, er, ir, nr), (mo, po, eo, io_, no) = out
    same, n, first = capi.compare_models(mr, mo, rtol_value=H.VALUE_RTOL)
    assert n == 12 and same == n, first
    np.testing.assert_allclose(po, pr, rtol=0, atol=1e-9)
    assert nr == no
    np.testing.assert_allclose(io_, ir, rtol=1e-6)
    for k in ("binary_logloss", "auc"):
        np.testing.assert_allclose(eo["v"][k], er["v"][k], rtol=1e-9)


def test_loaded_multiclass_model_predicts_like_the_reference_cpu():
    """a multiclass model the REFERENCE trained, loaded here from its text: K trees per iteration, softmax outputs"""
    if not H.have_ref():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(5)
    X = rng.standard_normal((3000, 6))
    y = (np.argmax(X[:, :3] + 0.5 * rng.standard_normal((3000, 3)), axis=1)).astype(np.float32)
    ref, ours = capi.reference(), capi.ours()
    _, br, _ = capi.train(ref, X, y, "objective=multiclass num_class=3 num_leaves=7 verbosity=-1 num_threads=2", 5)
    text = ref.model_string(br)
    bo = ours.load_model_from_string(text)
    for raw in (True, False):
        np.testing.assert_allclose(ours.predict_mat(bo, X, raw=raw, num_class=3), ref.predict_mat(br, X, raw=raw, num_class=3),
                                   rtol=0, atol=1e-15)
    k = C.c_int(0)
    ours.check(ours.lib.LGBM_BoosterGetNumClasses(bo, C.byref(k))); assert k.value == 3
    ours.check(ours.lib.LGBM_BoosterGetCurrentIteration(bo, C.byref(k))); assert k.value == 5
    assert ours.model_string(bo) == text


def test_subset_and_leaf_index_prediction_cpu():
    """LGBM_DatasetGetSubset (python's Dataset.subset / lgb.cv) and predict_type = leaf index on a loaded model: CPU-only
    checks against the reference -- the subset's bins equal the reference's subset's, the leaf indices equal its."""
    if not (H.have_ref() and os.path.exists(H.REF_SHIM_SO)):
        pytest.skip("oracle/_ref not built")
    os.environ["LGBM_B200_HOST_BINNING"] = "1"
    try:
        C.CDLL(H.REF_SO, mode=C.RTLD_GLOBAL); shim = C.CDLL(H.REF_SHIM_SO)
        rng = np.random.default_rng(8)
        X = rng.standard_normal((5000, 9)); X[::7, 3] = np.nan; X[:, 5] = rng.integers(0, 6, size=5000)
        y = (np.nan_to_num(X[:, 0]) + X[:, 1] ** 2 > 1).astype(np.float32)
        w = (rng.random(5000) + 0.5).astype(np.float32)
        idx = np.sort(rng.choice(5000, size=1700, replace=False)).astype(np.int32)
        ref, ours = capi.reference(), capi.ours()
        p = "max_bin=63 categorical_feature=5 verbosity=-1"
        subs = {}
        for name, lib in (("ref", ref), ("ours", ours)):
            ds = lib.dataset_from_mat(X, p)
            lib.set_field(ds, "label", y); lib.set_field(ds, "weight", w)
            sub = C.c_void_p()
            lib.check(lib.lib.LGBM_DatasetGetSubset(ds, idx.ctypes.data_as(C.POINTER(C.c_int32)), C.c_int32(len(idx)), C.c_char_p(p.encode()), C.byref(sub)))
            n = C.c_int(); lib.check(lib.lib.LGBM_DatasetGetNumData(sub, C.byref(n))); assert n.value == len(idx)
            subs[name] = sub
        nf = shim.ref_num_inner_features(subs["ref"])
        bins = np.empty((len(idx), nf), dtype=np.int32); real = np.empty(nf, dtype=np.int32)
        ours.check(ours.lib.LGBM_B200_DatasetDumpBins(subs["ours"], bins.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
        for f in range(nf):
            want = np.empty(len(idx), dtype=np.int32)
            shim.ref_feature_bin_column(subs["ref"], C.c_int(f), want.ctypes.data_as(C.POINTER(C.c_int32)))
            assert np.array_equal(bins[:, f], want), f"inner feature {f}: subset bins differ"
        # the subset's label came along
        out_len = C.c_int(); out_ptr = C.c_void_p(); out_type = C.c_int()
        ours.check(ours.lib.LGBM_DatasetGetField(subs["ours"], C.c_char_p(b"label"), C.byref(out_len), C.byref(out_ptr), C.byref(out_type)))
        got = np.ctypeslib.as_array(C.cast(out_ptr, C.POINTER(C.c_float)), shape=(out_len.value,))
        assert np.array_equal(got, y[idx])
        # leaf indices from a model the reference trained
        _, br, _ = capi.train(ref, X, y, "objective=binary num_leaves=15 categorical_feature=5 verbosity=-1 num_threads=2", 5)
        bo = ours.load_model_from_string(ref.model_string(br))
        def leaves(lib, b):
            out = np.empty(len(X) * 5, dtype=np.float64); n = C.c_int64()
            lib.check(lib.lib.LGBM_BoosterPredictForMat(b, X.ctypes.data_as(C.c_void_p), C.c_int(1), C.c_int32(X.shape[0]), C.c_int32(X.shape[1]),
                                                        C.c_int(1), C.c_int(2), C.c_int(0), C.c_int(-1), C.c_char_p(b""), C.byref(n),
                                                        out.ctypes.data_as(C.POINTER(C.c_double))))
            assert n.value == len(out)
            return out
        assert np.array_equal(leaves(ref, br), leaves(ours, bo))
    finally:
        os.environ.pop("LGBM_B200_HOST_BINNING", None)


def test_dataset_parameter_errors_cpu():
    """Dataset-level parameters are checked where the reference checks them (no device needed to build a Dataset)"""
    ours = capi.ours()
    X = np.random.default_rng(1).standard_normal((500, 4))
    with pytest.raises(capi.LGBMError, match="max_bin_by_feature"):
        ours.dataset_from_mat(X, "max_bin_by_feature=10,20 verbosity=-1")            # one entry per column is needed
    with pytest.raises(capi.LGBMError, match="max_bin_by_feature"):
        ours.dataset_from_mat(X, "max_bin_by_feature=10,20,1,5 verbosity=-1")        # every entry > 1
    with pytest.raises(capi.LGBMError, match="column name"):
        ours.dataset_from_mat(X, "categorical_feature=name:a,b verbosity=-1")        # names need a file header
    with pytest.raises(capi.LGBMError, match="max_bin"):
        ours.dataset_from_mat(X, "max_bin=70000 verbosity=-1")
    ds = ours.dataset_from_mat(np.floor(np.abs(X) * 3), "categorical_feature=0,2 max_bin_by_feature=4,8,16,32 verbosity=-1")
    n = C.c_int(); ours.check(ours.lib.LGBM_DatasetGetNumFeature(ds, C.byref(n))); assert n.value == 4
    sub = C.c_void_p()
    bad = np.array([3, 2, 1], dtype=np.int32)
    with pytest.raises(capi.LGBMError, match="sorted"):
        ours.check(ours.lib.LGBM_DatasetGetSubset(ds, bad.ctypes.data_as(C.POINTER(C.c_int32)), C.c_int32(3), C.c_char_p(b""), C.byref(sub)))


def test_dataset_parameters_cannot_change_after_construction_cpu():
    ours = capi.ours()
    ok = lambda old, new: ours.lib.LGBM_DatasetUpdateParamChecking(C.c_char_p(old.encode()), C.c_char_p(new.encode()))
    assert ok("max_bin=255 categorical_feature=1,2", "num_leaves=63 learning_rate=0.2") == 0
    for old, new, what in (("max_bin=255", "max_bin=63", "max_bin"), ("categorical_feature=1,2", "categorical_feature=1", "categorical_feature"),
                           ("max_bin_by_feature=4,8", "max_bin_by_feature=4,9", "max_bin_by_feature"), ("", "zero_as_missing=true", "zero_as_missing")):
        assert ok(old, new) != 0
        assert what in ours.lib.LGBM_GetLastError().decode()
