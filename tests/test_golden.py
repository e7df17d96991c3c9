"""Committed golden vectors (tests/golden/golden_v1.npz, generated from the unmodified reference by
tests/golden/make_golden.py).  CPU part: product binning and the oracle's tree growth against the fixture (no
oracle/_ref needed).  GPU part: short training runs through the C API against the reference's model text."""
import ctypes as C
import os

import numpy as np
import pytest

import capi
import helpers as H
from lightgbmv1_b200 import _native as N

G = np.load(os.path.join(H.ROOT, "tests", "golden", "golden_v1.npz"))


def _bounds_split(metas, flat):
    out, pos = [], 0
    for m in metas:
        out.append(flat[pos:pos + m[0]]); pos += m[0]
    return out


def test_binning_matches_golden():
    ours = capi.ours()
    X = G["X"]
    ds = ours.dataset_from_mat(X, str(G["bin_params"]))
    metas = G["bin_metas"]
    nf = C.c_int(); ours.check(ours.lib.LGBM_B200_DatasetNumInnerFeatures(ds, C.byref(nf)))
    assert nf.value == len(metas)
    bounds = _bounds_split(metas, G["bin_bounds"])
    bins = np.empty((X.shape[0], nf.value), dtype=np.int32); real = np.empty(nf.value, dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(ds, bins.ctypes.data_as(C.POINTER(C.c_int32)),
                                                  real.ctypes.data_as(C.POINTER(C.c_int32))))
    for f in range(nf.value):
        m = (C.c_int32 * 5)(); b = (C.c_double * 300)()
        ours.check(ours.lib.LGBM_B200_DatasetFeatureInfo(ds, C.c_int(f), m, b, C.c_int(300)))
        assert list(m) == list(metas[f])
        assert np.array_equal(np.array(b[:m[0]]), bounds[f], equal_nan=True)
        stored = G["bin_stored"][:, f].astype(np.int32)
        want = np.where(stored == 0, m[1], stored - 1 + (1 if m[1] == 0 else 0))
        assert np.array_equal(bins[:, f], want)


def _tree_arrays(text):
    return capi.parse_trees(text)[0]


def test_oracle_tree_matches_golden_tree():
    orc = H.load_oracle()
    metas = [tuple(m) for m in G["bin_metas"]]
    bins = np.ascontiguousarray(G["bin_stored"])
    bounds = _bounds_split(G["bin_metas"], G["bin_bounds"])
    g, h = G["tree_g"], G["tree_h"]
    tree = _tree_arrays(str(G["tree_model"]))
    tp = N.make_tree_params(num_leaves=15, min_data_in_leaf=20)
    fm = N.make_feature_meta(metas)
    recs = (N.SplitRecord * 14)()
    ns = orc.orc_train_tree(H.fptr(bins, C.c_uint8), C.c_int(bins.shape[0]), C.c_int(bins.shape[1]), fm, C.byref(tp),
                            H.fptr(g, C.c_float), H.fptr(h, C.c_float), None, C.c_int(0), None, recs, None)
    assert ns == int(tree["num_leaves"][0]) - 1
    for i in range(ns):
        assert recs[i].real_feature == int(tree["split_feature"][i])
        assert bounds[recs[i].feature][recs[i].threshold] == float(tree["threshold"][i])
        assert recs[i].left_count + recs[i].right_count == int(tree["internal_count"][i])
    lv = {0: 0.0}
    for i in range(ns):
        lv[recs[i].leaf] = recs[i].left_output; lv[recs[i].right_leaf] = recs[i].right_output
    for leaf in range(ns + 1):
        assert lv[leaf] == float(tree["leaf_value"][leaf])


@pytest.mark.gpu
def test_gpu_tree_matches_golden_tree():
    metas = [tuple(m) for m in G["bin_metas"]]
    bins = np.ascontiguousarray(G["bin_stored"])
    tree = _tree_arrays(str(G["tree_model"]))
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=15, min_data_in_leaf=20))
    L.set_gradients(G["tree_g"], G["tree_h"])
    L.begin_tree()
    recs = L.train_tree()
    assert len(recs) == int(tree["num_leaves"][0]) - 1
    bounds = _bounds_split(G["bin_metas"], G["bin_bounds"])
    for i, r in enumerate(recs):
        assert r.real_feature == int(tree["split_feature"][i])
        assert bounds[r.feature][r.threshold] == float(tree["threshold"][i])
        assert r.left_count + r.right_count == int(tree["internal_count"][i])
    L.close()


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["binary", "rank"])
def test_gpu_training_matches_golden_models(which):
    ours = capi.ours()
    X = G["X"]
    params = str(G[f"{which}_params"])
    if which == "binary":
        ds, b, done = capi.train(ours, X, G["y"], params, 6)
    else:
        ds, b, done = capi.train(ours, X, G["rank_label"], params, 5, group=G["rank_sizes"])
    text = ours.model_string(b)
    same, n, first = capi.compare_models(str(G[f"{which}_model"]), text)
    assert same == n and n == (6 if which == "binary" else 5), first
    pred = ours.get_predict(b, X.shape[0])
    assert np.max(np.abs(pred - G[f"{which}_pred"])) <= 1e-6
