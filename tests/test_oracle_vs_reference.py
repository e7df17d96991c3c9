"""Pins the test oracle (oracle/oracle.c) and the product's host-side binning against the UNMODIFIED reference
built by oracle/Makefile (oracle/_ref/lib_lightgbm.so + libref_shim.so).  CPU only.

Everything here is bit-exact: the oracle restates the reference's sequential double/float arithmetic, and the
binning is integer/ordering work on doubles.
"""
import ctypes as C
import os

import numpy as np
import pytest

import capi
import helpers as H
from lightgbmv1_b200 import _native as N

pytestmark = pytest.mark.skipif(not (os.path.exists(H.REF_SO) and os.path.exists(H.REF_SHIM_SO)),
                                reason="oracle/_ref not built (make -C oracle ref)")


@pytest.fixture(scope="module")
def shim():
    C.CDLL(H.REF_SO, mode=C.RTLD_GLOBAL)
    return C.CDLL(H.REF_SHIM_SO)


@pytest.fixture(scope="module")
def orc():
    return H.load_oracle()


@pytest.fixture(scope="module")
def ref():
    return capi.reference()


@pytest.fixture(scope="module")
def ours():
    return capi.ours()


def mixed_matrix(seed, n, dtype=np.float32):
    """Columns that exercise every binning branch: gaussian, half zeros, 3 levels, NaNs, all-negative, constant,
    heavy ties, >70% in one non-zero bin, sparse-ish, integers."""
    rng = np.random.default_rng(seed)
    cols = [
        rng.standard_normal(n),
        np.maximum(0, rng.standard_normal(n)),
        rng.choice([0.0, 1.087, 2.173], size=n),
        np.where(rng.random(n) < 0.1, np.nan, rng.standard_normal(n)),
        -np.abs(rng.standard_normal(n)) - 0.5,
        np.full(n, 3.0),
        np.round(rng.standard_normal(n), 1),
        np.where(rng.random(n) < 0.8, 5.0, rng.standard_normal(n)),
        np.where(rng.random(n) < 0.9, 0.0, rng.random(n)),
        rng.integers(0, 40, size=n).astype(np.float64),
        np.where(rng.random(n) < 0.05, np.nan, np.maximum(0, rng.standard_normal(n))),
        rng.standard_normal(n) * 1e-3,
    ]
    return np.stack(cols, axis=1).astype(dtype)


def feature_infos(lib_handle, fn_num, fn_info, ds):
    nf = fn_num(ds)
    out = []
    for f in range(nf):
        meta = (C.c_int32 * 5)()
        bounds = (C.c_double * 300)()
        fn_info(ds, C.c_int(f), meta, bounds, C.c_int(300))
        out.append((list(meta), np.array(bounds[:meta[0]])))
    return out


@pytest.mark.parametrize("params", [
    "max_bin=255 min_data_in_leaf=20 verbosity=-1",
    "max_bin=63 min_data_in_bin=1 min_data_in_leaf=1 verbosity=-1",
    "max_bin=15 verbosity=-1",
    "max_bin=255 zero_as_missing=true verbosity=-1",
    "max_bin=255 use_missing=false verbosity=-1",
    "max_bin=255 enable_bundle=false verbosity=-1",
    "max_bin=255 max_bin_by_feature=255,7,3,31,100,2,16,200,9,40,63,5 verbosity=-1",
    "max_bin=255 bin_construct_sample_cnt=1500 data_random_seed=7 verbosity=-1",
    "max_bin=255 feature_pre_filter=false min_data_in_leaf=3000 verbosity=-1",
])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_binning_matches_reference(shim, ref, ours, params, dtype):
    n = 6000
    X = mixed_matrix(3, n, dtype)
    dr = ref.dataset_from_mat(X, params)
    do = ours.dataset_from_mat(X, params)
    nf_o = C.c_int()
    ours.check(ours.lib.LGBM_B200_DatasetNumInnerFeatures(do, C.byref(nf_o)))
    nf_r = shim.ref_num_inner_features(dr)
    assert nf_o.value == nf_r
    for f in range(nf_r):
        mr = (C.c_int32 * 5)(); br = (C.c_double * 300)()
        mo = (C.c_int32 * 5)(); bo = (C.c_double * 300)()
        shim.ref_feature_info(dr, C.c_int(f), mr, br, C.c_int(300))
        ours.check(ours.lib.LGBM_B200_DatasetFeatureInfo(do, C.c_int(f), mo, bo, C.c_int(300)))
        assert list(mr) == list(mo), f"inner feature {f}: (num_bin, mfb, default_bin, missing, real) differ"
        a = np.array(br[:mr[0]]); b = np.array(bo[:mo[0]])
        assert np.array_equal(a, b, equal_nan=True), f"inner feature {f}: bin upper bounds differ"
    # bins of every row (BinIterator::Get space) and the stored encoding the device consumes
    bins = np.empty((n, nf_r), dtype=np.int32)
    real = np.empty(nf_r, dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(do, bins.ctypes.data_as(C.POINTER(C.c_int32)),
                                                  real.ctypes.data_as(C.POINTER(C.c_int32))))
    for f in range(nf_r):
        col = np.empty(n, dtype=np.uint8)
        shim.ref_feature_stored_column(dr, C.c_int(f), col.ctypes.data_as(C.POINTER(C.c_uint8)))
        mr = (C.c_int32 * 5)(); br = (C.c_double * 300)()
        shim.ref_feature_info(dr, C.c_int(f), mr, br, C.c_int(300))
        mfb = mr[1]
        want_bin = np.where(col == 0, mfb, col.astype(np.int32) - 1 + (1 if mfb == 0 else 0))
        assert np.array_equal(bins[:, f], want_bin), f"inner feature {f}: binned values differ"
    ref.dataset_free(dr); ours.dataset_free(do)


def test_aligned_dataset_uses_reference_mappers(shim, ref, ours):
    X = mixed_matrix(5, 4000)
    X2 = mixed_matrix(6, 1500)
    p = "max_bin=63 verbosity=-1"
    dr = ref.dataset_from_mat(X, p); dr2 = ref.dataset_from_mat(X2, p, reference=dr)
    do = ours.dataset_from_mat(X, p); do2 = ours.dataset_from_mat(X2, p, reference=do)
    nf = shim.ref_num_inner_features(dr2)
    bins = np.empty((1500, nf), dtype=np.int32); real = np.empty(nf, dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(do2, bins.ctypes.data_as(C.POINTER(C.c_int32)),
                                                  real.ctypes.data_as(C.POINTER(C.c_int32))))
    for f in range(nf):
        col = np.empty(1500, dtype=np.uint8)
        shim.ref_feature_stored_column(dr2, C.c_int(f), col.ctypes.data_as(C.POINTER(C.c_uint8)))
        mr = (C.c_int32 * 5)(); br = (C.c_double * 300)()
        shim.ref_feature_info(dr2, C.c_int(f), mr, br, C.c_int(300))
        want = np.where(col == 0, mr[1], col.astype(np.int32) - 1 + (1 if mr[1] == 0 else 0))
        assert np.array_equal(bins[:, f], want)


def _case(seed, n, F, kinds=None, max_bin=255):
    rng = np.random.default_rng(seed)
    metas = H.random_feature_metas(rng, F, max_bin=max_bin, kinds=kinds)
    bins = H.random_bins(rng, n, metas)
    g = rng.standard_normal(n).astype(np.float32)
    h = (rng.random(n) * 0.25 + 0.01).astype(np.float32)
    return rng, metas, bins, g, h


def test_oracle_histogram_bit_exact(shim, orc):
    n, F = 20_000, 9
    rng, metas, bins, g, h = _case(1, n, F)
    fm = N.make_feature_meta(metas)
    bases, total = H.hist_bases(metas)
    idx = np.sort(rng.choice(n, size=n // 3, replace=False)).astype(np.int32)
    for use_idx in (False, True):
        rows = idx if use_idx else None
        m = len(idx) if use_idx else n
        out = np.zeros(2 * total)
        orc.orc_hist_build(H.fptr(bins, C.c_uint8), C.c_int(F), fm, None if rows is None else H.fptr(rows, C.c_int32),
                           C.c_int(m), H.fptr(g, C.c_float), H.fptr(h, C.c_float), H.fptr(out, C.c_double))
        og = (g[rows] if use_idx else g).copy(); oh = (h[rows] if use_idx else h).copy()
        for f in range(F):
            col = np.ascontiguousarray(bins[:, f])
            ns = H.nslots(metas[f])
            want = np.zeros(2 * ns)
            shim.ref_dense_hist(H.fptr(col, C.c_uint8), C.c_int(n), C.c_int(ns),
                                None if rows is None else H.fptr(rows, C.c_int32), C.c_int(m),
                                H.fptr(og, C.c_float), H.fptr(oh, C.c_float), H.fptr(want, C.c_double))
            got = out[2 * bases[f]: 2 * (bases[f] + ns)]
            assert np.array_equal(got, want), f"feature {f}"


@pytest.mark.parametrize("kind", range(8))
def test_oracle_partition_bit_exact(shim, orc, kind):
    n, F = 9000, 3
    rng, metas, bins, g, h = _case(20 + kind, n, F, kinds=[kind], max_bin=30)
    fm = N.make_feature_meta(metas)
    idx = np.sort(rng.choice(n, size=n // 2, replace=False)).astype(np.int32)
    for feature in range(F):
        nb = metas[feature][0]
        col = np.ascontiguousarray(bins[:, feature])
        for threshold in sorted({0, 1, nb // 2, max(nb - 2, 0), nb - 1}):
            for dl in (0, 1):
                lte = np.empty(len(idx), np.int32); gt = np.empty(len(idx), np.int32)
                wl = np.empty(len(idx), np.int32); wg = np.empty(len(idx), np.int32)
                a = orc.orc_partition(H.fptr(bins, C.c_uint8), C.c_int(F), fm, C.c_int(feature), C.c_int(threshold),
                                      C.c_int(dl), H.fptr(idx, C.c_int32), C.c_int(len(idx)), H.fptr(lte, C.c_int32),
                                      H.fptr(gt, C.c_int32))
                b = shim.ref_dense_split(H.fptr(col, C.c_uint8), C.c_int(n), C.byref(fm[feature]), C.c_int(threshold),
                                         C.c_int(dl), H.fptr(idx, C.c_int32), C.c_int(len(idx)), H.fptr(wl, C.c_int32),
                                         H.fptr(wg, C.c_int32))
                assert a == b
                assert np.array_equal(lte[:a], wl[:b]) and np.array_equal(gt[:len(idx) - a], wg[:len(idx) - b])


PARAM_SETS = [
    dict(min_data_in_leaf=20, min_sum_hessian_in_leaf=1e-3),
    dict(min_data_in_leaf=1, min_sum_hessian_in_leaf=100.0),
    dict(min_data_in_leaf=50, min_sum_hessian_in_leaf=5.0, lambda_l1=0.5, lambda_l2=2.0, min_gain_to_split=0.1),
    dict(min_data_in_leaf=5, lambda_l2=1.0, max_delta_step=1.5),
    dict(min_data_in_leaf=5, lambda_l1=0.1, path_smooth=10.0),
]


@pytest.mark.parametrize("pi", range(len(PARAM_SETS)))
@pytest.mark.parametrize("kind", range(8))
def test_oracle_find_best_threshold_bit_exact(shim, orc, pi, kind):
    n, F = 8000, 4
    rng, metas, bins, g, h = _case(40 + kind, n, F, kinds=[kind], max_bin=50)
    params = N.make_tree_params(num_leaves=8, **PARAM_SETS[pi])
    fm = N.make_feature_meta(metas)
    bases, total = H.hist_bases(metas)
    hist = np.zeros(2 * total)
    orc.orc_hist_build(H.fptr(bins, C.c_uint8), C.c_int(F), fm, None, C.c_int(n), H.fptr(g, C.c_float),
                       H.fptr(h, C.c_float), H.fptr(hist, C.c_double))
    sg = float(np.sum(g.astype(np.float64))); sh = float(np.sum(h.astype(np.float64)))
    orc.orc_hist_fix(fm, C.c_int(F), C.c_double(sg), C.c_double(sh), H.fptr(hist, C.c_double))
    for f in range(F):
        data = np.ascontiguousarray(hist[2 * (bases[f] + 1): 2 * (bases[f] + H.nslots(metas[f]))]).copy()
        data = np.concatenate([data, np.zeros(8)])  # the reference may read one pair past a 1-slot feature
        for parent_output in (0.0, -0.37):
            a = N.SplitInfo(); b = N.SplitInfo(); sa = C.c_int(); sb = C.c_int()
            orc.orc_find_best_threshold(H.fptr(data, C.c_double), C.byref(fm[f]), C.byref(params), C.c_double(sg),
                                        C.c_double(sh), C.c_int(n), C.c_double(parent_output), C.byref(a), C.byref(sa))
            shim.ref_find_best_threshold(H.fptr(data, C.c_double), C.byref(fm[f]), C.byref(params), C.c_double(sg),
                                         C.c_double(sh), C.c_int(n), C.c_double(parent_output), C.byref(b), C.byref(sb))
            assert sa.value == sb.value
            for name, _ in N.SplitInfo._fields_:
                if name in ("feature", "real_feature"):
                    continue
                va, vb = getattr(a, name), getattr(b, name)
                if np.isfinite(b.gain) or name in ("gain", "default_left"):
                    assert va == vb or (np.isnan(va) and np.isnan(vb)), f"feature {f} field {name}: {va} vs {vb}"


@pytest.mark.parametrize("weighted", [False, True])
def test_oracle_gradients_bit_exact(shim, orc, weighted):
    n = 20_011
    rng = np.random.default_rng(9)
    score = rng.standard_normal(n) * 2
    weight = (rng.random(n) + 0.5).astype(np.float32) if weighted else None
    wp = None if weight is None else H.fptr(weight, C.c_float)
    g = np.empty(n, np.float32); h = np.empty(n, np.float32); wg = np.empty(n, np.float32); wh = np.empty(n, np.float32)
    # binary (is_unbalance exercises the label weights)
    label = (rng.random(n) < 0.3).astype(np.float32)
    npos = int(label.sum()); nneg = n - npos
    orc.orc_grad_binary(C.c_int(n), H.fptr(score, C.c_double), H.fptr(label, C.c_float), wp, C.c_double(0.7),
                        C.c_double(1.0), C.c_double(nneg / npos), H.fptr(g, C.c_float), H.fptr(h, C.c_float))
    assert shim.ref_gradients(b"objective=binary sigmoid=0.7 is_unbalance=true verbosity=-1", C.c_int(n),
                              H.fptr(label, C.c_float), wp, None, C.c_int(0), H.fptr(score, C.c_double),
                              H.fptr(wg, C.c_float), H.fptr(wh, C.c_float)) == 0
    assert np.array_equal(g, wg) and np.array_equal(h, wh)
    # l2
    label2 = rng.standard_normal(n).astype(np.float32)
    orc.orc_grad_l2(C.c_int(n), H.fptr(score, C.c_double), H.fptr(label2, C.c_float), wp, H.fptr(g, C.c_float), H.fptr(h, C.c_float))
    assert shim.ref_gradients(b"objective=regression verbosity=-1", C.c_int(n), H.fptr(label2, C.c_float), wp, None,
                              C.c_int(0), H.fptr(score, C.c_double), H.fptr(wg, C.c_float), H.fptr(wh, C.c_float)) == 0
    assert np.array_equal(g, wg) and np.array_equal(h, wh)
    # lambdarank
    sizes = []
    left = n
    while left > 0:
        s = int(min(left, max(1, rng.poisson(40))))
        sizes.append(s); left -= s
    sizes = np.array(sizes, dtype=np.int32)
    bounds = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    label3 = rng.choice(5, size=n, p=[.52, .32, .13, .02, .01]).astype(np.float32)
    score3 = np.round(score, 2)  # ties in score exercise the stable sort
    gains = np.array([0.0] + [float((1 << i) - 1) for i in range(1, 31)])
    for norm in (1, 0):
        orc.orc_grad_lambdarank(C.c_int(len(sizes)), H.fptr(bounds, C.c_int32), H.fptr(score3, C.c_double),
                                H.fptr(label3, C.c_float), wp, C.c_double(1.0), C.c_int(20), C.c_int(norm),
                                H.fptr(gains, C.c_double), C.c_int(len(gains)), H.fptr(g, C.c_float), H.fptr(h, C.c_float))
        p = f"objective=lambdarank lambdarank_norm={'true' if norm else 'false'} verbosity=-1".encode()
        assert shim.ref_gradients(p, C.c_int(n), H.fptr(label3, C.c_float), wp, H.fptr(sizes, C.c_int32),
                                  C.c_int(len(sizes)), H.fptr(score3, C.c_double), H.fptr(wg, C.c_float),
                                  H.fptr(wh, C.c_float)) == 0
        assert np.array_equal(g, wg) and np.array_equal(h, wh)


@pytest.mark.parametrize("pi", range(len(PARAM_SETS)))
def test_oracle_tree_matches_reference_tree(shim, orc, ref, pi):
    """One boosting iteration with given gradients: the reference grows a tree through its C API; the oracle grows
    one from the reference's own binned columns.  Same splits, same counts, same leaf values."""
    n = 12_000
    X = mixed_matrix(11, n)[:, [0, 1, 2, 3, 6, 9, 10]]
    rng = np.random.default_rng(12)
    g = (rng.standard_normal(n) + 0.8 * np.sign(np.nan_to_num(X[:, 0])) - 0.5 * (X[:, 2] > 1)).astype(np.float32)
    h = (rng.random(n) * 0.25 + 0.05).astype(np.float32)
    ps = PARAM_SETS[pi]
    pstr = " ".join(f"{k}={v}" for k, v in ps.items())
    params = f"objective=none num_leaves=31 learning_rate=1.0 verbosity=-1 num_threads=2 force_col_wise=true {pstr}"
    ds = ref.dataset_from_mat(X, params)
    ref.set_field(ds, "label", np.zeros(n, np.float32))
    b = ref.booster_create(ds, params)
    ref.update_custom(b, g, h)
    tree = capi.parse_trees(ref.model_string(b))[0]
    # the reference's own bins + metas
    nf = shim.ref_num_inner_features(ds)
    metas = []; cols = []; bounds_all = []
    for f in range(nf):
        m = (C.c_int32 * 5)(); bd = (C.c_double * 300)()
        shim.ref_feature_info(ds, C.c_int(f), m, bd, C.c_int(300))
        metas.append((m[0], m[1], m[2], m[3], m[4])); bounds_all.append(np.array(bd[:m[0]]))
        col = np.empty(n, np.uint8)
        shim.ref_feature_stored_column(ds, C.c_int(f), col.ctypes.data_as(C.POINTER(C.c_uint8)))
        cols.append(col)
    bins = np.ascontiguousarray(np.stack(cols, axis=1))
    tp = N.make_tree_params(num_leaves=31, **ps)
    if tp.path_smooth > 1e-15 and tp.min_data_in_leaf < 2:
        tp.min_data_in_leaf = 2
    fm = N.make_feature_meta(metas)
    recs = (N.SplitRecord * 30)()
    ns = orc.orc_train_tree(H.fptr(bins, C.c_uint8), C.c_int(n), C.c_int(nf), fm, C.byref(tp), H.fptr(g, C.c_float),
                            H.fptr(h, C.c_float), None, C.c_int(0), None, recs, None)
    assert ns == int(tree["num_leaves"][0]) - 1
    # replay: node i of the model is the i-th split
    sf = [int(v) for v in tree["split_feature"]]
    th = [float(v) for v in tree["threshold"]]
    ic = [int(v) for v in tree["internal_count"]]
    dt = [int(v) for v in tree["decision_type"]]
    for i in range(ns):
        r = recs[i]
        assert r.real_feature == sf[i]
        assert bounds_all[r.feature][r.threshold] == th[i]
        assert r.left_count + r.right_count == ic[i]
        assert ((dt[i] & 2) != 0) == bool(r.default_left)
    # leaf values / counts: replay the leaf ids
    leaf_value = {0: 0.0}; leaf_count = {0: n}
    for i in range(ns):
        r = recs[i]
        leaf_value[r.leaf] = r.left_output; leaf_value[r.right_leaf] = r.right_output
        leaf_count[r.leaf] = r.left_count; leaf_count[r.right_leaf] = r.right_count
    lv = [float(v) for v in tree["leaf_value"]]; lc = [int(v) for v in tree["leaf_count"]]
    for leaf in range(ns + 1):
        assert leaf_count[leaf] == lc[leaf]
        want = lv[leaf]
        got = leaf_value[leaf] if not np.isnan(leaf_value[leaf]) else 0.0
        if abs(got) <= 1e-35:
            got = 0.0
        assert got == want, f"leaf {leaf}: {got!r} vs {want!r}"


def sparse_matrix(seed, n, dtype=np.float32):
    """One-hot blocks (exclusive by construction), sparse numeric columns that collide in a few rows, a column whose
    most frequent value is not zero, and two dense columns: the reference bundles the sparse ones into shared columns."""
    rng = np.random.default_rng(seed)
    cols = []
    for levels in (6, 9):
        cat = rng.integers(0, levels, size=n)
        block = [(cat == v).astype(np.float64) * (1.0 + 0.25 * v) for v in range(levels)]
        for _ in range(2):                                 # two rows where a second member is set as well: a dense bundle
            r = int(rng.integers(0, n))                    # tolerates n / 10000 such rows and keeps only the last pushed value
            block[(int(cat[r]) + 1 + int(rng.integers(0, levels - 1))) % levels][r] = 1.0
        cols += block
    owner = rng.integers(0, 40, size=n)                    # sparse numeric columns on disjoint rows ...
    sparse = [np.where(owner == k, rng.integers(1, 6, size=n).astype(np.float64), 0.0) for k in range(8)]
    for k in range(0, 8, 2):                               # ... except one row per pair, where both are set: the bundle
        r = int(rng.integers(0, n))                        # tolerates n / 10000 such rows and drops one of the two values
        sparse[k][r] = 3.0; sparse[k + 1][r] = 2.0
    cols += sparse
    cols.append(np.where(rng.random(n) < 0.04, rng.standard_normal(n), 0.0))
    cols.append(np.where(rng.random(n) < 0.97, 7.0, rng.integers(0, 4, size=n).astype(np.float64)))   # mfb != zero bin
    cols.append(np.where(rng.random(n) < 0.02, np.nan, 0.0) + np.where(rng.random(n) < 0.03, 2.0, 0.0))
    cols.append(rng.standard_normal(n))
    cols.append(np.maximum(0, rng.standard_normal(n)))
    X = np.stack(cols, axis=1).astype(dtype)
    return X[:, rng.permutation(X.shape[1])]


@pytest.mark.parametrize("seed,n,params", [(1, 30_000, "max_bin=255 verbosity=-1"),
                                           (2, 42_000, "max_bin=63 min_data_in_bin=1 verbosity=-1"),
                                           (3, 25_000, "max_bin=255 zero_as_missing=true verbosity=-1")])
def test_bundled_features_bin_like_the_reference(shim, ref, ours, seed, n, params, monkeypatch):
    # exclusive feature bundling: the reference shares one column between several sparse features, and rows where two of
    # them are non-default lose all but the last pushed; this library keeps a column per feature and applies that rule
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "1")
    X = sparse_matrix(seed, n)
    dr = ref.dataset_from_mat(X, params); do = ours.dataset_from_mat(X, params)
    nf = shim.ref_num_inner_features(dr)
    assert shim.ref_num_groups(dr) < nf, "the case does not bundle anything"
    bins = np.empty((n, nf), dtype=np.int32); real = np.empty(nf, dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(do, bins.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
    lost = 0
    for f in range(nf):
        mr = (C.c_int32 * 5)(); br = (C.c_double * 300)()
        shim.ref_feature_info(dr, C.c_int(f), mr, br, C.c_int(300))
        assert mr[4] == real[f], "inner feature order"
        want = np.empty(n, dtype=np.int32)
        shim.ref_feature_bin_column(dr, C.c_int(f), want.ctypes.data_as(C.POINTER(C.c_int32)))
        assert np.array_equal(bins[:, f], want), f"inner feature {f} (column {real[f]}): bins differ"
    # the override must actually have happened somewhere: without it the matrices differ
    do2 = ours.dataset_from_mat(X, params + " enable_bundle=false")
    nf2 = C.c_int(); ours.check(ours.lib.LGBM_B200_DatasetNumInnerFeatures(do2, C.byref(nf2)))
    bins2 = np.empty((n, nf2.value), dtype=np.int32); real2 = np.empty(nf2.value, dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(do2, bins2.ctypes.data_as(C.POINTER(C.c_int32)), real2.ctypes.data_as(C.POINTER(C.c_int32))))
    by_real = {int(r): i for i, r in enumerate(real2)}
    for f in range(nf):
        lost += int(np.sum(bins[:, f] != bins2[:, by_real[int(real[f])]]))
    assert lost > 0, "no conflicting row in the case: the override rule is not exercised"
    # an aligned (validation) dataset is not bundled (Dataset::CreateValid): same bins as without bundling
    Xv = sparse_matrix(seed, n + 5000)[n:]
    dvr = ref.dataset_from_mat(Xv, params, reference=dr); dvo = ours.dataset_from_mat(Xv, params, reference=do)
    assert shim.ref_num_groups(dvr) == nf
    bv = np.empty((len(Xv), nf), dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(dvo, bv.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
    for f in range(nf):
        want = np.empty(len(Xv), dtype=np.int32)
        shim.ref_feature_bin_column(dvr, C.c_int(f), want.ctypes.data_as(C.POINTER(C.c_int32)))
        assert np.array_equal(bv[:, f], want), f"validation data, inner feature {f}: bins differ"
    ref.dataset_free(dr); ours.dataset_free(do); ours.dataset_free(do2)


def categorical_matrix(seed, n, dtype=np.float64):
    """columns 0, 2, 4, 5 are categorical: 3 categories (one-hot search), 60 skewed categories with NaN and negative
    values, 400 categories of which many are rare, a sparse one (mostly category 0); 1, 3, 6 are numerical."""
    rng = np.random.default_rng(seed)
    c0 = rng.integers(0, 3, size=n).astype(np.float64)
    c2 = np.floor(rng.exponential(8.0, size=n)).clip(0, 59)
    c2 = np.where(rng.random(n) < 0.03, np.nan, c2)
    c2 = np.where(rng.random(n) < 0.01, -2.0, c2)
    c4 = np.floor(rng.pareto(1.2, size=n) * 3).clip(0, 399)
    c5 = np.where(rng.random(n) < 0.9, 0.0, rng.integers(1, 12, size=n).astype(np.float64))
    X = np.stack([c0, rng.standard_normal(n), c2, np.maximum(0, rng.standard_normal(n)), c4, c5,
                  rng.standard_normal(n)], axis=1)
    return X.astype(dtype)


CAT_PARAMS = "categorical_feature=0,2,4,5 max_bin=255 verbosity=-1"


@pytest.mark.parametrize("seed,n,extra", [(1, 20_000, ""), (2, 9_000, " max_bin=63 min_data_in_bin=5"), (3, 15_000, " use_missing=false")])
def test_categorical_features_bin_like_the_reference(shim, ref, ours, seed, n, extra, monkeypatch):
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "1")
    X = categorical_matrix(seed, n)
    p = CAT_PARAMS + extra
    dr = ref.dataset_from_mat(X, p); do = ours.dataset_from_mat(X, p)
    nf = shim.ref_num_inner_features(dr)
    nfo = C.c_int(); ours.check(ours.lib.LGBM_B200_DatasetNumInnerFeatures(do, C.byref(nfo)))
    assert nfo.value == nf
    bins = np.empty((n, nf), dtype=np.int32); real = np.empty(nf, dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(do, bins.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
    for f in range(nf):
        mr = (C.c_int32 * 5)(); br = (C.c_double * 600)()
        mo = (C.c_int32 * 5)(); bo = (C.c_double * 600)()
        shim.ref_feature_info(dr, C.c_int(f), mr, br, C.c_int(600))
        ours.check(ours.lib.LGBM_B200_DatasetFeatureInfo(do, C.c_int(f), mo, bo, C.c_int(600)))
        assert list(mr) == list(mo), f"inner feature {f}: (num_bin, mfb, default_bin, missing, real) differ"
        # BinToValue: upper bounds of a numerical feature, the categories of a categorical one
        assert np.array_equal(np.array(br[:mr[0]]), np.array(bo[:mo[0]]), equal_nan=True), f"inner feature {f}: BinToValue differs"
        want = np.empty(n, dtype=np.int32)
        shim.ref_feature_bin_column(dr, C.c_int(f), want.ctypes.data_as(C.POINTER(C.c_int32)))
        assert np.array_equal(bins[:, f], want), f"inner feature {f} (column {real[f]}): bins differ"
    ref.dataset_free(dr); ours.dataset_free(do)


@pytest.mark.parametrize("fmt,dtype,ptr", [("csr", np.float64, np.int32), ("csr", np.float32, np.int64), ("csc", np.float64, np.int32)])
def test_sparse_input_bins_like_the_reference(shim, ref, ours, fmt, dtype, ptr, monkeypatch):
    # LGBM_DatasetCreateFromCSR / CSC: the reference samples and pushes a sparse source to the same dataset as the dense
    # matrix; this library densifies.  Every row's bins must equal the reference's own sparse-built dataset.
    sp = pytest.importorskip("scipy.sparse")
    monkeypatch.setenv("LGBM_B200_HOST_BINNING", "1")
    n = 20_000
    X = sparse_matrix(4, n, dtype=np.float64)
    X = np.nan_to_num(X)                      # (a stored NaN is kept by both; absent = 0)
    X[::97, 3] = np.nan
    S = (sp.csr_matrix(X) if fmt == "csr" else sp.csc_matrix(X)).astype(dtype)
    S.indptr = S.indptr.astype(ptr)
    # (a CSC source is pushed column by column on parallel threads in the reference, so WHICH member of a bundle survives
    # a conflicting row depends on thread timing there: the CSC case runs without bundling)
    p = "max_bin=255 verbosity=-1" + (" enable_bundle=false" if fmt == "csc" else "")
    dr = ref.dataset_from_sparse(S, p); do = ours.dataset_from_sparse(S, p)
    nf = shim.ref_num_inner_features(dr)
    nfo = C.c_int(); ours.check(ours.lib.LGBM_B200_DatasetNumInnerFeatures(do, C.byref(nfo)))
    assert nfo.value == nf
    bins = np.empty((n, nf), dtype=np.int32); real = np.empty(nf, dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(do, bins.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
    for f in range(nf):
        mr = (C.c_int32 * 5)(); br = (C.c_double * 300)()
        shim.ref_feature_info(dr, C.c_int(f), mr, br, C.c_int(300))
        assert mr[4] == real[f]
        want = np.empty(n, dtype=np.int32)
        shim.ref_feature_bin_column(dr, C.c_int(f), want.ctypes.data_as(C.POINTER(C.c_int32)))
        assert np.array_equal(bins[:, f], want), f"inner feature {f} (column {real[f]}): bins differ"
    # an aligned dataset from a sparse source
    Sv = sp.csr_matrix(np.nan_to_num(sparse_matrix(5, 3000, dtype=np.float64)))
    dvr = ref.dataset_from_sparse(Sv, p, reference=dr); dvo = ours.dataset_from_sparse(Sv, p, reference=do)
    bv = np.empty((3000, nf), dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(dvo, bv.ctypes.data_as(C.POINTER(C.c_int32)), real.ctypes.data_as(C.POINTER(C.c_int32))))
    for f in range(nf):
        want = np.empty(3000, dtype=np.int32)
        shim.ref_feature_bin_column(dvr, C.c_int(f), want.ctypes.data_as(C.POINTER(C.c_int32)))
        assert np.array_equal(bv[:, f], want)
