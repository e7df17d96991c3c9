"""GPU parity tests, kernel by kernel, through the C ABI (include/lgbcu.h) against the CPU oracle.

Bars (north_star): bin/partition/leaf assignment bit-exact; histogram sums <= 1e-6 relative; split feature /
threshold identical; gains and outputs to 1e-9 relative (the split search is all-double on both sides).
"""
import ctypes as C

import numpy as np
import pytest

import helpers as H
from lightgbmv1_b200 import _native as N

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orc():
    return H.load_oracle()


def _make_case(seed, n, F, kinds=None, max_bin=255, weights=False):
    rng = np.random.default_rng(seed)
    metas = H.random_feature_metas(rng, F, max_bin=max_bin, kinds=kinds)
    bins = H.random_bins(rng, n, metas)
    g = rng.standard_normal(n).astype(np.float32)
    h = (rng.random(n).astype(np.float32) * 0.25 + 0.01).astype(np.float32)
    return rng, metas, bins, g, h


def _oracle_hist(orc, bins, metas, idx, g, h, fix=True):
    fm = N.make_feature_meta(metas)
    _, total = H.hist_bases(metas)
    out = np.zeros(2 * total, dtype=np.float64)
    n = len(idx) if idx is not None else bins.shape[0]
    orc.orc_hist_build(H.fptr(bins, C.c_uint8), C.c_int(bins.shape[1]), fm,
                       None if idx is None else H.fptr(idx, C.c_int32), C.c_int(n), H.fptr(g, C.c_float),
                       H.fptr(h, C.c_float), H.fptr(out, C.c_double))
    if fix:
        rows = idx if idx is not None else np.arange(bins.shape[0])
        sg = float(np.sum(g[rows].astype(np.float64))); sh = float(np.sum(h[rows].astype(np.float64)))
        orc.orc_hist_fix(fm, C.c_int(len(metas)), C.c_double(sg), C.c_double(sh), H.fptr(out, C.c_double))
    return out


def _compare_hist(metas, got, want, g, h, skip_slot0=True):
    bases, total = H.hist_bases(metas)
    mask = np.ones(total, dtype=bool)
    if skip_slot0:
        # slot 0 of each group collects the most-frequent bin's rows; the reference never reads it and for
        # mfb > 0 features its FixHistogram'd twin lives at slot 1+mfb (which IS compared)
        for b in bases:
            mask[b] = False
    floor_g = 1e-6 * float(np.max(np.abs(g))) if len(g) else 1.0
    floor_h = 1e-6 * float(np.max(np.abs(h))) if len(h) else 1.0
    eg = H.rel_err(got[0::2][mask], want[0::2][mask], floor_g)
    eh = H.rel_err(got[1::2][mask], want[1::2][mask], floor_h)
    return max(eg, eh)


@pytest.mark.parametrize("mode", [N.HIST_GLOBAL_ATOMIC, N.HIST_WARP_PRIVATE, N.HIST_AUTO])
@pytest.mark.parametrize("n,F", [(50_000, 28), (4097, 17), (33, 5), (1, 3), (20_000, 137)])
def test_root_histogram(orc, mode, n, F):
    rng, metas, bins, g, h = _make_case(1 + n + F, n, F)
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=4, min_data_in_leaf=1))
    L.set_hist_mode(mode)
    L.set_gradients(g, h)
    L.begin_tree()
    L.hist_build(0)
    got = L.hist_read(0)
    want = _oracle_hist(orc, bins, metas, None, g, h)
    assert _compare_hist(metas, got, want, g, h) <= H.HIST_RTOL
    sg, sh, cnt = L.get_leaf_sums(0)
    assert cnt == n
    assert abs(sg - float(np.sum(g.astype(np.float64)))) <= H.VALUE_RTOL * float(np.sum(np.abs(g.astype(np.float64))) + 1e-30)
    assert abs(sh - float(np.sum(h.astype(np.float64)))) <= H.VALUE_RTOL * float(np.sum(h.astype(np.float64)))
    L.close()


@pytest.mark.parametrize("mode", [N.HIST_GLOBAL_ATOMIC, N.HIST_WARP_PRIVATE])
def test_indexed_histogram_with_bag(orc, mode):
    n, F = 60_000, 28
    rng, metas, bins, g, h = _make_case(7, n, F)
    bag = np.sort(rng.choice(n, size=n // 3, replace=False)).astype(np.int32)
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=4, min_data_in_leaf=1))
    L.set_hist_mode(mode)
    L.set_gradients(g, h)
    L.begin_tree(bag_indices=bag)
    L.hist_build(0)
    got = L.hist_read(0)
    want = _oracle_hist(orc, bins, metas, bag, g, h)
    assert _compare_hist(metas, got, want, g[bag], h[bag]) <= H.HIST_RTOL
    L.close()


def test_histogram_is_deterministic_and_mode_independent():
    # fixed point: any kernel, any launch, same bits
    n, F = 30_000, 28
    rng, metas, bins, g, h = _make_case(11, n, F)
    outs = []
    for mode in (N.HIST_GLOBAL_ATOMIC, N.HIST_WARP_PRIVATE, N.HIST_WARP_PRIVATE):
        L = N.CudaLearner()
        L.set_dataset(bins, metas)
        L.set_tree_params(N.make_tree_params(num_leaves=4))
        L.set_hist_mode(mode)
        L.set_gradients(g, h)
        L.begin_tree()
        L.hist_build(0)
        outs.append(L.hist_read(0))
        L.close()
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[1], outs[2])


def test_column_upload_matches_row_upload():
    n, F = 10_000, 19
    rng, metas, bins, g, h = _make_case(13, n, F)
    res = []
    for use_cols in (False, True):
        L = N.CudaLearner()
        if use_cols:
            L.set_dataset_columns([np.ascontiguousarray(bins[:, f]) for f in range(F)], metas)
        else:
            L.set_dataset(bins, metas)
        L.set_tree_params(N.make_tree_params(num_leaves=4))
        L.set_gradients(g, h)
        L.begin_tree()
        L.hist_build(0)
        res.append(L.hist_read(0))
        L.close()
    assert np.array_equal(res[0], res[1])


def _orc_find(orc, hist, metas, params, sg, sh, n, is_root=1, used=None):
    fm = N.make_feature_meta(metas)
    out = N.SplitInfo()
    sp = np.zeros(len(metas), dtype=np.int8)
    hist = hist.copy()
    orc.orc_find_best_split(H.fptr(hist, C.c_double), fm, C.c_int(len(metas)),
                            None if used is None else H.fptr(used, C.c_int8), C.byref(params), C.c_double(sg),
                            C.c_double(sh), C.c_int(n), C.c_int(is_root), C.c_double(0.0), C.byref(out),
                            H.fptr(sp, C.c_int8))
    return out, sp


PARAM_SETS = [
    dict(min_data_in_leaf=20, min_sum_hessian_in_leaf=1e-3),
    dict(min_data_in_leaf=1, min_sum_hessian_in_leaf=100.0),
    dict(min_data_in_leaf=50, min_sum_hessian_in_leaf=5.0, lambda_l1=0.5, lambda_l2=2.0, min_gain_to_split=0.1),
    dict(min_data_in_leaf=5, lambda_l2=1.0, max_delta_step=1.5),
    dict(min_data_in_leaf=5, lambda_l1=0.1, path_smooth=10.0),
]


@pytest.mark.parametrize("pi", range(len(PARAM_SETS)))
@pytest.mark.parametrize("kinds", [None, [0], [1], [2], [3], [4], [5], [6], [7]])
def test_find_best_split_root(orc, pi, kinds):
    n, F = 20_000, 12
    rng, metas, bins, g, h = _make_case(100 + pi, n, F, kinds=kinds, max_bin=64 if kinds else 255)
    params = N.make_tree_params(num_leaves=8, **PARAM_SETS[pi])
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(params)
    L.set_gradients(g, h)
    L.begin_tree()
    L.hist_build(0)
    got = L.find_best_split(0)
    sg, sh, cnt = L.get_leaf_sums(0)
    hist = _oracle_hist(orc, bins, metas, None, g, h, fix=False)
    # oracle side uses its own double sums of the float gradients (LeafSplits::Init)
    osg = float(np.sum(g.astype(np.float64))); osh = float(np.sum(h.astype(np.float64)))
    want, _ = _orc_find(orc, hist, metas, params, osg, osh, n)
    assert (got.feature, got.threshold, got.default_left) == (want.feature, want.threshold, want.default_left)
    assert (got.left_count, got.right_count) == (want.left_count, want.right_count)
    if want.feature >= 0 and np.isfinite(want.gain):
        for a, b in [(got.gain, want.gain), (got.left_output, want.left_output), (got.right_output, want.right_output),
                     (got.left_sum_hessian, want.left_sum_hessian), (got.right_sum_hessian, want.right_sum_hessian)]:
            assert abs(a - b) <= H.VALUE_RTOL * max(abs(b), 1e-12) + 1e-12
        for a, b in [(got.left_sum_gradient, want.left_sum_gradient), (got.right_sum_gradient, want.right_sum_gradient)]:
            assert abs(a - b) <= H.VALUE_RTOL * float(np.sum(np.abs(g)))
    L.close()


@pytest.mark.parametrize("kind", range(8))
@pytest.mark.parametrize("default_left", [0, 1])
def test_partition_bit_exact(orc, kind, default_left):
    n, F = 30_011, 9
    rng, metas, bins, g, h = _make_case(200 + kind, n, F, kinds=[kind], max_bin=40)
    fm = N.make_feature_meta(metas)
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=8))
    L.set_gradients(g, h)
    for feature in (0, F - 1):
        nb = metas[feature][0]
        for threshold in {0, 1, nb // 2, max(nb - 2, 0)}:
            L.begin_tree()
            left = L.partition(0, feature, threshold, default_left, 1)
            idx = np.arange(n, dtype=np.int32)
            lte = np.empty(n, dtype=np.int32); gt = np.empty(n, dtype=np.int32)
            want_left = orc.orc_partition(H.fptr(bins, C.c_uint8), C.c_int(F), fm, C.c_int(feature), C.c_int(threshold),
                                          C.c_int(default_left), H.fptr(idx, C.c_int32), C.c_int(n),
                                          H.fptr(lte, C.c_int32), H.fptr(gt, C.c_int32))
            assert left == want_left
            assert np.array_equal(L.leaf_indices(0), lte[:want_left])
            assert np.array_equal(L.leaf_indices(1), gt[:n - want_left])
    L.close()


def test_partition_nested_keeps_order(orc):
    # split the right child again: ranges, ping-pong buffers and stability across two levels
    n, F = 50_000, 6
    rng, metas, bins, g, h = _make_case(300, n, F, kinds=[0, 2, 4])
    fm = N.make_feature_meta(metas)
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=8))
    L.set_gradients(g, h)
    L.begin_tree()
    t0 = metas[0][0] // 2
    L.partition(0, 0, t0, 1, 1)
    right = L.leaf_indices(1).copy()
    t1 = metas[1][0] // 3
    left2 = L.partition(1, 1, t1, 0, 2)
    lte = np.empty(len(right), dtype=np.int32); gt = np.empty(len(right), dtype=np.int32)
    want = orc.orc_partition(H.fptr(bins, C.c_uint8), C.c_int(F), fm, C.c_int(1), C.c_int(t1), C.c_int(0),
                             H.fptr(right, C.c_int32), C.c_int(len(right)), H.fptr(lte, C.c_int32), H.fptr(gt, C.c_int32))
    assert left2 == want
    assert np.array_equal(L.leaf_indices(1), lte[:want])
    assert np.array_equal(L.leaf_indices(2), gt[:len(right) - want])
    L.close()


def test_partition_of_an_empty_leaf(orc):
    # a leaf with no local rows (one rank of a data-parallel run can hold none of a leaf's rows): the partition must
    # finish, report 0 rows on both sides and leave the other leaves alone -- step-wise API and whole-tree growth
    n, F = 5_000, 4
    rng, metas, bins, g, h = _make_case(310, n, F, kinds=[1])
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=8))
    L.set_gradients(g, h)
    L.begin_tree()
    nb = metas[0][0]
    left = L.partition(0, 0, nb - 1, 1, 1)      # threshold = last bin: every row goes left, leaf 1 is empty
    assert left == n and len(L.leaf_indices(1)) == 0
    left2 = L.partition(1, 1, 0, 1, 2)          # split the empty leaf
    assert left2 == 0 and len(L.leaf_indices(1)) == 0 and len(L.leaf_indices(2)) == 0
    assert np.array_equal(L.leaf_indices(0), np.arange(n, dtype=np.int32))
    L.close()


def _orc_train(orc, bins, metas, params, g, h, bag=None, used=None):
    fm = N.make_feature_meta(metas)
    recs = (N.SplitRecord * max(params.num_leaves - 1, 1))()
    leaf_of_row = np.empty(bins.shape[0], dtype=np.int32)
    ns = orc.orc_train_tree(H.fptr(bins, C.c_uint8), C.c_int(bins.shape[0]), C.c_int(bins.shape[1]), fm, C.byref(params),
                            H.fptr(g, C.c_float), H.fptr(h, C.c_float),
                            None if bag is None else H.fptr(bag, C.c_int32), C.c_int(0 if bag is None else len(bag)),
                            None if used is None else H.fptr(used, C.c_int8), recs, H.fptr(leaf_of_row, C.c_int32))
    return [recs[i] for i in range(ns)], leaf_of_row


def _check_tree(L, got, want, leaf_of_row, n, bag=None):
    """Split records must be identical.  The one tolerated deviation (north_star: "except on gain near-ties") is
    a pair of candidates whose gains agree to H.VALUE_RTOL relative: then the records may differ from that split on, and
    only the GPU tree's own consistency is checked past it."""
    k = 0
    while k < min(len(got), len(want)) and H.records_to_tuples([got[k]]) == H.records_to_tuples([want[k]]):
        a, b = got[k], want[k]
        assert a.real_feature == b.real_feature
        assert abs(a.gain - b.gain) <= 1e-6 * max(abs(b.gain), 1e-6)
        for x, y in [(a.left_output, b.left_output), (a.right_output, b.right_output),
                     (a.left_sum_hessian, b.left_sum_hessian), (a.right_sum_hessian, b.right_sum_hessian)]:
            assert abs(x - y) <= H.VALUE_RTOL * max(abs(y), 1e-9)
        k += 1
    exact = k == len(got) == len(want)
    if not exact:
        assert k < len(got) and k < len(want), "one tree stopped early"
        a, b = got[k], want[k]
        assert abs(a.gain - b.gain) <= H.VALUE_RTOL * max(abs(b.gain), 1e-9), \
            f"split {k} differs and is not a near-tie: {H.records_to_tuples([a])} vs {H.records_to_tuples([b])}"
    # leaf assignment: every bagged row in exactly one leaf, ascending inside a leaf (stable partition);
    # bit-exact against the oracle when the trees are identical
    rows = np.arange(n, dtype=np.int32) if bag is None else bag
    seen = []
    cnt = {0: len(rows)}
    for r in got:
        cnt[r.leaf] = r.left_count; cnt[r.right_leaf] = r.right_count
    for leaf in range(len(got) + 1):
        idx = L.leaf_indices(leaf)
        assert len(idx) == cnt[leaf]
        assert np.all(np.diff(idx) > 0)
        if exact:
            assert np.array_equal(idx, np.nonzero(leaf_of_row == leaf)[0].astype(np.int32))
        seen.append(idx)
    assert np.array_equal(np.sort(np.concatenate(seen)), np.sort(rows))
    return exact


@pytest.mark.parametrize("mode", [N.HIST_AUTO, N.HIST_GLOBAL_ATOMIC, N.HIST_WARP_PRIVATE])
@pytest.mark.parametrize("pi", range(len(PARAM_SETS)))
def test_train_tree_matches_oracle(orc, mode, pi):
    n, F = 40_000, 28
    rng, metas, bins, g, h = _make_case(400 + pi, n, F)
    # learnable signal so that trees are deep
    g = (g + 0.8 * np.sign(bins[:, 0].astype(np.float32) - 20) + 0.5 * (bins[:, 3] > 7)).astype(np.float32)
    params = N.make_tree_params(num_leaves=31, **PARAM_SETS[pi])
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(params)
    L.set_hist_mode(mode)
    L.set_gradients(g, h)
    L.begin_tree()
    got = L.train_tree()
    want, leaf_of_row = _orc_train(orc, bins, metas, params, g, h)
    assert len(want) >= 3
    _check_tree(L, got, want, leaf_of_row, n)
    L.close()


def test_train_tree_bagging_colsample_depth(orc):
    n, F = 30_000, 20
    rng, metas, bins, g, h = _make_case(500, n, F)
    g = (g + 0.6 * np.sign(bins[:, 2].astype(np.float32) - 10)).astype(np.float32)
    bag = np.sort(rng.choice(n, size=int(n * 0.8), replace=False)).astype(np.int32)
    used = (rng.random(F) < 0.8).astype(np.int8); used[2] = 1
    params = N.make_tree_params(num_leaves=63, max_depth=5, min_data_in_leaf=50, min_sum_hessian_in_leaf=5.0)
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(params)
    L.set_gradients(g, h)
    L.begin_tree(bag_indices=bag, feature_used=used)
    got = L.train_tree()
    want, leaf_of_row = _orc_train(orc, bins, metas, params, g, h, bag=bag, used=used)
    assert len(want) >= 3
    _check_tree(L, got, want, leaf_of_row, n, bag=bag)
    L.close()


def test_train_tree_exact_without_missing(orc):
    # no missing-value features => no reverse/forward ties => the whole tree must be identical
    n, F = 60_000, 28
    rng, metas, bins, g, h = _make_case(450, n, F, kinds=[0, 1, 6, 0])
    g = (g + 0.8 * np.sign(bins[:, 0].astype(np.float32) - 20) + 0.5 * (bins[:, 3] > 7)).astype(np.float32)
    params = N.make_tree_params(num_leaves=63, min_data_in_leaf=20, min_sum_hessian_in_leaf=1e-3)
    L = N.CudaLearner()
    L.set_dataset(bins, metas); L.set_tree_params(params); L.set_gradients(g, h)
    L.begin_tree()
    got = L.train_tree()
    want, leaf_of_row = _orc_train(orc, bins, metas, params, g, h)
    assert len(want) == 62
    assert _check_tree(L, got, want, leaf_of_row, n)
    L.close()


def test_train_tree_stops_without_gain(orc):
    n, F = 5000, 4
    rng, metas, bins, g, h = _make_case(600, n, F)
    params = N.make_tree_params(num_leaves=15, min_data_in_leaf=1, min_gain_to_split=1e9)
    L = N.CudaLearner()
    L.set_dataset(bins, metas); L.set_tree_params(params); L.set_gradients(g, h)
    L.begin_tree()
    assert L.train_tree() == []
    want, _ = _orc_train(orc, bins, metas, params, g, h)
    assert want == []
    L.close()


@pytest.mark.parametrize("weighted", [False, True])
def test_gradients_binary_and_l2(orc, weighted):
    n = 100_003
    rng = np.random.default_rng(5)
    score = rng.standard_normal(n) * 3
    label = (rng.random(n) < 0.4).astype(np.float32)
    weight = (rng.random(n) + 0.5).astype(np.float32) if weighted else None
    metas = [(4, 0, 0, 0)]
    bins = np.zeros((n, 1), dtype=np.uint8)
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_label(label)
    L.set_weight(weight)
    L.set_score(score)
    wg = np.empty(n, np.float32); wh = np.empty(n, np.float32)
    # binary
    L.set_objective(N.OBJ_BINARY, 1.0, 1.0, 1.7)
    L.get_gradients()
    g, h = L.copy_gradients()
    orc.orc_grad_binary(C.c_int(n), H.fptr(score, C.c_double), H.fptr(label, C.c_float),
                        None if weight is None else H.fptr(weight, C.c_float), C.c_double(1.0), C.c_double(1.0),
                        C.c_double(1.7), H.fptr(wg, C.c_float), H.fptr(wh, C.c_float))
    # device exp() is within 1 ulp of libm's: allow 1 float ulp on the rounded result
    assert np.max(np.abs(g - wg) / np.maximum(np.abs(wg), 1e-30)) <= 2.5e-7
    assert np.max(np.abs(h - wh) / np.maximum(np.abs(wh), 1e-30)) <= 2.5e-7
    assert np.mean(g == wg) > 0.999
    # l2: exact
    label2 = rng.standard_normal(n).astype(np.float32)
    L.set_label(label2)
    L.set_objective(N.OBJ_L2)
    L.get_gradients()
    g, h = L.copy_gradients()
    orc.orc_grad_l2(C.c_int(n), H.fptr(score, C.c_double), H.fptr(label2, C.c_float),
                    None if weight is None else H.fptr(weight, C.c_float), H.fptr(wg, C.c_float), H.fptr(wh, C.c_float))
    assert np.array_equal(g, wg) and np.array_equal(h, wh)
    L.close()


def test_add_score_by_partition(orc):
    n, F = 20_000, 8
    rng, metas, bins, g, h = _make_case(700, n, F)
    g = (g + 0.7 * np.sign(bins[:, 1].astype(np.float32) - 9)).astype(np.float32)
    params = N.make_tree_params(num_leaves=15, min_data_in_leaf=10)
    L = N.CudaLearner()
    L.set_dataset(bins, metas); L.set_tree_params(params); L.set_gradients(g, h)
    score0 = rng.standard_normal(n)
    L.set_score(score0)
    L.begin_tree()
    recs = L.train_tree()
    nl = len(recs) + 1
    vals = rng.standard_normal(nl)
    L.add_score_by_partition(vals)
    L.add_score_const(0.25)
    got = L.copy_score()
    want_tree, leaf_of_row = _orc_train(orc, bins, metas, params, g, h)
    want = score0 + vals[leaf_of_row] + 0.25
    assert np.array_equal(got, want)
    L.close()


def test_errors_are_reported_not_fatal():
    L = N.CudaLearner()
    with pytest.raises(N.LgbCuError):
        L.begin_tree()          # no dataset
    with pytest.raises(N.LgbCuError):
        L.set_dataset(np.zeros((4, 2), np.uint8), [(70000, 0, 0, 0)])   # > 16-bit bins
    L.close()


def test_lambdarank_gradients_match_oracle(orc):
    rng = np.random.default_rng(21)
    sizes = np.clip(rng.poisson(90, size=400), 1, 700).astype(np.int32)
    sizes[5] = 1; sizes[17] = 650
    n = int(sizes.sum())
    bounds = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    label = rng.choice(5, size=n, p=[.52, .32, .13, .02, .01]).astype(np.float32)
    score = np.round(rng.standard_normal(n) * 2, 2)   # ties exercise the stable sort
    weight = (rng.random(n) + 0.5).astype(np.float32)
    gains = np.array([0.0] + [float((1 << i) - 1) for i in range(1, 31)])
    for w in (None, weight):
        for norm in (1, 0):
            L = N.CudaLearner()
            L.set_dataset(np.zeros((n, 1), np.uint8), [(4, 0, 0, 0)])
            L.set_label(label); L.set_weight(w); L.set_query(bounds)
            L.set_score(score)
            L.set_objective(N.OBJ_LAMBDARANK, 1.0, 20, norm)
            L.get_gradients()
            g, h = L.copy_gradients()
            wg = np.empty(n, np.float32); wh = np.empty(n, np.float32)
            orc.orc_grad_lambdarank(C.c_int(len(sizes)), H.fptr(bounds, C.c_int32), H.fptr(score, C.c_double),
                                    H.fptr(label, C.c_float), None if w is None else H.fptr(w, C.c_float), C.c_double(1.0),
                                    C.c_int(20), C.c_int(norm), H.fptr(gains, C.c_double), C.c_int(len(gains)),
                                    H.fptr(wg, C.c_float), H.fptr(wh, C.c_float))
            # per-document float accumulation is replayed in the reference's order: bit-exact except that the
            # normalisation factor sums the pair lambdas in a different (block-reduction) order: <= 1 float ulp
            assert np.max(np.abs(g - wg) / np.maximum(np.abs(wg), 1e-30)) <= 2.5e-7
            assert np.max(np.abs(h - wh) / np.maximum(np.abs(wh), 1e-30)) <= 2.5e-7
            if norm == 0:
                assert np.array_equal(g, wg) and np.array_equal(h, wh)
            L.close()


def test_wide_feature_matrix_like_epsilon(orc):
    # config #3 shape in miniature: 2000 features (125 feature tiles), every feature 255 bins
    n, F = 6000, 2000
    rng = np.random.default_rng(31)
    metas = [(255, 127, 127, 0)] * F
    bins = rng.integers(0, 256, size=(n, F), dtype=np.int64)
    bins = np.where(bins == 128, 0, bins).astype(np.uint8)  # stored 0 = most frequent bin, 128 would be its own slot
    g = rng.standard_normal(n).astype(np.float32)
    g += (0.7 * np.sign(bins[:, 1234].astype(np.float32) - 100)).astype(np.float32)
    h = (rng.random(n) * 0.25 + 0.01).astype(np.float32)
    params = N.make_tree_params(num_leaves=15, min_data_in_leaf=20)
    L = N.CudaLearner()
    L.set_dataset(bins, metas); L.set_tree_params(params); L.set_gradients(g, h)
    L.begin_tree()
    L.hist_build(0)
    got = L.hist_read(0)
    want = _oracle_hist(orc, bins, metas, None, g, h)
    assert _compare_hist(metas, got, want, g, h) <= H.HIST_RTOL
    L.begin_tree()
    recs = L.train_tree()
    want_recs, leaf_of_row = _orc_train(orc, bins, metas, params, g, h)
    assert len(want_recs) >= 3 and recs[0].feature == 1234
    _check_tree(L, recs, want_recs, leaf_of_row, n)
    L.close()
