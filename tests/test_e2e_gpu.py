"""End-to-end parity through the drop-in C API: the same LGBM_* calls are issued against the reference's
lib_lightgbm.so (CPU, tree_learner=serial, force_col_wise) and against lib_lightgbm_b200.so (CUDA).

Bars (north_star): identical split feature / threshold per node except on gain near-ties; training loss / AUC
within 1e-5; the model text is loadable by the unmodified reference.
"""
import os

import numpy as np
import pytest

import capi
import helpers as H

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]

REF_EXTRA = " force_col_wise=true num_threads=4"


def logloss(p, y):
    p = np.clip(p, 1e-15, 1 - 1e-15)
    return float(-np.mean(y * np.log(p) + (1 - y) * np.log(1 - p)))


def auc(p, y):
    order = np.argsort(p, kind="mergesort")
    ranks = np.empty(len(p)); ranks[order] = np.arange(1, len(p) + 1)
    # average ranks for ties
    sp = p[order]
    i = 0
    while i < len(sp):
        j = i
        while j + 1 < len(sp) and sp[j + 1] == sp[i]:
            j += 1
        if j > i:
            ranks[order[i:j + 1]] = (i + j + 2) / 2.0
        i = j + 1
    npos = y.sum(); nneg = len(y) - npos
    return float((ranks[y > 0].sum() - npos * (npos + 1) / 2) / (npos * nneg))


def higgs_like(seed, n, F=28):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, F)).astype(np.float32)
    for c in range(21, min(25, F)):
        X[:, c] = np.maximum(0, X[:, c])
    for c in range(25, F):
        X[:, c] = rng.choice(np.array([0.0, 1.087, 2.173], dtype=np.float32), size=n)
    w = np.random.default_rng(seed + 1).standard_normal(F)
    y = ((X @ w + 0.5 * X[:, 0] * X[:, 1] + rng.standard_normal(n)) > 0).astype(np.float32)
    return X, y


def run_both(X, y, params, iters, **kw):
    ref, ours = capi.reference(), capi.ours()
    dsr, br, nr = capi.train(ref, X, y, params + REF_EXTRA, iters, **kw)
    dso, bo, no = capi.train(ours, X, y, params, iters, **kw)
    return ref, ours, br, bo, nr, no


def assert_same_model(ref, ours, br, bo, min_identical_frac=1.0):
    tr, to = ref.model_string(br), ours.model_string(bo)
    hr, ho = capi.header_lines(tr), capi.header_lines(to)
    for k in ("version", "num_class", "num_tree_per_iteration", "label_index", "max_feature_idx", "objective",
              "feature_names", "feature_infos"):
        assert hr.get(k) == ho.get(k), k
    same, n, first = capi.compare_models(tr, to)
    assert n > 0
    assert same >= int(np.ceil(min_identical_frac * n)), f"{same}/{n} trees identical; first mismatch: {first}"
    return tr, to


def test_binary_identical_trees_and_metrics():
    X, y = higgs_like(0, 60_000)
    params = "objective=binary num_leaves=63 learning_rate=0.1 min_data_in_leaf=0 min_sum_hessian_in_leaf=100 verbosity=-1"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 20)
    assert nr == no == 20
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, len(y)), ours.get_predict(bo, len(y))
    assert abs(logloss(pr, y) - logloss(po, y)) <= 1e-5
    assert abs(auc(pr, y) - auc(po, y)) <= 1e-5
    assert np.max(np.abs(pr - po)) <= 1e-9


def test_reference_loads_our_model_and_predicts_the_same():
    X, y = higgs_like(2, 20_000)
    params = "objective=binary num_leaves=31 verbosity=-1"
    ours = capi.ours(); ref = capi.reference()
    dso, bo, _ = capi.train(ours, X, y, params, 8)
    text = ours.model_string(bo)
    loaded = ref.load_model_from_string(text)
    Xt, _ = higgs_like(3, 5000)
    Xt[::7, 3] = np.nan
    for raw in (True, False):
        a = ref.predict_mat(loaded, Xt, raw=raw)
        b = ours.predict_mat(bo, Xt, raw=raw)
        assert np.max(np.abs(a - b)) <= 1e-12
    # training scores kept on the device == predictions of the final model on the training matrix
    tr = ours.get_predict(bo, len(y))
    pm = ours.predict_mat(bo, X)
    assert np.max(np.abs(tr - pm)) <= 1e-9


def test_regression_l2_with_weights():
    rng = np.random.default_rng(5)
    n = 30_000
    X = rng.standard_normal((n, 12)).astype(np.float64)
    X[:, 3] = np.where(rng.random(n) < 0.1, np.nan, X[:, 3])
    y = (2 * X[:, 0] + np.nan_to_num(X[:, 3]) ** 2 + np.sin(3 * X[:, 5]) + 0.1 * rng.standard_normal(n)).astype(np.float32)
    w = (rng.random(n) + 0.5).astype(np.float32)
    params = "objective=regression num_leaves=31 learning_rate=0.2 lambda_l2=1.0 min_data_in_leaf=20 verbosity=-1"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 15, weight=w)
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, n), ours.get_predict(bo, n)
    assert abs(np.mean(w * (pr - y) ** 2) - np.mean(w * (po - y) ** 2)) <= 1e-5 * max(1.0, np.mean(w * (pr - y) ** 2))


def test_regression_l2_constant_hessian_no_missing():
    rng = np.random.default_rng(6)
    n = 40_000
    X = rng.standard_normal((n, 10)).astype(np.float32)
    y = (X[:, 0] * X[:, 1] + np.abs(X[:, 2]) + 0.1 * rng.standard_normal(n)).astype(np.float32)
    params = "objective=regression num_leaves=63 learning_rate=0.1 verbosity=-1"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 12)
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, n), ours.get_predict(bo, n)
    assert np.max(np.abs(pr - po)) <= 1e-9


def test_bagging_feature_fraction_like_config1():
    # examples/binary_classification/train.conf shape: 7000 x 28, 63 leaves, feature_fraction 0.8,
    # bagging_fraction 0.8 / freq 5, min_data_in_leaf 50, min_sum_hessian_in_leaf 5, weights
    X, y = higgs_like(7, 7000)
    w = (np.random.default_rng(8).random(7000) + 0.5).astype(np.float32)
    params = ("objective=binary num_leaves=63 learning_rate=0.1 feature_fraction=0.8 bagging_fraction=0.8 bagging_freq=5 "
              "min_data_in_leaf=50 min_sum_hessian_in_leaf=5.0 max_bin=255 verbosity=-1")
    ref, ours, br, bo, nr, no = run_both(X, y, params, 30, weight=w)
    assert nr == no
    # leaves of ~100 rows leave most of the 255 bins empty: thresholds tie exactly unless FixHistogram's 1e-15
    # residue and the kEpsilon-seeded sequential sums are replayed in the reference's order -- they are
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, 7000), ours.get_predict(bo, 7000)
    assert abs(logloss(pr, y) - logloss(po, y)) <= 1e-5
    assert abs(auc(pr, y) - auc(po, y)) <= 1e-5


def test_lambdarank_small():
    rng = np.random.default_rng(9)
    sizes = np.clip(rng.poisson(60, size=200), 1, 300).astype(np.int32)
    n = int(sizes.sum())
    X = rng.standard_normal((n, 20)).astype(np.float32)
    X[:, 10:] = np.round(np.abs(X[:, 10:]) * 3)
    latent = X[:, 0] + 0.5 * X[:, 11] + 0.5 * rng.standard_normal(n)
    y = np.digitize(latent, np.quantile(latent, [0.52, 0.84, 0.97, 0.99])).astype(np.float32)
    params = "objective=lambdarank num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 verbosity=-1"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10, group=sizes)
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, n), ours.get_predict(bo, n)
    assert np.max(np.abs(pr - po)) <= 1e-6


def test_custom_objective_host_gradients():
    X, y = higgs_like(11, 20_000, F=10)
    params = "objective=none num_leaves=31 learning_rate=0.1 verbosity=-1"
    ref, ours = capi.reference(), capi.ours()
    out = []
    for lib, extra in ((ref, REF_EXTRA), (ours, "")):
        ds = lib.dataset_from_mat(X, params + extra)
        lib.set_field(ds, "label", y)
        b = lib.booster_create(ds, params + extra)
        score = np.zeros(len(y))
        for it in range(5):
            p = 1 / (1 + np.exp(-score))
            lib.update_custom(b, (p - y).astype(np.float32), (p * (1 - p)).astype(np.float32))
            score = lib.get_predict(b, len(y))
        out.append((lib, b))
    same, n, first = capi.compare_models(out[0][0].model_string(out[0][1]), out[1][0].model_string(out[1][1]))
    assert same == n == 5, first


def test_init_score_and_stop_when_unsplittable():
    X, y = higgs_like(13, 3000, F=6)
    init = np.full(len(y), 0.3)
    params = "objective=binary num_leaves=15 min_data_in_leaf=1400 verbosity=-1"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 6, init_score=init)
    assert nr == no
    assert_same_model(ref, ours, br, bo)


def test_errors_surface_through_last_error():
    ours = capi.ours()
    X, y = higgs_like(14, 500, F=4)
    with pytest.raises(capi.LGBMError):
        ours.dataset_from_mat(X, "max_bin=1")
    ds = ours.dataset_from_mat(X, "verbosity=-1")
    ours.set_field(ds, "label", y)
    with pytest.raises(capi.LGBMError):
        ours.booster_create(ds, "objective=no_such_objective")
    with pytest.raises(capi.LGBMError):
        ours.booster_create(ds, "objective=binary boosting=gbdt tree_learner=voting")
    with pytest.raises(capi.LGBMError):
        ours.set_field(ds, "label", y[:10])


def test_more_than_256_leaves_uses_the_general_control_path():
    # trees of more than 256 leaves take the sequential commit/select path of the control step
    X, y = higgs_like(21, 60_000)
    params = "objective=binary num_leaves=400 min_data_in_leaf=5 min_sum_hessian_in_leaf=1e-3 learning_rate=0.1 verbosity=-1"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 3)
    assert_same_model(ref, ours, br, bo)


@pytest.mark.parametrize("lookahead", ["1", "3", "16"])
def test_model_does_not_depend_on_the_lookahead(lookahead, monkeypatch):
    # look-ahead growth expands leaves early but commits them in the reference's order: any K gives the same trees
    X, y = higgs_like(22, 80_000)
    params = "objective=binary num_leaves=63 min_data_in_leaf=20 learning_rate=0.1 max_depth=9 verbosity=-1"
    monkeypatch.setenv("LGBCU_LOOKAHEAD", lookahead)
    ref, ours, br, bo, nr, no = run_both(X, y, params, 5)
    assert_same_model(ref, ours, br, bo)


def test_training_on_a_row_subset():
    """LGBM_DatasetGetSubset (python's Dataset.subset / lgb.cv folds): both libraries train on the same rows of the same
    parent dataset -- the parent's bin mappers, the subset's labels and weights -- and give the same trees"""
    import ctypes as C
    X, y = higgs_like(51, 40_000)
    w = (np.random.default_rng(52).random(len(y)) + 0.5).astype(np.float32)
    idx = np.sort(np.random.default_rng(53).choice(len(y), size=26_000, replace=False)).astype(np.int32)
    params = "objective=binary num_leaves=31 min_data_in_leaf=20 verbosity=-1"
    out = {}
    for name, lib, extra in (("ref", capi.reference(), REF_EXTRA), ("ours", capi.ours(), "")):
        ds = lib.dataset_from_mat(X, params + extra)
        lib.set_field(ds, "label", y); lib.set_field(ds, "weight", w)
        sub = C.c_void_p()
        lib.check(lib.lib.LGBM_DatasetGetSubset(ds, idx.ctypes.data_as(C.POINTER(C.c_int32)), C.c_int32(len(idx)),
                                                C.c_char_p((params + extra).encode()), C.byref(sub)))
        b = lib.booster_create(sub, params + extra)
        for _ in range(8):
            lib.update(b)
        out[name] = (lib, b)
    (ref, br), (ours, bo) = out["ref"], out["ours"]
    assert_same_model(ref, ours, br, bo)
    assert np.max(np.abs(ref.get_predict(br, len(idx)) - ours.get_predict(bo, len(idx)))) <= 1e-9
