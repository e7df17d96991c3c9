"""ctypes binding of the C-ABI library ``lib_lightgbm_b200.so`` (include/lgbcu.h, include/lgbm_b200.h).

There is no CPU fallback: if the shared library is missing the import fails loudly, and if it loads but no
CUDA device is usable every compute entry point returns -1 with a message (raised here as ``LgbCuError``).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib_lightgbm_b200.so")


class LgbCuError(RuntimeError):
    pass


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make lib` (or __graft_entry__.build()). "
            "lightgbmv1_b200 has no CPU fallback.")
    return C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)


lib = _load()


class FeatureMeta(C.Structure):
    _fields_ = [("num_bin", C.c_int32), ("most_freq_bin", C.c_int32), ("default_bin", C.c_int32),
                ("missing_type", C.c_int32), ("real_index", C.c_int32), ("bin_type", C.c_int32),
                ("penalty", C.c_double)]


class TreeParams(C.Structure):
    _fields_ = [("num_leaves", C.c_int32), ("max_depth", C.c_int32), ("min_data_in_leaf", C.c_int32),
                ("reserved", C.c_int32), ("min_sum_hessian_in_leaf", C.c_double), ("lambda_l1", C.c_double),
                ("lambda_l2", C.c_double), ("min_gain_to_split", C.c_double), ("max_delta_step", C.c_double),
                ("path_smooth", C.c_double)]


class SplitInfo(C.Structure):
    _fields_ = [("feature", C.c_int32), ("real_feature", C.c_int32), ("threshold", C.c_int32),
                ("default_left", C.c_int32), ("left_count", C.c_int32), ("right_count", C.c_int32),
                ("gain", C.c_double), ("left_output", C.c_double), ("right_output", C.c_double),
                ("left_sum_gradient", C.c_double), ("left_sum_hessian", C.c_double),
                ("right_sum_gradient", C.c_double), ("right_sum_hessian", C.c_double)]


class SplitRecord(C.Structure):
    _fields_ = [("leaf", C.c_int32), ("right_leaf", C.c_int32), ("feature", C.c_int32), ("real_feature", C.c_int32),
                ("threshold", C.c_int32), ("default_left", C.c_int32), ("left_count", C.c_int32),
                ("right_count", C.c_int32), ("gain", C.c_float), ("reserved", C.c_int32),
                ("left_output", C.c_double), ("right_output", C.c_double), ("left_sum_hessian", C.c_double),
                ("right_sum_hessian", C.c_double), ("left_sum_gradient", C.c_double),
                ("right_sum_gradient", C.c_double)]


MISSING_NONE, MISSING_ZERO, MISSING_NAN = 0, 1, 2
OBJ_NONE, OBJ_BINARY, OBJ_L2, OBJ_LAMBDARANK = 0, 1, 2, 3
HIST_AUTO, HIST_GLOBAL_ATOMIC, HIST_WARP_PRIVATE = 0, 1, 2

# every symbol include/lgbcu.h declares (tests check that the library exports all of them)
LGBCU_SYMBOLS = [
    "lgbcu_last_error", "lgbcu_multiply_score", "lgbcu_valid_multiply_score", "lgbcu_set_cat_params", "lgbcu_set_monotone", "lgbcu_tree_cat_sets", "lgbcu_set_tree_cat_sets", "lgbcu_device_count", "lgbcu_launch_count", "lgbcu_create", "lgbcu_destroy",
    "lgbcu_synchronize", "lgbcu_set_dataset", "lgbcu_set_dataset_columns", "lgbcu_set_tree_params",
    "lgbcu_set_hist_mode", "lgbcu_set_label", "lgbcu_set_weight", "lgbcu_set_query", "lgbcu_set_objective",
    "lgbcu_set_label_gain", "lgbcu_get_gradients", "lgbcu_set_gradients", "lgbcu_copy_gradients",
    "lgbcu_set_score", "lgbcu_add_score_const", "lgbcu_copy_score", "lgbcu_add_score_by_partition", "lgbcu_add_score_by_tree",
    "lgbcu_begin_tree", "lgbcu_hist_build", "lgbcu_hist_read", "lgbcu_num_total_bin", "lgbcu_hist_subtract",
    "lgbcu_find_best_split", "lgbcu_partition", "lgbcu_leaf_indices", "lgbcu_set_leaf_sums",
    "lgbcu_get_leaf_sums", "lgbcu_train_tree", "lgbcu_comm_unique_id", "lgbcu_comm_init", "lgbcu_comm_destroy", "lgbcu_comm_release_all", "lgbcu_bagging_reset", "lgbcu_bagging", "lgbcu_bagging_balanced", "lgbcu_copy_bag", "lgbcu_goss_scores", "lgbcu_goss_apply", "lgbcu_set_num_class", "lgbcu_set_class", "lgbcu_set_ova_label_weights", "lgbcu_renew_leaf_outputs", "lgbcu_bin_matrix", "lgbcu_binned_copy_to_host", "lgbcu_binned_device", "lgbcu_binned_free", "lgbcu_set_dataset_binned", "lgbcu_comm_allreduce_sum_f64",
    "lgbcu_profile_enable", "lgbcu_profile_read", "lgbcu_partition_rows_noalloc", "lgbcu_profile_phases", "lgbcu_pin_host_buffer", "lgbcu_unpin_host_buffer", "lgbcu_valid_create", "lgbcu_valid_set_label",
    "lgbcu_valid_set_weight", "lgbcu_valid_set_query", "lgbcu_valid_set_score", "lgbcu_valid_add_const", "lgbcu_valid_add_tree",
    "lgbcu_valid_copy_score", "lgbcu_eval_set_ndcg", "lgbcu_eval",
]

lib.lgbcu_last_error.restype = C.c_char_p
lib.lgbcu_launch_count.restype = C.c_int64


def check(rc: int) -> None:
    if rc != 0:
        raise LgbCuError(lib.lgbcu_last_error().decode())


def device_count() -> int:
    n = C.c_int(0)
    lib.lgbcu_device_count(C.byref(n))
    return n.value


def launch_count() -> int:
    return int(lib.lgbcu_launch_count())


def _ptr(a: np.ndarray, ctype):
    return a.ctypes.data_as(C.POINTER(ctype))


def make_feature_meta(metas) -> C.Array:
    """metas: iterable of dicts / tuples (num_bin, most_freq_bin, default_bin, missing_type[, real_index, penalty])."""
    metas = list(metas)
    arr = (FeatureMeta * len(metas))()
    for i, m in enumerate(metas):
        if isinstance(m, dict):
            arr[i].num_bin = m["num_bin"]; arr[i].most_freq_bin = m["most_freq_bin"]
            arr[i].default_bin = m["default_bin"]; arr[i].missing_type = m["missing_type"]
            arr[i].real_index = m.get("real_index", i); arr[i].penalty = m.get("penalty", 1.0)
        else:
            arr[i].num_bin, arr[i].most_freq_bin, arr[i].default_bin, arr[i].missing_type = m[:4]
            arr[i].real_index = m[4] if len(m) > 4 else i
            arr[i].penalty = m[5] if len(m) > 5 else 1.0
    return arr


def make_tree_params(num_leaves=31, max_depth=-1, min_data_in_leaf=20, min_sum_hessian_in_leaf=1e-3, lambda_l1=0.0,
                     lambda_l2=0.0, min_gain_to_split=0.0, max_delta_step=0.0, path_smooth=0.0) -> TreeParams:
    p = TreeParams()
    p.num_leaves = num_leaves; p.max_depth = max_depth; p.min_data_in_leaf = min_data_in_leaf
    p.min_sum_hessian_in_leaf = min_sum_hessian_in_leaf; p.lambda_l1 = lambda_l1; p.lambda_l2 = lambda_l2
    p.min_gain_to_split = min_gain_to_split; p.max_delta_step = max_delta_step; p.path_smooth = path_smooth
    return p


class CudaLearner:
    """Python mirror of one LgbCuContext: the device-resident tree learner + score + objective."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        check(lib.lgbcu_create(C.c_int(device), C.byref(self._h)))
        self.num_data = 0
        self.num_features = 0
        self.params = None

    def close(self):
        if self._h:
            lib.lgbcu_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- dataset / params
    def set_dataset(self, bins: np.ndarray, metas) -> None:
        bins = np.ascontiguousarray(bins, dtype=np.uint8)
        assert bins.ndim == 2
        self._metas = metas if isinstance(metas, C.Array) else make_feature_meta(metas)
        # byte columns: one per feature, two (low byte first) for a feature with more than 256 stored values
        self.num_data, self.num_features = bins.shape[0], len(self._metas)
        fs = sum(2 if 1 + m.num_bin - (1 if m.most_freq_bin == 0 else 0) > 256 else 1 for m in self._metas)
        if fs != bins.shape[1]:
            raise LgbCuError(f"set_dataset: {len(self._metas)} feature descriptors need {fs} byte columns, the matrix has {bins.shape[1]}")
        check(lib.lgbcu_set_dataset(self._h, C.c_int32(bins.shape[0]), C.c_int32(self.num_features),
                                    _ptr(bins, C.c_uint8), self._metas))

    def set_dataset_columns(self, columns, metas) -> None:
        cols = [np.ascontiguousarray(c, dtype=np.uint8) for c in columns]
        self._metas = metas if isinstance(metas, C.Array) else make_feature_meta(metas)
        self.num_data, self.num_features = len(cols[0]), len(cols)
        arr = (C.POINTER(C.c_uint8) * len(cols))(*[_ptr(c, C.c_uint8) for c in cols])
        check(lib.lgbcu_set_dataset_columns(self._h, C.c_int32(self.num_data), C.c_int32(self.num_features), arr,
                                            self._metas))

    def set_tree_params(self, params: TreeParams) -> None:
        self.params = params
        check(lib.lgbcu_set_tree_params(self._h, C.byref(params)))

    def set_hist_mode(self, mode: int) -> None:
        check(lib.lgbcu_set_hist_mode(self._h, C.c_int(mode)))

    def num_total_bin(self) -> int:
        n = C.c_int32()
        check(lib.lgbcu_num_total_bin(self._h, C.byref(n)))
        return n.value

    # ---- metadata / objective / score
    def set_label(self, label) -> None:
        a = np.ascontiguousarray(label, dtype=np.float32)
        check(lib.lgbcu_set_label(self._h, _ptr(a, C.c_float), C.c_int32(len(a))))

    def set_weight(self, weight) -> None:
        if weight is None:
            check(lib.lgbcu_set_weight(self._h, None, C.c_int32(0)))
            return
        a = np.ascontiguousarray(weight, dtype=np.float32)
        check(lib.lgbcu_set_weight(self._h, _ptr(a, C.c_float), C.c_int32(len(a))))

    def set_query(self, boundaries) -> None:
        a = np.ascontiguousarray(boundaries, dtype=np.int32)
        check(lib.lgbcu_set_query(self._h, _ptr(a, C.c_int32), C.c_int32(len(a) - 1)))

    def set_objective(self, kind: int, p0=0.0, p1=0.0, p2=0.0) -> None:
        check(lib.lgbcu_set_objective(self._h, C.c_int(kind), C.c_double(p0), C.c_double(p1), C.c_double(p2)))

    def set_label_gain(self, gains) -> None:
        a = np.ascontiguousarray(gains, dtype=np.float64)
        check(lib.lgbcu_set_label_gain(self._h, _ptr(a, C.c_double), C.c_int32(len(a))))

    def get_gradients(self) -> None:
        check(lib.lgbcu_get_gradients(self._h))

    def set_gradients(self, grad, hess) -> None:
        g = np.ascontiguousarray(grad, dtype=np.float32)
        h = np.ascontiguousarray(hess, dtype=np.float32)
        check(lib.lgbcu_set_gradients(self._h, _ptr(g, C.c_float), _ptr(h, C.c_float), C.c_int32(len(g))))

    def copy_gradients(self):
        g = np.empty(self.num_data, dtype=np.float32)
        h = np.empty(self.num_data, dtype=np.float32)
        check(lib.lgbcu_copy_gradients(self._h, _ptr(g, C.c_float), _ptr(h, C.c_float), C.c_int32(self.num_data)))
        return g, h

    def set_score(self, score) -> None:
        if score is None:
            check(lib.lgbcu_set_score(self._h, None, C.c_int32(0)))
            return
        a = np.ascontiguousarray(score, dtype=np.float64)
        check(lib.lgbcu_set_score(self._h, _ptr(a, C.c_double), C.c_int32(len(a))))

    def add_score_const(self, v: float) -> None:
        check(lib.lgbcu_add_score_const(self._h, C.c_double(v)))

    def copy_score(self) -> np.ndarray:
        a = np.empty(self.num_data, dtype=np.float64)
        check(lib.lgbcu_copy_score(self._h, _ptr(a, C.c_double), C.c_int32(self.num_data)))
        return a

    def add_score_by_partition(self, leaf_values) -> None:
        a = np.ascontiguousarray(leaf_values, dtype=np.float64)
        check(lib.lgbcu_add_score_by_partition(self._h, _ptr(a, C.c_double), C.c_int32(len(a))))

    def add_score_by_tree(self, split_feature_inner, threshold_bin, decision_type, left_child, right_child,
                          leaf_value, rows=None, rows_tag=0) -> None:
        lv = np.ascontiguousarray(leaf_value, dtype=np.float64)
        sf = np.ascontiguousarray(split_feature_inner, dtype=np.int32)
        tb = np.ascontiguousarray(threshold_bin, dtype=np.int32)
        dt = np.ascontiguousarray(decision_type, dtype=np.int8)
        lc = np.ascontiguousarray(left_child, dtype=np.int32)
        rc = np.ascontiguousarray(right_child, dtype=np.int32)
        r = None if rows is None else np.ascontiguousarray(rows, dtype=np.int32)
        check(lib.lgbcu_add_score_by_tree(self._h, C.c_int32(len(lv)), _ptr(sf, C.c_int32), _ptr(tb, C.c_int32),
                                          _ptr(dt, C.c_int8), _ptr(lc, C.c_int32), _ptr(rc, C.c_int32),
                                          _ptr(lv, C.c_double), None if r is None else _ptr(r, C.c_int32),
                                          C.c_int32(0 if r is None else len(r)), C.c_int64(rows_tag)))

    # ---- tree steps
    def begin_tree(self, bag_indices=None, feature_used=None) -> None:
        bi = None if bag_indices is None else np.ascontiguousarray(bag_indices, dtype=np.int32)
        fu = None if feature_used is None else np.ascontiguousarray(feature_used, dtype=np.int8)
        check(lib.lgbcu_begin_tree(self._h, None if bi is None else _ptr(bi, C.c_int32),
                                   C.c_int32(0 if bi is None else len(bi)),
                                   None if fu is None else _ptr(fu, C.c_int8)))

    # ---- bagging on the device
    DEVICE_BAG = C.cast(C.c_void_p(-1), C.POINTER(C.c_int32))

    def bagging_reset(self, seed: int) -> None:
        check(lib.lgbcu_bagging_reset(self._h, C.c_int32(seed)))

    def bagging(self, fraction: float) -> int:
        n = C.c_int32()
        check(lib.lgbcu_bagging(self._h, C.c_double(fraction), C.byref(n)))
        return n.value

    def copy_bag(self, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.int32)
        check(lib.lgbcu_copy_bag(self._h, _ptr(out, C.c_int32), C.c_int32(n)))
        return out

    def begin_tree_device_bag(self, feature_used=None) -> None:
        fu = None if feature_used is None else np.ascontiguousarray(feature_used, dtype=np.int8)
        check(lib.lgbcu_begin_tree(self._h, self.DEVICE_BAG, C.c_int32(0), None if fu is None else _ptr(fu, C.c_int8)))

    def hist_build(self, leaf: int) -> None:
        check(lib.lgbcu_hist_build(self._h, C.c_int32(leaf)))

    def hist_read(self, leaf: int) -> np.ndarray:
        out = np.empty(2 * self.num_total_bin(), dtype=np.float64)
        check(lib.lgbcu_hist_read(self._h, C.c_int32(leaf), _ptr(out, C.c_double), C.c_int64(len(out))))
        return out

    def hist_subtract(self, parent_leaf: int, smaller_leaf: int) -> None:
        check(lib.lgbcu_hist_subtract(self._h, C.c_int32(parent_leaf), C.c_int32(smaller_leaf)))

    def find_best_split(self, leaf: int) -> SplitInfo:
        s = SplitInfo()
        check(lib.lgbcu_find_best_split(self._h, C.c_int32(leaf), C.byref(s)))
        return s

    def partition(self, leaf, feature, threshold, default_left, right_leaf) -> int:
        n = C.c_int32()
        check(lib.lgbcu_partition(self._h, C.c_int32(leaf), C.c_int32(feature), C.c_int32(threshold),
                                  C.c_int32(int(default_left)), C.c_int32(right_leaf), C.byref(n)))
        return n.value

    def leaf_indices(self, leaf: int) -> np.ndarray:
        n = C.c_int32()
        check(lib.lgbcu_leaf_indices(self._h, C.c_int32(leaf), None, C.c_int32(0), C.byref(n)))
        out = np.empty(max(n.value, 1), dtype=np.int32)
        check(lib.lgbcu_leaf_indices(self._h, C.c_int32(leaf), _ptr(out, C.c_int32), C.c_int32(len(out)), C.byref(n)))
        return out[:n.value]

    def set_leaf_sums(self, leaf, sum_g, sum_h, output=0.0) -> None:
        check(lib.lgbcu_set_leaf_sums(self._h, C.c_int32(leaf), C.c_double(sum_g), C.c_double(sum_h),
                                      C.c_double(output)))

    def get_leaf_sums(self, leaf):
        g, h, n = C.c_double(), C.c_double(), C.c_int32()
        check(lib.lgbcu_get_leaf_sums(self._h, C.c_int32(leaf), C.byref(g), C.byref(h), C.byref(n)))
        return g.value, h.value, n.value

    def train_tree(self):
        L = self.params.num_leaves
        recs = (SplitRecord * max(L - 1, 1))()
        n = C.c_int32()
        check(lib.lgbcu_train_tree(self._h, recs, C.byref(n)))
        return [recs[i] for i in range(n.value)]

    def synchronize(self) -> None:
        check(lib.lgbcu_synchronize(self._h))

    # ---- comm / profile
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = (C.c_uint8 * 128)()
        check(lib.lgbcu_comm_unique_id(buf))
        return bytes(buf)

    def comm_init(self, uid: bytes, rank: int, world: int) -> None:
        buf = (C.c_uint8 * 128).from_buffer_copy(uid)
        check(lib.lgbcu_comm_init(self._h, buf, C.c_int32(rank), C.c_int32(world)))

    def profile_enable(self, on: bool) -> None:
        check(lib.lgbcu_profile_enable(self._h, C.c_int(1 if on else 0)))

    def profile_phases(self):
        a = (C.c_double * 8)()
        check(lib.lgbcu_profile_phases(self._h, a))
        return list(a)

    def profile_read(self, reset=False):
        ms, calls, nbytes = C.c_double(), C.c_int64(), C.c_double()
        check(lib.lgbcu_profile_read(self._h, C.byref(ms), C.byref(calls), C.byref(nbytes), C.c_int(1 if reset else 0)))
        return ms.value, calls.value, nbytes.value
