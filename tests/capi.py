"""One ctypes driver for the LGBM_* C API, usable with BOTH libraries: the reference's lib_lightgbm.so
(oracle/_ref, test infrastructure) and the product's lib_lightgbm_b200.so.  The tests read like the
reference's own tests/c_api_test/test_.py because the entry points are the same."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "lib_lightgbm.so")
OURS_SO = os.environ.get("LGBM_B200_LIB", os.path.join(ROOT, "lightgbmv1_b200", "lib_lightgbm_b200.so"))  # (override: debug builds, tools/build_dbg.sh)

DTYPE_F32, DTYPE_F64, DTYPE_I32 = 0, 1, 2


class LGBMError(RuntimeError):
    pass


class LGBMLib:
    def __init__(self, path, is_reference):
        self.lib = C.CDLL(path)
        self.is_reference = is_reference
        self.lib.LGBM_GetLastError.restype = C.c_char_p

    def check(self, rc):
        if rc != 0:
            raise LGBMError(self.lib.LGBM_GetLastError().decode())

    def dataset_from_mat(self, X, params="", reference=None):
        X = np.ascontiguousarray(X)
        assert X.dtype in (np.float32, np.float64) and X.ndim == 2
        h = C.c_void_p()
        self.check(self.lib.LGBM_DatasetCreateFromMat(
            X.ctypes.data_as(C.c_void_p), C.c_int(DTYPE_F32 if X.dtype == np.float32 else DTYPE_F64),
            C.c_int32(X.shape[0]), C.c_int32(X.shape[1]), C.c_int(1), C.c_char_p(params.encode()),
            reference, C.byref(h)))
        return h

    def dataset_from_sparse(self, S, params="", reference=None):
        """S: scipy.sparse csr_matrix or csc_matrix (LGBM_DatasetCreateFromCSR / LGBM_DatasetCreateFromCSC)"""
        csr = S.format == "csr"
        assert csr or S.format == "csc"
        ptr = np.ascontiguousarray(S.indptr); idx = np.ascontiguousarray(S.indices, dtype=np.int32)
        val = np.ascontiguousarray(S.data)
        assert ptr.dtype in (np.int32, np.int64) and val.dtype in (np.float32, np.float64)
        h = C.c_void_p()
        fn = self.lib.LGBM_DatasetCreateFromCSR if csr else self.lib.LGBM_DatasetCreateFromCSC
        self.check(fn(ptr.ctypes.data_as(C.c_void_p), C.c_int(DTYPE_I32 if ptr.dtype == np.int32 else 3),
                      idx.ctypes.data_as(C.POINTER(C.c_int32)), val.ctypes.data_as(C.c_void_p),
                      C.c_int(DTYPE_F32 if val.dtype == np.float32 else DTYPE_F64), C.c_int64(len(ptr)), C.c_int64(len(val)),
                      C.c_int64(S.shape[1] if csr else S.shape[0]), C.c_char_p(params.encode()), reference, C.byref(h)))
        return h

    def predict_csr(self, b, S, raw=False, num_class=1):
        ptr = np.ascontiguousarray(S.indptr); idx = np.ascontiguousarray(S.indices, dtype=np.int32)
        val = np.ascontiguousarray(S.data, dtype=np.float64)
        out = np.empty(S.shape[0] * num_class, dtype=np.float64); n = C.c_int64()
        self.check(self.lib.LGBM_BoosterPredictForCSR(
            b, ptr.ctypes.data_as(C.c_void_p), C.c_int(DTYPE_I32 if ptr.dtype == np.int32 else 3),
            idx.ctypes.data_as(C.POINTER(C.c_int32)), val.ctypes.data_as(C.c_void_p), C.c_int(DTYPE_F64), C.c_int64(len(ptr)),
            C.c_int64(len(val)), C.c_int64(S.shape[1]), C.c_int(1 if raw else 0), C.c_int(0), C.c_int(-1), C.c_char_p(b""),
            C.byref(n), out.ctypes.data_as(C.POINTER(C.c_double))))
        assert n.value == len(out)
        return out

    def set_field(self, ds, name, arr):
        arr = np.ascontiguousarray(arr)
        t = {np.dtype(np.float32): DTYPE_F32, np.dtype(np.float64): DTYPE_F64, np.dtype(np.int32): DTYPE_I32}[arr.dtype]
        self.check(self.lib.LGBM_DatasetSetField(ds, C.c_char_p(name.encode()), arr.ctypes.data_as(C.c_void_p),
                                                 C.c_int(len(arr)), C.c_int(t)))

    def dataset_free(self, ds):
        self.check(self.lib.LGBM_DatasetFree(ds))

    def booster_create(self, ds, params):
        h = C.c_void_p()
        self.check(self.lib.LGBM_BoosterCreate(ds, C.c_char_p(params.encode()), C.byref(h)))
        return h

    def booster_free(self, b):
        self.check(self.lib.LGBM_BoosterFree(b))

    def update(self, b):
        fin = C.c_int(0)
        self.check(self.lib.LGBM_BoosterUpdateOneIter(b, C.byref(fin)))
        return fin.value

    def update_custom(self, b, grad, hess):
        fin = C.c_int(0)
        g = np.ascontiguousarray(grad, dtype=np.float32); h = np.ascontiguousarray(hess, dtype=np.float32)
        self.check(self.lib.LGBM_BoosterUpdateOneIterCustom(b, g.ctypes.data_as(C.POINTER(C.c_float)),
                                                            h.ctypes.data_as(C.POINTER(C.c_float)), C.byref(fin)))
        return fin.value

    def get_predict(self, b, n, data_idx=0, out=None):
        """`out`: a caller-owned float64 buffer of n entries (e.g. one whose pages are already mapped)"""
        if out is None:
            out = np.empty(n, dtype=np.float64)
        assert out.dtype == np.float64 and out.size == n and out.flags.c_contiguous
        ln = C.c_int64(0)
        self.check(self.lib.LGBM_BoosterGetPredict(b, C.c_int(data_idx), C.byref(ln), out.ctypes.data_as(C.POINTER(C.c_double))))
        assert ln.value == n
        return out

    def add_valid(self, b, ds):
        self.check(self.lib.LGBM_BoosterAddValidData(b, ds))

    def eval_names(self, b):
        n = C.c_int(0)
        self.check(self.lib.LGBM_BoosterGetEvalCounts(b, C.byref(n)))
        bufs = [C.create_string_buffer(255) for _ in range(n.value)]
        ptrs = (C.c_char_p * n.value)(*map(C.addressof, bufs))
        got, need = C.c_int(0), C.c_size_t(0)
        self.check(self.lib.LGBM_BoosterGetEvalNames(b, C.c_int(n.value), C.byref(got), C.c_size_t(255), C.byref(need), ptrs))
        return [bufs[i].value.decode() for i in range(got.value)]

    def get_eval(self, b, data_idx):
        n = C.c_int(0)
        self.check(self.lib.LGBM_BoosterGetEvalCounts(b, C.byref(n)))
        out = np.zeros(max(n.value, 1), dtype=np.float64)
        ln = C.c_int(0)
        self.check(self.lib.LGBM_BoosterGetEval(b, C.c_int(data_idx), C.byref(ln), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out[:ln.value]

    def predict_mat(self, b, X, raw=False, num_iteration=-1, num_class=1):
        X = np.ascontiguousarray(X)
        out = np.empty(X.shape[0] * num_class, dtype=np.float64)
        ln = C.c_int64(0)
        self.check(self.lib.LGBM_BoosterPredictForMat(
            b, X.ctypes.data_as(C.c_void_p), C.c_int(DTYPE_F32 if X.dtype == np.float32 else DTYPE_F64),
            C.c_int32(X.shape[0]), C.c_int32(X.shape[1]), C.c_int(1), C.c_int(1 if raw else 0), C.c_int(0),
            C.c_int(num_iteration), C.c_char_p(b""), C.byref(ln), out.ctypes.data_as(C.POINTER(C.c_double))))
        assert ln.value == out.size
        return out

    def model_string(self, b):
        ln = C.c_int64(0)
        buf = C.create_string_buffer(1 << 20)
        self.check(self.lib.LGBM_BoosterSaveModelToString(b, C.c_int(0), C.c_int(-1), C.c_int(0), C.c_int64(len(buf)),
                                                          C.byref(ln), buf))
        if ln.value > len(buf):
            buf = C.create_string_buffer(ln.value)
            self.check(self.lib.LGBM_BoosterSaveModelToString(b, C.c_int(0), C.c_int(-1), C.c_int(0),
                                                              C.c_int64(len(buf)), C.byref(ln), buf))
        return buf.value.decode()

    def load_model_from_string(self, s):
        """reference only: proves our model text is loadable by the unmodified reference"""
        h = C.c_void_p(); it = C.c_int(0)
        self.check(self.lib.LGBM_BoosterLoadModelFromString(C.c_char_p(s.encode()), C.byref(it), C.byref(h)))
        return h


def reference():
    return LGBMLib(REF_SO, True)


def ours():
    return LGBMLib(OURS_SO, False)


def parse_trees(model_text):
    """-> list of dicts (one per tree) with the per-node arrays of the model file"""
    trees = []
    cur = None
    for line in model_text.splitlines():
        if line.startswith("Tree="):
            cur = {}
            trees.append(cur)
        elif line.startswith("end of trees"):
            cur = None
        elif cur is not None and "=" in line:
            k, v = line.split("=", 1)
            cur[k] = v.split(" ") if v else []
    return trees


def header_lines(model_text):
    out = {}
    for line in model_text.splitlines():
        if line.startswith("Tree="):
            break
        if "=" in line:
            k, v = line.split("=", 1)
            out[k] = v
    return out


def train(lib, X, y, params, iters, weight=None, group=None, init_score=None, ds_params=None):
    ds = lib.dataset_from_mat(X, ds_params if ds_params is not None else params)
    lib.set_field(ds, "label", np.asarray(y, dtype=np.float32))
    if weight is not None:
        lib.set_field(ds, "weight", np.asarray(weight, dtype=np.float32))
    if group is not None:
        lib.set_field(ds, "group", np.asarray(group, dtype=np.int32))
    if init_score is not None:
        lib.set_field(ds, "init_score", np.asarray(init_score, dtype=np.float64))
    b = lib.booster_create(ds, params)
    done = 0
    for _ in range(iters):
        if lib.update(b):
            break
        done += 1
    return ds, b, done


def compare_models(text_a, text_b, rtol_value=1e-9, allow_near_ties=True, stop_at_first=True):
    """Structure (split_feature, threshold, decision_type, children, leaf_count) must be identical tree by tree;
    leaf values / weights within rtol.  Returns (num_trees_identical, num_trees, first_mismatch_description).
    stop_at_first=True counts the identical PREFIX (a differing tree changes the gradients of every later one);
    False counts every identical tree (a near-tie that picks an equivalent threshold leaves the later trees equal)."""
    ta, tb = parse_trees(text_a), parse_trees(text_b)
    n = min(len(ta), len(tb))
    same = 0
    first = None
    for i in range(n):
        a, b = ta[i], tb[i]
        ok = True
        for k in ("num_leaves", "num_cat", "split_feature", "threshold", "decision_type", "left_child", "right_child",
                  "leaf_count", "internal_count", "cat_boundaries", "cat_threshold"):
            if a.get(k) != b.get(k):
                ok = False
                if first is None:
                    first = f"tree {i} key {k}"
                break
        if ok:
            for k in ("leaf_value", "leaf_weight", "split_gain", "internal_value", "internal_weight"):
                va = np.array(a.get(k, []), dtype=np.float64); vb = np.array(b.get(k, []), dtype=np.float64)
                tol = 1e-5 if k in ("split_gain", "internal_value", "internal_weight") else rtol_value
                if va.shape != vb.shape or not np.allclose(va, vb, rtol=tol, atol=1e-12):
                    ok = False
                    if first is None:
                        first = f"tree {i} key {k} max diff {np.max(np.abs(va - vb)) if va.shape == vb.shape else 'shape'}"
                    break
        if ok:
            same += 1
        elif stop_at_first:
            break  # later trees depend on this one
    return same, n, first
