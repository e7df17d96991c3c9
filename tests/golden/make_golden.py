"""Generates tests/golden/golden_v1.npz from the UNMODIFIED reference (oracle/_ref/lib_lightgbm.so + shim, built by
oracle/Makefile from /root/reference).  Run in the build container:  python tests/golden/make_golden.py
The fixture lets the GPU box (where /root/reference does not exist) check binning, single trees and short training
runs against the reference even if oracle/_ref did not travel."""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import capi  # noqa: E402
import helpers as H  # noqa: E402


def mixed(seed, n):
    rng = np.random.default_rng(seed)
    cols = [rng.standard_normal(n), np.maximum(0, rng.standard_normal(n)), rng.choice([0.0, 1.087, 2.173], size=n),
            np.where(rng.random(n) < 0.1, np.nan, rng.standard_normal(n)), np.round(rng.standard_normal(n), 1),
            rng.integers(0, 40, size=n).astype(np.float64), np.where(rng.random(n) < 0.8, 5.0, rng.standard_normal(n)),
            -np.abs(rng.standard_normal(n)) - 0.5]
    return np.stack(cols, axis=1).astype(np.float32)


def main():
    C.CDLL(H.REF_SO, mode=C.RTLD_GLOBAL)
    shim = C.CDLL(H.REF_SHIM_SO)
    ref = capi.reference()
    out = {}
    n = 3000
    X = mixed(1, n)
    rng = np.random.default_rng(2)
    y = ((np.nan_to_num(X[:, 0]) + 0.7 * (X[:, 2] > 1) - 0.5 * np.nan_to_num(X[:, 3]) + 0.3 * rng.standard_normal(n)) > 0).astype(np.float32)
    out["X"], out["y"] = X, y
    # ---- binning
    p_bin = "max_bin=63 min_data_in_leaf=20 verbosity=-1"
    ds = ref.dataset_from_mat(X, p_bin)
    nf = shim.ref_num_inner_features(ds)
    metas, bounds, cols = [], [], []
    for f in range(nf):
        m = (C.c_int32 * 5)(); b = (C.c_double * 300)()
        shim.ref_feature_info(ds, C.c_int(f), m, b, C.c_int(300))
        metas.append(list(m)); bounds.append(np.array(b[:m[0]]))
        col = np.empty(n, np.uint8)
        shim.ref_feature_stored_column(ds, C.c_int(f), col.ctypes.data_as(C.POINTER(C.c_uint8)))
        cols.append(col)
    out["bin_params"] = np.array(p_bin)
    out["bin_metas"] = np.array(metas, dtype=np.int32)
    out["bin_bounds"] = np.concatenate(bounds)
    out["bin_stored"] = np.stack(cols, axis=1)
    # ---- one tree from given gradients (custom objective), learning_rate 1
    g = (rng.standard_normal(n) + 0.8 * np.sign(np.nan_to_num(X[:, 0])) - 0.5 * (X[:, 2] > 1)).astype(np.float32)
    h = (rng.random(n) * 0.25 + 0.05).astype(np.float32)
    p_tree = "objective=none num_leaves=15 learning_rate=1.0 min_data_in_leaf=20 max_bin=63 verbosity=-1 num_threads=2 force_col_wise=true"
    ds2 = ref.dataset_from_mat(X, p_tree); ref.set_field(ds2, "label", np.zeros(n, np.float32))
    b2 = ref.booster_create(ds2, p_tree); ref.update_custom(b2, g, h)
    out["tree_g"], out["tree_h"], out["tree_params"] = g, h, np.array(p_tree)
    out["tree_model"] = np.array(ref.model_string(b2).split("feature_importances")[0])
    # ---- short training runs
    p_binary = "objective=binary num_leaves=15 learning_rate=0.1 min_data_in_leaf=20 max_bin=63 verbosity=-1"
    _, bb, _ = capi.train(ref, X, y, p_binary + " num_threads=2 force_col_wise=true", 6)
    out["binary_params"] = np.array(p_binary)
    out["binary_model"] = np.array(ref.model_string(bb).split("feature_importances")[0])
    out["binary_pred"] = ref.get_predict(bb, n)
    sizes = np.clip(rng.poisson(50, size=60), 1, 200).astype(np.int32)
    sizes[-1] += n - int(sizes.sum()) if int(sizes.sum()) < n else 0
    while int(sizes.sum()) > n:
        sizes[np.argmax(sizes)] -= 1
    lab = np.digitize(np.nan_to_num(X[:, 0]) + 0.5 * rng.standard_normal(n), [0.2, 1.0, 1.8, 2.4]).astype(np.float32)
    p_rank = "objective=lambdarank num_leaves=15 learning_rate=0.1 min_data_in_leaf=20 max_bin=63 verbosity=-1"
    _, br, _ = capi.train(ref, X, lab, p_rank + " num_threads=2 force_col_wise=true", 5, group=sizes)
    out["rank_params"], out["rank_sizes"], out["rank_label"] = np.array(p_rank), sizes, lab
    out["rank_model"] = np.array(ref.model_string(br).split("feature_importances")[0])
    out["rank_pred"] = ref.get_predict(br, n)
    np.savez_compressed(os.path.join(HERE, "golden_v1.npz"), **out)
    print("wrote", os.path.join(HERE, "golden_v1.npz"), os.path.getsize(os.path.join(HERE, "golden_v1.npz")), "bytes")


if __name__ == "__main__":
    main()
