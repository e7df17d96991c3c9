"""Shared test plumbing: oracle loader (TEST INFRASTRUCTURE), synthetic binned data, tolerances."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "lib_lightgbm.so")
REF_SHIM_SO = os.path.join(ROOT, "oracle", "_ref", "libref_shim.so")

from lightgbmv1_b200._native import (FeatureMeta, SplitInfo, SplitRecord, TreeParams, make_feature_meta,  # noqa: E402
                                     make_tree_params)

# tolerances (north_star): histogram sums 1e-6 relative; integer / index work bit-exact
HIST_RTOL = 1e-6
# Histograms are exact 64-bit fixed-point sums (per-row rounding <= 2^-39 of max|g| at 10^7 rows): sums, gains and
# leaf values derived from them agree with the reference's float -> double sums to about 1e-10; VALUE_RTOL is the bar.
VALUE_RTOL = 1e-9


def load_oracle() -> C.CDLL:
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
            os.path.join(ROOT, "oracle", "oracle.c")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"], stdout=subprocess.DEVNULL)
    lib = C.CDLL(ORACLE_SO)
    lib.orc_total_bin.restype = C.c_int
    lib.orc_partition.restype = C.c_int
    lib.orc_train_tree.restype = C.c_int
    return lib


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def fptr(a, ct):
    return a.ctypes.data_as(C.POINTER(ct))


def random_feature_metas(rng, num_features, max_bin=255, kinds=None):
    """Feature metas covering the (missing_type, most_freq_bin, default_bin) cases of DenseBin::SplitInner."""
    metas = []
    for f in range(num_features):
        kind = kinds[f % len(kinds)] if kinds else int(rng.integers(0, 8))
        nb = int(rng.integers(3, max_bin + 1))
        if kind == 0:      # dense continuous, no missing: mfb = default bin somewhere in the middle
            d = int(rng.integers(1, nb - 1)); m = (nb, d, d, 0)
        elif kind == 1:    # non-negative feature: zero is bin 0 = mfb (offset 1)
            m = (nb, 0, 0, 0)
        elif kind == 2:    # NaN missing, mfb = default in the middle
            d = int(rng.integers(1, nb - 1)); m = (nb, d, d, 2)
        elif kind == 3:    # NaN missing, mfb = 0
            m = (nb, 0, 0, 2)
        elif kind == 4:    # zero-as-missing, mfb == default
            d = int(rng.integers(1, nb - 1)); m = (nb, d, d, 1)
        elif kind == 5:    # zero-as-missing, mfb != default (sparse-rate >= 0.7 case)
            d = int(rng.integers(1, nb - 1)); mf = int(rng.integers(0, nb)); mf = mf if mf != d else (d + 1) % nb
            m = (nb, mf, d, 1)
        elif kind == 6:    # low cardinality (3 levels)
            m = (3, 0, 0, 0)
        else:              # NaN missing where the NaN bin itself is the most frequent one
            m = (nb, nb - 1, int(rng.integers(0, nb - 1)), 2)
        metas.append(m)
    return metas


def random_bins(rng, n, metas, skew=True):
    """Stored values (0 = most-frequent bin, else bin + 1 - [mfb==0]) for random bin draws."""
    F = len(metas)
    out = np.zeros((n, F), dtype=np.uint8)
    for f, m in enumerate(metas):
        nb, mfb = m[0], m[1]
        if skew:
            p = rng.random(nb) + 0.05
            p[mfb] += 1.5
            p /= p.sum()
            b = rng.choice(nb, size=n, p=p)
        else:
            b = rng.integers(0, nb, size=n)
        stored = np.where(b == mfb, 0, b + 1 - (1 if mfb == 0 else 0))
        out[:, f] = stored.astype(np.uint8)
    return out


def nslots(m):
    return 1 + m[0] - (1 if m[1] == 0 else 0)


def hist_bases(metas):
    base, acc = [], 0
    for m in metas:
        base.append(acc)
        acc += nslots(m)
    return base, acc


def rel_err(a, b, floor):
    """max |a-b| / max(|b|, floor)"""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0


def records_to_tuples(recs):
    return [(r.leaf, r.right_leaf, r.feature, r.threshold, r.default_left, r.left_count, r.right_count) for r in recs]
