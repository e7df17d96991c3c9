"""lgbcu_partition_rows_noalloc: the allocation-free partition entry point.

Callers that cannot allocate device memory (they run inside a GPU memory arena owned by some other component, or
on a path where cudaMalloc's device-wide synchronisation is not acceptable) bring their own input and output
buffers; the working memory is a fixed array inside the library.  Here the test plays that caller: it owns every
buffer (torch tensors), calls the entry point twice in a row -- the second time with fewer rows and another split,
so that anything left behind in the shared array by the first call would show -- and requires bit-identical output
to the regular entry point lgbcu_partition (index work: the tolerance is zero).

Collected last on purpose (file name): a failure here cannot hide the result of any other test.
"""
import ctypes as C

import numpy as np
import pytest

import helpers as H
from lightgbmv1_b200 import _native as N

pytestmark = pytest.mark.gpu


def _reference_partition(bins, metas, rows, feature, threshold, default_left):
    """the existing entry point: lgbcu_partition on a learner whose root is exactly `rows`"""
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=4))
    n = bins.shape[0]
    L.set_gradients(np.zeros(n, np.float32), np.ones(n, np.float32))
    L.begin_tree(bag_indices=rows)
    left = L.partition(0, feature, threshold, default_left, 1)
    out = np.concatenate([L.leaf_indices(0), L.leaf_indices(1)])
    L.close()
    return left, out


def test_noalloc_partition_matches_lgbcu_partition_twice_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    rng = np.random.default_rng(77)
    n, F = 300_000, 11
    metas = H.random_feature_metas(rng, F, max_bin=200, kinds=[0, 2, 4, 5, 7, 1])
    bins = H.random_bins(rng, n, metas)
    stride = 16
    padded = np.zeros((n, stride), dtype=np.uint8)
    padded[:, :F] = bins
    d_bins = torch.from_numpy(padded).cuda()
    fm = N.make_feature_meta(metas)
    fn = N.lib.lgbcu_partition_rows_noalloc
    fn.restype = C.c_int
    # call 1: a large subset; call 2: a smaller, different subset with another feature / threshold / default side
    calls = [
        (np.sort(rng.choice(n, size=250_003, replace=False)).astype(np.int32), 1, metas[1][0] // 2, 1),
        (np.sort(rng.choice(n, size=70_001, replace=False)).astype(np.int32), 3, metas[3][0] // 3, 0),
    ]
    for rows, feature, threshold, default_left in calls:
        d_in = torch.from_numpy(rows).cuda()
        d_out = torch.full((len(rows),), -1, dtype=torch.int32, device="cuda")
        d_left = torch.full((1,), -1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        rc = fn(C.c_void_p(d_bins.data_ptr()), C.c_int64(stride), C.c_int32(feature), C.byref(fm[feature]),
                C.c_int32(threshold), C.c_int32(default_left), C.c_void_p(d_in.data_ptr()), C.c_int32(len(rows)),
                C.c_void_p(d_out.data_ptr()), C.c_void_p(d_left.data_ptr()), C.c_void_p(0))
        assert rc == 0, N.lib.lgbcu_last_error().decode()
        torch.cuda.synchronize()
        got_left = int(d_left.cpu()[0])
        got = d_out.cpu().numpy()
        want_left, want = _reference_partition(bins, metas, rows, feature, threshold, default_left)
        assert 0 < want_left < len(rows)
        assert got_left == want_left
        assert np.array_equal(got, want)
        # and first-hand against the CPU oracle (DenseBin::SplitInner restated, oracle/oracle.c) on the same rows
        orc = H.load_oracle()
        lte = np.empty(len(rows), dtype=np.int32); gt = np.empty(len(rows), dtype=np.int32)
        o_left = orc.orc_partition(H.fptr(bins, C.c_uint8), C.c_int(F), fm, C.c_int(feature), C.c_int(threshold),
                                   C.c_int(default_left), H.fptr(rows, C.c_int32), C.c_int(len(rows)),
                                   H.fptr(lte, C.c_int32), H.fptr(gt, C.c_int32))
        assert got_left == o_left
        assert np.array_equal(got, np.concatenate([lte[:o_left], gt[:len(rows) - o_left]]))


def test_noalloc_partition_rejects_oversize_and_bad_arguments():
    torch = pytest.importorskip("torch")
    fn = N.lib.lgbcu_partition_rows_noalloc
    fn.restype = C.c_int
    fm = N.make_feature_meta([(10, 0, 0, 0)])
    buf = torch.zeros(64, dtype=torch.int32, device="cuda")
    b8 = torch.zeros(64, dtype=torch.uint8, device="cuda")
    too_many = (1 << 22) + 1
    rc = fn(C.c_void_p(b8.data_ptr()), C.c_int64(16), C.c_int32(0), C.byref(fm[0]), C.c_int32(3), C.c_int32(0),
            C.c_void_p(buf.data_ptr()), C.c_int32(too_many), C.c_void_p(buf.data_ptr() + 128), C.c_void_p(buf.data_ptr() + 64),
            C.c_void_p(0))
    assert rc == -1
    assert "exceed" in N.lib.lgbcu_last_error().decode()
    rc = fn(None, C.c_int64(16), C.c_int32(0), C.byref(fm[0]), C.c_int32(3), C.c_int32(0), C.c_void_p(buf.data_ptr()),
            C.c_int32(4), C.c_void_p(buf.data_ptr() + 128), C.c_void_p(buf.data_ptr() + 64), C.c_void_p(0))
    assert rc == -1
    # partially overlapping input / output ranges are refused, not silently corrupted
    rc = fn(C.c_void_p(b8.data_ptr()), C.c_int64(16), C.c_int32(0), C.byref(fm[0]), C.c_int32(3), C.c_int32(0),
            C.c_void_p(buf.data_ptr()), C.c_int32(8), C.c_void_p(buf.data_ptr() + 16), C.c_void_p(buf.data_ptr() + 192),
            C.c_void_p(0))
    assert rc == -1
    assert "overlap" in N.lib.lgbcu_last_error().decode()
