"""Device-side bagging (lgbcu_bagging) against a numpy restatement of GBDT::BaggingHelper (src/boosting/gbdt.cpp:
162-181): one LCG per 1024-row block (include/LightGBM/utils/random.h:20-112), seeded bagging_seed + block; a row is in
the bag when NextFloat() < bagging_fraction; the generators keep their state from one re-bagging to the next.  The list
must be bit-identical: in-bag rows ascending, then out-of-bag rows ascending."""
import numpy as np
import pytest

import helpers as H
from lightgbmv1_b200 import _native as N

pytestmark = pytest.mark.gpu


class BlockLCG:
    def __init__(self, n, seed):
        self.n = n
        self.nblock = (n + 1023) // 1024
        self.x = (np.uint64(seed) + np.arange(self.nblock, dtype=np.uint64)) & np.uint64(0xffffffff)

    def draw(self, fraction):
        in_bag = np.zeros(self.nblock * 1024, dtype=bool)
        x = self.x.copy()
        rows_in_block = np.minimum(1024, self.n - 1024 * np.arange(self.nblock))
        for i in range(1024):   # a generator draws once per existing row of its block, in row order
            live = i < rows_in_block
            nx = (np.uint64(214013) * x + np.uint64(2531011)) & np.uint64(0xffffffff)
            x = np.where(live, nx, x)
            r = ((x >> np.uint64(16)) & np.uint64(0x7fff)).astype(np.float32) / np.float32(32768.0)
            in_bag[i::1024][:self.nblock] = live & (r.astype(np.float64) < fraction)
        self.x = x
        in_bag = in_bag.reshape(self.nblock, 1024).reshape(-1)[:self.n]
        rows = np.arange(self.n, dtype=np.int32)
        return np.concatenate([rows[in_bag], rows[~in_bag]]), int(in_bag.sum())


@pytest.mark.parametrize("n,seed,fraction", [(7000, 3, 0.8), (100_000, 12345, 0.5), (1024, 0, 0.999), (1025, 7, 0.1), (300_001, 99, 0.632)])
def test_device_bag_is_bit_identical_to_the_reference_generator(n, seed, fraction):
    rng = np.random.default_rng(n)
    metas = H.random_feature_metas(rng, 3, kinds=[0])
    bins = H.random_bins(rng, n, metas)
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.bagging_reset(seed)
    ref = BlockLCG(n, seed)
    for _ in range(3):   # three re-baggings in a row: the generator states carry over
        cnt = L.bagging(fraction)
        want, want_cnt = ref.draw(fraction)
        assert cnt == want_cnt
        assert np.array_equal(L.copy_bag(n), want)
    L.close()


def test_tree_on_device_bag_equals_tree_on_uploaded_bag():
    rng = np.random.default_rng(5)
    n, F = 40_000, 12
    metas = H.random_feature_metas(rng, F)
    bins = H.random_bins(rng, n, metas)
    g = rng.standard_normal(n).astype(np.float32)
    h = (rng.random(n).astype(np.float32) * 0.25 + 0.01).astype(np.float32)
    recs = []
    for device in (True, False):
        L = N.CudaLearner()
        L.set_dataset(bins, metas)
        L.set_tree_params(N.make_tree_params(num_leaves=31, min_data_in_leaf=20))
        L.set_gradients(g, h)
        L.bagging_reset(11)
        cnt = L.bagging(0.7)
        if device:
            L.begin_tree_device_bag()
        else:
            L.begin_tree(bag_indices=L.copy_bag(n)[:cnt])
        recs.append(H.records_to_tuples(L.train_tree()))
        L.close()
    assert recs[0] == recs[1] and len(recs[0]) >= 5
