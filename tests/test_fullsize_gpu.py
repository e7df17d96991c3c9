"""BASELINE.json full-size configs, checked through size-independent properties (the oracle would take minutes at
these sizes): histogram linearity / checksums, partition = stable disjoint cover, tree bookkeeping consistency, and
invariance of the whole tree to the histogram code path (fixed-point sums are order independent, so any kernel must
give the same bits)."""
import numpy as np
import pytest

import helpers as H
from lightgbmv1_b200 import _native as N

pytestmark = pytest.mark.gpu


def _synthetic_bins(n, F, seed, metas):
    rng = np.random.default_rng(seed)
    bins = np.empty((n, F), dtype=np.uint8)
    for f, m in enumerate(metas):
        b = rng.integers(0, m[0], size=n, dtype=np.int32)
        bins[:, f] = np.where(b == m[1], 0, b + 1 - (1 if m[1] == 0 else 0)).astype(np.uint8)
    g = rng.standard_normal(n).astype(np.float32)
    g += (0.5 * np.sign(bins[:, 0].astype(np.float32) - 100) + 0.3 * (bins[:, 1] > 50)).astype(np.float32)
    h = (rng.random(n).astype(np.float32) * 0.25 + 0.01).astype(np.float32)
    return bins, g, h


def _check_tree_properties(L, recs, n_rows, metas, bins):
    cnt = {0: n_rows}
    for r in recs:
        assert cnt[r.leaf] == r.left_count + r.right_count
        cnt[r.leaf] = r.left_count; cnt[r.right_leaf] = r.right_count
        assert r.left_count > 0 and r.right_count > 0
        assert 0 <= r.threshold < metas[r.feature][0]
    seen = np.zeros(n_rows, dtype=np.uint8)
    for leaf in range(len(recs) + 1):
        idx = L.leaf_indices(leaf)
        assert len(idx) == cnt[leaf]
        assert np.all(np.diff(idx) > 0)                    # stable partition keeps rows ascending
        seen[idx] += 1
    assert np.all(seen == 1)                                # disjoint cover
    # the first split really separates the rows by its threshold (stored-value semantics, no missing here)
    r0 = recs[0]
    m = metas[r0.feature]
    stored = bins[:, r0.feature].astype(np.int32)
    bin_of = np.where(stored == 0, m[1], stored - 1 + (1 if m[1] == 0 else 0))
    left_rows = np.nonzero(bin_of <= r0.threshold)[0]
    assert len(left_rows) == r0.left_count


@pytest.mark.parametrize("n,F,leaves,name", [(10_500_000, 28, 255, "higgs"), (400_000, 2000, 63, "epsilon"),
                                             (2_270_000, 137, 255, "mslr")])
def test_full_size_properties(n, F, leaves, name):
    metas = [(3, 0, 0, 0) if f % 9 == 8 else ((130, 0, 0, 0) if f % 7 == 6 else (255, 127, 127, 0)) for f in range(F)]
    bins, g, h = _synthetic_bins(n, F, 1, metas)
    L = N.CudaLearner()
    L.set_dataset(bins, metas)
    L.set_tree_params(N.make_tree_params(num_leaves=leaves, min_data_in_leaf=0, min_sum_hessian_in_leaf=100.0))
    L.set_gradients(g, h)
    # ---- histogram: every feature's bins sum to the leaf totals (checksum of checksums), both code paths identical
    L.begin_tree()
    L.hist_build(0)
    hist = L.hist_read(0)
    sg, sh, cnt = L.get_leaf_sums(0)
    assert cnt == n
    G = float(np.sum(g.astype(np.float64))); Hs = float(np.sum(h.astype(np.float64)))
    assert abs(sg - G) <= H.VALUE_RTOL * float(np.sum(np.abs(g.astype(np.float64))))
    assert abs(sh - Hs) <= H.VALUE_RTOL * Hs
    bases, total = H.hist_bases(metas)
    for f in (0, 6, 8, F - 1):
        ns = H.nslots(metas[f])
        hg = hist[2 * bases[f]: 2 * (bases[f] + ns): 2]; hh = hist[2 * bases[f] + 1: 2 * (bases[f] + ns): 2]
        # slot 0 (most-frequent bin rows) is duplicated at slot 1+mfb for mfb > 0 features: count it once
        dup = 1 + metas[f][1] if metas[f][1] > 0 else None
        tg = hg.sum() - (hg[dup] if dup is not None else 0.0); th = hh.sum() - (hh[dup] if dup is not None else 0.0)
        assert abs(tg - sg) <= H.VALUE_RTOL * float(np.sum(np.abs(g.astype(np.float64))))
        assert abs(th - sh) <= H.VALUE_RTOL * sh
    if name == "higgs":
        L.set_hist_mode(N.HIST_GLOBAL_ATOMIC)
        L.begin_tree(); L.hist_build(0)
        assert np.array_equal(L.hist_read(0), hist)
        L.set_hist_mode(N.HIST_AUTO)
    # ---- whole tree: bookkeeping + partition properties; linearity: scaling the gradients by 2 gives the same tree
    L.begin_tree()
    recs = L.train_tree()
    assert len(recs) == leaves - 1
    _check_tree_properties(L, recs, n, metas, bins)
    # (power-of-two scaling is exact in float and in the fixed-point histograms; min_sum_hessian scales with it)
    L.set_tree_params(N.make_tree_params(num_leaves=leaves, min_data_in_leaf=0, min_sum_hessian_in_leaf=200.0))
    L.set_gradients(2 * g, 2 * h)
    L.begin_tree()
    recs2 = L.train_tree()
    assert [t[:7] for t in H.records_to_tuples(recs2)[:32]] == [t[:7] for t in H.records_to_tuples(recs)[:32]]
    L.close()
