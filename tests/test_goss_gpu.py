"""boosting=goss against the unmodified reference (oracle/_ref): GOSS::Bagging / BaggingHelper (src/boosting/goss.hpp:
103-177).  The sample is drawn per thread chunk (Threading::BlockInfoForceSize), so both arms get the same num_threads;
learning_rate is large so that sampling starts after 1 / learning_rate iterations.  Identical trees are the bar:
the sampled rows, the multiplier applied to the small-gradient rows and the out-of-bag score update all feed them."""
import os

import numpy as np
import pytest

import capi
import helpers as H
from test_e2e_gpu import assert_same_model, higgs_like, logloss, run_both

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(H.REF_SO), reason="oracle/_ref not built")]

BASE = ("objective=binary boosting=goss num_leaves=31 learning_rate=0.5 min_data_in_leaf=20 "
        "min_sum_hessian_in_leaf=1e-3 verbosity=-1 num_threads=4")


@pytest.mark.parametrize("n,top,other", [(50_000, 0.2, 0.1),     # top + other <= 0.5: the reference trains on a row subset copy
                                         (50_000, 0.4, 0.3),     # > 0.5: bagging by indices
                                         (5_000, 0.2, 0.1),      # fewer chunks than threads would like (1024-row minimum)
                                         (70_001, 0.05, 0.45)])
def test_goss_identical_trees(n, top, other):
    X, y = higgs_like(11, n)
    params = BASE + f" top_rate={top} other_rate={other}"
    ref, ours, br, bo, nr, no = run_both(X, y, params, 10)
    assert nr == no == 10
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, len(y)), ours.get_predict(bo, len(y))
    assert np.max(np.abs(pr - po)) <= 1e-9
    assert abs(logloss(pr, y) - logloss(po, y)) <= 1e-5


def test_goss_rejects_bagging_like_the_reference():
    X, y = higgs_like(3, 3000)
    ours = capi.ours()
    ds = ours.dataset_from_mat(X, BASE)
    ours.set_field(ds, "label", y)
    with pytest.raises(Exception, match="Cannot use bagging in GOSS"):
        ours.booster_create(ds, BASE + " bagging_fraction=0.5 bagging_freq=1")
    ours.dataset_free(ds)


@pytest.mark.parametrize("pos,neg,freq", [(0.5, 0.8, 2), (1.0, 0.3, 1), (0.25, 1.0, 3)])
def test_balanced_bagging_identical_trees(pos, neg, freq):
    """pos_/neg_bagging_fraction (GBDT::BalancedBaggingHelper, src/boosting/gbdt.cpp:183-207): the per-block generators
    are the plain bagging ones, the fraction a row is tested against depends on its label."""
    X, y = higgs_like(21, 40_000)
    params = ("objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 min_sum_hessian_in_leaf=1e-3 "
              f"verbosity=-1 pos_bagging_fraction={pos} neg_bagging_fraction={neg} bagging_freq={freq} bagging_seed=7")
    ref, ours, br, bo, nr, no = run_both(X, y, params, 12)
    assert nr == no == 12
    assert_same_model(ref, ours, br, bo)
    pr, po = ref.get_predict(br, len(y)), ours.get_predict(bo, len(y))
    assert np.max(np.abs(pr - po)) <= 1e-9
