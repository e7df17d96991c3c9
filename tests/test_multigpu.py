"""tree_learner=data (A11).  The oracle is the reference's OWN data-parallel learner: 2 machines over localhost sockets,
contiguous row shards binned against a common reference dataset (tests/golden/make_dp_golden.py wrote rank 0's model
to tests/golden/dp_reference_2rank_model.txt).
  (gpu, 1 GPU)      one rank with the data-parallel count semantics (LGBCU_FORCE_DP_COUNTS) on the unsharded matrix,
                    same bin mappers, must reproduce the reference's 2-machine model;
  (gpu, >= 2 GPUs)  the 2-rank NCCL run must reproduce it too -- and equal the 1-rank run bit for bit (the histogram
                    all-reduce is an integer sum); with shards SORTED by a feature (skewed classes, leaves with no
                    rows on one rank -- a case the reference itself gets wrong, see make_dp_golden.dp_case) and
                    is_unbalance (label weights from the class counts of ALL ranks, binary_objective.hpp:72-75) the
                    2-rank run must still equal the 1-rank run;
  (cpu, gloo, world_size 2) the host-side shard plumbing: shards built against a common reference dataset carry the
                    same bin mappers and together hold exactly the bins of the unsharded matrix."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import capi
import helpers as H
from lightgbmv1_b200 import _native as N

GOLDEN = os.path.join(H.ROOT, "tests", "golden", "dp_reference_2rank_model.txt")
SCRIPT = os.path.join(H.ROOT, "tools", "run_dp.py")


def _torchrun(nproc, port, out, extra="", sorted_shards=False, env=None, max_bin=0):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), SCRIPT, "--out", out]
    if max_bin:
        cmd += ["--max-bin", str(max_bin)]
    if extra:
        cmd += ["--extra", extra]
    if sorted_shards:
        cmd += ["--sorted-shards"]
    subprocess.check_call(cmd, env=dict(os.environ, **(env or {})), timeout=900)


def _assert_same(a, b, what):
    same, n, first = capi.compare_models(a, b, rtol_value=H.VALUE_RTOL)
    assert n > 0 and same == n, f"{what}: {same}/{n} trees identical; first mismatch: {first}"


@pytest.mark.gpu
def test_one_rank_with_dp_counts_reproduces_the_reference_two_machine_model(tmp_path):
    one = str(tmp_path / "one.txt")
    subprocess.check_call([sys.executable, SCRIPT, "--out", one], env=dict(os.environ, LGBCU_FORCE_DP_COUNTS="1"), timeout=600)
    _assert_same(open(GOLDEN).read(), open(one).read(), "reference 2-machine socket run vs 1 rank with data-parallel counts")


@pytest.mark.gpu
def test_two_rank_model_equals_reference_two_machine_model_and_one_rank_model(tmp_path):
    if N.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    one, two = str(tmp_path / "one.txt"), str(tmp_path / "two.txt")
    subprocess.check_call([sys.executable, SCRIPT, "--out", one], env=dict(os.environ, LGBCU_FORCE_DP_COUNTS="1"), timeout=600)
    _torchrun(2, 29511, two)
    a, b = open(one).read(), open(two).read()
    _assert_same(open(GOLDEN).read(), b, "reference 2-machine socket run vs 2-rank NCCL run")
    same, n, first = capi.compare_models(a, b, rtol_value=0.0)
    assert same == n, (same, n, first)
    assert a.split("parameters:")[0] == b.split("parameters:")[0]


@pytest.mark.gpu
@pytest.mark.parametrize("sorted_shards", [False, True])
def test_two_rank_feature_ownership_mode(tmp_path, sorted_shards):
    """The reference's own scheme (data_parallel_tree_learner.cpp:176-247): every rank owns a block of features, gets
    the global histogram of that block only (grouped ncclReduce), scans it, and the per-leaf winners are all-gathered
    and reduced with SplitInfo::operator> on every rank.  Chosen by shape for wide data; forced here.  Same model as
    the all-reduce mode, i.e. the reference's 2-machine model and the 1-rank model."""
    if N.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    one, two = str(tmp_path / "one.txt"), str(tmp_path / "two.txt")
    cmd = [sys.executable, SCRIPT, "--out", one] + (["--sorted-shards"] if sorted_shards else [])
    subprocess.check_call(cmd, env=dict(os.environ, LGBCU_FORCE_DP_COUNTS="1"), timeout=600)
    _torchrun(2, 29513 + int(sorted_shards), two, sorted_shards=sorted_shards, env={"LGBCU_DP_OWNERSHIP": "1"})
    a, b = open(one).read(), open(two).read()
    if not sorted_shards:
        _assert_same(open(GOLDEN).read(), b, "reference 2-machine socket run vs 2-rank NCCL run with feature ownership")
    same, n, first = capi.compare_models(a, b, rtol_value=0.0)
    assert same == n, (same, n, first)


@pytest.mark.gpu
@pytest.mark.parametrize("form", ["0", "1", "2"])   # NCCL all-reduce / one-step pull (the 2-rank default) / two-step (slice reduce + gather)
@pytest.mark.parametrize("sorted_shards", [False, True])
def test_two_rank_one_shot_exchange(tmp_path, sorted_shards, form):
    """No collective on the histograms: every rank adds into a local pool exported with CUDA IPC, publishes each round
    with a flag in its peers' memory, and the split search sums the ranks' local slots over NVLink while it loads them
    (LGBCU_DP_ONESHOT=1).  Integer sums: the model must equal the NCCL all-reduce mode's bit for bit -- the reference's
    2-machine model and the 1-rank model."""
    if N.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    one, two = str(tmp_path / "one.txt"), str(tmp_path / "two.txt")
    cmd = [sys.executable, SCRIPT, "--out", one] + (["--sorted-shards"] if sorted_shards else [])
    subprocess.check_call(cmd, env=dict(os.environ, LGBCU_FORCE_DP_COUNTS="1"), timeout=600)
    _torchrun(2, 29515 + int(sorted_shards) + 2 * int(form), two, sorted_shards=sorted_shards, env={"LGBCU_DP_ONESHOT": form})
    a, b = open(one).read(), open(two).read()
    if not sorted_shards:
        _assert_same(open(GOLDEN).read(), b, "reference 2-machine socket run vs 2-rank run with the one-shot exchange")
    same, n, first = capi.compare_models(a, b, rtol_value=0.0)
    assert same == n, (same, n, first)


@pytest.mark.gpu
@pytest.mark.parametrize("env", [{}, {"LGBCU_DP_OWNERSHIP": "1"}, {"LGBCU_DP_ONESHOT": "0"}])
def test_two_rank_wide_features_equal_one_rank(tmp_path, env):
    """max_bin = 1000 under tree_learner=data: the wide histogram kernel adds into the exchanged pool, the general split
    search runs on the reduced histograms (in every exchange form) -- integer sums, so the 2-rank model must equal the
    1-rank model with the data-parallel count semantics bit for bit."""
    if N.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    one, two = str(tmp_path / "one.txt"), str(tmp_path / "two.txt")
    subprocess.check_call([sys.executable, SCRIPT, "--out", one, "--max-bin", "1000"], env=dict(os.environ, LGBCU_FORCE_DP_COUNTS="1"), timeout=600)
    _torchrun(2, 29531 + len(env) + 2 * ("LGBCU_DP_ONESHOT" in env), two, env=env, max_bin=1000)
    a, b = open(one).read(), open(two).read()
    same, n, first = capi.compare_models(a, b, rtol_value=0.0)
    assert same == n, (same, n, first)
    assert "max_bin: 1000" in a


@pytest.mark.gpu
@pytest.mark.parametrize("extra,env", [("categorical_feature=25,26,27,0 max_cat_to_onehot=2 cat_smooth=1 min_data_per_group=10", {}),
                                       ("categorical_feature=25,26,27,0 max_cat_to_onehot=2 cat_smooth=1 min_data_per_group=10", {"LGBCU_DP_ONESHOT": "0"}),
                                       ("monotone_constraints=1,-1,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0 monotone_penalty=0.5", {}),
                                       ("monotone_constraints=1,-1,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0", {"LGBCU_DP_OWNERSHIP": "1"})])
def test_two_rank_categorical_and_monotone_equal_one_rank(tmp_path, extra, env):
    """Categorical features and monotone constraints under tree_learner=data: every rank runs the general split search on
    the same reduced histograms (a categorical split's bin set stays in the rank's own pool; feature ownership is not
    used with categorical data), so the 2-rank model equals the 1-rank model with the data-parallel count semantics."""
    if N.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    one, two = str(tmp_path / "one.txt"), str(tmp_path / "two.txt")
    subprocess.check_call([sys.executable, SCRIPT, "--out", one, "--extra", extra], env=dict(os.environ, LGBCU_FORCE_DP_COUNTS="1"), timeout=600)
    _torchrun(2, 29541 + len(extra) % 7 + 3 * len(env), two, extra=extra, env=env)
    a, b = open(one).read(), open(two).read()
    same, n, first = capi.compare_models(a, b, rtol_value=0.0)
    assert same == n, (same, n, first)
    if "categorical" in extra:
        assert any(int(t["num_cat"][0]) > 0 for t in capi.parse_trees(b)), "no categorical split in the case"


@pytest.mark.gpu
def test_two_rank_sorted_shards_is_unbalance_equals_one_rank(tmp_path):
    if N.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    one, two = str(tmp_path / "one.txt"), str(tmp_path / "two.txt")
    subprocess.check_call([sys.executable, SCRIPT, "--out", one, "--extra", "is_unbalance=true", "--sorted-shards"],
                          env=dict(os.environ, LGBCU_FORCE_DP_COUNTS="1"), timeout=600)
    _torchrun(2, 29512, two, extra="is_unbalance=true", sorted_shards=True)
    same, n, first = capi.compare_models(open(one).read(), open(two).read(), rtol_value=0.0)
    assert same == n, (same, n, first)


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import torch
    ours = capi.ours()
    rng = np.random.default_rng(0)
    n, F = 6000, 7
    X = rng.standard_normal((n, F)).astype(np.float32)
    X[:, 3] = np.maximum(0, X[:, 3])
    ref_ds = ours.dataset_from_mat(X[:2000], "max_bin=63 verbosity=-1")
    lo, hi = rank * n // world, (rank + 1) * n // world
    ds = ours.dataset_from_mat(np.ascontiguousarray(X[lo:hi]), "max_bin=63 verbosity=-1", reference=ref_ds)
    nf = C.c_int(); ours.check(ours.lib.LGBM_B200_DatasetNumInnerFeatures(ds, C.byref(nf)))
    bins = np.empty((hi - lo, nf.value), dtype=np.int32); real = np.empty(nf.value, dtype=np.int32)
    ours.check(ours.lib.LGBM_B200_DatasetDumpBins(ds, bins.ctypes.data_as(C.POINTER(C.c_int32)),
                                                  real.ctypes.data_as(C.POINTER(C.c_int32))))
    # the rendezvous payload travels like the NCCL id would: rank 0's 128 bytes to everyone
    uid = torch.arange(128, dtype=torch.uint8) if rank == 0 else torch.zeros(128, dtype=torch.uint8)
    dist.broadcast(uid, 0)
    assert uid.tolist() == list(range(128))
    gathered = [None] * world
    dist.all_gather_object(gathered, (lo, hi, real.tolist(), int(bins.sum()), bins.tobytes()))
    if rank == 0:
        full = ours.dataset_from_mat(X, "max_bin=63 verbosity=-1", reference=ref_ds)
        fb = np.empty((n, nf.value), dtype=np.int32); fr = np.empty(nf.value, dtype=np.int32)
        ours.check(ours.lib.LGBM_B200_DatasetDumpBins(full, fb.ctypes.data_as(C.POINTER(C.c_int32)),
                                                      fr.ctypes.data_as(C.POINTER(C.c_int32))))
        ok = True
        for (a, b, r, _, raw) in gathered:
            ok = ok and r == fr.tolist()
            ok = ok and np.array_equal(np.frombuffer(raw, dtype=np.int32).reshape(b - a, nf.value), fb[a:b])
        q.put(ok)
    dist.barrier()
    dist.destroy_process_group()


def test_row_shards_share_mappers_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, 29533, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


@pytest.mark.gpu
def test_entry_points_leave_the_callers_current_device_alone():
    """A host application may be working on another GPU in the same thread (torch does): dataset creation (device
    binning), booster creation and training select their own device and put the caller's back."""
    if N.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch
    torch.cuda.set_device(1)
    rng = np.random.default_rng(0)
    X = rng.standard_normal((20_000, 6)).astype(np.float32)
    y = (X[:, 0] + 0.3 * rng.standard_normal(20_000) > 0).astype(np.float32)
    ours = capi.ours()
    params = "objective=binary num_leaves=15 verbosity=-1 gpu_device_id=0"
    ds = ours.dataset_from_mat(X, params)
    assert torch.cuda.current_device() == 1
    ours.set_field(ds, "label", y)
    b = ours.booster_create(ds, params)
    assert torch.cuda.current_device() == 1
    for _ in range(3):
        ours.update(b)
    ours.get_predict(b, len(y))
    assert torch.cuda.current_device() == 1
    ours.booster_free(b); ours.dataset_free(ds)
    assert torch.cuda.current_device() == 1
