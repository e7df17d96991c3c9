"""Golden model of the reference's OWN data-parallel learner (tree_learner=data over localhost sockets, 2 machines):
    python tests/golden/make_dp_golden.py            (run in the container that has oracle/_ref built)
Two processes each hold one contiguous row shard of the same synthetic matrix; both shards are binned against a common
`reference` dataset (built from the first REF_ROWS rows BEFORE the network is initialised, so its bin mappers are
found on one machine), then LGBM_NetworkInit + LGBM_BoosterCreate(tree_learner=data num_machines=2) + ITERS x
LGBM_BoosterUpdateOneIter (src/treelearner/data_parallel_tree_learner.cpp:125-256).  Rank 0's model text is the golden
the 2-rank NCCL run -- and the 1-rank run with data-parallel count semantics -- must reproduce
(tests/test_multigpu.py).  The generator below is shared with the test (dp_case)."""
import ctypes as C
import os
import sys
import multiprocessing as mp

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

N_ROWS, N_FEAT, REF_ROWS, ITERS = 200_000, 28, 50_000, 8
# boost_from_average=false: the reference averages the machines' LOCAL initial scores (gbdt.cpp:312-335), which a
# single-rank rerun on the unsharded matrix cannot reproduce; everything after the initial score is what is compared
BASE = "objective=binary num_leaves=63 learning_rate=0.1 min_data_in_leaf=20 max_bin=255 boost_from_average=false verbosity=-1"
OUT = os.path.join(HERE, "dp_reference_2rank_model.txt")


def dp_case(sorted_shards=False):
    """the matrix, labels and reference-dataset rows of the golden run.  sorted_shards=True (NOT used for the golden)
    sorts the rows by a feature, so that class balance and feature ranges differ between the machines and some leaves
    have no rows at all on one of them.  The reference cannot serve as the oracle for that case: with zero local rows
    Dataset::ConstructHistograms returns before it clears the histogram buffer (include/LightGBM/dataset.h:508-510),
    the stale contents of the pool slot enter the reduce-scatter, and its model depends on them (observed here: split
    gains of 4.7e6 on 200 000 rows).  tests/test_multigpu.py uses the sorted case to compare this library's 2-rank run
    with its own 1-rank run instead."""
    rng = np.random.default_rng(0)
    X = rng.standard_normal((N_ROWS, N_FEAT)).astype(np.float32)
    X[:, 21:25] = np.maximum(0, X[:, 21:25])
    X[:, 25:28] = np.array([0.0, 1.087, 2.173], dtype=np.float32)[rng.integers(0, 3, size=(N_ROWS, 3))]
    w = np.random.default_rng(1).standard_normal(N_FEAT)
    y = ((X @ w + 0.5 * X[:, 0] * X[:, 1] + rng.standard_normal(N_ROWS)) > 0).astype(np.float32)
    if sorted_shards:
        order = np.argsort(X[:, 3], kind="stable")
        X, y = np.ascontiguousarray(X[order]), np.ascontiguousarray(y[order])
    # the common reference dataset must see the whole range: a fixed stride sample
    ref_rows = np.arange(0, N_ROWS, N_ROWS // REF_ROWS)[:REF_ROWS]
    return X, y, ref_rows


def worker(rank, world, port0, q):
    import capi
    ref = capi.reference()
    X, y, ref_rows = dp_case()
    params = BASE + " force_col_wise=true num_threads=2"
    ref_ds = ref.dataset_from_mat(np.ascontiguousarray(X[ref_rows]), params)
    lo, hi = rank * N_ROWS // world, (rank + 1) * N_ROWS // world
    ds = ref.dataset_from_mat(np.ascontiguousarray(X[lo:hi]), params, reference=ref_ds)
    ref.set_field(ds, "label", np.ascontiguousarray(y[lo:hi]))
    machines = ",".join(f"127.0.0.1:{port0 + r}" for r in range(world))
    ref.check(ref.lib.LGBM_NetworkInit(C.c_char_p(machines.encode()), C.c_int(port0 + rank), C.c_int(120), C.c_int(world)))
    b = ref.booster_create(ds, params + f" tree_learner=data num_machines={world}")
    for _ in range(ITERS):
        ref.update(b)
    text = ref.model_string(b)
    ref.check(ref.lib.LGBM_NetworkFree())
    q.put((rank, text))


def main():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port0 = 12400 + 2 * (os.getpid() % 700)   # a fresh pair of ports per run: a port in TIME_WAIT makes the reference wait
    procs = [ctx.Process(target=worker, args=(r, 2, port0, q)) for r in range(2)]
    for p in procs:
        p.start()
    texts = dict(q.get(timeout=600) for _ in procs)
    for p in procs:
        p.join()
    a, b = texts[0].split("feature_importances")[0], texts[1].split("feature_importances")[0]
    assert a == b, "the two machines disagree on the model"
    open(OUT, "w").write(texts[0])
    print("wrote", OUT, len(texts[0]), "bytes")


if __name__ == "__main__":
    main()
