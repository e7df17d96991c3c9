"""Data-parallel training driver (one process per GPU; launched with torchrun).  Trains `iters` iterations on a
row shard with tree_learner=data over NCCL and writes the model text of rank 0 to --out.  With WORLD_SIZE=1 it
trains on the full matrix (single GPU) using the SAME bin mappers, so the two runs can be compared."""
import argparse, ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=200_000)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--leaves", type=int, default=31)
ap.add_argument("--out", default="")
args = ap.parse_args()
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))

rng = np.random.default_rng(0)
n, F = args.rows, 28
X = rng.standard_normal((n, F)).astype(np.float32)
X[:, 21:25] = np.maximum(0, X[:, 21:25])
w = np.random.default_rng(1).standard_normal(F)
y = ((X @ w + 0.5 * X[:, 0] * X[:, 1] + rng.standard_normal(n)) > 0).astype(np.float32)
ours = capi.ours(); lib = ours.lib
base = f"objective=binary num_leaves={args.leaves} learning_rate=0.1 min_data_in_leaf=20 boost_from_average=false verbosity=-1"
ref_ds = ours.dataset_from_mat(X[:50_000], base)
if world > 1:
    import torch, torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl")
    lo, hi = rank * n // world, (rank + 1) * n // world
    ds = ours.dataset_from_mat(np.ascontiguousarray(X[lo:hi]), base, reference=ref_ds)
    ours.set_field(ds, "label", np.ascontiguousarray(y[lo:hi]))
    uid = np.zeros(128, dtype=np.uint8)
    if rank == 0:
        ours.check(lib.LGBM_NetworkGetNCCLUniqueId(uid.ctypes.data_as(C.POINTER(C.c_uint8))))
    t = torch.from_numpy(uid).cuda(); dist.broadcast(t, 0); uid = t.cpu().numpy()
    ours.check(lib.LGBM_NetworkInitNCCL(uid.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_int(rank), C.c_int(world)))
    params = base + f" tree_learner=data num_machines={world} gpu_device_id={local}"
else:
    ds = ours.dataset_from_mat(X, base, reference=ref_ds)
    ours.set_field(ds, "label", y)
    params = base
b = ours.booster_create(ds, params)
for _ in range(args.iters):
    ours.update(b)
text = ours.model_string(b)
if rank == 0 and args.out:
    open(args.out, "w").write(text)
if world > 1:
    # every rank must hold the identical model
    import hashlib
    h = int(hashlib.sha256(text.split("feature_importances")[0].encode()).hexdigest()[:15], 16)
    tt = torch.tensor([h], dtype=torch.int64, device="cuda")
    lst = [torch.zeros_like(tt) for _ in range(world)]
    dist.all_gather(lst, tt)
    assert all(int(x) == h for x in lst), "ranks disagree on the model"
    dist.barrier()
print(f"rank {rank}/{world} done", flush=True)
