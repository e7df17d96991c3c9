"""Data-parallel training driver (one process per GPU; launched with torchrun).  Trains the case of
tests/golden/make_dp_golden.py -- the same matrix, shards, common reference dataset, parameters and iteration count
as the reference's own 2-machine socket run -- with tree_learner=data over NCCL and writes rank 0's model text to
--out.  With WORLD_SIZE=1 it trains on the full matrix (single GPU) using the SAME bin mappers; with
LGBCU_FORCE_DP_COUNTS=1 that run uses the data-parallel count semantics and must give the same model."""
import argparse, ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np
import capi
import make_dp_golden as G

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="")
ap.add_argument("--extra", default="", help="extra parameters, e.g. is_unbalance=true")
ap.add_argument("--max-bin", type=int, default=0, help="replace the case's max_bin (e.g. 1000: wide features)")
ap.add_argument("--iters", type=int, default=G.ITERS)
ap.add_argument("--sorted-shards", action="store_true", help="rows sorted by a feature: skewed shards, empty local leaves")
args = ap.parse_args()
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))

X, y, ref_rows = G.dp_case(sorted_shards=args.sorted_shards)
n = len(y)
ours = capi.ours(); lib = ours.lib
base = G.BASE + (" " + args.extra if args.extra else "")
if args.max_bin:
    base = base.replace("max_bin=255", f"max_bin={args.max_bin}")
ref_ds = ours.dataset_from_mat(np.ascontiguousarray(X[ref_rows]), base + f" gpu_device_id={local}")
if world > 1:
    import torch, torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl")
    lo, hi = rank * n // world, (rank + 1) * n // world
    ds = ours.dataset_from_mat(np.ascontiguousarray(X[lo:hi]), base + f" gpu_device_id={local}", reference=ref_ds)
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
    ours.booster_free(b)
    ours.check(lib.LGBM_NetworkFree())
    dist.destroy_process_group()
print(f"rank {rank}/{world} done", flush=True)
