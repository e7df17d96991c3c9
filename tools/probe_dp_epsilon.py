"""Scratch probe (torchrun, one process per GPU): ms per iteration of the Epsilon-shaped configuration (400K x 2000)
with tree_learner=data, rows sharded over the ranks -- run once with LGBCU_DP_OWNERSHIP=0 (all-reduce + replicated
search) and once with =1 (feature ownership).  usage: torchrun ... probe_dp_epsilon.py [rows] [features] [iterations]"""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import capi, workloads as W
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
feats = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 8
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
import torch, torch.distributed as dist
torch.cuda.set_device(local)
if world > 1: dist.init_process_group("nccl")
ours = capi.ours(); lib = ours.lib
X, y = W.epsilon_shaped(rows, feats)
base = W.EPSILON_PARAMS + f" num_threads={max(1, (os.cpu_count() or 4) // world)} gpu_device_id={local} boost_from_average=false"
ref_ds = ours.dataset_from_mat(np.ascontiguousarray(X[:50_000]), base)
lo, hi = rank * rows // world, (rank + 1) * rows // world
ds = ours.dataset_from_mat(np.ascontiguousarray(X[lo:hi]), base, reference=ref_ds)
ours.set_field(ds, "label", np.ascontiguousarray(y[lo:hi]))
params = base
if world > 1:
    uid = np.zeros(128, dtype=np.uint8)
    if rank == 0: ours.check(lib.LGBM_NetworkGetNCCLUniqueId(uid.ctypes.data_as(C.POINTER(C.c_uint8))))
    t = torch.from_numpy(uid).cuda(); dist.broadcast(t, 0); uid = t.cpu().numpy()
    ours.check(lib.LGBM_NetworkInitNCCL(uid.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_int(rank), C.c_int(world)))
    params = base + f" tree_learner=data num_machines={world}"
b = ours.booster_create(ds, params)
for _ in range(3): ours.update(b)
ours.get_predict(b, hi - lo)
if world > 1: dist.barrier()
t0 = time.perf_counter()
for _ in range(iters): ours.update(b)
ours.get_predict(b, hi - lo)
ms = 1e3 * (time.perf_counter() - t0) / iters
import hashlib
h = hashlib.sha256(ours.model_string(b).split("feature_importances")[0].encode()).hexdigest()[:12]
print(f"rank {rank}/{world} ownership={os.environ.get('LGBCU_DP_OWNERSHIP', 'auto')} rows/rank {hi - lo} x {feats}: {ms:.2f} ms/iter model {h}", flush=True)
if world > 1:
    dist.barrier(); ours.booster_free(b); ours.check(lib.LGBM_NetworkFree()); dist.destroy_process_group()
