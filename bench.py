#!/usr/bin/env python
"""bench.py -- boosting iterations/s of the B200 tree learner on the Higgs-shaped workload (BASELINE.json
configs[1]: 10.5M x 28 dense, max_bin=255, num_leaves=255, binary logloss, lr 0.1, min_data_in_leaf=0,
min_sum_hessian_in_leaf=100), through the drop-in C API (LGBM_DatasetCreateFromMat / LGBM_BoosterCreate /
LGBM_BoosterUpdateOneIter).

  python bench.py --gpus N --steps K --warmup W              our arm (CUDA, sm_100a)
  python bench.py --impl reference --gpus N --steps K ...    the reference's own CPU learner (oracle/_ref) on the
                                                             same matrix, same parameters, all host threads

A "step" is one boosting iteration (gradients -> 254 x {histogram, split search, partition} -> score update).
`value`  : iterations/s with everything resident in HBM (that is how the API works: the matrix is uploaded once by
           LGBM_BoosterCreate); K synchronous calls, each ending in a stream sync + device->host read of the
           split records, bracketed by barrier + device sync, max over ranks.
`e2e`    : the same K iterations INCLUDING LGBM_BoosterCreate (host->device upload of the binned matrix + labels)
           and a final LGBM_BoosterGetPredict (device->host read of all training scores), divided by K:
           what a C-API user pays from host buffers.  h2d/d2h bytes are per step (total / K).
`roofline`: the histogram kernel k_hist, algorithmic bytes per SURVEY 8(d)
           (n*(G+8) [+4n if indexed] + total_bin*16 per call) over their CUDA-event time on the launch stream,
           measured in a separate profiled pass after the timed region.
`cpu_baseline`: the reference CPU learner timed here on the box's host cores on the same workload (few iterations).

Multi-GPU (torchrun, one rank per GPU): STRONG scaling of the same 10.5M-row workload, rows sharded contiguously,
tree_learner=data, one NCCL int64 all-reduce of the smaller leaf's histogram per split.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# rank 0 must print ONE JSON line on stdout.  Libraries (NCCL's version banner, LightGBM's log) write to fd 1 as well,
# so everything but that line is sent to stderr: fd 1 is pointed at fd 2 and the line goes to the saved descriptor.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

PARAMS = ("objective=binary max_bin=255 num_leaves=255 learning_rate=0.1 min_data_in_leaf=0 "
          "min_sum_hessian_in_leaf=100 verbosity=-1")
N_ROWS, N_FEAT = 10_500_000, 28


def higgs_shaped(n, seed=0):
    """SURVEY.md 8(d) config #2 generator: 21 gaussian, 4 half-zero, 3 three-level columns."""
    rng = np.random.default_rng(seed)
    X = np.empty((n, N_FEAT), dtype=np.float32)
    chunk = 1 << 20
    w = np.random.default_rng(seed + 1).standard_normal(N_FEAT).astype(np.float32)
    y = np.empty(n, dtype=np.float32)
    levels = np.array([0.0, 1.087, 2.173], dtype=np.float32)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        b = rng.standard_normal((e - s, N_FEAT), dtype=np.float32)
        b[:, 21:25] = np.maximum(0, b[:, 21:25])
        b[:, 25:28] = levels[rng.integers(0, 3, size=(e - s, 3))]
        X[s:e] = b
        y[s:e] = ((b @ w + 0.5 * b[:, 0] * b[:, 1] + rng.standard_normal(e - s, dtype=np.float32)) > 0)
    return X, y


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index=0):
        self.rows = []
        self.proc = None
        self.idx = gpu_index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def run_reference(args):
    """The reference's own CPU implementation of the path (oracle/_ref), all host threads, same config."""
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    import capi
    if not os.path.exists(capi.REF_SO):
        emit({"impl": "reference", "unavailable": "oracle/_ref/lib_lightgbm.so not built (make -C oracle ref)"})
        return 0
    cores = os.cpu_count() or 1
    n = args.rows
    X, y = higgs_shaped(n)
    ref = capi.reference()
    params = PARAMS + f" force_col_wise=true num_threads={cores}"
    ds = ref.dataset_from_mat(X, params)
    ref.set_field(ds, "label", y)
    b = ref.booster_create(ds, params)
    for _ in range(args.warmup):
        ref.update(b)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ref.update(b)
    dt = time.perf_counter() - t0
    v = args.steps / dt
    line = {
        "impl": "reference", "metric": "boosting iters/sec", "value": v, "unit": "iters/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"Higgs-shaped {n}x{N_FEAT} dense, max_bin=255, num_leaves=255, binary, lr 0.1 (configs[1])",
                   "params": PARAMS, "l2_note": "inputs larger than L2"},
        "cpu_baseline": {"value": v, "unit": "iters/s", "cores": cores, "kind": "reference",
                         "sample": f"full workload, {args.steps} iterations after {args.warmup} warm-up, force_col_wise"},
        "e2e": {"value": v, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)
    return 0


def run_ours(args):
    rank, world, local = dist_env()
    import capi
    ours = capi.ours()
    lib = ours.lib
    lib.LGBM_B200_LaunchCount.restype = C.c_int64
    use_dist = world > 1
    if use_dist:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl")
    n = args.rows
    X, y = higgs_shaped(n)
    ds_params = PARAMS
    if use_dist:
        # identical bin mappers on every rank: a reference dataset from the first 200k rows, then this rank's shard
        ref_ds = ours.dataset_from_mat(X[:200_000], ds_params)
        lo, hi = rank * n // world, (rank + 1) * n // world
        Xs, ys = np.ascontiguousarray(X[lo:hi]), np.ascontiguousarray(y[lo:hi])
        ds = ours.dataset_from_mat(Xs, ds_params, reference=ref_ds)
        ours.set_field(ds, "label", ys)
        def net_init():
            # a fresh NCCL id per communicator (= per booster): rank 0 creates it, torch.distributed carries it
            uid = np.zeros(128, dtype=np.uint8)
            if rank == 0:
                ours.check(lib.LGBM_NetworkGetNCCLUniqueId(uid.ctypes.data_as(C.POINTER(C.c_uint8))))
            t = torch.from_numpy(uid).cuda()
            dist.broadcast(t, 0)
            uid = t.cpu().numpy()
            ours.check(lib.LGBM_NetworkInitNCCL(uid.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_int(rank), C.c_int(world)))
        params = PARAMS + f" tree_learner=data num_machines={world} gpu_device_id={local}"
        local_rows = hi - lo
    else:
        ds = ours.dataset_from_mat(X, ds_params)
        ours.set_field(ds, "label", y)
        params = PARAMS + f" gpu_device_id={local}"
        local_rows = n

        def net_init():
            pass
    del X

    def barrier():
        if use_dist:
            torch.cuda.synchronize()
            dist.barrier()

    # one-time process cost (CUDA context + module load) is paid before any timing
    from lightgbmv1_b200 import _native as _N
    _N.CudaLearner(local).close()

    # ---- e2e pass: BoosterCreate (H2D of the binned matrix + labels) + K iterations + GetPredict (D2H scores)
    net_init()
    barrier()
    t0 = time.perf_counter()
    b = ours.booster_create(ds, params)
    for _ in range(args.steps):
        ours.update(b)
    scores = ours.get_predict(b, local_rows)
    barrier()
    e2e_dt = time.perf_counter() - t0
    row_stride = (N_FEAT + 15) // 16 * 16
    h2d_total = local_rows * (N_FEAT + 4)  # binned matrix (as passed by the host) + float labels
    d2h_total = local_rows * 8
    ours.booster_free(b)

    # ---- timed pass: W warm-up + exactly K iterations, device-resident
    net_init()
    b = ours.booster_create(ds, params)
    for _ in range(args.warmup):
        ours.update(b)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = lib.LGBM_B200_LaunchCount()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ours.update(b)  # every call ends with a stream sync + D2H of the split records
    barrier()
    dt = time.perf_counter() - t0
    launches = lib.LGBM_B200_LaunchCount() - launches0
    clocks = sampler.stop() if rank == 0 else None
    if use_dist:
        tt = torch.tensor([dt, e2e_dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt, e2e_dt = float(tt[0]), float(tt[1])

    # ---- roofline pass (profiled, outside the timed region): histogram kernels only
    hist_ms, hist_calls, hist_bytes = C.c_double(), C.c_int64(), C.c_double()
    ours.check(lib.LGBM_B200_BoosterProfile(b, 1, 1, C.byref(hist_ms), C.byref(hist_calls), C.byref(hist_bytes)))
    prof_iters = 3
    hist_launches = 0
    ph = (C.c_double * 8)()
    for _ in range(prof_iters):
        ours.update(b)
        ours.check(lib.LGBM_B200_BoosterPhases(b, ph))
        hist_launches += int(ph[7])   # rounds the tree needed = k_hist launches that did work (each covers up to 16 leaves)
    ours.check(lib.LGBM_B200_BoosterProfile(b, 0, 0, C.byref(hist_ms), C.byref(hist_calls), C.byref(hist_bytes)))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # DRAM traffic of the histogram kernel per launch: from the committed ncu pass over one tree of this very workload
    # (profiles/hist_traffic.json, written by tools/summarize_traffic.py from the ncu csv next to it)
    traffic = None
    try:
        traffic = float(json.load(open(os.path.join(ROOT, "profiles", "hist_traffic.json")))["dram_bytes_per_launch"])
    except Exception:
        pass
    achieved = (hist_bytes.value / 1e9) / (hist_ms.value / 1e3) if hist_ms.value > 0 else 0.0

    if use_dist:
        ours.booster_free(b)
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0
    # ---- cpu_baseline (N=1 only): the reference CPU learner on the same workload, few iterations
    cpu = None
    if world == 1 and not args.no_cpu_baseline and os.path.exists(capi.REF_SO):
        cores = os.cpu_count() or 1
        Xc, yc = higgs_shaped(n)
        ref = capi.reference()
        rp = PARAMS + f" force_col_wise=true num_threads={cores}"
        rds = ref.dataset_from_mat(Xc, rp)
        ref.set_field(rds, "label", yc)
        rb = ref.booster_create(rds, rp)
        ref.update(rb)
        k = 3
        t0 = time.perf_counter()
        for _ in range(k):
            ref.update(rb)
        cdt = time.perf_counter() - t0
        cpu = {"value": k / cdt, "unit": "iters/s", "cores": cores, "kind": "reference",
               "sample": f"same {n}x{N_FEAT} matrix and parameters, {k} iterations after 1 warm-up, force_col_wise, all host threads"}
    line = {
        "metric": "boosting iters/sec", "value": args.steps / dt, "unit": "iters/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "int64 fixed-point histograms / f64 split search",
        "data": "synthetic",
        "config": {"workload": f"Higgs-shaped {n}x{N_FEAT} dense, max_bin=255, num_leaves=255, binary, lr 0.1 (configs[1])",
                   "params": PARAMS, "parallelism": f"data-parallel x{world}" if world > 1 else "single GPU",
                   "l2_note": "inputs larger than L2 (binned matrix 336 MB + 168 MB of gradients/scores per pass)"},
        "clocks": clocks,
        "e2e": {"value": args.steps / e2e_dt, "unit": "iters/s", "h2d_bytes_per_step": h2d_total / args.steps,
                "d2h_bytes_per_step": d2h_total / args.steps,
                "what": "LGBM_BoosterCreate (upload) + K x LGBM_BoosterUpdateOneIter + LGBM_BoosterGetPredict, / K"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_hist (block-shared sub-histogram, 32-bit shared atomic pairs, DSMEM cluster reduce; global-atomic path for leaves < 256 rows)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None,
                     "traffic": traffic, "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                     "launches": hist_launches, "leaf_histograms": int(hist_calls.value), "ms": hist_ms.value, "bytes": hist_bytes.value,
                     "algorithmic_bytes_per_launch": hist_bytes.value / max(1, hist_launches),
                     "note": "achieved/traffic are per-launch averages over every histogram launch of %d iterations (one launch per "
                             "round of tree growth, each building the histograms of up to 16 leaves; ~25 rounds per tree); the "
                             "root launch alone runs at ~1 TB/s, see profiles/" % prof_iters},
        "cpu_baseline": cpu,
        "train_check": {"mean_score": float(np.mean(scores))},
    }
    emit(line)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=N_ROWS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
