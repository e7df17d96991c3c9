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
`e2e`    : the same K iterations from HOST buffers: LGBM_DatasetCreateFromMat on the float32 matrix (bin mappers on a
           host sample, host->device copy of the raw matrix, value->bin on the device), LGBM_DatasetSetField(label),
           LGBM_BoosterCreate, K x LGBM_BoosterUpdateOneIter and a final LGBM_BoosterGetPredict (device->host read of all
           training scores), divided by K: what a C-API user pays.  h2d/d2h bytes are per step (total / K).  The
           reference arm's `e2e` is timed from the same call (its `value` stays the iterations alone).  The
           data-parallel rendezvous (LGBM_NetworkInitNCCL) and the common reference dataset of the shards are made
           once per process, as the reference's Network::Init is.
`roofline`: the histogram kernel k_hist ALONE (CUDA events around the kernel, not around the all-reduce that follows
           it in data-parallel runs), algorithmic bytes per SURVEY 8(d) (n*(G+8) [+4n if indexed] + total_bin*16 per
           call), measured in a separate profiled pass after the timed region.
`cpu_baseline`: the reference CPU learner timed here on the box's host cores on the same workload (few iterations).
`train_check`: N = 1: the model the reference CPU learner grows in those iterations against the model this library
           grows from the same start -- trees compared / identical, first mismatch, largest score difference.
           N > 1: the N-rank model against the 1-rank model grown on the unsharded matrix with the data-parallel count
           semantics (must be identical: the histogram all-reduce is an integer sum) and the agreement of all ranks.

Multi-GPU (torchrun, one rank per GPU): STRONG scaling of the same 10.5M-row workload, rows sharded contiguously,
tree_learner=data, one NCCL int64 all-reduce of the round's fresh histograms per round of tree growth.
"""
import argparse
import ctypes as C
import hashlib
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

import workloads as W  # noqa: E402

PARAMS = W.HIGGS_PARAMS
N_ROWS, N_FEAT = 10_500_000, 28
higgs_shaped = W.higgs_shaped
PARITY_ITERS = 10   # iterations compared tree by tree with the reference CPU learner (N = 1)
DP_CHECK_ITERS = 5  # iterations of the N-rank == 1-rank check (N > 1)


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


def workload_name(n):
    return f"Higgs-shaped {n}x{N_FEAT} dense, max_bin=255, num_leaves=255, binary, lr 0.1 (configs[1])"


def config_of(n):
    """identical in both arms (the driver compares the dicts)"""
    return {"workload": workload_name(n), "params": PARAMS,
            "l2_note": "inputs larger than L2 (binned matrix ~300 MB + gradients / scores per pass)"}


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
    # The reference has two histogram layouts on the CPU (column-wise DenseBin scans / row-wise multi-value bins) and
    # picks one by timing both on the first iteration; both are timed here explicitly and the FASTER one is the arm's
    # value -- on 16 cores and this shape that is row-wise.
    per_mode = {}
    for mode in ("row", "col"):
        params = PARAMS + f" force_{mode}_wise=true num_threads={cores}"
        ds = ref.dataset_from_mat(X, params)
        ref.set_field(ds, "label", y)
        b = ref.booster_create(ds, params)
        for _ in range(args.warmup):
            ref.update(b)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ref.update(b)
        per_mode[mode] = args.steps / (time.perf_counter() - t0)
        ref.booster_free(b); ref.dataset_free(ds)
    best = max(per_mode, key=per_mode.get)
    v = per_mode[best]
    dt = args.steps / v
    params = PARAMS + f" force_{best}_wise=true num_threads={cores}"
    # e2e: the same K iterations from the host matrix (LGBM_DatasetCreateFromMat .. LGBM_BoosterGetPredict), / K
    score_out = np.empty(n, dtype=np.float64)
    score_out.fill(0.0)
    t0 = time.perf_counter()
    ds = ref.dataset_from_mat(X, params)
    ref.set_field(ds, "label", y)
    b = ref.booster_create(ds, params)
    for _ in range(args.steps):
        ref.update(b)
    ref.get_predict(b, n, out=score_out)
    e2e_dt = time.perf_counter() - t0
    line = {
        "impl": "reference", "metric": "boosting iters/sec", "value": v, "unit": "iters/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(n),
        "cpu_baseline": {"value": v, "unit": "iters/s", "cores": cores, "kind": "reference",
                         "sample": f"full workload, {args.steps} iterations after {args.warmup} warm-up, force_{best}_wise (the faster of the reference's two layouts)",
                         "col_wise_value": per_mode["col"], "row_wise_value": per_mode["row"]},
        # the contract wants the line's own value here (no copies on the host arm); the same from-the-matrix pipeline the
        # other arm's e2e times is reported beside it
        "e2e": {"value": v, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "from_mat": {"value": args.steps / e2e_dt, "unit": "iters/s",
                             "what": "LGBM_DatasetCreateFromMat + SetField(label) + LGBM_BoosterCreate + K x "
                                     "LGBM_BoosterUpdateOneIter + LGBM_BoosterGetPredict, / K (all on the host cores)"}},
    }
    emit(line)
    return 0


CONFIG5_ROWS = 100_000_000


def gen_rows_threaded(n, seed, threads):
    """config #2's generator, 1M-row chunks drawn by a few python threads (numpy releases the GIL)"""
    from concurrent.futures import ThreadPoolExecutor
    X = np.empty((n, N_FEAT), dtype=np.float32)
    y = np.empty(n, dtype=np.float32)
    w = np.random.default_rng(1).standard_normal(N_FEAT).astype(np.float32)
    levels = np.array([0.0, 1.087, 2.173], dtype=np.float32)
    chunk = 1 << 20
    starts = list(range(0, n, chunk))

    def fill(i):
        s_, e_ = starts[i], min(n, starts[i] + chunk)
        rng = np.random.default_rng([seed, i])
        b = rng.standard_normal((e_ - s_, N_FEAT), dtype=np.float32)
        b[:, 21:25] = np.maximum(0, b[:, 21:25])
        b[:, 25:28] = levels[rng.integers(0, 3, size=(e_ - s_, 3))]
        X[s_:e_] = b
        y[s_:e_] = ((b @ w + 0.5 * b[:, 0] * b[:, 1] + rng.standard_normal(e_ - s_, dtype=np.float32)) > 0)

    with ThreadPoolExecutor(max_workers=max(1, threads)) as ex:
        list(ex.map(fill, range(len(starts))))
    return X, y


def run_config5(ours, lib, rank, world, local, host_threads, total_rows, barrier, allmax):
    """BASELINE.json configs[4]: Higgs-shaped 100M x 28 rows sharded over the ranks (each rank generates only its
    shard), tree_learner=data; the common bin mappers come from a 200k-row sample every rank draws identically."""
    t_all = time.perf_counter()
    rows = total_rows // world
    Xr, _ = gen_rows_threaded(200_000, 555, host_threads)
    dsp = PARAMS + f" num_threads={host_threads} gpu_device_id={local}"
    ref_ds = ours.dataset_from_mat(Xr, dsp)
    Xs, ys = gen_rows_threaded(rows, 1000 + rank, host_threads)
    t_gen = time.perf_counter() - t_all
    t0 = time.perf_counter()
    ds = ours.dataset_from_mat(Xs, dsp, reference=ref_ds)
    ours.set_field(ds, "label", ys)
    t_ds = time.perf_counter() - t0
    del Xs
    params = dsp + (f" tree_learner=data num_machines={world}" if world > 1 else "")
    b = ours.booster_create(ds, params)
    for _ in range(2):
        ours.update(b)
    barrier()
    t0 = time.perf_counter()
    k = 5
    for _ in range(k):
        ours.update(b)
    barrier()
    dt = allmax(time.perf_counter() - t0)
    hist_ms, hist_calls, hist_bytes = C.c_double(), C.c_int64(), C.c_double()
    ours.check(lib.LGBM_B200_BoosterProfile(b, 1, 1, C.byref(hist_ms), C.byref(hist_calls), C.byref(hist_bytes)))
    ours.update(b)
    ph = (C.c_double * 8)()
    ours.check(lib.LGBM_B200_BoosterPhases(b, ph))
    ours.check(lib.LGBM_B200_BoosterProfile(b, 0, 0, C.byref(hist_ms), C.byref(hist_calls), C.byref(hist_bytes)))
    ours.booster_free(b); ours.dataset_free(ds); ours.dataset_free(ref_ds)
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak))
    except Exception:
        pass
    ach = (hist_bytes.value / 1e9) / (hist_ms.value / 1e3) if hist_ms.value > 0 else 0.0
    return {"workload": f"Higgs-shaped {total_rows}x{N_FEAT}, rows sharded over {world} rank(s), tree_learner=data (configs[4])",
            "rows_total": rows * world, "rows_per_gpu": rows, "iterations_timed": k, "ms_per_iter": 1e3 * dt / k,
            "iters_per_s": k / dt, "hist_frac_of_hbm_peak": ach / peak, "hist_kernel_ms_per_iter": ph[1],
            "hist_plus_allreduce_ms_per_iter": ph[0], "find_split_ms_per_iter": ph[2], "partition_ms_per_iter": ph[4],
            "rounds_per_tree": ph[7], "generate_s": t_gen, "dataset_create_s": t_ds,
            "dataset_create_h2d_gb_per_s": rows * N_FEAT * 4 / 1e9 / t_ds if t_ds > 0 else None}


def model_hash(text):
    return int(hashlib.sha256(text.split("feature_importances")[0].encode()).hexdigest()[:15], 16)


def run_ours(args):
    rank, world, local = dist_env()
    import capi
    ours = capi.ours()
    lib = ours.lib
    lib.LGBM_B200_LaunchCount.restype = C.c_int64
    use_dist = world > 1
    cores = os.cpu_count() or 1
    host_threads = max(1, cores // world)   # torchrun pins OMP_NUM_THREADS=1: ask for this rank's share of the cores
    if use_dist:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl")
    n = args.rows
    X, y = higgs_shaped(n)
    ds_params = PARAMS + f" num_threads={host_threads}"
    ds_full = None
    if use_dist:
        # identical bin mappers on every rank: a reference dataset from the first 200k rows, then this rank's shard
        ref_ds = ours.dataset_from_mat(X[:200_000], ds_params + f" gpu_device_id={local}")
        lo, hi = rank * n // world, (rank + 1) * n // world
        Xs, ys = np.ascontiguousarray(X[lo:hi]), np.ascontiguousarray(y[lo:hi])
        ds = ours.dataset_from_mat(Xs, ds_params + f" gpu_device_id={local}", reference=ref_ds)
        ours.set_field(ds, "label", ys)
        if rank == 0:   # the unsharded matrix with the same mappers, for the N-rank == 1-rank check
            ds_full = ours.dataset_from_mat(X, ds_params + f" gpu_device_id={local}", reference=ref_ds)
            ours.set_field(ds_full, "label", y)
        # the rendezvous is made ONCE per process (the reference's Network::Init, src/network/network.cpp:45-58):
        # rank 0 creates the NCCL id, torch.distributed carries it; every booster of this process trains over it
        uid = np.zeros(128, dtype=np.uint8)
        if rank == 0:
            ours.check(lib.LGBM_NetworkGetNCCLUniqueId(uid.ctypes.data_as(C.POINTER(C.c_uint8))))
        t = torch.from_numpy(uid).cuda()
        dist.broadcast(t, 0)
        uid = t.cpu().numpy()
        ours.check(lib.LGBM_NetworkInitNCCL(uid.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_int(rank), C.c_int(world)))
        params = ds_params + f" tree_learner=data num_machines={world} gpu_device_id={local}"
        local_rows = hi - lo
    else:
        ds = ours.dataset_from_mat(X, ds_params + f" gpu_device_id={local}")
        ours.set_field(ds, "label", y)
        params = ds_params + f" gpu_device_id={local}"
        local_rows = n
    X_e2e = None if use_dist else X

    def barrier():
        if use_dist:
            torch.cuda.synchronize()
            dist.barrier()

    # one-time process cost (CUDA context + module load, NCCL communicator) is paid before any timing
    from lightgbmv1_b200 import _native as _N
    _N.CudaLearner(local).close()
    # ... and so is one untimed pass over the e2e path itself (dataset from the matrix, booster, W iterations, scores
    # back): the device pool and the page-locked staging the timed pass reuses are grown here, as the W warm-up
    # iterations do for the device-resident number
    score_out = np.empty(local_rows, dtype=np.float64)
    score_out.fill(0.0)   # (maps the pages: np.zeros would not)
    if use_dist:
        ds_w = ours.dataset_from_mat(Xs, ds_params + f" gpu_device_id={local}", reference=ref_ds)
        ours.set_field(ds_w, "label", ys)
    else:
        ds_w = ours.dataset_from_mat(X_e2e, ds_params + f" gpu_device_id={local}")
        ours.set_field(ds_w, "label", y)
    b = ours.booster_create(ds_w, params)
    for _ in range(max(1, args.warmup)):
        ours.update(b)
    ours.get_predict(b, local_rows, out=score_out)
    ours.booster_free(b)
    ours.dataset_free(ds_w)

    # ---- e2e pass, from host buffers: DatasetCreateFromMat (H2D of the raw float matrix, binned on the device) +
    # SetField(label) + BoosterCreate + K iterations + GetPredict (D2H of all scores into a host buffer the caller owns
    # -- allocated and touched before the clock starts, in both arms: mapping fresh pages is the OS's time, not the path's)
    barrier()
    t0 = time.perf_counter()
    if use_dist:
        ds_e = ours.dataset_from_mat(Xs, ds_params + f" gpu_device_id={local}", reference=ref_ds)
        ours.set_field(ds_e, "label", ys)
    else:
        ds_e = ours.dataset_from_mat(X_e2e, ds_params + f" gpu_device_id={local}")
        ours.set_field(ds_e, "label", y)
    b = ours.booster_create(ds_e, params)
    for _ in range(args.steps):
        ours.update(b)
    scores = ours.get_predict(b, local_rows, out=score_out)
    barrier()
    e2e_dt = time.perf_counter() - t0
    h2d_total = local_rows * (N_FEAT * 4 + 4)  # the raw float32 matrix + float labels
    d2h_total = local_rows * 8
    ours.booster_free(b)
    ours.dataset_free(ds_e)
    del scores

    # ---- timed pass: W warm-up + exactly K iterations, device-resident
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
    phases = np.zeros(8)
    ph = (C.c_double * 8)()
    for _ in range(prof_iters):
        ours.update(b)
        ours.check(lib.LGBM_B200_BoosterPhases(b, ph))
        hist_launches += int(ph[7])   # rounds the tree needed = k_hist launches that did work (each covers up to 16 leaves)
        phases += np.array(list(ph))
    ours.check(lib.LGBM_B200_BoosterProfile(b, 0, 0, C.byref(hist_ms), C.byref(hist_calls), C.byref(hist_bytes)))
    ours.booster_free(b)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = (hist_bytes.value / 1e9) / (hist_ms.value / 1e3) if hist_ms.value > 0 else 0.0

    # ---- train_check
    check = {}
    check_params = params + " boost_from_average=false"
    if use_dist:
        # (a) the N-rank model; (b) rank 0 alone on the unsharded matrix, same mappers, data-parallel count semantics
        b = ours.booster_create(ds, check_params)
        for _ in range(DP_CHECK_ITERS):
            ours.update(b)
        text_dp = ours.model_string(b)
        ours.booster_free(b)
        h = model_hash(text_dp)
        th = torch.tensor([h], dtype=torch.int64, device="cuda")
        lst = [torch.zeros_like(th) for _ in range(world)]
        dist.all_gather(lst, th)
        ranks_agree = all(int(x) == h for x in lst)
        if rank == 0:
            os.environ["LGBCU_FORCE_DP_COUNTS"] = "1"
            b1 = ours.booster_create(ds_full, ds_params + f" tree_learner=serial gpu_device_id={local} boost_from_average=false")
            del os.environ["LGBCU_FORCE_DP_COUNTS"]
            for _ in range(DP_CHECK_ITERS):
                ours.update(b1)
            text_1 = ours.model_string(b1)
            ours.booster_free(b1)
            same, ntrees, first = capi.compare_models(text_1, text_dp, rtol_value=0.0)
            check = {"kind": f"{world}-rank data-parallel model vs 1-rank model on the unsharded matrix (data-parallel count "
                             "semantics, same bin mappers, boost_from_average=false)",
                     "trees_compared": ntrees, "trees_identical": same, "first_mismatch": first,
                     "dp_equals_single": bool(same == ntrees and ntrees == DP_CHECK_ITERS), "all_ranks_agree": bool(ranks_agree)}
        config5 = None
        if not args.no_config5:
            ours.dataset_free(ds)
            if ds_full is not None:
                ours.dataset_free(ds_full)

            def allmax(v):
                t5 = torch.tensor([v], dtype=torch.float64, device="cuda")
                dist.all_reduce(t5, op=dist.ReduceOp.MAX)
                return float(t5[0])
            try:
                config5 = run_config5(ours, lib, rank, world, local, host_threads, args.config5_rows, barrier, allmax)
            except Exception as ex:  # the headline line must still be printed
                config5 = {"error": str(ex)[:300]}
        dist.barrier()
        ours.check(lib.LGBM_NetworkFree())
        dist.destroy_process_group()
    if rank != 0:
        return 0
    # ---- cpu_baseline (N=1 only): the reference CPU learner on the same workload, few iterations; its model is kept
    cpu = None
    if world == 1 and not args.no_cpu_baseline and os.path.exists(capi.REF_SO):
        ref = capi.reference()
        rp = PARAMS + f" force_col_wise=true num_threads={cores}"
        rds = ref.dataset_from_mat(X, rp)
        ref.set_field(rds, "label", y)
        rb = ref.booster_create(rds, rp)
        ref.update(rb)
        k = PARITY_ITERS - 1
        t0 = time.perf_counter()
        for _ in range(k):
            ref.update(rb)
        cdt = time.perf_counter() - t0
        cpu = {"value": k / cdt, "unit": "iters/s", "cores": cores, "kind": "reference",
               "sample": f"same {n}x{N_FEAT} matrix and parameters, {k} iterations after 1 warm-up, force_col_wise, all host threads"}
        text_ref = ref.model_string(rb)
        pred_ref = ref.get_predict(rb, n)
        ref.booster_free(rb)
        # the reference's other histogram layout (row-wise multi-value bins), same dataset, same threads
        try:
            rw = ref.booster_create(rds, PARAMS + f" force_row_wise=true num_threads={cores}")
            ref.update(rw)
            t0 = time.perf_counter()
            for _ in range(4):
                ref.update(rw)
            cpu["row_wise_value"] = 4 / (time.perf_counter() - t0)
            cpu["row_wise_sample"] = "4 iterations after 1 warm-up, force_row_wise"
            cpu["col_wise_value"] = cpu["value"]
            if cpu["row_wise_value"] > cpu["value"]:   # the baseline is the reference's faster layout
                cpu["value"] = cpu["row_wise_value"]
                cpu["sample"] = (f"same {n}x{N_FEAT} matrix and parameters, all host threads; the faster of the reference's two "
                                 f"layouts: force_row_wise, 4 iterations after 1 warm-up (force_col_wise, whose model train_check "
                                 f"compares: {k} iterations, col_wise_value)")
            ref.booster_free(rw)
        except Exception as ex:
            cpu["row_wise_value"] = None; cpu["row_wise_sample"] = str(ex)[:200]
        ref.dataset_free(rds)
        b = ours.booster_create(ds, params)
        for _ in range(PARITY_ITERS):
            ours.update(b)
        text_ours = ours.model_string(b)
        pred_ours = ours.get_predict(b, n)
        ours.booster_free(b)
        same, ntrees, first = capi.compare_models(text_ref, text_ours, rtol_value=1e-9)
        p0 = np.clip(pred_ref, 1e-15, 1 - 1e-15); p1 = np.clip(pred_ours, 1e-15, 1 - 1e-15)
        ll = lambda p: float(-np.mean(y * np.log(p) + (1 - y) * np.log(1 - p)))
        check = {"kind": "reference CPU learner (oracle/_ref) vs this library, same matrix, same start",
                 "trees_compared": ntrees, "trees_identical": same, "first_mismatch": first,
                 "max_abs_score_diff": float(np.max(np.abs(pred_ref - pred_ours))),
                 "logloss_reference": ll(p0), "logloss_ours": ll(p1)}
    if world == 1:
        config5 = None
        if not args.no_config5:
            ours.dataset_free(ds)
            X = y = X_e2e = None
            try:
                config5 = run_config5(ours, lib, 0, 1, local, host_threads, args.config5_rows, lambda: None, lambda v: v)
            except Exception as ex:
                config5 = {"error": str(ex)[:300]}
    line = {
        "metric": "boosting iters/sec", "value": args.steps / dt, "unit": "iters/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "int64 fixed-point histograms / f64 split search",
        "data": "synthetic",
        "config": config_of(n),
        "parallelism": f"data-parallel x{world} (rows sharded, NCCL)" if world > 1 else "single GPU",
        "clocks": clocks,
        "e2e": {"value": args.steps / e2e_dt, "unit": "iters/s", "h2d_bytes_per_step": h2d_total / args.steps,
                "d2h_bytes_per_step": d2h_total / args.steps,
                "what": "LGBM_DatasetCreateFromMat (host sample -> bin mappers, H2D of the float32 matrix, value->bin on the device) + "
                        "SetField(label) + LGBM_BoosterCreate + K x LGBM_BoosterUpdateOneIter + LGBM_BoosterGetPredict, / K"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_hist (block-shared sub-histogram of 32 features, lane = feature, one byte load per (row, lane), "
                                               "32-bit shared atomic pairs, DSMEM cluster reduce; global-atomic path for leaves < 256 rows)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None,
                     "traffic": None, "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                     "launches": hist_launches, "leaf_histograms": int(hist_calls.value), "ms": hist_ms.value, "bytes": hist_bytes.value,
                     "algorithmic_bytes_per_launch": hist_bytes.value / max(1, hist_launches),
                     "note": "achieved is the average over every histogram launch of %d iterations (one launch per round of tree "
                             "growth, each building the histograms of up to 16 leaves), CUDA events around the kernel alone; "
                             "DRAM traffic per launch is in the ncu captures under profiles/, not measurable in the run" % prof_iters},
        "phases_ms_per_iter": {"hist_plus_allreduce": phases[0] / prof_iters, "hist_kernel": phases[1] / prof_iters,
                               "find_split_and_control": phases[2] / prof_iters, "partition": phases[4] / prof_iters,
                               "rounds": phases[7] / prof_iters},
        "cpu_baseline": cpu,
        "train_check": check,
        "extra": {"config5": config5},
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
    ap.add_argument("--no-config5", action="store_true", help="skip the 100M-row data-parallel workload (extra.config5)")
    ap.add_argument("--config5-rows", type=int, default=CONFIG5_ROWS)
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
