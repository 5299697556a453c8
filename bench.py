#!/usr/bin/env python
"""bench.py -- headline benchmark of the frame2frame hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload tvl1|f2f]

Workloads (BASELINE.json configs):
  tvl1 : configs[1] -- TV-L1 flow on synthetic 1080p gray frame pairs, 5-scale
         pyramid, default warps; frame pairs sharded over the GPUs (no
         collective on the data path).  A step = one batch of PAIRS_PER_STEP
         frame pairs per GPU.  metric = flow fields / s.
  f2f  : configs[2] -- online F2F denoising of a synthetic 1080p sequence
         (see bench_f2f.py); selected automatically once that path is built.

One JSON line on stdout (rank 0).  `value` is timed with the frames resident in
HBM; `e2e` is the same metric through the host-buffer C-ABI call
(f2f_tvl1_flow_batch_host: H2D of the frames and D2H of the flows inside the
timed region, from/to pinned host memory).  `--impl reference` times the
reference's own CPU code (oracle/_ref, tvl1flow_lib.c compiled in place) on the
host cores.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W = 1080, 1920
NSCALES = 5
NSLOTS = 16                  # frame pairs advancing in lock-step on a GPU (one "group")
PAIRS_PER_STEP = 32          # per GPU: two groups, so that the host-buffer call's upload / solve / download pipeline
                             # (f2f_tvl1_flow_batch_host double-buffers successive groups of ONE call) is exercised;
                             # 32 pairs x 2 frames x 8.3 MB = 531 MB > L2 (126 MB)
ALG_BYTES_PER_PX_ITER = 64   # read rho_c,I1wx,I1wy,grad,u1,u2,p11..p22; write u1,u2,p11..p22


from frame2frame_b200.benchutil import ClockSampler, peaks  # noqa: E402
import bench_f2f  # noqa: E402


_REAL_STDOUT = None


def quiet_stdout():
    """File descriptor 1 -> stderr for the duration of the run; emit() writes the JSON line to the real stdout."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(obj):
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.dup2(_REAL_STDOUT, 1)
    print(json.dumps(obj), flush=True)
    if _REAL_STDOUT is not None:
        os.dup2(2, 1)


def make_pairs(n, seq):
    """n consecutive synthetic 1080p frame pairs (I0 = frame t+1, I1 = frame t: the
    backward flow tvl1flow.sh:11-14 computes)."""
    import numpy as np
    from frame2frame_b200 import synth
    nf = min(n, 8) + 1                      # cycle through 8 distinct pairs; noise differs per seq
    frames = [synth.noisy_frame(H, W, t, seq=seq) for t in range(nf)]
    I0 = np.stack([frames[(k % (nf - 1)) + 1] for k in range(n)])
    I1 = np.stack([frames[(k % (nf - 1))] for k in range(n)])
    return I0, I1


def reference_arm(args, rank, world):
    """The reference's own CPU implementation (tvl1flow_lib.c built in place, OpenMP, all host
    threads) on a bounded sample of the same workload: one 1080p frame pair per step."""
    if rank != 0:
        return
    import numpy as np
    from oracle.tvl1 import OracleTvl1, RefTvl1
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)      # torchrun exports OMP_NUM_THREADS=1
    if RefTvl1.available():
        impl, kind = RefTvl1(), "reference"
        RefTvl1.set_threads(cores)
        run = lambda a, b: impl.flow(a, b, nscales=NSCALES)
    else:
        impl, kind = OracleTvl1(), "port"
        run = lambda a, b: impl.flow(a, b, nscales=NSCALES)[0]
    I0, I1 = make_pairs(max(1, min(args.steps + args.warmup, 8)), 0)
    for k in range(args.warmup):
        run(I0[k % len(I0)], I1[k % len(I0)])
    t0 = time.perf_counter()
    for k in range(args.steps):
        run(I0[(args.warmup + k) % len(I0)], I1[(args.warmup + k) % len(I0)])
    dt = time.perf_counter() - t0
    v = args.steps / dt
    emit({
        "impl": "reference", "metric": "TV-L1 flow fields/sec @1080p gray", "value": v, "unit": "fields/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: tvl1flow 1080p gray pairs, 5 scales, 5 warps, eps 0.01 "
                               "(bounded sample: 1 pair per step)"},
        "cpu_baseline": {"value": v, "unit": "fields/s", "cores": cores, "kind": kind,
                         "sample": "%d x 1 pair 1920x1080, OpenMP %d threads" % (args.steps, cores)},
        "e2e": {"value": v, "unit": "fields/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def cpu_baseline(seconds_budget=25.0):
    """Reference CPU path on this box's host cores, bounded sample (rank 0, N=1)."""
    from oracle.tvl1 import OracleTvl1, RefTvl1
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)      # torchrun exports OMP_NUM_THREADS=1
    if RefTvl1.available():
        impl, kind = RefTvl1(), "reference"
        RefTvl1.set_threads(cores)
        run = lambda a, b: impl.flow(a, b, nscales=NSCALES)
    else:
        impl, kind = OracleTvl1(), "port"
        run = lambda a, b: impl.flow(a, b, nscales=NSCALES)[0]
    I0, I1 = make_pairs(2, 0)
    n, t0 = 0, time.perf_counter()
    while True:
        run(I0[n % 2], I1[n % 2])
        n += 1
        dt = time.perf_counter() - t0
        if n >= 2 and (dt > seconds_budget * 0.6 or n >= 4):
            break
        if dt > seconds_budget:
            break
    return {"value": n / dt, "unit": "fields/s", "cores": cores, "kind": kind,
            "sample": "%d pair(s) 1920x1080 in %.1f s, OpenMP %d threads" % (n, dt, cores)}


def ours_tvl1(args, rank, world, local_rank, with_cpu=True):
    import numpy as np
    import torch
    import torch.distributed as dist
    from frame2frame_b200 import _lib
    from frame2frame_b200.tvl1 import make_params, tvl1flow

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    lib = _lib.lib()                                   # raises if the CUDA library is missing
    n = PAIRS_PER_STEP
    I0h, I1h = make_pairs(n, rank)
    # pinned host copies for the e2e leg
    pI0 = torch.from_numpy(I0h).pin_memory()
    pI1 = torch.from_numpy(I1h).pin_memory()
    pU = torch.empty((n, 2, H, W), dtype=torch.float32).pin_memory()
    dI0, dI1 = pI0.to(dev), pI1.to(dev)
    p = make_params(W, H, nscales=NSCALES)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        return tvl1flow(dI0, dI1, nslots=NSLOTS, nscales=NSCALES)

    def step_host():
        _lib.check(lib.f2f_tvl1_flow_batch_host(pI0.data_ptr(), pI1.data_ptr(), pU.data_ptr(), n, W, H, C.byref(p),
                                                None, NSLOTS), "f2f_tvl1_flow_batch_host")

    # ---- device-resident timing (`value`) ----
    for _ in range(args.warmup):
        step_device()
    prof = (C.c_double * 5)()
    lib.f2f_profile_read(prof)
    launches_before = prof[4]
    lib.f2f_profile_enable(1)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        flow = step_device()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    lib.f2f_profile_read(prof)
    lib.f2f_profile_enable(0)
    iter_ms, slot_iters0, px0, launches0, launches_after = [prof[i] for i in range(5)]
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * n * args.steps / (ms_max * 1e-3)

    # ---- end to end through the host-buffer C ABI (`e2e`) ----
    step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * n * args.steps / float(t.item())
    checksum = float(pU.double().abs().mean())

    if rank != 0:
        return
    pk, pk_kind = peaks()
    alg_bytes = ALG_BYTES_PER_PX_ITER * px0 * slot_iters0
    achieved = alg_bytes / (iter_ms * 1e-3) / 1e9 if iter_ms > 0 else None
    line = {
        "metric": "TV-L1 flow fields/sec @1080p gray", "value": value, "unit": "fields/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: tvl1flow on synthetic 1080p gray frame pairs, 5-scale pyramid, 5 warps, "
                               "eps 0.01, frame pairs sharded over GPUs",
                   "pairs_per_step_per_gpu": n, "pairs_in_lock_step": NSLOTS, "frame": "%dx%d" % (W, H),
                   "l2": "inputs larger than L2 (531 MB of frames + 3.4 GB of solver state per step)"},
        "e2e": {"value": e2e_value, "unit": "fields/s", "h2d_bytes_per_step": 2 * n * H * W * 4,
                "d2h_bytes_per_step": 2 * n * H * W * 4, "api": "f2f_tvl1_flow_batch_host (C ABI, pinned host buffers)",
                "flow_abs_mean": checksum},
        "gpu_launches": int(launches_after - launches_before - 0) if launches_after else None,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "k_iter_strip<4> (fused primal-dual iteration, finest scale)",
                     "achieved": achieved, "peak": pk["hbm_gbs"], "peak_source": pk_kind, "unit": "GB/s",
                     "frac": (achieved / pk["hbm_gbs"]) if achieved else None,
                     "frac_of_nominal_8TBs": (achieved / 8000.0) if achieved else None,
                     "alg_bytes_per_px_iter": ALG_BYTES_PER_PX_ITER,
                     "px_iters_per_step": px0 * slot_iters0 / args.steps,
                     "launches_per_step": launches0 / args.steps,
                     "avg_launch_us": iter_ms * 1e3 / launches0 if launches0 else None,
                     "traffic": bench_f2f._traffic("k_iter_strip")[0], "traffic_source": bench_f2f._traffic("k_iter_strip")[1]},
    }
    if world == 1 and not args.no_cpu and with_cpu:
        line["cpu_baseline"] = cpu_baseline()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="f2f", choices=["f2f", "tvl1", "batch"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-tvl1", action="store_true", help="f2f workload: skip the secondary TV-L1 measurement")
    ap.add_argument("--no-batch", action="store_true", help="f2f workload: skip the secondary batch-mode (NCCL) measurement")
    ap.add_argument("--no-precise", action="store_true", help="f2f workload: skip the secondary precise-mode measurement")
    ap.add_argument("--no-config0", action="store_true", help="reference arm: skip the non-extrapolated configs[0] point")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries the JSON line and nothing else: libraries that write to file descriptor 1 on their own (NCCL's
    # version banner under torchrun) are sent to stderr for the duration of the run
    quiet_stdout()
    bench_f2f.emit = emit
    if args.impl == "reference":
        if args.workload == "f2f":
            bench_f2f.reference_arm(args, rank, world)
        else:
            reference_arm(args, rank, world)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback in the product path)")
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        if args.workload == "tvl1":
            line = ours_tvl1(args, rank, world, local_rank)
        elif args.workload == "batch":
            line = bench_f2f.ours_batch(args, rank, world, local_rank, ClockSampler, peaks)
        else:
            line = bench_f2f.ours(args, rank, world, local_rank, ClockSampler, peaks)
            keys = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "scaling", "config", "e2e", "roofline",
                    "cpu_baseline", "clocks", "gpu_launches")
            if not args.no_tvl1:
                # the metric's second half (TV-L1 flow fields/sec, configs[1]) at every N, >= 5 timed steps
                targs = argparse.Namespace(**vars(args))
                targs.steps, targs.warmup = max(5, min(args.steps, 8)), 3
                t = ours_tvl1(targs, rank, world, local_rank, with_cpu=True)
                if rank == 0:
                    line["tvl1"] = {k: t[k] for k in keys if k in t}
            if not args.no_batch:
                # configs[3]: batch-mode fine-tune with the NCCL gradient all-reduce (the one path with a collective)
                bargs = argparse.Namespace(**vars(args))
                bargs.steps, bargs.warmup = 5, 2
                t = bench_f2f.ours_batch(bargs, rank, world, local_rank, ClockSampler, peaks)
                if rank == 0:
                    line["batch"] = {k: t[k] for k in keys + ("model_tflops_per_gpu", "comm") if k in t}
        if rank == 0:
            emit(line)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
