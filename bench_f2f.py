"""F2F workload of bench.py: BASELINE.json configs[2] -- online frame-to-frame denoising of a
synthetic 1080p gray sequence (sigma = 25), DnCNN-S-25 layout, --iter 20 per frame.

A step = ONE frame of the online loop (blind_denoising.py:187-256): TV-L1 backward flow of the
noisy pair on the device (tvl1flow.sh:11-14), warp + masks once, 20 x (train forward, masked L1,
backward, Adam), eval forward.  With N > 1 GPUs every rank denoises its own chunk of the sequence
(the reference's --first/--last sharding, SURVEY 8e: weights carry over from frame to frame, so
the online loop itself does not shard; no collective on the data path) -> weak scaling.
"""
import ctypes as C
import json
import os
import time

H, W = 1080, 1920
ITERS = 20
FLOP_PER_PX_FWD = 1108224
FLOP_PER_PX_TRAIN = 3323520
CONV64_FLOP_PER_PX = 2 * 9 * 64 * 64


def emit(obj):                      # bench.py replaces this with its clean-stdout writer
    print(json.dumps(obj), flush=True)


LOOKAHEAD = 16                     # frame pairs per batched TV-L1 call (OnlineF2F.run_sequence)
FLOW_KW = dict(nscales=5)          # tvl1flow defaults of main.c:25-35 with the 5-scale pyramid of configs[1]
METRIC = "F2F denoised frames/sec @1080p gray"


def _frames(n, seq, synth):
    import numpy as np
    return [synth.noisy_frame(H, W, t, seq=seq) / np.float32(255.0) for t in range(n)]


def _oracle_frame_seconds(torch, cores, train_iters=1):
    """Reference CPU path (oracle/dncnn_oracle.py: stock PyTorch FP32 on the host cores) on a bounded
    sample: `train_iters` fine-tuning iterations + one eval forward of one 1080p frame;
    -> (seconds per train iteration, seconds per eval forward)."""
    import numpy as np
    from frame2frame_b200 import synth
    from oracle import dncnn_oracle as do
    from oracle import f2f_oracle as fo
    torch.set_num_threads(cores)
    model = do.load_numpy_state(do.OracleDnCNN(residual=False), synth.pretrained_state_dict())
    opt = do.make_adam(model)
    f = _frames(2, 0, synth)
    flow = synth.backward_flow(H, W, 1)
    t0 = time.perf_counter()
    warped, _, mask = fo.warp_and_mask(f[0], flow, f64=True)
    tc, tw, tm = (torch.from_numpy(a)[None, None] for a in (f[1], warped, mask))
    crit = torch.nn.L1Loss(reduction="sum")
    model.train()
    for _ in range(train_iters):
        loss = crit(tm * (tc - model(tc)), tm * tw)
        opt.zero_grad()
        loss.backward()
        opt.step()
    t1 = time.perf_counter()
    model.eval()
    with torch.no_grad():
        _ = tc - model(tc)
    t2 = time.perf_counter()
    return (t1 - t0) / train_iters, t2 - t1


def _oracle_flow_seconds(cores):
    """The frame's backward TV-L1 flow on the host cores: the reference's own tvl1flow_lib.c (oracle/_ref, OpenMP) when
    it is built, else the C restatement -- one 1080p pair with the parameters the device path uses (FLOW_KW)."""
    from frame2frame_b200 import synth
    from oracle.tvl1 import OracleTvl1, RefTvl1
    os.environ["OMP_NUM_THREADS"] = str(cores)          # torchrun exports OMP_NUM_THREADS=1
    f = [synth.noisy_frame(H, W, t) for t in range(2)]
    if RefTvl1.available():
        RefTvl1.set_threads(cores)
        impl = RefTvl1()
        t0 = time.perf_counter()
        impl.flow(f[1], f[0], **FLOW_KW)
    else:
        impl = OracleTvl1()
        t0 = time.perf_counter()
        impl.flow(f[1], f[0], **FLOW_KW)
    return time.perf_counter() - t0


def cpu_baseline():
    import torch
    cores = os.cpu_count() or 1
    t_it, t_ev = _oracle_frame_seconds(torch, cores, 1)
    t_fl = _oracle_flow_seconds(cores)
    sec = ITERS * t_it + t_ev + t_fl
    return {"value": 1.0 / sec, "unit": "frames/s", "cores": cores, "kind": "port",
            "sample": "1 fine-tuning iteration (%.1f s) + 1 eval forward (%.1f s) + the TV-L1 flow (%.1f s, tvl1flow_lib.c) of "
                      "one 1920x1080 frame, torch %s CPU FP32 %d threads; extrapolated to iter=20 per frame"
                      % (t_it, t_ev, t_fl, torch.__version__, cores)}


def reference_arm(args, rank, world):
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    for _ in range(min(args.warmup, 1)):
        _oracle_frame_seconds(torch, cores, 1)
    t_it = t_ev = t_fl = 0.0
    for _ in range(args.steps):
        a, b = _oracle_frame_seconds(torch, cores, 1)
        t_it += a
        t_ev += b
        t_fl += _oracle_flow_seconds(cores)
    sec = (ITERS * t_it + t_ev + t_fl) / args.steps
    v = 1.0 / sec
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[2]: online F2F denoise, 1080p gray sigma=25 synthetic seq, DnCNN-25, iter=20 per frame "
                               "(bounded sample per step: 1 fine-tuning iteration + 1 eval forward + the frame's TV-L1 flow, extrapolated to iter=20)"},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
                         "sample": "%d x (1 iteration + 1 eval forward + TV-L1 flow) of a 1920x1080 frame, torch CPU FP32 %d threads; the "
                                   "reference's DnCNN submodule is un-vendored, so its loop is restated with the same stock "
                                   "torch modules (oracle/dncnn_oracle.py)" % (args.steps, cores)},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def ours(args, rank, world, local_rank, ClockSampler, peaks):
    import numpy as np
    import torch
    import torch.distributed as dist
    from frame2frame_b200 import _lib, synth
    from frame2frame_b200.blind_denoising import OnlineF2F, load_network, psnr

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    lib = _lib.lib()                                   # raises if the CUDA library is missing
    h16 = "fp16" if lib.f2f_dncnn_storage_dtype() else "bf16"
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", "dncnn_s25_synth.npz")
    engine = OnlineF2F(load_network(gold, dev))
    nf = 4                                             # distinct noisy frames of this rank's chunk, cycled
    host = [torch.from_numpy(f)[None, None].pin_memory() for f in _frames(nf, rank, synth)]
    devf = [h.to(dev) for h in host]
    out_host = torch.empty((1, 1, H, W), dtype=torch.float32).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run_device(k0, nsteps):
        """nsteps frames of the online loop, frames resident in HBM; the backward flows of up to LOOKAHEAD
        consecutive pairs are computed in one batched TV-L1 call inside this (timed) region."""
        fr = [devf[(k0 + j) % nf] for j in range(nsteps + 1)]
        den = None
        for _, den in engine.run_sequence(fr, ITERS, lookahead=LOOKAHEAD, flow_kw=FLOW_KW):
            pass
        return den

    def run_host(k0, nsteps):
        """The same from pinned host frames to pinned host outputs: every frame is uploaded once (H2D) and every
        denoised frame read back (D2H) inside the timed region."""
        fr = [host[(k0 + j) % nf] for j in range(nsteps + 1)]
        outs = [out_host] * (nsteps + 1)
        for _ in engine.run_sequence(fr, ITERS, lookahead=LOOKAHEAD, flow_kw=FLOW_KW, out=outs):
            pass
        torch.cuda.current_stream().synchronize()

    run_device(0, args.warmup)
    prof = (C.c_double * 5)()
    lib.f2f_profile_read(prof)
    launches_before = prof[4]
    lib.f2f_profile_dncnn(1)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    den = run_device(args.warmup, args.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    dn = (C.c_double * 4)()
    lib.f2f_profile_dncnn_read(dn)
    lib.f2f_profile_dncnn(0)
    lib.f2f_profile_read(prof)
    launches = prof[4] - launches_before
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * args.steps / (ms_max * 1e-3)

    # ---- end to end: host frames in, denoised frame out, through OnlineF2F.step_frame ----
    run_host(0, 1)
    barrier()
    t0 = time.perf_counter()
    run_host(1, args.steps)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * args.steps / float(t.item())
    kk = (1 + args.steps) % nf
    quality = psnr(synth.clean_frame_u8(H, W, kk) / np.float32(255.0), out_host[0, 0].numpy())

    if rank != 0:
        return
    pk, pk_kind = peaks()
    P = H * W
    conv_ms, conv_n, wg_ms, wg_n = (dn[i] for i in range(4))
    conv_tf = CONV64_FLOP_PER_PX * P * conv_n / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else None
    wg_tf = CONV64_FLOP_PER_PX * P * wg_n / (wg_ms * 1e-3) / 1e12 if wg_ms > 0 else None
    peak_tf = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
    frame_flop = (ITERS * FLOP_PER_PX_TRAIN + FLOP_PER_PX_FWD) * P
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": h16, "data": "synthetic",
        "config": {"workload": "configs[2]: online F2F denoise, 1080p gray sigma=25 synthetic seq, DnCNN-25 layout "
                               "(weights pre-trained on synthetic textures: tests/golden/dncnn_s25_synth.npz), iter=20 per "
                               "frame, TV-L1 backward flows computed on the device inside the timed region (up to %d consecutive "
                               "pairs per batched call); 1 frame per step per GPU, " % LOOKAHEAD +
                               "ranks denoise independent chunks (no collective)",
                   "frame": "%dx%d" % (W, H), "iter": ITERS, "precision": h16.upper() + " operands/activations/gradients, FP32 accumulate, "
                   "FP32 master weights, Adam and BatchNorm statistics",
                   "l2": "inputs larger than L2: every layer streams 265 MB activations (4.8 GB workspace per step)"},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": H * W * 4, "d2h_bytes_per_step": H * W * 4,
                "api": "OnlineF2F.run_sequence (frame2frame_b200/blind_denoising.py) from/to pinned host frames; "
                       "(steps+1)/steps frames uploaded per step (the first frame of the run is only a `prev`)",
                "psnr_db_last_frame": float(quality)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "model_tflops": frame_flop * value / world / 1e12,
        "roofline": {"bound": "tensor", "kernel": "k_conv64 / k_conv64t (tcgen05 implicit-GEMM 3x3 64->64 conv: the forward and "
                                                  "input-gradient launches)",
                     "achieved": conv_tf, "peak": peak_tf, "peak_source": pk_kind + " bf16_tflops_sustained", "unit": "TFLOP/s",
                     "frac": conv_tf / peak_tf if conv_tf else None, "flop_per_launch": CONV64_FLOP_PER_PX * P,
                     "launches_per_step": conv_n / args.steps, "avg_launch_us": conv_ms * 1e3 / conv_n if conv_n else None,
                     "share_of_step": conv_ms / ms if ms else None,
                     "wgrad64": {"achieved": wg_tf, "frac": wg_tf / peak_tf if wg_tf else None,
                                 "avg_launch_us": wg_ms * 1e3 / wg_n if wg_n else None, "share_of_step": wg_ms / ms if ms else None},
                     # dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full (profiles/r01_ncu_conv64_wgrad64.txt):
                     # forward k_conv64t<1,1> 266 + 218 MB, input gradient k_conv64t<0,2> 531 + 235 MB (algorithmic 531 / 796 MB;
                     # the writes still dirty in the 126 MB L2 at kernel end are not counted by ncu); mean over the 15 + 15 launches
                     "traffic": 625e6,
                     "traffic_note": "mean of forward (484 MB) and input-gradient (766 MB) launches, ncu --set full, "
                                     "profiles/r01_ncu_conv64_wgrad64.txt; wgrad64 1041 MB (algorithmic 1061 MB)"},
    }
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline()
    return line


def ours_batch(args, rank, world, local_rank, ClockSampler, peaks):
    """BASELINE.json configs[3]: batch-mode F2F fine-tune over 8 sequences at 1080p, DnCNN gradients
    all-reduced with NCCL (frame2frame_b200/batch_f2f.py).  A step = one frame of EVERY sequence:
    warp + masks once per sequence, 20 x (per local sequence: train forward, masked L1, backward;
    then ONE gradient all-reduce + BatchNorm-statistics all-gather + Adam), eval forward per
    sequence.  Total work is fixed (8 sequences) -> strong scaling over 1/2/4/8 GPUs."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from frame2frame_b200 import _lib, synth
    from frame2frame_b200.batch_f2f import BatchF2F
    from frame2frame_b200.blind_denoising import load_network

    NSEQ = 8
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    h16 = "fp16" if _lib.lib().f2f_dncnn_storage_dtype() else "bf16"
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", "dncnn_s25_synth.npz")
    eng = BatchF2F(load_network(gold, dev), NSEQ)
    nf = 3
    frames = {s: [torch.from_numpy(f)[None, None].to(dev) for f in _frames(nf, s, synth)] for s in eng.local}
    flow = torch.from_numpy(synth.backward_flow(H, W, 1))[None].to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(k):
        currs = [frames[s][(k + 1) % nf] for s in eng.local]
        prevs = [frames[s][k % nf] for s in eng.local]
        eng.finetune(currs, prevs, [flow] * len(currs), ITERS)
        return [eng.denoise(c) for c in currs]

    # end-to-end leg: the same step from pinned host frames (H2D of curr and prev of every local sequence) to
    # pinned host results (D2H of every denoised frame) inside the timed region
    host = {s: [f.cpu().pin_memory() for f in frames[s]] for s in eng.local}
    out_host = [torch.empty((1, 1, H, W), dtype=torch.float32).pin_memory() for _ in eng.local]

    def step_host(k):
        currs = [host[s][(k + 1) % nf].to(dev, non_blocking=True) for s in eng.local]
        prevs = [host[s][k % nf].to(dev, non_blocking=True) for s in eng.local]
        eng.finetune(currs, prevs, [flow] * len(currs), ITERS)
        for o, c in zip(out_host, currs):
            o.copy_(eng.denoise(c), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for k in range(args.warmup):
        step(k)
    lib = _lib.lib()
    prof = (C.c_double * 5)()
    lib.f2f_profile_read(prof)
    launches_before = prof[4]
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(args.steps):
        step(args.warmup + k)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    lib.f2f_profile_read(prof)
    launches = prof[4] - launches_before
    step_host(args.warmup + args.steps)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        step_host(args.warmup + k)
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = NSEQ * args.steps / float(t.item())
    if rank != 0:
        return
    value = NSEQ * args.steps / (ms_max * 1e-3)
    P = H * W
    frame_flop = (ITERS * FLOP_PER_PX_TRAIN + FLOP_PER_PX_FWD) * P
    pk, pk_kind = peaks()
    peak_tf = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
    tf = frame_flop * value / world / 1e12
    return {
        "metric": "batch-mode F2F fine-tuned frames/sec @1080p gray", "value": value, "unit": "frames/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": h16, "data": "synthetic",
        "config": {"workload": "configs[3]: batch-mode F2F fine-tune, 8 sequences x 1080p gray sigma=25, iter=20 per frame, one "
                               "Adam step per inner iteration on the summed loss; sequences sharded over ranks, flat 556032-float "
                               "gradient all-reduced with NCCL, BatchNorm batch statistics all-gathered (ordered EMA)",
                   "frame": "%dx%d" % (W, H), "iter": ITERS, "sequences": NSEQ,
                   "l2": "inputs larger than L2 (265 MB per activation)"},
        "clocks": clocks, "model_tflops_per_gpu": tf,
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": 2 * NSEQ * H * W * 4,
                "d2h_bytes_per_step": NSEQ * H * W * 4,
                "api": "BatchF2F.finetune + denoise (frame2frame_b200/batch_f2f.py) from/to pinned host frames"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "kernel": "whole train step (model FLOPs)", "achieved": tf, "peak": peak_tf,
                     "peak_source": pk_kind + " bf16_tflops_sustained", "unit": "TFLOP/s", "frac": tf / peak_tf, "traffic": None},
    }
