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


def workload_config():
    """`config` of BOTH arms (the driver compares them): the workload, nothing about how an arm computes it."""
    return {"workload": "configs[2]: online F2F denoise, 1080p gray sigma=25 synthetic seq, DnCNN-25 layout (weights pre-trained on "
                        "synthetic textures: tests/golden/dncnn_s25_synth.npz), iter=20 per frame, TV-L1 backward flow of every "
                        "frame pair (5 scales, 5 warps, main.c defaults) computed inside the timed region; 1 frame per step per "
                        "GPU, ranks denoise independent chunks (no collective)",
            "frame": "%dx%d" % (W, H), "iter": ITERS,
            "l2": "inputs larger than L2: every layer streams 265 MB activations (4.8 GB workspace per step)"}


def _sample(torch, cores):
    """One bounded sample of a frame's work on the host cores: 1 fine-tuning iteration + 1 eval forward + the TV-L1 flow.
    -> (seconds measured, seconds per frame extrapolated linearly to iter = ITERS, parts)"""
    t_it, t_ev = _oracle_frame_seconds(torch, cores, 1)
    t_fl = _oracle_flow_seconds(cores)
    return t_it + t_ev + t_fl, ITERS * t_it + t_ev + t_fl, (t_it, t_ev, t_fl)


def cpu_baseline():
    import torch
    cores = os.cpu_count() or 1
    meas, frame, (t_it, t_ev, t_fl) = _sample(torch, cores)
    return {"value": 1.0 / frame, "unit": "frames/s", "cores": cores, "kind": "port",
            "sample": "1 fine-tuning iteration (%.1f s) + 1 eval forward (%.1f s) + the TV-L1 flow (%.1f s, tvl1flow_lib.c) of "
                      "one 1920x1080 frame, torch %s CPU FP32 %d threads; extrapolated to iter=20 per frame"
                      % (t_it, t_ev, t_fl, torch.__version__, cores)}


C0_H, C0_W, C0_FRAMES = 540, 960, 2      # BASELINE.json configs[0] (the reference's own CPU-runnable case), first frames


def config0_reference(torch, cores):
    """A NON-extrapolated reference point: the FP32 CPU loop on the first C0_FRAMES frames of configs[0] (960x540,
    analytic flows, iter = 20 each, eval forward, PSNR) -- every iteration really runs.  -> dict"""
    import numpy as np
    from frame2frame_b200 import synth
    from frame2frame_b200.blind_denoising import psnr
    from oracle import dncnn_oracle as do
    torch.set_num_threads(cores)
    model = do.load_numpy_state(do.OracleDnCNN(residual=False), synth.pretrained_state_dict())
    opt = do.make_adam(model)
    fr = [synth.noisy_frame(C0_H, C0_W, t) / np.float32(255.0) for t in range(C0_FRAMES + 1)]
    t0 = time.perf_counter()
    ps = []
    for t in range(1, C0_FRAMES + 1):
        den = do.finetune_frame(model, opt, fr[t], fr[t - 1], synth.backward_flow(C0_H, C0_W, t), ITERS)
        ps.append(float(psnr(synth.clean_frame_u8(C0_H, C0_W, t) / np.float32(255.0), den)))
    dt = time.perf_counter() - t0
    return {"workload": "configs[0]: %d frames 960x540, analytic flows, iter=20, every iteration executed (no extrapolation)" % C0_FRAMES,
            "value": C0_FRAMES / dt, "unit": "frames/s", "seconds": dt, "cores": cores, "psnr_db": ps}


def reference_arm(args, rank, world):
    """The reference's CPU loop on the host cores.  A step = ONE bounded sample of a frame's work (1 iteration + 1 eval
    forward + the flow); `ms_per_step` is that measured time, `value` the frames/s it implies for iter = 20
    (`extrapolation` says how), so steps x ms_per_step is what this run really took."""
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    for _ in range(min(args.warmup, 1)):
        _sample(torch, cores)
    meas = frame = 0.0
    parts = [0.0, 0.0, 0.0]
    for _ in range(args.steps):
        m, f, p = _sample(torch, cores)
        meas += m
        frame += f
        parts = [a + b for a, b in zip(parts, p)]
    meas, frame = meas / args.steps, frame / args.steps
    parts = [v / args.steps for v in parts]
    v = 1.0 / frame
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": meas * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(),
        "extrapolation": {"frame_ms": frame * 1e3, "formula": "iter * t_iteration + t_eval + t_flow", "iter": ITERS,
                          "t_iteration_s": parts[0], "t_eval_s": parts[1], "t_flow_s": parts[2],
                          "note": "value = 1000 / frame_ms; ms_per_step is the measured sample (1 iteration + eval + flow)"},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
                         "sample": "%d x (1 iteration + 1 eval forward + TV-L1 flow) of a 1920x1080 frame, torch CPU FP32 %d threads; the "
                                   "reference's DnCNN submodule is un-vendored, so its loop is restated with the same stock "
                                   "torch modules (oracle/dncnn_oracle.py); flow by the reference's tvl1flow_lib.c (oracle/_ref)"
                                   % (args.steps, cores)},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if args.gpus == 1 and not getattr(args, "no_config0", False):
        line["config0"] = config0_reference(torch, cores)
    emit(line)


def _traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed ncu --set full summary
    (profiles/r02_traffic.json, written by tools/ncu_summary.py from the .ncu-rep) -- None when no capture is committed."""
    fn = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_traffic.json")
    try:
        d = json.load(open(fn))
        return d.get(kernel), "profiles/r02_traffic.json (%s)" % d.get("source", "ncu --set full")
    except Exception:
        return None, "no ncu --set full capture committed for this kernel"


def _golden(name):
    try:
        return json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", name)))
    except Exception:
        return None


def parity_leg(torch, dev, gold, nframes):
    """The CUDA loop from the pretrained weights on the first frames of a committed FP32-oracle record (analytic
    flows): per-frame PSNR difference against it, and the device time per frame.  Not part of any timed region."""
    import numpy as np
    from frame2frame_b200 import synth
    from frame2frame_b200.blind_denoising import OnlineF2F, load_network, psnr
    w, h = (int(v) for v in gold["frame"].split("x"))
    n = min(nframes, len(gold["frames"]))
    fixture = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests", "golden", "dncnn_s25_synth.npz")
    eng = OnlineF2F(load_network(fixture, dev))
    host = [torch.from_numpy(synth.noisy_frame(h, w, t) / np.float32(255.0))[None, None].pin_memory() for t in range(n + 1)]
    flows = {t: torch.from_numpy(synth.backward_flow(h, w, t))[None].to(dev) for t in range(1, n + 1)}
    outs = [torch.empty((1, 1, h, w), dtype=torch.float32).pin_memory() for _ in range(n + 1)]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in eng.run_sequence(host, gold["iters"], flows=flows, out=outs):
        pass
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    d = []
    for t in range(1, n + 1):
        clean = synth.clean_frame_u8(h, w, t) / np.float32(255.0)
        d.append(float(psnr(clean, outs[t][0, 0].numpy())) - gold["frames"][t - 1]["psnr_oracle"])
    return d, n / dt


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

    # ---- `value`: frames resident in HBM, no instrumentation inside the timed region ----
    run_device(0, args.warmup)
    prof = (C.c_double * 5)()
    lib.f2f_profile_read(prof)
    launches_before = prof[4]
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run_device(args.warmup, args.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    lib.f2f_profile_read(prof)
    launches = prof[4] - launches_before
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * args.steps / (ms_max * 1e-3)

    # ---- `e2e`: host frames in, denoised frames out, through OnlineF2F.run_sequence ----
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

    # ---- roofline: per-launch CUDA events around every 64->64 convolution / weight-gradient launch, in a SEPARATE pass
    # of the same loop after the timed regions (the event records sit between the PDL-chained launches, so they must
    # not be inside `value`) ----
    psteps = min(args.steps, 3)
    lib.f2f_profile_dncnn(1)
    torch.cuda.synchronize()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    run_device(args.warmup + args.steps, psteps)
    p1.record()
    torch.cuda.synchronize()
    pms = p0.elapsed_time(p1)
    dn = (C.c_double * 4)()
    lib.f2f_profile_dncnn_read(dn)
    lib.f2f_profile_dncnn(0)

    if rank != 0:
        return
    pk, pk_kind = peaks()
    P = H * W
    conv_ms, conv_n, wg_ms, wg_n = (dn[i] for i in range(4))
    conv_tf = CONV64_FLOP_PER_PX * P * conv_n / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else None
    wg_tf = CONV64_FLOP_PER_PX * P * wg_n / (wg_ms * 1e-3) / 1e12 if wg_ms > 0 else None
    peak_tf = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
    frame_flop = (ITERS * FLOP_PER_PX_TRAIN + FLOP_PER_PX_FWD) * P
    traffic, traffic_src = _traffic("k_conv64t")
    wg_traffic, _ = _traffic("k_wgrad64")
    mode = h16.upper() + " operands/activations/gradients, FP32 accumulate, FP32 master weights, Adam and BatchNorm statistics"
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": h16, "data": "synthetic", "config": workload_config(),
        "mode": {"name": h16, "precision": mode,
                 "note": "the mode that produced `value`, `e2e` and `psnr_delta_db`; see DESIGN.md section 2 for its parity record"},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": H * W * 4, "d2h_bytes_per_step": H * W * 4,
                "api": "OnlineF2F.run_sequence (frame2frame_b200/blind_denoising.py) from/to pinned host frames; "
                       "(steps+1)/steps frames uploaded per step (the first frame of the run is only a `prev`)",
                "psnr_db_last_frame": float(quality)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "model_tflops": frame_flop * value / world / 1e12,
        "mfu": frame_flop * value / world / 1e12 / peak_tf,
        "roofline": {"bound": "tensor", "kernel": "dnt::k_conv64t (tcgen05 implicit-GEMM 3x3 64->64 conv: the forward and "
                                                  "input-gradient launches)",
                     "achieved": conv_tf, "peak": peak_tf, "peak_source": pk_kind + " bf16_tflops_sustained", "unit": "TFLOP/s",
                     "frac": conv_tf / peak_tf if conv_tf else None, "flop_per_launch": CONV64_FLOP_PER_PX * P,
                     "launches_per_step": conv_n / psteps, "avg_launch_us": conv_ms * 1e3 / conv_n if conv_n else None,
                     "share_of_step": conv_ms / pms if pms else None,
                     "measured": "CUDA events around every launch, %d steps of the same loop run after the timed regions" % psteps,
                     "wgrad64": {"achieved": wg_tf, "frac": wg_tf / peak_tf if wg_tf else None,
                                 "avg_launch_us": wg_ms * 1e3 / wg_n if wg_n else None, "share_of_step": wg_ms / pms if pms else None,
                                 "traffic": wg_traffic},
                     "traffic": traffic, "traffic_source": traffic_src},
    }
    # ---- parity against the committed FP32-oracle records (outside every timed region) ----
    g1080, g0 = _golden("f2f_1080p_ref.json"), _golden("f2f_config0_ref.json")
    if g1080 is not None:
        d, _ = parity_leg(torch, dev, g1080, 2)
        line["psnr_delta_db"] = {"vs": "tests/golden/f2f_1080p_ref.json (FP32 CPU oracle loop, 1920x1080, analytic flows)",
                                 "per_frame": d, "max_abs": float(np.max(np.abs(d))), "mode": h16}
    if world == 1 and g0 is not None:
        d, fps = parity_leg(torch, dev, g0, C0_FRAMES)
        line["config0"] = {"workload": "configs[0]: %d frames 960x540, analytic flows, iter=20, from/to pinned host frames" % C0_FRAMES,
                           "value": fps, "unit": "frames/s", "psnr_delta_db": d,
                           "note": "the reference arm times the same frames on the host cores without extrapolation (its `config0`)"}
    # ---- the contract-grade mode: FP32-class split operands on the tensor cores (f2f_dncnn_set_precise(2)) ----
    if not getattr(args, "no_precise", False) and lib.f2f_dncnn_storage_dtype():
        from frame2frame_b200 import dncnn
        dncnn.set_precise(2)
        try:
            peng = OnlineF2F(load_network(gold, dev))
            psteps2 = 3

            def prun(k0, nsteps):
                fr = [devf[(k0 + j) % nf] for j in range(nsteps + 1)]
                for _ in peng.run_sequence(fr, ITERS, lookahead=LOOKAHEAD, flow_kw=FLOW_KW):
                    pass
            prun(0, 1)
            torch.cuda.synchronize()
            q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            q0.record()
            prun(1, psteps2)
            q1.record()
            torch.cuda.synchronize()
            pfps = psteps2 / (q0.elapsed_time(q1) * 1e-3)
            blk = {"mode": "precise 2: every 64->64 operand as an FP16 hi + lo pair (22 bits), three tcgen05 products per term, "
                           "FP32 activations / gradients between layers (csrc/dncnn_precise2.cuh)",
                   "value": pfps, "unit": "frames/s", "steps": psteps2, "n_gpus": 1,
                   "model_tflops_fp32_equivalent": frame_flop * pfps / 1e12,
                   "note": "same workload as `value` (device-resident frames, flows inside), one GPU"}
            del peng
            if g1080 is not None:
                d, _ = parity_leg(torch, dev, g1080, 2)
                blk["psnr_delta_db"] = {"vs": "tests/golden/f2f_1080p_ref.json", "per_frame": d, "max_abs": float(np.max(np.abs(d)))}
            line["precise"] = blk
        finally:
            dncnn.set_precise(0)
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
        "comm": {"backend": "nccl" if world > 1 else "none (one rank)", "ranks": world,
                 "per_inner_iteration": "one all-reduce(SUM) of the flat 556032-float gradient + one all-gather of 1920 BatchNorm "
                                        "statistics per local sequence, issued after the backward pass of the last local sequence"},
        "roofline": {"bound": "tensor", "kernel": "whole train step (model FLOPs)", "achieved": tf, "peak": peak_tf,
                     "peak_source": pk_kind + " bf16_tflops_sustained", "unit": "TFLOP/s", "frac": tf / peak_tf, "traffic": None},
    }
