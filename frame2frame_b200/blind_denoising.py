"""Model-blind video denoising by frame-to-frame fine-tuning -- the reference's
``blind_denoising.py`` (CLI flags :264-288, ``blind_denoising(**args)`` :125-259, ``psnr``
:36-41) on B200.

Same flags, same inputs (frames ``--input %03d.png``/``.tiff``, flows ``--flow %03d.flo`` from
frame i to frame i-1, reference frames ``--ref``, DnCNN-S-25 ``--network`` state dict with the
``module.`` prefix), same outputs (denoised frames, one PSNR per line in ``--output_psnr``,
``torch.save([model, optimizer])`` to ``--output_network``).  Underneath, every inner iteration
``out = curr - model(curr); loss = WarpedLoss(out, prev, flow); backward; Adam`` is ONE call into
libf2f_b200.so (f2f_step_train), the warp / validity / occlusion mask is computed once per frame
instead of once per iteration on the host, and nothing runs on the CPU but file I/O.

``OnlineF2F`` is the engine (frames and flows as tensors); ``blind_denoising()`` wraps it with
the reference's file handling.  ``--flow`` may be omitted: the backward flow is then computed on
the device by the TV-L1 kernels with the script's parameters (what tvl1flow.sh:11-14 does in a
separate process: lambda 0.2, nscales 100 -> clamp, fscale 2; ``--flow_*`` flags override them).
"""
import argparse
import os

import numpy as np
import torch
import torch.nn as nn

from ._lib import check, lib
from .dncnn import Adam, DnCNN
from .frame_io import OrderedSink, Prefetcher
from .tvl1 import read_flo, tvl1flow
from .tvl1flow_cli import SCRIPT_PARAMS
from .warped_loss import WarpedLoss, warp_mask

LR, WEIGHT_DECAY, BETAS, EPS = 5e-5, 0.00001, (0.9, 0.999), 1e-08   # blind_denoising.py:158-165


def psnr(img1, img2, peak=1):
    """blind_denoising.py:36-41 (img2 is NOT clipped)."""
    x = (np.array(img1).squeeze() - np.array(img2).squeeze()).flatten()
    return 10 * np.log10(peak ** 2 / np.mean(x ** 2))


def load_network(model_fn, device):
    """blind_denoising.py:142-147: DnCNN(channels=1, num_of_layers=17) in DataParallel([0]) + state dict.
    Accepts a torch ``.pth`` state dict or an ``.npz`` of arrays with the same keys."""
    net = DnCNN(channels=1, num_of_layers=17, residual=False)      # the paper script's DnCNN returns the noise
    model = nn.DataParallel(net, device_ids=[device.index or 0]).to(device)
    if model_fn.endswith(".npz"):
        sd = {k: torch.from_numpy(np.asarray(v, dtype=np.float32 if v.dtype.kind == "f" else v.dtype))
              for k, v in np.load(model_fn).items()}
    else:
        sd = torch.load(model_fn, map_location="cpu", weights_only=False)
        if isinstance(sd, (list, tuple)):            # an --output_network file: torch.save([model, optimizer]), :258 / :148-151
            sd = sd[0]
        if isinstance(sd, nn.Module):
            sd = sd.state_dict()
    if not any(k.startswith("module.") for k in sd):
        sd = {"module." + k: v for k, v in sd.items()}
    model.load_state_dict(sd)
    return model


class OnlineF2F:
    """The per-frame body of blind_denoising.py:187-256 with everything resident on the device."""

    def __init__(self, model, optimizer=None, align_corners=False):
        self.model = model
        self.net = model.module if isinstance(model, nn.DataParallel) else model
        if getattr(self.net, "channels", 1) != 1:
            raise NotImplementedError("the online loop is the paper script's: DnCNN(channels=1) on gray frames "
                                      "(blind_denoising.py:144)")
        self.dev = self.net.flat_params.device
        self.criterion = WarpedLoss(align_corners=align_corners).cuda(self.dev)
        self.optimizer = optimizer or Adam(model.parameters(), lr=LR, betas=BETAS, eps=EPS, weight_decay=WEIGHT_DECAY,
                                           amsgrad=False)
        self._scratch = None

    def _buffers(self, H, W):
        if self._scratch is None or self._scratch[0] != (H, W):
            mk = lambda: torch.empty((1, 1, H, W), dtype=torch.float32, device=self.dev)
            self._scratch = ((H, W), mk(), mk(), torch.zeros((), dtype=torch.float64, device=self.dev))
        return self._scratch[1:]

    def finetune(self, curr, prev, flow, iters=20, losses=None):
        """blind_denoising.py:211-222.  curr, prev: [1,1,H,W] in [0,1]; flow [1,2,H,W] (frame i -> i-1)."""
        net = self.net
        curr = curr.to(self.dev, torch.float32).contiguous()
        _, _, H, W = curr.shape
        self.model.train()
        warped, mask = self.criterion.precompute(prev, flow)          # once per frame
        out, dout, loss = self._buffers(H, W)
        st, group = self.optimizer.fused_state(net)
        ws = net._ws_train.get(1, H, W, True, self.dev)
        b1, b2 = group["betas"]
        with torch.cuda.device(self.dev):
            stream = torch.cuda.current_stream(self.dev).cuda_stream
            for _ in range(iters):
                check(lib().f2f_step_train(net.flat_params.data_ptr(), net.flat_grads.data_ptr(), st["exp_avg"].data_ptr(),
                                           st["exp_avg_sq"].data_ptr(), net._running.data_ptr(), net._nbt.data_ptr(),
                                           curr.data_ptr(), warped.data_ptr(), mask.data_ptr(), out.data_ptr(),
                                           dout.data_ptr(), loss.data_ptr(), ws.data_ptr(), ws.numel(), H, W,
                                           self.criterion.KIND, st["step"] + 1, group["lr"], b1, b2, group["eps"],
                                           group["weight_decay"], stream), "f2f_step_train")
                self.optimizer.advance()
                if losses is not None:
                    losses.append(float(loss.item()))
        net._fwd_id += 1
        net._attach_grads()

    def denoise(self, curr):
        """blind_denoising.py:226-230: eval-mode ``curr - model(curr)``."""
        self.model.eval()
        with torch.no_grad():
            x = curr.to(self.dev, torch.float32).contiguous()
            res = self.net.residual
            self.net.residual = True
            try:
                return self.model(x)
            finally:
                self.net.residual = res

    def step_frame(self, curr, prev, flow=None, iters=20, losses=None, flow_kw=None):
        """Fine-tune on (curr, prev) and return the denoised curr.  flow=None: TV-L1 backward flow
        (I0 = curr, I1 = prev, on 0..255 values like tvl1flow.sh:11-14) computed on the device."""
        if flow is None:
            flow = tvl1flow(curr[0, 0].to(self.dev) * 255.0, prev[0, 0].to(self.dev) * 255.0, **(flow_kw or {}))[None]
        self.finetune(curr, prev, flow, iters, losses)
        return self.denoise(curr)

    def flows_ahead(self, frames, flow_kw=None):
        """Backward flows of consecutive frames in ONE batched TV-L1 call: frames = [f_0 .. f_L] ([1,1,H,W] in
        [0,1], on the device) -> flow[L,2,H,W], flow[k] from f_{k+1} to f_k (I0 = f_{k+1}, I1 = f_k on 0..255
        values, tvl1flow.sh:11-14).  Frame pairs are independent (the reference runs one tvl1flow process per pair
        over the whole sequence before denoising starts), so they advance in lock-step as slots of one launch
        sequence instead of 2 300 latency-bound launches per pair."""
        seq = torch.cat([f.to(self.dev, torch.float32).reshape(1, f.shape[-2], f.shape[-1]) for f in frames]) * 255.0
        return tvl1flow(seq[1:], seq[:-1], **(flow_kw or {}))

    def run_sequence(self, frames, iters=20, lookahead=16, flows=None, flow_kw=None, out=None, losses=None):
        """The frame loop of blind_denoising.py:187-256 over frames[0..n-1] (host or device tensors [1,1,H,W]):
        yields (i, denoised frame i) for i = 1..n-1.  flows: optional sequence with flows[i] = flow from frame i to
        i-1 ([1,2,H,W]); when None they are computed on the device `lookahead` pairs at a time (flows_ahead).
        out: optional pinned host tensors to receive the denoised frames (asynchronous D2H on the current stream)."""
        n = len(frames)
        i = 1
        prev = frames[0].to(self.dev, non_blocking=True)
        while i < n:
            L = min(max(1, lookahead), n - i)
            window = [prev] + [frames[k].to(self.dev, non_blocking=True) for k in range(i, i + L)]
            wflows = None if flows is not None else self.flows_ahead(window, flow_kw)
            for k in range(L):
                curr = window[k + 1]
                flow = flows[i + k] if flows is not None else wflows[k][None]
                self.finetune(curr, window[k], flow, iters, losses)
                den = self.denoise(curr)
                if out is not None:
                    out[i + k].copy_(den, non_blocking=True)
                yield i + k, den
            prev = window[L]
            i += L


# ------------------------------------------------------------------ file handling
def _is_tiff(path):
    return path[-4:] == "tiff" or path[-3:] == "tif"


def imread_gray(path):
    """skimage.io.imread(path, as_gray=True) for 8-bit gray PNGs / tifffile.imread for TIFFs
    (blind_denoising.py:170-174): ndarray [H,W]; colour images use the .2125/.7154/.0721 luma
    skimage applies and come back in [0,1] like skimage does."""
    from PIL import Image
    im = Image.open(path)
    if _is_tiff(path):
        return np.array(im)
    a = np.array(im)
    if a.ndim == 3:
        a = a[..., :3].astype(np.float64) / 255.0
        return a @ np.array([0.2125, 0.7154, 0.0721])
    return a


def imsave(path, arr):
    from PIL import Image
    Image.fromarray(arr).save(path)


def _to_batch(im):
    """blind_denoising.py:177-180: [H,W] -> [1,1,H,W] / 255."""
    if len(im.shape) < 4:
        im = np.expand_dims(np.expand_dims(im, 0), 0)
        im = im / 255.
    return im


def blind_denoising(**args):
    """Main function (blind_denoising.py:125-259).  args: the CLI parameters below."""
    dev = torch.device("cuda", int(args.get("gpu", 0) or 0))
    # everything below (pinned-buffer copies, the `ready` events, the prefetch workers' pin_memory) acts on the CURRENT
    # device and its current stream: make that this run's GPU, not cuda:0 (--gpu N, torchrun ranks with LOCAL_RANK >= 1)
    torch.cuda.set_device(dev)
    if args.get("precise"):
        from .dncnn import set_precise
        set_precise(int(args["precise"]))
    model_fn = args["network"]
    if not os.path.isabs(model_fn):
        model_fn = os.path.join(os.path.abspath(os.path.dirname(__file__)), model_fn)
    print("Loading model a pre-trained DnCNN\n")
    model = load_network(model_fn, dev)
    engine = OnlineF2F(model, align_corners=bool(args.get("align_corners", False)))

    first, last = args["first"], args["last"]

    def read_frame(k):                   # frame k of the run = file index first + k (blind_denoising.py:170-180,192-203)
        return torch.FloatTensor(_to_batch(imread_gray(args["input"] % (first + k)))).pin_memory()

    def read_flow(k):                    # --flow files: flow from frame i to frame i-1 (blind_denoising.py:206-209)
        return torch.Tensor(np.expand_dims(read_flo(args["flow"] % (first + k)), 0)).permute(0, 3, 1, 2)

    plot_psnr = open(args["output_psnr"], "w")

    def finish_frame(i, den_host, ready):
        """blind_denoising.py:233-253 for frame i, on the writer thread: save, score, log -- in frame order."""
        ready.synchronize()
        print(i)
        den_np = den_host.numpy()
        if _is_tiff(args["ref"]):
            imsave(args["output"] % (i), np.squeeze(255. * den_np).astype(np.float32))
        else:
            imsave(args["output"] % (i), np.squeeze(255. * np.clip(den_np, 0., 1.)).astype(np.uint8))
        ref = _to_batch(imread_gray(args["ref"] % (i)))
        quant = psnr(ref, den_np)
        plot_psnr.write(str(quant) + "\n")
        print(quant)

    lookahead = int(args.get("lookahead", 16) or 1)
    n = last - first + 1
    # without --flow files the backward flows come from the device TV-L1 with the parameters the reference pipeline
    # produces its .flo files with (tvl1flow.sh:11-18: lambda 0.2, nscales 100 -> clamp, fscale 2), unless overridden
    flow_kw = dict(SCRIPT_PARAMS)
    flow_kw.update({k[len("flow_"):]: v for k, v in args.items() if k.startswith("flow_") and v is not None})
    # decode ahead of the loop and encode/score behind it on host threads (frame_io.py); the loop itself never
    # waits on a PNG.  Each frame gets its own pinned result buffer, alive until the writer is done with it.
    with Prefetcher(read_frame, n, depth=lookahead + 2, workers=4) as frames, \
            Prefetcher(read_flow, n, depth=2, workers=1) as flow_files, OrderedSink(depth=8) as sink:
        flows = flow_files if args.get("flow") else None
        for k, den in engine.run_sequence(frames, iters=args["iter"], lookahead=lookahead, flows=flows, flow_kw=flow_kw):
            host = torch.empty(den.shape, dtype=den.dtype, pin_memory=True)
            with torch.cuda.device(dev):
                host.copy_(den, non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(torch.cuda.current_stream(dev))       # the stream the copy was enqueued on
            sink.submit(finish_frame, first + k, host, ready)
    torch.save([model, engine.optimizer], args["output_network"])
    plot_psnr.close()


# ------------------------------------------------------------------ frame chunks over GPUs
def chunk_bounds(first, last, rank, world):
    """The online loop does not shard inside a stream (weights, Adam and BatchNorm state carry from frame to frame),
    so several GPUs take independent --first/--last chunks, each restarting from the pretrained weights -- the
    reference's own way of splitting a sequence (blind_denoising.py:279-282; SURVEY 8e, configs[4]).  Frames
    first+1 .. last are the denoised ones (blind_denoising.py:187); rank `rank` of `world` takes a contiguous block
    of them and starts ONE frame earlier: that frame is only its first `prev`.  -> (first_r, last_r), or None when
    there are fewer frames than ranks and this rank has none."""
    n = max(last - first, 0)
    base, extra = divmod(n, world)
    count = base + (1 if rank < extra else 0)
    if count == 0:
        return None
    lo = first + 1 + rank * base + min(rank, extra)
    return lo - 1, lo + count - 1


def _chunk_name(path, rank):
    root, ext = os.path.splitext(path)
    return "%s.chunk%03d%s" % (root, rank, ext)


def run_chunked(args, rank, world, run=None, barrier=None, local_rank=None):
    """One rank's share of blind_denoising(**args) under `world` processes (torchrun): its chunk of frames on its own
    GPU, output frames under the usual names, its network under `output_network` + '.chunkRRR', and its PSNR lines in
    a part file that rank 0 concatenates in frame order into `output_psnr` once every rank is done."""
    run = blind_denoising if run is None else run
    part = _chunk_name(args["output_psnr"], rank)
    b = chunk_bounds(args["first"], args["last"], rank, world)
    if b is None:
        open(part, "w").close()
    else:
        run(**dict(args, first=b[0], last=b[1], output_psnr=part, output_network=_chunk_name(args["output_network"], rank),
                   gpu=rank if local_rank is None else local_rank))
    if barrier is not None:
        barrier()
    if rank == 0:
        with open(args["output_psnr"], "w") as out:
            for r in range(world):
                fn = _chunk_name(args["output_psnr"], r)
                with open(fn) as f:
                    out.write(f.read())
                os.remove(fn)


def main(argv=None):
    parser = argparse.ArgumentParser(description="Blind_denoising")
    parser.add_argument("--input", type=str, default="", help="path to input frames (C type)")
    parser.add_argument("--ref", type=str, default="", help="path to reference frames (C type), against which the psnr is going to be computed")
    parser.add_argument("--flow", type=str, default="", help="path to optical flow (C type); empty = TV-L1 on the device")
    parser.add_argument("--output", type=str, default="./%03d.png", help="path to output image (C type)")
    parser.add_argument("--output_psnr", type=str, default="plot_psnr.txt", help="path to output psnr")
    parser.add_argument("--output_network", type=str, default="final.pth", help="path to output network")
    parser.add_argument("--first", type=int, default=1, help="index first frame")
    parser.add_argument("--last", type=int, default=300, help="index last frame")
    parser.add_argument("--iter", type=int, default=20, help="number of time the learning is done on a given frame")
    parser.add_argument("--network", type=str, default="dncnn/logs/DnCNN-S-25/net.pth", help="path to the network")
    parser.add_argument("--gpu", type=int, default=0, help="(added) CUDA device index")
    parser.add_argument("--align_corners", action="store_true", help="(added) sample with align_corners=True")
    parser.add_argument("--precise", type=int, nargs="?", const=2, default=0, choices=[0, 1, 2],
                        help="(added) FP32-class arithmetic throughout the network, the reference's precision: 2 = split FP16 "
                             "operand pairs on the tensor cores (~5x the default mode's time), 1 = FP32 on the CUDA cores (~40x); "
                             "default 0 = 16-bit tensor-core operands")
    parser.add_argument("--lookahead", type=int, default=16, help="(added) frame pairs per batched TV-L1 call when --flow is empty")
    # (added) TV-L1 parameters of the on-device flow used when --flow is empty; defaults = tvl1flow/tvl1flow.sh:11-18
    for name, typ in (("tau", float), ("lam", float), ("theta", float), ("nscales", int), ("fscale", int), ("zfactor", float),
                      ("nwarps", int), ("epsilon", float)):
        parser.add_argument("--flow_" + name, type=typ, default=None,
                            help="(added) TV-L1 %s of the on-device flow (default %s, tvl1flow.sh)" % (name, SCRIPT_PARAMS[name]))
    argspar = parser.parse_args(argv)
    print("\n### Model-blind Video Denoising Via Frame-to-frame Training ###")
    print("> Parameters:")
    for p, v in zip(argspar.__dict__.keys(), argspar.__dict__.values()):
        print("\t{}: {}".format(p, v))
    print("\n")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        # launched with torchrun: one process per GPU, every rank an independent chunk of the sequence; the only
        # exchange is the barrier before rank 0 assembles the PSNR file (gloo: nothing on the data path)
        import torch.distributed as dist
        dist.init_process_group("gloo")
        try:
            run_chunked(vars(argspar), int(os.environ["RANK"]), world, barrier=dist.barrier,
                        local_rank=int(os.environ.get("LOCAL_RANK", "0")))
        finally:
            dist.destroy_process_group()
        return
    blind_denoising(**vars(argspar))


if __name__ == "__main__":
    main()
