"""Batch-mode frame-to-frame fine-tuning over several sequences (BASELINE.json configs[3]).

The reference fine-tunes ONE sequence online (blind_denoising.py:187-256).  Batch mode advances S
independent sequences frame by frame together and takes one Adam step per inner iteration on the
SUM of the per-sequence losses -- the loss is already a sum over pixels (nn.L1Loss(size_average=
False), blind_denoising.py:47,121), so the batch gradient is the sum of the per-sequence
gradients.  B = 1 semantics are kept per sequence (the occlusion mask uses the sequence's own flow,
BatchNorm statistics are per frame, blind_denoising.py:92-103,212), i.e. it is exactly the
single-process loop "for each sequence: forward, loss, backward (accumulating); then step".

Multi-GPU (one process per GPU, torch.distributed): sequences are split into contiguous blocks over
the ranks; per inner iteration there is ONE exchange step,
  * the flat 556 032-float gradient buffer is summed over ranks with a single all-reduce (NCCL over
    NVLink), or -- ``deterministic=True`` -- all-gathered and summed in sequence order, which makes
    the result bit-identical for every world size;
  * the BatchNorm batch statistics of every sequence (15 layers x (mean, unbiased var) x 64) are
    all-gathered and the running-stat EMA (momentum 0.1, net.py:17 defaults) is applied in sequence
    order on every rank, so running statistics equal the single-process loop's for any world size;
then every rank takes the identical Adam step (blind_denoising.py:163-165).

The collective/host logic below works on CPU tensors too (gloo), which is how tests/ covers the
N > 1 path without GPUs; the compute always goes through libf2f_b200.so (no CPU fallback).
"""
import ctypes as C

import torch
import torch.distributed as dist

BN_MOMENTUM = 0.1     # torch.nn.BatchNorm2d default, net.py:17
NBN, NCH = 15, 64


def shard_sequences(nseq, rank, world):
    """Contiguous block of sequence indices owned by `rank` (blocks differ by at most one)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    base, rem = divmod(nseq, world)
    lo = rank * base + min(rank, rem)
    return list(range(lo, lo + base + (1 if rank < rem else 0)))


def _world(group):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def _staged(t, group):
    """gloo moves host memory: stage CUDA tensors through the host for it (NCCL takes them as they are)."""
    return t.is_cuda and dist.get_backend(group) == "gloo"


def exchange_gradients(local_grads, nseq, group=None, deterministic=False):
    """Sum of the per-sequence gradients over all ranks.

    local_grads: [n_local, P] (deterministic) -- one row per local sequence, in sequence order -- or
    a single [P] tensor already summed over the local sequences.  Returns a [P] tensor, identical on
    every rank.  deterministic=True reproduces the single-process left-to-right sum bit for bit."""
    rank, world = _world(group)
    if not deterministic:
        g = local_grads if local_grads.dim() == 1 else local_grads.sum(0)
        if world > 1:
            if _staged(g, group):
                h = g.cpu()
                dist.all_reduce(h, op=dist.ReduceOp.SUM, group=group)
                g.copy_(h)
            else:
                dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
        return g
    if local_grads.dim() != 2:
        raise ValueError("deterministic exchange needs one gradient row per local sequence")
    rows = gather_rows(local_grads, nseq, group)
    total = rows[0].clone()
    for s in range(1, nseq):
        total += rows[s]
    return total


def gather_rows(local_rows, nseq, group=None):
    """All-gather of per-sequence rows ([n_local, ...] on each rank, contiguous block sharding) into
    [nseq, ...] in sequence order on every rank."""
    rank, world = _world(group)
    if world == 1:
        if local_rows.shape[0] != nseq:
            raise ValueError("expected %d rows, got %d" % (nseq, local_rows.shape[0]))
        return local_rows
    nmax = (nseq + world - 1) // world
    pad = torch.zeros((nmax,) + tuple(local_rows.shape[1:]), dtype=local_rows.dtype, device=local_rows.device)
    pad[:local_rows.shape[0]] = local_rows
    if _staged(pad, group):
        hpad = pad.cpu()
        hout = [torch.empty_like(hpad) for _ in range(world)]
        dist.all_gather(hout, hpad, group=group)
        out = [h.to(pad.device) for h in hout]
    else:
        out = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(out, pad, group=group)
    rows = []
    for r in range(world):
        rows.append(out[r][:len(shard_sequences(nseq, r, world))])
    return torch.cat(rows, 0)


def batch_stats_from_sums(sums, n):
    """Epilogue sums of one train-mode forward ([15, 2, 64] float64: sum z, sum z^2 over n = H*W
    values) -> [15, 2, 64] float32 (batch mean, UNBIASED batch variance), SURVEY A.4."""
    mean = sums[:, 0] / n
    var = (sums[:, 1] / n - mean * mean).clamp_min(0.0)
    unbiased = var * (n / (n - 1.0)) if n > 1 else var
    return torch.stack([mean, unbiased], 1).to(torch.float32)


def apply_bn_ema(running, nbt, stats_all, momentum=BN_MOMENTUM):
    """running [15, 2, 64] float32, nbt [15] int64, stats_all [nseq, 15, 2, 64] float32 in sequence
    order: the train-mode running-stat update of nn.BatchNorm2d, once per sequence, in order."""
    m = torch.tensor(momentum, dtype=torch.float32, device=running.device)
    keep = torch.tensor(1.0, dtype=torch.float32, device=running.device) - m
    for s in range(stats_all.shape[0]):
        running.mul_(keep).add_(stats_all[s] * m)
    nbt += stats_all.shape[0]
    return running


class BatchF2F:
    """S sequences, one shared DnCNN; see the module docstring.  `model` is a frame2frame_b200 DnCNN
    (optionally inside DataParallel) on this rank's GPU; every rank must hold identical weights."""

    def __init__(self, model, nseq, optimizer=None, group=None, deterministic=False, align_corners=False):
        import torch.nn as nn
        from .blind_denoising import BETAS, EPS, LR, WEIGHT_DECAY
        from .dncnn import Adam
        from .warped_loss import WarpedLoss
        self.model = model
        self.net = model.module if isinstance(model, nn.DataParallel) else model
        self.dev = self.net.flat_params.device
        self.group = group
        self.rank, self.world = _world(group)
        self.nseq = nseq
        self.local = shard_sequences(nseq, self.rank, self.world)
        self.deterministic = deterministic
        self.criterion = WarpedLoss(align_corners=align_corners).cuda(self.dev)
        self.optimizer = optimizer or Adam(model.parameters(), lr=LR, betas=BETAS, eps=EPS, weight_decay=WEIGHT_DECAY,
                                           amsgrad=False)
        self._shape = None

    def _buffers(self, H, W):
        from ._lib import lib
        if self._shape != (H, W, lib().f2f_dncnn_get_precise()):
            n = len(self.local)
            mk = lambda *s, dt=torch.float32: torch.empty(s, dtype=dt, device=self.dev)
            self._out, self._dout = mk(1, 1, H, W), mk(1, 1, H, W)
            self._loss = torch.zeros(max(n, 1), dtype=torch.float64, device=self.dev)
            self._grows = mk(max(n, 1), self.net.flat_params.numel()) if self.deterministic else None
            self._scr_running = torch.empty_like(self.net._running)
            self._scr_nbt = torch.empty_like(self.net._nbt)
            lay = (C.c_size_t * 8)()
            lib().f2f_dncnn_workspace_layout(1, H, W, 1, lay)
            self._stats_off = int(lay[7])
            self._shape = (H, W, lib().f2f_dncnn_get_precise())

    def finetune(self, currs, prevs, flows, iters=20, losses=None):
        """One frame of every LOCAL sequence: currs/prevs [1,1,H,W] lists, flows [1,2,H,W] list (frame
        i -> i-1), in the order of self.local.  `losses` receives the global loss per iteration."""
        from ._lib import check, lib
        net, L = self.net, lib()
        if not (len(currs) == len(prevs) == len(flows) == len(self.local)):
            raise ValueError("expected %d local sequences" % len(self.local))
        currs = [c.to(self.dev, torch.float32).contiguous() for c in currs]
        H, W = (currs[0].shape[-2:] if currs else self._shape[:2])
        self._buffers(H, W)
        self.model.train()
        pre = [self.criterion.precompute(p, f) for p, f in zip(prevs, flows)]          # once per frame
        st, group = self.optimizer.fused_state(net)
        ws = net._ws_train.get(1, H, W, True, self.dev)
        sums = ws[self._stats_off:self._stats_off + NBN * 128 * 8].view(torch.float64).view(NBN, 2, NCH)
        b1, b2 = group["betas"]
        stream = torch.cuda.current_stream(self.dev).cuda_stream
        P = net.flat_params.numel()
        n = len(self.local)
        stats_local = torch.empty((n, NBN, 2, NCH), dtype=torch.float32, device=self.dev)
        with torch.cuda.device(self.dev):
            for _ in range(iters):
                for j in range(n):
                    warped, mask = pre[j]
                    # the kernels' in-place running-stat update goes to a scratch copy; the real update is
                    # the ordered EMA below, identical for every world size
                    self._scr_running.copy_(net._running)
                    self._scr_nbt.copy_(net._nbt)
                    check(L.f2f_dncnn_forward(net.flat_params.data_ptr(), self._scr_running.data_ptr(),
                                              self._scr_nbt.data_ptr(), currs[j].data_ptr(), self._out.data_ptr(),
                                              ws.data_ptr(), ws.numel(), 1, H, W, 1, 1, stream), "f2f_dncnn_forward")
                    stats_local[j] = batch_stats_from_sums(sums, float(H * W))
                    check(L.f2f_masked_loss(self._out.data_ptr(), warped.data_ptr(), mask.data_ptr(), self._dout.data_ptr(),
                                            self._loss[j:].data_ptr(), H * W, self.criterion.KIND, stream), "f2f_masked_loss")
                    gdst = self._grows[j] if self.deterministic else net.flat_grads
                    check(L.f2f_dncnn_backward(net.flat_params.data_ptr(), gdst.data_ptr(), self._dout.data_ptr(),
                                               ws.data_ptr(), ws.numel(), 1, H, W, 1,
                                               0 if (self.deterministic or j == 0) else 1, stream), "f2f_dncnn_backward")
                # ---- the exchange step ----
                if self.deterministic:
                    total = exchange_gradients(self._grows[:n], self.nseq, self.group, True)
                    net.flat_grads.copy_(total)
                else:
                    if n == 0:
                        net.flat_grads.zero_()
                    exchange_gradients(net.flat_grads, self.nseq, self.group, False)
                stats_all = gather_rows(stats_local, self.nseq, self.group)
                apply_bn_ema(net._running, net._nbt, stats_all)
                check(L.f2f_adam_step(net.flat_params.data_ptr(), net.flat_grads.data_ptr(), st["exp_avg"].data_ptr(),
                                      st["exp_avg_sq"].data_ptr(), P, st["step"] + 1, group["lr"], b1, b2, group["eps"],
                                      group["weight_decay"], stream), "f2f_adam_step")
                self.optimizer.advance()
                if losses is not None:
                    per_seq = gather_rows(self._loss[:n].reshape(n, 1), self.nseq, self.group)   # in sequence order
                    tot = per_seq[0, 0].clone()
                    for q in range(1, self.nseq):
                        tot += per_seq[q, 0]
                    losses.append(float(tot.item()))
        net._fwd_id += 1
        net._attach_grads()

    def denoise(self, curr):
        """Eval-mode ``curr - model(curr)`` (blind_denoising.py:226-230) with the shared weights."""
        self.model.eval()
        with torch.no_grad():
            x = curr.to(self.dev, torch.float32).contiguous()
            res = self.net.residual
            self.net.residual = True
            try:
                return self.model(x)
            finally:
                self.net.residual = res
