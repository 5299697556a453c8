"""DnCNN on B200: host-side mirror of the reference network and optimiser.

Reference interface this replaces
  * ``DnCNN(channels, num_of_layers=17)`` of lib/frame2frame/dncnn/net.py:6-39 -- same
    constructor, same ``.dncnn`` ``nn.Sequential`` (so ``state_dict()`` has the same 92 keys
    ``dncnn.{0,2,5,...}.weight``, ``dncnn.{3,6,...}.{weight,bias,running_mean,running_var,
    num_batches_tracked}`` and DnCNN-S-25 ``.pth`` files load, with the ``module.`` prefix when
    wrapped in ``nn.DataParallel`` as blind_denoising.py:144-147 does), ``.train()`` /
    ``.eval()`` BatchNorm semantics, 4-D and 5-D inputs (net.py:23-37);
  * ``torch.optim.Adam`` as configured at blind_denoising.py:163-165.

What runs underneath is libf2f_b200.so: all parameters live in ONE flat FP32 buffer (the modules'
parameters are views into it), the forward and backward passes are one C-ABI call each
(f2f_dncnn_forward / f2f_dncnn_backward: tcgen05 implicit-GEMM convolutions with BatchNorm
statistics and ReLU fused into their prologues/epilogues), gradients land in one flat buffer
(ready for a single NCCL all-reduce), and Adam is one kernel over the flat buffers.
There is no CPU fallback: tensors must be on a CUDA device.
"""
import torch
import torch.nn as nn

from ._lib import check, lib

FEATURES = 64
NBN = 15
NPARAMS = 556032


def _stream():
    return torch.cuda.current_stream().cuda_stream


def set_precise(mode):
    """Precision mode of every DnCNN in the process (f2f_dncnn_set_precise): 0 / False = the default 16-bit tensor-core
    operands; 1 / True = FP32 on the CUDA cores, the arithmetic of the reference's FP32 torch loop up to summation order
    (~40x slower); 2 = FP32-class split operands (hi + lo FP16 pairs, three products) on the tensor cores (~5x slower).
    Workspaces are re-planned on the next call."""
    lib().f2f_dncnn_set_precise(int(mode))


def get_precise():
    return int(lib().f2f_dncnn_get_precise())


class _Workspace:
    """Device workspace of f2f_dncnn_forward/backward, cached per (B, H, W, train, device)."""

    def __init__(self):
        self.key = None
        self.buf = None

    def get(self, B, H, W, train, device, channels=1):
        key = (B, H, W, bool(train), str(device), lib().f2f_dncnn_get_precise(), channels)
        if self.key != key:
            n = lib().f2f_dncnn_workspace_bytes_c(B, H, W, int(train), channels)
            if n == 0:
                raise ValueError("bad DnCNN geometry B=%d H=%d W=%d" % (B, H, W))
            self.buf = None
            self.buf = torch.empty(n, dtype=torch.uint8, device=device)
            self.key = key
        return self.buf


class _DnCNNFunction(torch.autograd.Function):
    """y = net(x); parameter gradients are written straight into the model's flat gradient buffer
    (and exposed as the ``.grad`` views of the modules' parameters)."""

    @staticmethod
    def forward(ctx, x, flat, model):
        B, C, H, W = x.shape
        train = model.training
        ws = model._ws_train if train else model._ws_eval
        buf = ws.get(B, H, W, train, x.device, C)
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            check(lib().f2f_dncnn_forward_c(flat.data_ptr(), model._running.data_ptr(), model._nbt.data_ptr(),
                                            x.data_ptr(), y.data_ptr(), buf.data_ptr(), buf.numel(), B, H, W,
                                            int(train), int(model.residual), C, _stream()), "f2f_dncnn_forward")
        ctx.model, ctx.geom, ctx.train = model, (B, H, W), train
        if train:
            model._fwd_id += 1
            ctx.fwd_id = model._fwd_id
        return y

    @staticmethod
    def backward(ctx, dy):
        model = ctx.model
        if not ctx.train:
            raise RuntimeError("DnCNN backward needs a forward in train() mode (activations are not kept in eval())")
        if ctx.fwd_id != model._fwd_id:
            raise RuntimeError("DnCNN backward: the workspace was overwritten by a later forward pass")
        B, H, W = ctx.geom
        buf = model._ws_train.get(B, H, W, True, dy.device, model.channels)
        accumulate = model._grads_attached()
        dy = dy.contiguous().float()
        with torch.cuda.device(dy.device):
            check(lib().f2f_dncnn_backward_c(model._flat.data_ptr(), model._flat_grad.data_ptr(), dy.data_ptr(),
                                             buf.data_ptr(), buf.numel(), B, H, W, int(model.residual), int(accumulate),
                                             model.channels, _stream()), "f2f_dncnn_backward")
        model._attach_grads()
        return None, None, None


class DnCNN(nn.Module):
    """lib/frame2frame/dncnn/net.py:6-39.  ``residual=True`` returns ``x - dncnn(x)`` like that file
    (net.py:31); ``residual=False`` returns the noise estimate like the ``dncnn.models.DnCNN`` the
    paper script imports (blind_denoising.py:25,218 subtracts it itself)."""

    def __init__(self, channels=1, num_of_layers=17, residual=True):
        super().__init__()
        if channels not in (1, 3) or num_of_layers != 17:
            raise NotImplementedError("the B200 kernels are built for the 17-layer DnCNN with 1 (gray) or 3 (RGB) image "
                                      "channels, got channels=%d, num_of_layers=%d" % (channels, num_of_layers))
        self.channels = channels
        kernel_size, padding, features = 3, 1, FEATURES
        layers = [nn.Conv2d(channels, features, kernel_size, padding=padding, bias=False), nn.ReLU(inplace=True)]
        for _ in range(num_of_layers - 2):
            layers.append(nn.Conv2d(features, features, kernel_size, padding=padding, bias=False))
            layers.append(nn.BatchNorm2d(features))
            layers.append(nn.ReLU(inplace=True))
        layers.append(nn.Conv2d(features, channels, kernel_size, padding=padding, bias=False))
        self.dncnn = nn.Sequential(*layers)
        self.residual = residual
        self._ws_train, self._ws_eval = _Workspace(), _Workspace()
        self._fwd_id = 0
        self._flatten()

    # ---- flat storage -------------------------------------------------------------------
    def _ordered(self):
        """(parameters, bn modules) in flat-buffer order = state_dict order."""
        params, bns = [], []
        for m in self.dncnn:
            if isinstance(m, nn.Conv2d):
                params.append(m.weight)
            elif isinstance(m, nn.BatchNorm2d):
                params += [m.weight, m.bias]
                bns.append(m)
        return params, bns

    def _flatten(self):
        params, bns = self._ordered()
        dev = params[0].device
        with torch.no_grad():
            flat = torch.cat([p.detach().reshape(-1).float() for p in params]).to(dev).contiguous()
            assert flat.numel() == NPARAMS + 1152 * (self.channels - 1)
            running = torch.stack([torch.stack([b.running_mean.float(), b.running_var.float()]) for b in bns]).to(dev).contiguous()
            nbt = torch.stack([b.num_batches_tracked.reshape(()).long() for b in bns]).to(dev).contiguous()
        flat.requires_grad_(True)
        self.__dict__["_flat"] = flat
        self.__dict__["_flat_grad"] = torch.zeros_like(flat, requires_grad=False)
        self.__dict__["_running"] = running
        self.__dict__["_nbt"] = nbt
        self.__dict__["_grad_views"] = []
        off = 0
        for p in params:
            n = p.numel()
            p.data = flat.detach()[off:off + n].view(p.shape)
            self._grad_views.append(self._flat_grad[off:off + n].view(p.shape))
            p.grad = None
            off += n
        for k, b in enumerate(bns):
            b.running_mean = running[k, 0]
            b.running_var = running[k, 1]
            b.num_batches_tracked = nbt[k]

    def _apply(self, fn, *args, **kwargs):
        super()._apply(fn, *args, **kwargs)
        self._flatten()
        return self

    def load_state_dict(self, state_dict, *args, **kwargs):
        r = super().load_state_dict(state_dict, *args, **kwargs)
        self._flatten()      # assign=True or dtype changes would detach the views
        return r

    def __getstate__(self):
        # torch.save([model, optimizer]) (blind_denoising.py:258): keep the modules, drop device scratch
        d = self.__dict__.copy()
        for k in ("_ws_train", "_ws_eval", "_flat", "_flat_grad", "_grad_views", "_running", "_nbt"):
            d.pop(k, None)
        return d

    def __setstate__(self, d):
        self.__dict__.update(d)
        self.__dict__["_ws_train"], self.__dict__["_ws_eval"] = _Workspace(), _Workspace()
        self._fwd_id = 0
        self._flatten()

    def _grads_attached(self):
        """True when every parameter's .grad is still the view into the flat gradient buffer
        (i.e. no zero_grad(set_to_none=True) since the last backward): accumulate like autograd does."""
        params, _ = self._ordered()
        return all(p.grad is not None and p.grad.data_ptr() == g.data_ptr() for p, g in zip(params, self._grad_views))

    def _attach_grads(self):
        params, _ = self._ordered()
        for p, g in zip(params, self._grad_views):
            p.grad = g

    @property
    def flat_params(self):
        return self._flat

    @property
    def flat_grads(self):
        return self._flat_grad

    # ---- forward ------------------------------------------------------------------------
    def forward(self, x):
        if not x.is_cuda or not self._flat.is_cuda:
            raise RuntimeError("DnCNN (B200) needs CUDA tensors and a model moved to a CUDA device: no CPU fallback")
        shape5 = None
        if x.ndim == 5:                         # net.py:25-29: video batch -> (b t) c h w
            shape5 = x.shape
            x = x.reshape(-1, *x.shape[2:])
        x = x.contiguous().float()
        if x.shape[1] != self.channels:
            raise ValueError("expected a %d-channel input, got %s" % (self.channels, tuple(x.shape)))
        y = _DnCNNFunction.apply(x, self._flat, self)
        if shape5 is not None:
            y = y.reshape(shape5)
        return y


class Adam(torch.optim.Optimizer):
    """torch.optim.Adam(lr, betas, eps, weight_decay, amsgrad=False) semantics
    (blind_denoising.py:163-165) with the update done by f2f_adam_step.  When the parameters of
    a group are views into one flat buffer (a ``DnCNN`` above) the whole group is ONE launch."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if amsgrad:
            raise NotImplementedError("amsgrad=True is not provided (the reference uses amsgrad=False)")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False))
        self._flat_state = {}

    @staticmethod
    def _contiguous_run(tensors):
        """(base tensor index order) if `tensors` tile one contiguous float32 region in order, else None."""
        ptr = tensors[0].data_ptr()
        for t in tensors:
            if t.dtype != torch.float32 or not t.is_contiguous() or t.data_ptr() != ptr:
                return False
            ptr += t.numel() * 4
        return True

    def __setstate__(self, state):
        # torch.optim.Optimizer.__getstate__ keeps defaults / state / param_groups only (torch.save([model, optimizer]),
        # blind_denoising.py:258): the flat moment buffers are rebuilt from the per-parameter state on first use
        super().__setstate__(state)
        self.__dict__["_flat_state"] = {}

    def _flat_group_state(self, gi, ps):
        st = self._flat_state.get(gi)
        if st is None:
            n = sum(p.numel() for p in ps)
            dev = ps[0].device
            st = self._flat_state[gi] = {
                "step": 0,
                "exp_avg": torch.zeros(n, dtype=torch.float32, device=dev),
                "exp_avg_sq": torch.zeros(n, dtype=torch.float32, device=dev)}
            off = 0
            for p in ps:
                k = p.numel()
                old = self.state.get(p) or {}
                views = {"exp_avg": st["exp_avg"][off:off + k].view(p.shape),
                         "exp_avg_sq": st["exp_avg_sq"][off:off + k].view(p.shape)}
                step = torch.tensor(0.0)
                if "exp_avg" in old:                  # resumed (unpickled / load_state_dict): keep the moments and the step
                    views["exp_avg"].copy_(old["exp_avg"].to(dev))
                    views["exp_avg_sq"].copy_(old["exp_avg_sq"].to(dev))
                    step = torch.as_tensor(old.get("step", 0.0), dtype=torch.float32).cpu().clone()
                    st["step"] = int(step.item())
                self.state[p] = {"step": step, **views}
                off += k
        return st

    def fused_state(self, model):
        """For the one-call training step (f2f_step_train): the flat Adam state of `model`'s
        parameters; the caller passes st["step"] + 1 to the kernel and then calls advance()."""
        group = self.param_groups[0]
        ps = list(group["params"])
        if len(self.param_groups) != 1 or not self._contiguous_run([p.data for p in ps]) \
                or ps[0].data_ptr() != model.flat_params.data_ptr():
            raise RuntimeError("fused step needs an optimizer built from model.parameters() of one DnCNN")
        return self._flat_group_state(0, ps), group

    def advance(self):
        st = self._flat_state[0]
        st["step"] += 1
        for p in self.param_groups[0]["params"]:
            self.state[p]["step"] += 1

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            b1, b2 = group["betas"]
            flat = len(ps) == len(group["params"]) and self._contiguous_run([p.data for p in ps]) \
                and self._contiguous_run([p.grad for p in ps])
            if flat:
                n = sum(p.numel() for p in ps)
                st = self._flat_group_state(gi, ps)
                st["step"] += 1
                for p in ps:
                    self.state[p]["step"] += 1
                with torch.cuda.device(ps[0].device):
                    check(lib().f2f_adam_step(ps[0].data_ptr(), ps[0].grad.data_ptr(), st["exp_avg"].data_ptr(),
                                              st["exp_avg_sq"].data_ptr(), n, st["step"], group["lr"], b1, b2,
                                              group["eps"], group["weight_decay"], _stream()), "f2f_adam_step")
            else:
                for p in ps:
                    if not p.is_cuda:
                        raise RuntimeError("Adam (B200) needs CUDA parameters: no CPU fallback")
                    st = self.state[p]
                    if not st:
                        st["step"] = torch.tensor(0.0)
                        st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                        st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["step"] += 1
                    g = p.grad.contiguous()
                    with torch.cuda.device(p.device):
                        check(lib().f2f_adam_step(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(),
                                                  st["exp_avg_sq"].data_ptr(), p.numel(), int(st["step"].item()),
                                                  group["lr"], b1, b2, group["eps"], group["weight_decay"], _stream()),
                              "f2f_adam_step")
        return loss
