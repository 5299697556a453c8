"""WarpedLoss on B200: host-side mirror of the reference's frame-to-frame loss.

Same classes, method names, argument meaning and results as
  * blind_denoising.py:44-122        -> ``WarpedLoss``      (paper script)
  * lib/frame2frame/warped_loss.py:117-248 -> ``LibWarpedLoss`` (package flavour,
    loss_type="warp": ``dist_crit`` "l2" / "l1", ``in_mask``)
but every tensor op runs in libf2f_b200.so (f2f_warp_mask, f2f_masked_loss):
the warp, validity mask, occlusion mask, dilation and frame are ONE kernel, and
since target, flow and masks do not change over the inner iterations of a frame
 ``precompute()`` lets the caller do that once per frame.
There is no CPU fallback: CPU tensors are moved to the module's device, as the
reference's ``.cuda()`` calls do (blind_denoising.py:64-65).
"""
import torch
import torch.nn as nn

from ._lib import check, lib


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _prep(t, device):
    return t.detach().to(device=device, dtype=torch.float32).contiguous()


def warp_mask(x, flo, in_mask=None, align_corners=False, occ_f64=True, want_valid=True, want_mask=True):
    """-> (warped, valid, mask) CUDA tensors shaped like x."""
    if not x.is_cuda:
        raise RuntimeError("warp_mask needs CUDA tensors (no CPU fallback)")
    B, C, H, W = x.shape
    if tuple(flo.shape) != (B, 2, H, W):
        raise ValueError("flow must be [B,2,H,W] = %s, got %s" % ((B, 2, H, W), tuple(flo.shape)))
    warped = torch.empty_like(x)
    valid = torch.empty_like(x) if want_valid else None
    mask = torch.empty_like(x) if want_mask else None
    with torch.cuda.device(x.device):
        check(lib().f2f_warp_mask(x.data_ptr(), flo.data_ptr(), in_mask.data_ptr() if in_mask is not None else None,
                                  warped.data_ptr(), valid.data_ptr() if want_valid else None,
                                  mask.data_ptr() if want_mask else None, B, C, H, W, int(align_corners),
                                  int(occ_f64), _stream()), "f2f_warp_mask")
    return warped, valid, mask


class _MaskedLoss(torch.autograd.Function):
    """loss(input) with warped/mask constant; gradient produced in the same pass."""

    @staticmethod
    def forward(ctx, inp, warped, mask, kind):
        x = inp.detach().contiguous()
        grad = torch.empty_like(x)
        loss = torch.empty((), dtype=torch.float64, device=x.device)
        with torch.cuda.device(x.device):
            check(lib().f2f_masked_loss(x.data_ptr(), warped.data_ptr(), mask.data_ptr(), grad.data_ptr(),
                                        loss.data_ptr(), x.numel(), kind, _stream()), "f2f_masked_loss")
        ctx.save_for_backward(grad)
        return loss.to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g.to(grad.dtype), None, None, None


class WarpedLoss(nn.Module):
    """blind_denoising.py:44-122.  ``align_corners=False`` reproduces the reference as it
    runs under torch >= 1.3; ``True`` is the convention its grid scaling was written for."""

    KIND = 0          # nn.L1Loss(size_average=False): sum
    OCC_F64 = True    # numpy float64 occlusion test (blind_denoising.py:88-94)

    def __init__(self, align_corners=False):
        super().__init__()
        self.align_corners = align_corners
        self._dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else None
        self._cache = None

    def cuda(self, device=None):
        if isinstance(device, torch.device):
            self._dev = device if device.index is not None else torch.device("cuda", torch.cuda.current_device())
        else:
            self._dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        return self

    def _device(self, *ts):
        for t in ts:
            if t.is_cuda:
                return t.device
        if self._dev is None:
            raise RuntimeError("WarpedLoss needs a CUDA device (no CPU fallback)")
        return self._dev

    def warp(self, x, flo):
        """(output, mask): bilinear warp of x by flo and its validity mask (:49-78)."""
        dev = self._device(x, flo)
        warped, valid, _ = warp_mask(_prep(x, dev), _prep(flo, dev), align_corners=self.align_corners,
                                     occ_f64=self.OCC_F64, want_mask=False)
        return warped, valid

    def occlusion_mask(self, warped, of, old_mask):
        """:81-113"""
        dev = self._device(warped, of, old_mask)
        B, C, H, W = warped.shape
        of, old = _prep(of, dev), _prep(old_mask, dev)
        out = torch.empty_like(old)
        with torch.cuda.device(dev):
            check(lib().f2f_occlusion_mask(of.data_ptr(), old.data_ptr(), out.data_ptr(), B, C, H, W,
                                           int(self.OCC_F64), _stream()), "f2f_occlusion_mask")
        return out

    def precompute(self, target, flow, in_mask=None):
        """Once per frame: (warped target, final mask).  forward() caches this by tensor identity."""
        dev = self._device(target, flow)
        warped, _, mask = warp_mask(_prep(target, dev), _prep(flow, dev),
                                    None if in_mask is None else _prep(in_mask, dev),
                                    align_corners=self.align_corners, occ_f64=self.OCC_F64, want_valid=False)
        return warped, mask

    def _warped_mask(self, target, flow, in_mask=None):
        key = (target.data_ptr(), flow.data_ptr(), target._version, flow._version,
               None if in_mask is None else (in_mask.data_ptr(), in_mask._version))
        if self._cache is None or self._cache[0] != key:
            self._cache = (key, self.precompute(target, flow, in_mask), (target, flow, in_mask))
        return self._cache[1]

    def forward(self, input, target, flow):
        """:115-122 -- scalar, differentiable wrt input only."""
        warped, mask = self._warped_mask(target, flow)
        self.loss = _MaskedLoss.apply(input.to(warped.device), warped, mask, self.KIND)
        return self.loss


class LibWarpedLoss(WarpedLoss):
    """lib/frame2frame/warped_loss.py:117-248 with loss_type="warp" (the stnls patch-search
    branch needs an absent third-party CUDA package and is out of scope)."""

    OCC_F64 = False   # float32 torch ops, warped_loss.py:184-190

    def __init__(self, dist_crit="l2", use_stnls=False, search=None, loss_type="warp", ws=9, ps=7, dist_mask=2e-1,
                 ps_scale=1., ps_final=1, align_corners=False):
        super().__init__(align_corners)
        if loss_type != "warp" or use_stnls:
            raise NotImplementedError("only loss_type='warp' without stnls is provided")
        if dist_crit not in ("l1", "l2"):
            raise ValueError("Uknown criterion [%s]" % dist_crit)      # warped_loss.py:248
        self.dist_crit = dist_crit
        self.loss_type = loss_type

    def forward(self, input, target, flow, step=0, in_mask=None):
        warped, mask = self._warped_mask(target, flow, in_mask)
        return _MaskedLoss.apply(input.to(warped.device), warped, mask, 1 if self.dist_crit == "l2" else 2)
