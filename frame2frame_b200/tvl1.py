"""TV-L1 optical flow on B200: host-side mirror of the reference's tvl1flow.

Reference interface this replaces: the ``tvl1flow`` executable run once per
frame pair by tvl1flow/tvl1flow.sh:9-19 (positional parameters of
tvl1flow/main.c:86-99) and its library entry point
``Dual_TVL1_optic_flow_multiscale`` (tvl1flow/tvl1flow_lib.c:356-372).

``tvl1flow(I0, I1)`` takes CUDA tensors ``[N,H,W]`` (or ``[H,W]``) and returns
``flow[N,2,H,W]`` with channel 0 = horizontal displacement, the layout
``read_flow(...).permute(2,0,1)`` has in blind_denoising.py:206-209.  All
arithmetic runs in libf2f_b200.so; torch only owns the memory and the stream.
"""
import ctypes as C

import numpy as np
import torch

from ._lib import Tvl1Params, check, lib


def make_params(nx, ny, tau=0.25, lam=0.15, theta=0.3, nscales=100, fscale=0, zfactor=0.5,
                nwarps=5, epsilon=0.01, verbose=0):
    """Parameters with main.c's validation (:102-141) and nscales clamp (:159-163) applied."""
    p = Tvl1Params(tau, lam, theta, nscales, fscale, zfactor, nwarps, epsilon, verbose)
    check(lib().f2f_tvl1_validate_params(C.byref(p), nx, ny), "f2f_tvl1_validate_params")
    return p


class Tvl1Workspace:
    """Device workspace for `nslots` concurrent frame pairs (reused across calls)."""

    def __init__(self, nx, ny, params, nslots, device):
        self.nbytes = lib().f2f_tvl1_workspace_bytes(nx, ny, C.byref(params), nslots)
        if self.nbytes == 0:
            raise ValueError("bad TV-L1 geometry/parameters for %dx%d" % (nx, ny))
        self.key = (nx, ny, params.nscales, params.nwarps, nslots, str(device))
        self.buf = torch.empty(self.nbytes, dtype=torch.uint8, device=device)


_ws_cache = {}


def tvl1flow(I0, I1, nslots=16, return_iters=False, **kw):
    """Flow from I0 to I1 for each pair.  I0, I1: float32 CUDA tensors [N,H,W] or [H,W]."""
    if not I0.is_cuda:
        raise RuntimeError("tvl1flow needs CUDA tensors: there is no CPU fallback in this package")
    squeeze = I0.dim() == 2
    if squeeze:
        I0, I1 = I0[None], I1[None]
    I0 = I0.contiguous().float()
    I1 = I1.contiguous().float()
    if I0.shape != I1.shape or I0.dim() != 3:
        raise ValueError("input images size mismatch %s != %s" % (tuple(I0.shape), tuple(I1.shape)))  # main.c:189-193
    n, ny, nx = I0.shape
    p = make_params(nx, ny, **kw)
    nslots = max(1, min(nslots, n))
    key = (nx, ny, p.nscales, p.nwarps, nslots, str(I0.device))
    ws = _ws_cache.get(key)
    if ws is None:
        _ws_cache.clear()
        ws = _ws_cache[key] = Tvl1Workspace(nx, ny, p, nslots, I0.device)
    flow = torch.empty((n, 2, ny, nx), dtype=torch.float32, device=I0.device)
    iters = np.zeros((n, p.nscales, p.nwarps), np.int32) if return_iters else None
    with torch.cuda.device(I0.device):
        st = torch.cuda.current_stream().cuda_stream
        check(lib().f2f_tvl1_flow_batch(I0.data_ptr(), I1.data_ptr(), flow.data_ptr(), n, nx, ny, C.byref(p),
                                        iters.ctypes.data if iters is not None else None,
                                        ws.buf.data_ptr(), ws.nbytes, st), "f2f_tvl1_flow_batch")
    if squeeze:
        flow = flow[0]
        iters = iters[0] if iters is not None else None
    return (flow, iters) if return_iters else flow


def tvl1flow_bidir(A, B, nslots=16, return_iters=False, **kw):
    """Forward AND backward flow of every pair (A[k], B[k]) in one pass with shared pyramids (tvl1flow.sh:11-18 computes
    them as two processes per frame): -> (flow A->B [N,2,H,W], flow B->A [N,2,H,W]), bit-identical to
    tvl1flow(A, B), tvl1flow(B, A).  A, B: float32 CUDA tensors [N,H,W] or [H,W]."""
    if not A.is_cuda:
        raise RuntimeError("tvl1flow_bidir needs CUDA tensors: there is no CPU fallback in this package")
    squeeze = A.dim() == 2
    if squeeze:
        A, B = A[None], B[None]
    A, B = A.contiguous().float(), B.contiguous().float()
    if A.shape != B.shape or A.dim() != 3:
        raise ValueError("input images size mismatch %s != %s" % (tuple(A.shape), tuple(B.shape)))
    n, ny, nx = A.shape
    p = make_params(nx, ny, **kw)
    nslots = max(2, min(nslots, 2 * n) & ~1)
    key = (nx, ny, p.nscales, p.nwarps, nslots, str(A.device))
    ws = _ws_cache.get(key)
    if ws is None:
        _ws_cache.clear()
        ws = _ws_cache[key] = Tvl1Workspace(nx, ny, p, nslots, A.device)
    flow = torch.empty((n, 2, 2, ny, nx), dtype=torch.float32, device=A.device)
    iters = np.zeros((n, 2, p.nscales, p.nwarps), np.int32) if return_iters else None
    if n:
        with torch.cuda.device(A.device):
            st = torch.cuda.current_stream().cuda_stream
            check(lib().f2f_tvl1_flow_bidir(A.data_ptr(), B.data_ptr(), flow.data_ptr(), n, nx, ny, C.byref(p),
                                            iters.ctypes.data if iters is not None else None,
                                            ws.buf.data_ptr(), ws.nbytes, st), "f2f_tvl1_flow_bidir")
    fwd, bwd = flow[:, 0], flow[:, 1]
    if squeeze:
        fwd, bwd = fwd[0], bwd[0]
        iters = iters[0] if iters is not None else None
    return (fwd, bwd, iters) if return_iters else (fwd, bwd)


def tvl1flow_host(I0, I1, max_slots=16, return_iters=False, **kw):
    """Same through host (numpy) buffers: H2D, batch, D2H inside the library."""
    I0 = np.ascontiguousarray(I0, np.float32)
    I1 = np.ascontiguousarray(I1, np.float32)
    if I0.ndim == 2:
        I0, I1 = I0[None], I1[None]
    n, ny, nx = I0.shape
    p = make_params(nx, ny, **kw)
    flow = np.empty((n, 2, ny, nx), np.float32)
    iters = np.zeros((n, p.nscales, p.nwarps), np.int32) if return_iters else None
    check(lib().f2f_tvl1_flow_batch_host(I0.ctypes.data, I1.ctypes.data, flow.ctypes.data, n, nx, ny, C.byref(p),
                                         iters.ctypes.data if iters is not None else None, max_slots),
          "f2f_tvl1_flow_batch_host")
    return (flow, iters) if return_iters else flow


# --- .flo files: byte-compatible with iio.c:2114-2129 / readFlowFile.py:18-34 ---
TAG_FLOAT = 202021.25


def write_flo(path, flow):
    """flow: [2,H,W] (tensor or array) -> Middlebury .flo (tag, int32 w, int32 h, interleaved u,v)."""
    f = flow.detach().cpu().numpy() if hasattr(flow, "detach") else np.asarray(flow)
    _, h, w = f.shape
    with open(path, "wb") as fh:
        np.array([TAG_FLOAT], np.float32).tofile(fh)
        np.array([w, h], np.int32).tofile(fh)
        np.ascontiguousarray(f.transpose(1, 2, 0), np.float32).tofile(fh)


def read_flo(path):
    """-> ndarray (h, w, 2) float32, as readFlowFile.read_flow returns."""
    with open(path, "rb") as fh:
        tag = np.fromfile(fh, np.float32, 1)[0]
        if tag != np.float32(TAG_FLOAT):
            raise ValueError("Flow number %r incorrect. Invalid .flo file" % tag)
        w, h = np.fromfile(fh, np.int32, 2)
        data = np.fromfile(fh, np.float32, 2 * int(w) * int(h))
    return data.reshape(int(h), int(w), 2)
