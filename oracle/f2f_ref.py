"""oracle/f2f_ref.py -- TEST INFRASTRUCTURE.  Loads the REFERENCE's own
blind_denoising.py (and lib/frame2frame/dncnn/net.py) in this container so that
golden vectors can be generated from it and the restatement in
oracle/f2f_oracle.py can be pinned against it.  Needs /root/reference; it is
never imported on the GPU box (tests read tests/golden/*.npz instead).

How: the reference imports an un-vendored `dncnn.models`,
`skimage`, `tifffile`; stub modules are injected for those (DnCNN's constructor
comes from the in-tree lib/frame2frame/dncnn/net.py:7-21 with SaoYan's
noise-estimate forward) and `.cuda()` is made a no-op so everything runs on CPU.
"""
import importlib.util
import os
import sys
import types
import warnings

import numpy as np

REF = os.environ.get("F2F_REFERENCE", "/root/reference")


def available():
    return os.path.exists(os.path.join(REF, "blind_denoising.py"))


_mod = None


def load():
    """-> module object of the reference's blind_denoising.py (CPU only)."""
    global _mod
    if _mod is not None:
        return _mod
    import torch
    import torch.nn as nn
    warnings.filterwarnings("ignore")
    spec = importlib.util.spec_from_file_location("_ref_dncnn_net", os.path.join(REF, "lib/frame2frame/dncnn/net.py"))
    net = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(net)

    class DnCNN(net.DnCNN):                 # SaoYan-style forward: returns the noise estimate
        def forward(self, x):
            return self.dncnn(x)

    pkg = types.ModuleType("dncnn")
    models = types.ModuleType("dncnn.models")
    models.DnCNN = DnCNN
    models.__all__ = ["DnCNN"]
    pkg.models = models
    sys.modules["dncnn"] = pkg
    sys.modules["dncnn.models"] = models
    for name in ("skimage", "skimage.io", "tifffile"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["skimage"].io = sys.modules["skimage.io"]
    torch.Tensor.cuda = lambda self, *a, **k: self
    nn.Module.cuda = lambda self, *a, **k: self
    sys.path.insert(0, REF)
    spec = importlib.util.spec_from_file_location("_ref_blind_denoising", os.path.join(REF, "blind_denoising.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    os.environ.pop("CUDA_VISIBLE_DEVICES", None)   # blind_denoising.py:33-34 sets it at import
    mod.RefDnCNN = DnCNN
    mod.LibDnCNN = net.DnCNN
    _mod = mod
    return mod


def step_inputs(H, W, t=1, seq=0, sigma=25.0):
    """Seeded inputs of one F2F step: curr, prev in [0,1], backward flow (analytic + smooth wobble)."""
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, ROOT)
    from frame2frame_b200 import synth
    curr = synth.noisy_frame(H, W, t, sigma, seq) / np.float32(255.0)
    prev = synth.noisy_frame(H, W, t - 1, sigma, seq) / np.float32(255.0)
    flow = synth.backward_flow(H, W, t)
    y, x = np.mgrid[0:H, 0:W].astype(np.float32)
    flow[0] += (0.3 * np.sin(y / 7.0) + 0.2 * np.cos(x / 5.0)).astype(np.float32)
    flow[1] += (0.25 * np.cos(y / 9.0) * np.sin(x / 6.0)).astype(np.float32)
    return curr.astype(np.float32), prev.astype(np.float32), flow.astype(np.float32)


def make_golden(gold_dir):
    """Writes tests/golden/f2f_step_ref.npz: outputs of the reference WarpedLoss."""
    import torch
    ref = load()
    out = {}
    for name, (H, W) in {"s": (37, 53), "m": (96, 128)}.items():
        curr, prev, flow = step_inputs(H, W)
        crit = ref.WarpedLoss()
        tprev = torch.from_numpy(prev)[None, None]
        tflow = torch.from_numpy(flow)[None]
        warped, valid = crit.warp(tprev, tflow)
        mask = crit.occlusion_mask(warped, tflow, valid)
        inp = torch.from_numpy(curr)[None, None].clone().requires_grad_(True)
        loss = crit(inp, tprev, tflow)
        loss.backward()
        out[name + "_hw"] = np.array([H, W], np.int32)
        out[name + "_warped"] = warped[0, 0].numpy()
        out[name + "_valid"] = valid[0, 0].numpy().astype(np.uint8)
        out[name + "_mask"] = mask[0, 0].numpy().astype(np.uint8)
        out[name + "_loss"] = np.array(loss.item(), np.float64)
        out[name + "_grad"] = inp.grad[0, 0].numpy().astype(np.int8)
    np.savez_compressed(os.path.join(gold_dir, "f2f_step_ref.npz"), **out)
    print("f2f_step_ref.npz:", {k: getattr(v, "shape", None) for k, v in out.items()})
