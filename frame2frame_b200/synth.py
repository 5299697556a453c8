"""Synthetic sequences, flows and DnCNN weights for parity tests and bench.py.

There is no network for datasets or checkpoints, so every input is generated
here, deterministically:

* frames: a textured background translating by (1.5, -0.75) px/frame with a
  25 %-area textured disc translating by (+2.0, -1.0) px/frame on top of it
  (so real occlusions and a motion boundary exist), plus N(0, sigma^2) noise,
  quantised to uint8 like the PNG inputs the reference reads
  (blind_denoising.py:173-180);
* flows: the analytic backward flow (frame t -> frame t-1) of that motion;
* weights: seeded DnCNN-S-25-layout state dict with the ``module.`` prefix the
  reference's DataParallel checkpoint has (blind_denoising.py:144-147).
"""
import math

import numpy as np

BG_V = (1.5, -0.75)      # background shift of the texture coordinates per frame
DISC_V = (2.0, -1.0)     # disc centre velocity, px / frame


def _disc(H, W, t):
    r = math.sqrt(0.25 * H * W / math.pi)
    cx = 0.45 * W + DISC_V[0] * t
    cy = 0.55 * H + DISC_V[1] * t
    return cx, cy, r


def clean_frame(H, W, t):
    """Noise-free frame t, float64 in roughly [20, 235]."""
    y, x = np.mgrid[0:H, 0:W].astype(np.float64)
    a, b = BG_V[0] * t, BG_V[1] * t
    xa, yb = x + a, y + b
    bg = (127.5 + 40 * np.sin(xa / 17) + 40 * np.cos(yb / 23)
          + 30 * np.sin((xa + yb) / 31) + 15 * np.sin(xa / 5) * np.cos(yb / 7))
    cx, cy, r = _disc(H, W, t)
    xd, yd = x - cx, y - cy
    fg = (127.5 + 50 * np.sin(xd / 9) * np.cos(yd / 11) + 35 * np.cos((xd - yd) / 13)
          + 15 * np.sin((xd + 2 * yd) / 4))
    return np.where(xd * xd + yd * yd <= r * r, fg, bg)


def noisy_frame(H, W, t, sigma=25.0, seq=0):
    """uint8-quantised noisy frame as float32 in 0..255 (what TV-L1 consumes;
    the F2F step consumes it divided by 255)."""
    rng = np.random.default_rng(1000 + seq * 1000 + t)
    f = clean_frame(H, W, t) + rng.normal(0.0, sigma, (H, W))
    return np.clip(np.round(f), 0, 255).astype(np.float32)


def clean_frame_u8(H, W, t):
    return np.clip(np.round(clean_frame(H, W, t)), 0, 255).astype(np.float32)


def backward_flow(H, W, t):
    """Analytic flow from frame t to frame t-1, float32 [2,H,W] (ch0 = x)."""
    y, x = np.mgrid[0:H, 0:W].astype(np.float64)
    cx, cy, r = _disc(H, W, t)
    inside = (x - cx) ** 2 + (y - cy) ** 2 <= r * r
    u = np.where(inside, -DISC_V[0], BG_V[0])
    v = np.where(inside, -DISC_V[1], BG_V[1])
    return np.stack([u, v]).astype(np.float32)


def dncnn_state_dict(seed=0, channels=1, num_of_layers=17, features=64, prefix="module.",
                     scale=1.0):
    """Seeded DnCNN weights in the reference checkpoint layout
    (lib/frame2frame/dncnn/net.py:7-21: conv at 0,2,5,...; BN at 3,6,...).
    Kaiming-normal convs, BN gamma=1, beta=0, running stats (0,1).  Returned as
    a dict of numpy arrays so it has no torch dependency; the host API converts.
    """
    rng = np.random.default_rng(seed)
    sd = {}

    def conv(idx, cout, cin):
        std = math.sqrt(2.0 / (cin * 9)) * scale
        sd["%sdncnn.%d.weight" % (prefix, idx)] = rng.normal(0, std, (cout, cin, 3, 3)).astype(np.float32)

    conv(0, features, channels)
    idx = 2
    for _ in range(num_of_layers - 2):
        conv(idx, features, features)
        b = idx + 1
        sd["%sdncnn.%d.weight" % (prefix, b)] = np.ones(features, np.float32)
        sd["%sdncnn.%d.bias" % (prefix, b)] = np.zeros(features, np.float32)
        sd["%sdncnn.%d.running_mean" % (prefix, b)] = np.zeros(features, np.float32)
        sd["%sdncnn.%d.running_var" % (prefix, b)] = np.ones(features, np.float32)
        sd["%sdncnn.%d.num_batches_tracked" % (prefix, b)] = np.zeros((), np.int64)
        idx += 3
    conv(idx, channels, features)
    return sd


def pretrained_state_dict(path=None):
    """The DnCNN-S-25-layout denoiser trained on synthetic textures by tools/pretrain_dncnn.py
    (stand-in for the un-vendored dncnn/logs/DnCNN-S-25/net.pth), `module.`-prefixed,
    as float32 numpy arrays."""
    import os
    if path is None:
        path = os.environ.get("F2F_PRETRAINED") or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                            "dncnn_s25_synth.npz")
    return {k: (np.asarray(v, np.float32) if v.dtype.kind == "f" else np.asarray(v)) for k, v in np.load(path).items()}
