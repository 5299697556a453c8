"""`lib/frame2frame` entry points on B200: what the reference's package exports for the DnCNN path --
``frame2frame.load_model(cfg)`` / ``extract_config(cfg)`` (lib/frame2frame/__init__.py:24-39, dncnn/io.py:22-86) and
``frame2frame.WarpedLoss`` (warped_loss.py:117) -- so that a harness written against the reference package can
call the CUDA path.  The reference's config plumbing (``dev_basics.configs.ExtractConfig``) is an un-vendored
dependency; its observable behaviour for this model is restated: keys, defaults, ``restrict`` filtering.

Differences, stated rather than hidden: ``channels`` is 1 (north_star's grayscale DnCNN) or 3 (the reference's default,
dncnn/io.py:64-67) -- other counts raise; ``net_name`` values of the FastDVDnet family raise ``ValueError`` like any
unknown name (that model is outside SURVEY section 8).
"""
import torch

from .dncnn import DnCNN
from .warped_loss import LibWarpedLoss as WarpedLoss  # noqa: F401  (warped_loss.py:117-248 constructor and forward)

ARCH_PAIRS = {"channels": 3, "num_of_layers": 17}                                  # dncnn/io.py:64-67
IO_PAIRS = {"pretrained_path": "", "pretrained_root": "", "pretrained_type": "",   # dncnn/io.py:69-76
            "pretrained_load": False, "net_name": "dncnn"}


def optional(cfg, key, default):
    """lib/frame2frame utils.optional: cfg[key] if present else default (dicts and attribute bags)."""
    if isinstance(cfg, dict):
        return cfg.get(key, default)
    return getattr(cfg, key, default)


def extract_config(cfg):
    """The model's own keys with defaults filled in (dncnn/io.py:17-18 `extract_config`, restrict=True)."""
    mtype = optional(cfg, "net_name", "dncnn")
    if mtype not in ("dncnn",):
        raise ValueError(f"Uknown model type [{mtype}]")                            # __init__.py:31,39 (sic)
    out = {"device": optional(cfg, "device", "cuda:0")}
    for pairs in (IO_PAIRS, ARCH_PAIRS):
        for k, v in pairs.items():
            out[k] = optional(cfg, k, v)
    return out


def load_pretrained(model, cfg):
    """dncnn/io.py:78-82.  Accepts a raw state dict, ``{"state_dict": ...}`` (Lightning) or ``{"model_state_dict": ...}``
    checkpoints; ``module.`` / ``net.`` prefixes are stripped."""
    if not cfg["pretrained_load"]:
        return
    import os
    path = os.path.join(cfg["pretrained_root"], cfg["pretrained_path"]) if cfg["pretrained_root"] else cfg["pretrained_path"]
    print("Loading model: ", path)
    sd = torch.load(path, map_location="cpu")
    for key in ("state_dict", "model_state_dict"):
        if isinstance(sd, dict) and key in sd:
            sd = sd[key]
    clean = {}
    for k, v in sd.items():
        for prefix in ("module.", "net."):
            if k.startswith(prefix):
                k = k[len(prefix):]
        clean[k] = v
    model.load_state_dict(clean)


def load_model(cfg):
    """frame2frame.load_model(cfg) (__init__.py:32-39 -> dncnn/io.py:22-60): DnCNN(channels, num_of_layers) with the
    in-tree forward convention ``x - dncnn(x)`` (net.py:31), optional checkpoint, moved to cfg.device."""
    c = extract_config(cfg)
    print("load_model:", c["net_name"])
    if int(c["channels"]) not in (1, 3):
        raise NotImplementedError("frame2frame_b200 implements the DnCNN for 1 (gray) or 3 (RGB) image channels; got channels=%r"
                                  % (c["channels"],))
    model = DnCNN(int(c["channels"]), int(c["num_of_layers"]), residual=True)
    load_pretrained(model, c)
    return model.to(c["device"])
