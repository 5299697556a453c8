"""``frame2frame`` -- the reference package's import name (setup.py:18-32 installs ``lib/frame2frame`` under it), served
by the B200 implementation: an unmodified harness that does

    import frame2frame
    net  = frame2frame.load_model(cfg)              # lib/frame2frame/__init__.py:32-39
    crit = frame2frame.WarpedLoss(dist_crit="l2")   # lib/frame2frame/warped_loss.py:117-131
    cfg  = frame2frame.extract_config(cfg)          # lib/frame2frame/__init__.py:24-31

gets the CUDA path of ``frame2frame_b200``.  Only what SURVEY section 8 puts in scope is here: the DnCNN loader and the
warped loss.  The reference's other exports (``lightning``, ``DnlsLoss``, ``get_loss_fxn``, ``fastdvdnet``) wrap absent
third-party packages and raise ``ImportError`` with that explanation when touched.
"""
from frame2frame_b200.lib_api import WarpedLoss, extract_config, load_model, optional  # noqa: F401
from . import dncnn  # noqa: F401


def extract_model_config(cfg):
    """lib/frame2frame/__init__.py:21-22"""
    return extract_config(cfg)


_OUT_OF_SCOPE = {
    "lightning": "pytorch_lightning / dev_basics / data_hub harness",
    "DnlsLoss": "stnls non-local search losses",
    "get_loss_fxn": "the reference's `.losses` module (absent from its own tree)",
    "fastdvdnet": "the reference's `.fastdvdnet` module (absent from its own tree)",
}


def __getattr__(name):
    if name in _OUT_OF_SCOPE:
        raise ImportError("frame2frame.%s is outside the hot path this build replaces (%s); see SURVEY.md section 2"
                          % (name, _OUT_OF_SCOPE[name]))
    raise AttributeError("module 'frame2frame' has no attribute %r" % name)
