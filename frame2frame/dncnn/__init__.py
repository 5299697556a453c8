"""``frame2frame.dncnn`` (lib/frame2frame/dncnn/__init__.py:1): ``load_model``, ``extract_config`` and the ``DnCNN``
module (dncnn/net.py:6-39) of the B200 implementation."""
from frame2frame_b200.dncnn import DnCNN  # noqa: F401
from frame2frame_b200.lib_api import extract_config, load_model  # noqa: F401
