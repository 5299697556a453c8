"""oracle/make_golden.py -- TEST INFRASTRUCTURE.  Regenerates tests/golden/*.npz
by executing the REFERENCE itself in this container:

  * TV-L1: oracle/_ref/libtvl1_ref.so (tvl1flow_lib.c compiled in place);
  * F2F:   /root/reference/blind_denoising.py imported with stub modules
           (see oracle/f2f_ref.py).

The fixtures are small (a few hundred KB) and committed; the GPU box has no
/root/reference, so the -m gpu tests read only these files.
    python oracle/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from frame2frame_b200 import synth  # noqa: E402
from oracle.tvl1 import OracleTvl1, RefTvl1  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def tvl1():
    ref, orc = RefTvl1(), OracleTvl1()
    out = {}
    cases = [("a", 68, 120, dict()), ("b", 97, 131, dict(nscales=3, nwarps=3)),
             ("c", 135, 240, dict(lam=0.2, nscales=100, fscale=2))]   # c = tvl1flow.sh:14 settings
    for name, H, W, kw in cases:
        I0 = synth.noisy_frame(H, W, 1)
        I1 = synth.noisy_frame(H, W, 0)
        u = ref.flow(I0, I1, **kw)
        _, iters, _ = orc.flow(I0, I1, **kw)
        out["%s_hw" % name] = np.array([H, W], np.int32)
        out["%s_flow" % name] = u
        out["%s_iters" % name] = iters
        out["%s_kw" % name] = np.array(repr(kw))
    np.savez_compressed(os.path.join(GOLD, "tvl1_ref.npz"), **out)
    print("tvl1_ref.npz:", {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    tvl1()
    try:
        from oracle import f2f_ref
        f2f_ref.make_golden(GOLD)
    except ImportError as e:
        print("f2f golden skipped:", e)
