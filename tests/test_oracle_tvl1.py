"""CPU tests: the TV-L1 oracle restatement against the reference's outputs.

* against tests/golden/tvl1_ref.npz (flows produced by the compiled reference,
  oracle/make_golden.py) -- bit-exact;
* against oracle/_ref/libtvl1_ref.so stage by stage when it is present.
"""
import ast
import os

import numpy as np
import pytest

from frame2frame_b200 import synth
from oracle.tvl1 import OracleTvl1, RefTvl1, clamp_scales

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "tvl1_ref.npz"))


@pytest.mark.parametrize("case", ["a", "b", "c"])
def test_oracle_matches_reference_golden(case):
    H, W = GOLD[case + "_hw"]
    kw = ast.literal_eval(str(GOLD[case + "_kw"]))
    u, iters, _ = OracleTvl1().flow(synth.noisy_frame(H, W, 1), synth.noisy_frame(H, W, 0), **kw)
    assert np.array_equal(u, GOLD[case + "_flow"])          # bit-exact
    assert np.array_equal(iters, GOLD[case + "_iters"])


def test_scale_clamp_matches_main_c():
    # main.c:159-163: N = 1 + log(hypot(nx,ny)/16)/log(1/zfactor)
    assert clamp_scales(1920, 1080, 0.5, 100, 0) == (8, 0)
    assert clamp_scales(3840, 2160, 0.5, 100, 0) == (9, 0)
    assert clamp_scales(960, 540, 0.5, 5, 0) == (5, 0)
    assert clamp_scales(120, 68, 0.5, 100, 7)[1] <= clamp_scales(120, 68, 0.5, 100, 7)[0]


@pytest.mark.skipif(not RefTvl1.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("hw", [(33, 47), (68, 120)])
def test_oracle_stages_match_reference_library(hw):
    H, W = hw
    o, r = OracleTvl1(), RefTvl1()
    I0, I1 = synth.noisy_frame(H, W, 3), synth.noisy_frame(H, W, 2)
    assert np.array_equal(o.gaussian(I0, 0.8), r.gaussian(I0, 0.8))
    assert np.array_equal(o.zoom_out(I0, 0.5), r.zoom_out(I0, 0.5))
    assert np.array_equal(o.zoom_out(I0, 0.7), r.zoom_out(I0, 0.7))
    a, b = o.normalize(I0 * 0.3 - 7, I1 * 0.3 - 7), r.normalize(I0 * 0.3 - 7, I1 * 0.3 - 7)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    a, b = o.centered_gradient(I0), r.centered_gradient(I0)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    z = o.zoom_out(I0, 0.5)
    zi = r.zoom_in(z, W, H) * (np.float32(1.0) / np.float32(0.5))
    assert np.array_equal(o.zoom_in_scaled(z, W, H, 0.5), zi)
    # one scale, from a non-zero initial flow
    u0 = np.full((H, W), 0.7, np.float32)
    v0 = np.full((H, W), -0.4, np.float32)
    g0, g1 = o.gaussian(I0, 0.8), o.gaussian(I1, 0.8)
    a = o.solve_scale(g0, g1, u0, v0, nwarps=2)
    b = r.solve_scale(g0, g1, u0, v0, nwarps=2)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.skipif(not RefTvl1.available(), reason="oracle/_ref not built")
def test_oracle_constant_images():
    # den == 0 branch of image_normalization (tvl1flow_lib.c:341-347), zero gradients
    I = np.full((40, 50), 3.0, np.float32)
    u, iters, _ = OracleTvl1().flow(I, I)
    assert np.array_equal(u, RefTvl1().flow(I, I))
    assert np.all(u == 0)
