"""CPU tests: the numpy restatement of the F2F step against the reference's outputs
(tests/golden/f2f_step_ref.npz, produced by the reference's own WarpedLoss under torch CPU),
and against the reference itself where /root/reference is present."""
import os

import numpy as np
import pytest

from oracle import f2f_oracle as fo
from oracle import f2f_ref

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "f2f_step_ref.npz"))


@pytest.mark.parametrize("case", ["s", "m"])
def test_oracle_matches_reference_golden(case):
    H, W = GOLD[case + "_hw"]
    curr, prev, flow = f2f_ref.step_inputs(H, W)
    warped, valid, mask = fo.warp_and_mask(prev, flow)
    assert np.array_equal(warped, GOLD[case + "_warped"])        # bit-exact
    assert np.array_equal(valid.astype(np.uint8), GOLD[case + "_valid"])
    assert np.array_equal(mask.astype(np.uint8), GOLD[case + "_mask"])
    loss, grad = fo.masked_loss(curr, warped, mask)
    assert abs(loss - float(GOLD[case + "_loss"])) <= 2e-6 * abs(loss)   # the reference sums in float32
    assert np.array_equal(grad.astype(np.int8), GOLD[case + "_grad"])
    assert 0.3 < mask.mean() < 0.99 and (1 - fo.occlusion(flow)).mean() < 0.999   # occlusions do fire


def test_zero_flow_is_not_identity_as_run():
    # with grid_sample's default align_corners=False the zero-flow warp is not the identity and
    # the validity mask loses a border; with align_corners=True it is.
    rng = np.random.default_rng(0)
    x = rng.random((20, 30), dtype=np.float32)
    z = np.zeros((2, 20, 30), np.float32)
    w, v = fo.warp(x, z, align_corners=False)
    assert not np.array_equal(w, x) and v[0].sum() == 0 and v[:, 0].sum() == 0
    w, v = fo.warp(x, z, align_corners=True)
    assert np.abs(w - x).max() < 1e-6 and v.all()


def test_occlusion_direction_quirk():
    # the test is du/dy + dv/dx, not a divergence
    H, W = 16, 16
    y, x = np.mgrid[0:H, 0:W].astype(np.float32)
    assert fo.occlusion(np.stack([y, 0 * y]).astype(np.float32)).all()          # u ramps along y: all occluded
    assert fo.occlusion(np.stack([x, 0 * x]).astype(np.float32))[1:-1, 1:-1].sum() == 0   # u ramps along x: none


@pytest.mark.skipif(not f2f_ref.available(), reason="/root/reference not present")
def test_oracle_matches_reference_live():
    import torch
    ref = f2f_ref.load()
    curr, prev, flow = f2f_ref.step_inputs(61, 83, t=2, seq=1)
    crit = ref.WarpedLoss()
    tp, tf = torch.from_numpy(prev)[None, None], torch.from_numpy(flow)[None]
    warped, valid = crit.warp(tp, tf)
    mask = crit.occlusion_mask(warped, tf, valid)
    w2, v2, m2 = fo.warp_and_mask(prev, flow)
    assert np.array_equal(w2, warped[0, 0].numpy())
    assert np.array_equal(v2, valid[0, 0].numpy()) and np.array_equal(m2, mask[0, 0].numpy())
    assert abs(fo.psnr(curr, prev) - ref.psnr(curr, prev)) < 1e-9
