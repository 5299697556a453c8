"""GPU parity tests of the fused warp + occlusion-mask kernel and the masked loss against the
oracle and the reference's golden outputs.  Bar: masks bit-exact, warped bit-exact (same fused
multiply-adds as ATen's CPU kernel), loss within float32 summation error, gradient exact."""
import os

import numpy as np
import pytest
import torch

from frame2frame_b200.warped_loss import LibWarpedLoss, WarpedLoss, warp_mask
from oracle import f2f_oracle as fo
from oracle import f2f_ref

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "f2f_step_ref.npz"))


@pytest.mark.parametrize("case", ["s", "m"])
def test_step_matches_reference_golden(cuda, case):
    H, W = GOLD[case + "_hw"]
    curr, prev, flow = f2f_ref.step_inputs(H, W)
    crit = WarpedLoss().cuda()
    tprev = torch.from_numpy(prev)[None, None].to(cuda)
    tflow = torch.from_numpy(flow)[None]                      # CPU tensor, as in blind_denoising.py:208
    warped, valid = crit.warp(tprev, tflow)
    mask = crit.occlusion_mask(warped, tflow, valid)
    assert np.array_equal(warped[0, 0].cpu().numpy(), GOLD[case + "_warped"])
    assert np.array_equal(valid[0, 0].cpu().numpy().astype(np.uint8), GOLD[case + "_valid"])
    assert np.array_equal(mask[0, 0].cpu().numpy().astype(np.uint8), GOLD[case + "_mask"])
    inp = torch.from_numpy(curr)[None, None].to(cuda).requires_grad_(True)
    loss = crit(inp, tprev, tflow)
    loss.backward()
    assert abs(loss.item() - float(GOLD[case + "_loss"])) <= 2e-6 * abs(loss.item())
    assert np.array_equal(inp.grad[0, 0].cpu().numpy().astype(np.int8), GOLD[case + "_grad"])


@pytest.mark.parametrize("hw", [(1080, 1920), (33, 31)])
def test_fused_kernel_vs_oracle(cuda, hw):
    H, W = hw
    curr, prev, flow = f2f_ref.step_inputs(H, W, t=3)
    for ac in (False, True):
        for f64 in (True, False):
            w, v, m = warp_mask(torch.from_numpy(prev)[None, None].to(cuda), torch.from_numpy(flow)[None].to(cuda),
                                align_corners=ac, occ_f64=f64)
            w2, v2, m2 = fo.warp_and_mask(prev, flow, f64=f64, align_corners=ac)
            assert np.array_equal(v[0, 0].cpu().numpy(), v2)
            assert np.array_equal(m[0, 0].cpu().numpy(), m2)
            assert np.array_equal(w[0, 0].cpu().numpy(), w2)


def test_lib_flavour_losses_and_in_mask(cuda):
    H, W = 64, 96
    curr, prev, flow = f2f_ref.step_inputs(H, W)
    rng = np.random.default_rng(1)
    in_mask = (rng.random((H, W)) > 0.3).astype(np.float32)
    for crit_name in ("l2", "l1"):
        crit = LibWarpedLoss(dist_crit=crit_name).cuda()
        inp = torch.from_numpy(curr)[None, None].to(cuda).requires_grad_(True)
        loss = crit(inp, torch.from_numpy(prev)[None, None].to(cuda), torch.from_numpy(flow)[None].to(cuda), 0,
                    torch.from_numpy(in_mask)[None, None].to(cuda))
        loss.backward()
        w2, v2, m2 = fo.warp_and_mask(prev, flow, f64=False)
        l2, g2 = fo.masked_loss(curr, w2, in_mask * m2, crit_name)
        assert abs(loss.item() - l2) <= 1e-5 * abs(l2)
        assert np.abs(inp.grad[0, 0].cpu().numpy() - g2).max() <= 1e-6 * max(1e-12, np.abs(g2).max())
    with pytest.raises(ValueError):
        LibWarpedLoss(dist_crit="huber")


def test_batch_quirks_and_empty(cuda):
    # B=2, C=3: flow of batch item 0 drives every occlusion test; only mask[0,0] is dilated
    H, W = 24, 40
    _, prev, flow = f2f_ref.step_inputs(H, W)
    x = torch.from_numpy(np.stack([prev, prev * 0.5, prev + 0.1]))[None].repeat(2, 1, 1, 1).contiguous().to(cuda)
    fl = torch.from_numpy(np.stack([flow, flow * 0.0])).to(cuda)
    w, v, m = warp_mask(x, fl)
    occ_d = fo.occlusion(flow)                       # dilated, from batch 0's flow
    raw = np.abs((np.pad(flow[0, 1:] - flow[0, :-1], ((0, 1), (0, 0)))).astype(np.float64)
                 + np.pad(flow[1, :, 1:] - flow[1, :, :-1], ((0, 0), (0, 1)))) > 0.75
    raw[0, :] = raw[-1, :] = True
    raw[:, 0] = raw[:, -1] = True
    for b in range(2):
        _, vb = fo.warp(prev, flow if b == 0 else flow * 0.0)
        for c in range(3):
            occ = occ_d if (b == 0 and c == 0) else raw.astype(np.uint8)
            assert np.array_equal(m[b, c].cpu().numpy(), vb * (1 - occ))
    e = torch.empty((0, 1, 8, 8), device=cuda)
    assert warp_mask(e, torch.empty((0, 2, 8, 8), device=cuda))[0].shape == (0, 1, 8, 8)
    with pytest.raises(ValueError):
        warp_mask(torch.zeros((1, 1, 8, 8), device=cuda), torch.zeros((1, 2, 8, 9), device=cuda))
