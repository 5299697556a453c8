"""The FP32 ("precise") mode of the DnCNN path (csrc/dncnn_precise.cuh, f2f_dncnn_set_precise): every kernel against
the stock-torch FP32 oracle at FP32 tolerances, and the online loop against the recorded FP32 CPU oracle run at
north_star's 0.05 dB.

Tolerances: forward / gradients relative L2 <= 2e-5 / 2e-4 (FP32 with another summation order through 17 layers and
train-mode BatchNorm); PSNR of the online loop within 0.05 dB per frame of tests/golden/f2f_config0_ref.json."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture()
def precise(cuda):
    from frame2frame_b200 import dncnn
    dncnn.set_precise(True)
    yield
    dncnn.set_precise(False)


def _models(cuda, residual):
    from frame2frame_b200 import synth
    from frame2frame_b200.dncnn import DnCNN
    from oracle import dncnn_oracle as do
    sd = synth.pretrained_state_dict()
    ref = do.load_numpy_state(do.OracleDnCNN(residual=residual), sd)
    net = torch.nn.DataParallel(DnCNN(residual=residual), device_ids=[0]).cuda()
    net.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    return net, ref


def _rel(a, b):
    a, b = a.detach().double().flatten().cpu(), b.detach().double().flatten().cpu()
    return float((a - b).norm() / b.norm())


@pytest.mark.parametrize("B,H,W", [(1, 67, 93), (2, 40, 56), (1, 5, 7), (1, 1, 40), (1, 130, 3), (1, 17, 16), (1, 16, 129)])
def test_precise_network_equals_fp32_oracle(cuda, precise, B, H, W):
    from frame2frame_b200 import synth
    from oracle import dncnn_oracle as do
    net, ref = _models(cuda, True)
    x = torch.from_numpy(np.stack([synth.noisy_frame(H, W, t) / 255.0 for t in range(B)])[:, None].astype(np.float32))
    dy = torch.from_numpy(np.random.default_rng(0).standard_normal(x.shape).astype(np.float32))
    net.eval(); ref.eval()
    with torch.no_grad():
        assert _rel(net(x.to(cuda)), ref(x)) <= 2e-5
    net.train(); ref.train()
    y, yr = net(x.to(cuda)), ref(x)
    tiny = H * W * B < 200                       # BatchNorm over a few dozen pixels is badly conditioned in any arithmetic
    assert _rel(y, yr) <= (2e-3 if tiny else 2e-5)
    (y * dy.to(cuda)).sum().backward()
    (yr * dy).sum().backward()
    m = net.module
    assert _rel(m.flat_grads, do.flat_grads(ref)) <= (5e-2 if tiny else 2e-4)
    rs = torch.stack([torch.stack([b.running_mean, b.running_var]) for b in ref.dncnn if isinstance(b, torch.nn.BatchNorm2d)])
    assert _rel(m._running, rs) <= 1e-5


def test_mode_switch_replans_the_workspace(cuda):
    """Both modes in one process, back and forth: the workspace is re-planned, results of the default mode are unchanged."""
    from frame2frame_b200 import dncnn, synth
    net, ref = _models(cuda, True)
    x = torch.from_numpy((synth.noisy_frame(48, 64, 0) / 255.0)[None, None].astype(np.float32)).to(cuda)
    net.eval()
    with torch.no_grad():
        a = net(x).clone()
        dncnn.set_precise(True)
        try:
            b = net(x).clone()
        finally:
            dncnn.set_precise(False)
        c = net(x).clone()
        r = ref.eval()(x.cpu())
    assert torch.equal(a, c)
    assert _rel(b, r) < _rel(a, r) and _rel(b, r) <= 2e-5


def test_precise_online_loop_meets_the_psnr_contract(cuda, precise):
    """BASELINE.json configs[0] in the precise mode against the recorded FP32 CPU oracle loop: |PSNR difference| <= 0.05 dB
    on EVERY frame (north_star), losses within 1e-4."""
    from frame2frame_b200 import synth
    from frame2frame_b200.blind_denoising import OnlineF2F, psnr
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "f2f_config0_ref.json")))
    W, H = (int(v) for v in gold["frame"].split("x"))
    ITERS, NF = gold["iters"], len(gold["frames"])
    net, _ = _models(cuda, False)
    eng = OnlineF2F(net)
    frames = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255.0))[None, None] for t in range(NF + 1)]
    flows = {t: torch.from_numpy(synth.backward_flow(H, W, t))[None] for t in range(1, NF + 1)}
    losses, diffs = [], []
    for t, den in eng.run_sequence(frames, ITERS, flows=flows, losses=losses):
        g = gold["frames"][t - 1]
        assert abs(losses[(t - 1) * ITERS] / g["loss_first"][0] - 1) <= 1e-4
        assert abs(losses[t * ITERS - 1] / g["loss_last"][0] - 1) <= 1e-4
        clean = synth.clean_frame_u8(H, W, t) / np.float32(255.0)
        diffs.append(float(psnr(clean, den[0, 0].cpu().numpy())) - g["psnr_oracle"])
    print("precise mode, PSNR - oracle per frame (dB):", " ".join("%+.4f" % d for d in diffs))
    assert np.abs(diffs).max() <= 0.05, diffs


@pytest.mark.parametrize("mode", ["default", "precise"])
def test_three_channel_network(cuda, mode):
    """channels = 3, the reference loader's default (lib/frame2frame/dncnn/io.py:64-67, net.py:13,19): forward (train and
    eval), gradients and running statistics against the stock-torch FP32 network with the same seeded weights, in both
    modes.  Bars: precise 2e-5 / 2e-4 like the gray network; default mode 3e-2 / 6e-2 (16-bit operands)."""
    from frame2frame_b200 import dncnn, synth
    from oracle import dncnn_oracle as do
    sd = synth.dncnn_state_dict(seed=3, channels=3)
    ref = do.load_numpy_state(do.OracleDnCNN(channels=3, residual=True), sd)
    net = torch.nn.DataParallel(dncnn.DnCNN(channels=3, residual=True), device_ids=[0]).cuda()
    net.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    B, H, W = 2, 37, 53
    rng = np.random.default_rng(5)
    x = torch.from_numpy(rng.uniform(0, 1, (B, 3, H, W)).astype(np.float32))
    dy = torch.from_numpy(rng.standard_normal((B, 3, H, W)).astype(np.float32))
    ty, tg = (2e-5, 2e-4) if mode == "precise" else (3e-2, 6e-2)
    dncnn.set_precise(mode == "precise")
    try:
        net.eval(); ref.eval()
        with torch.no_grad():
            assert _rel(net(x.to(cuda)), ref(x)) <= ty
        net.train(); ref.train()
        y, yr = net(x.to(cuda)), ref(x)
        assert _rel(y, yr) <= ty
        (y * dy.to(cuda)).sum().backward()
        (yr * dy).sum().backward()
        m = net.module
        g, gr = m.flat_grads, do.flat_grads(ref)
        assert g.numel() == 556032 + 2304
        assert _rel(g, gr) <= tg
        assert _rel(g[:1728], gr[:1728]) <= tg and _rel(g[-1728:], gr[-1728:]) <= tg      # conv0 / conv16 on their own
        with pytest.raises(ValueError):
            net(torch.zeros(1, 1, 8, 8, device=cuda))
    finally:
        dncnn.set_precise(False)
