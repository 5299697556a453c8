"""GPU parity tests of the DnCNN path (tcgen05 convolutions, weight gradient, BatchNorm
bookkeeping, Adam, the fused F2F training step) through the C ABI.

Bars (floating point, stated per test):
  * each kernel against the PyTorch FP32 reference of the SAME op on the kernel's own inputs
    (oracle/dncnn_bf16.py data flow): relative L2 error <= 2e-3 (BF16 rounding flips only);
  * whole network against the un-rounded FP32 oracle with the pretrained weights: <= 3e-2;
  * Adam against torch.optim.Adam: <= 2e-7 absolute after 25 steps;
  * online F2F fine-tuning against the FP32 CPU oracle loop: loss trajectory within 2 %, PSNR
    within 0.25 dB on the sequence average and 0.5 dB per frame at the test's 108x192 (north_star's
    0.05 dB is measured at 1080p by tools/parity_f2f.py; see test_online_f2f_psnr_parity).
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "dncnn_ref.npz"))


@pytest.fixture(scope="module")
def chk(cuda):
    from tools import check_dncnn
    return check_dncnn


@pytest.mark.parametrize("hw", [(16, 16), (40, 72), (135, 240)])
def test_wgrad64_vs_torch(chk, hw):
    chk.ok = True
    chk.check_wgrad(*hw)
    assert chk.ok


@pytest.mark.parametrize("hw", [(48, 64), (97, 131)])
def test_forward_layer_by_layer(chk, hw):
    chk.ok = True
    chk.check_layers(*hw)
    assert chk.ok


@pytest.mark.parametrize("hw", [(48, 64), (97, 131)])
def test_backward_layer_by_layer(chk, hw):
    chk.ok = True
    chk.check_backward_layers(*hw)
    assert chk.ok


def test_adam_matches_torch(chk):
    chk.ok = True
    chk.check_adam()
    assert chk.ok


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


@pytest.mark.parametrize("B,H,W", [(1, 67, 93), (2, 40, 56)])
def test_whole_network_vs_fp32_oracle_pretrained(cuda, B, H, W):
    from frame2frame_b200 import synth
    from oracle import dncnn_oracle as do
    net, ref = _models(cuda, True)
    x = torch.from_numpy(np.stack([synth.noisy_frame(H, W, t) / 255.0 for t in range(B)])[:, None].astype(np.float32))
    dy = torch.from_numpy(np.random.default_rng(0).standard_normal(x.shape).astype(np.float32))
    net.eval(); ref.eval()
    with torch.no_grad():
        assert _rel(net(x.to(cuda)), ref(x)) <= 3e-2
    net.train(); ref.train()
    y, yr = net(x.to(cuda)), ref(x)
    assert _rel(y, yr) <= 3e-2
    (y * dy.to(cuda)).sum().backward()
    (yr * dy).sum().backward()
    m = net.module
    assert _rel(m.flat_grads, do.flat_grads(ref)) <= 6e-2
    rs = torch.stack([torch.stack([b.running_mean, b.running_var]) for b in ref.dncnn if isinstance(b, torch.nn.BatchNorm2d)])
    nbt_ref = sum(int(b.num_batches_tracked) for b in ref.dncnn if isinstance(b, torch.nn.BatchNorm2d))
    assert _rel(m._running, rs) <= 1e-2 and int(m._nbt.sum()) == nbt_ref       # one train-mode forward each
    # 5-D video input (net.py:25-37) and the input-shape contract
    with torch.no_grad():
        v = net(x.to(cuda)[None])
    assert v.shape == (1, B, 1, H, W)
    with pytest.raises(ValueError):
        net(torch.zeros(1, 3, 8, 8, device=cuda))


@pytest.mark.parametrize("H,W", [(5, 7), (3, 130), (130, 3), (1, 40), (1, 7), (2, 40), (17, 16), (16, 129)])
def test_degenerate_frame_shapes(cuda, H, W):
    """Frames smaller than one tile, one pixel high, one tile plus one pixel: the reference takes any H x W
    (3x3 convolutions with padding 1, net.py:13-19); so must the tiled kernels (zero fill outside the frame,
    partial tiles masked out of the BatchNorm sums and the weight gradient).  BatchNorm over a few dozen pixels is
    badly conditioned -- for one-row frames the CPU emulation of the 16-bit data flow (oracle/dncnn_bf16.py) is
    itself 18-37 % away from the FP32 oracle's gradients -- so the bar is relative to what that data flow allows:
    at most twice the emulation's own distance from the FP32 oracle, plus a small absolute term."""
    from frame2frame_b200 import _lib, synth
    from oracle import dncnn_bf16 as emu
    from oracle import dncnn_oracle as do
    net, ref = _models(cuda, True)
    x = torch.from_numpy((synth.noisy_frame(H, W, 0) / 255.0)[None, None].astype(np.float32))
    dy = torch.from_numpy(np.random.default_rng(1).standard_normal(x.shape).astype(np.float32))
    net.eval(); ref.eval()
    with torch.no_grad():
        assert _rel(net(x.to(cuda)), ref(x)) <= 3e-2
    net.train(); ref.train()
    y, yr = net(x.to(cuda)), ref(x)
    (y * dy.to(cuda)).sum().backward()
    (yr * dy).sum().backward()
    emu.STORAGE = _lib.storage_dtype()
    flat = torch.cat([p.detach().reshape(-1) for p in ref.parameters()])
    e = emu.forward_backward(flat, x, dy, train=True, residual=True)
    assert _rel(y, yr) <= 2 * _rel(e["y"], yr) + 1e-3
    bound = 2 * _rel(e["grads"], do.flat_grads(ref)) + 2e-2
    if H * W < 200:
        bound = max(bound, 0.2)       # BatchNorm over < 200 values: the emulation's own distance varies 4-37 % with the weights
    assert _rel(net.module.flat_grads, do.flat_grads(ref)) <= bound


def test_cta_pair_convolutions_are_bit_identical(cuda):
    """The opt-in cta_group::2 variant of the 64->64 convolution kernels (f2f_set_conv_pair) against the default
    single-CTA kernels: same outputs bit for bit, forward (4 variants) and input gradient (2), including grids with
    an odd number of tiles and frames smaller than a tile."""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "check_pair.py"), "quick", "notime"], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0 and "ALL OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_reference_golden_finetune(cuda):
    """The reference's own loop (3 iterations, seeded random-init weights) from tests/golden/dncnn_ref.npz."""
    from frame2frame_b200 import synth
    from frame2frame_b200.blind_denoising import OnlineF2F
    from frame2frame_b200.dncnn import DnCNN
    from oracle import f2f_ref
    H, W = (int(v) for v in GOLD["ft_hw"])
    curr, prev, flow = f2f_ref.step_inputs(H, W)
    net = torch.nn.DataParallel(DnCNN(residual=False), device_ids=[0]).cuda()
    net.load_state_dict({k: torch.from_numpy(v) for k, v in synth.dncnn_state_dict(seed=0).items()})
    eng = OnlineF2F(net)
    losses = []
    den = eng.step_frame(torch.from_numpy(curr)[None, None], torch.from_numpy(prev)[None, None],
                         torch.from_numpy(flow)[None], 3, losses)
    # seeded RANDOM-init weights: a chaotic 17-layer stack, BF16 rounding differences grow with every update
    assert np.allclose(losses, GOLD["ft_losses"], rtol=6e-2)
    assert _rel(den[0, 0], torch.from_numpy(GOLD["ft_den"])) <= 0.15      # random-init 17-layer stack: chaotic


def test_autograd_path_equals_fused_step(cuda):
    """model(x) / WarpedLoss / backward / optimizer.step() written like blind_denoising.py:216-222
    gives the same parameters as the one-call f2f_step_train."""
    from frame2frame_b200 import synth
    from frame2frame_b200.blind_denoising import BETAS, EPS, LR, WEIGHT_DECAY, OnlineF2F
    from frame2frame_b200.dncnn import Adam
    from frame2frame_b200.warped_loss import WarpedLoss
    H, W = 56, 72
    f = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255))[None, None].to(cuda) for t in range(2)]
    flow = torch.from_numpy(synth.backward_flow(H, W, 1))[None]
    a, _ = _models(cuda, False)
    b, _ = _models(cuda, False)
    eng = OnlineF2F(a)
    la = []
    eng.finetune(f[1], f[0], flow, 3, la)
    crit = WarpedLoss().cuda()
    opt = Adam(b.parameters(), lr=LR, betas=BETAS, eps=EPS, weight_decay=WEIGHT_DECAY, amsgrad=False)
    b.train()
    opt.zero_grad()
    lb = []
    for _ in range(3):
        out_train = f[1] - b(f[1])
        loss = crit(out_train, f[0], flow)
        opt.zero_grad()
        loss.backward()
        opt.step()
        lb.append(float(loss.item()))
    assert np.allclose(la, lb, rtol=1e-5)
    assert _rel(b.module.flat_params, a.module.flat_params) <= 1e-6
    assert _rel(b.module._running, a.module._running) <= 1e-6


def test_online_f2f_psnr_parity(cuda):
    """Online fine-tuning (3 frames x 20 iterations from the pretrained weights) against the FP32 CPU oracle loop.

    Bars asserted here: the loss of EVERY inner iteration within 2 % of the oracle's; denoised PSNR within 0.25 dB on
    the sequence average and 0.5 dB per frame.  north_star asks for 0.05 dB; what was measured on B200
    (tools/parity_f2f.py, profiles/r01_parity_*.json): 1080p frame +0.018 dB, 540x960 frames -0.13 / -0.15 / -0.06 dB,
    108x192 frames up to +-0.35 dB.  The loop is sensitive to the arithmetic: sign() loss gradients and Adam's
    normalised steps amplify differences -- two runs of the FP32 oracle itself that differ only in the number of CPU
    threads differ by 0.003-0.03 dB at 540x960 (tools/chaos_floor.py, profiles/r01_reference_reproducibility_540.json)
    and an FP32 restatement with another summation order by 0.12 dB at this test's 108x192 (tools/exp_bf16_parity.py
    "none").  16-bit tensor-core operands (BF16 here; TF32 in the reference's own cuDNN path has FP16's mantissa) sit
    5-10x above that floor.  So the 0.05 dB bar holds at the metric's resolution on the frame measured, not at the
    small sizes a test can afford; DESIGN.md section 2 carries the table."""
    from frame2frame_b200 import synth
    from frame2frame_b200.blind_denoising import OnlineF2F, psnr
    from oracle import dncnn_oracle as do
    H, W, NF, ITERS = 108, 192, 3, 20
    net, ref = _models(cuda, False)
    opt = do.make_adam(ref)
    eng = OnlineF2F(net)
    frames = [synth.noisy_frame(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
    clean = [synth.clean_frame_u8(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
    diffs = []
    for t in range(1, NF + 1):
        flow = synth.backward_flow(H, W, t)
        lr, lg = [], []
        den_ref = do.finetune_frame(ref, opt, frames[t], frames[t - 1], flow, ITERS, lr)
        den = eng.step_frame(torch.from_numpy(frames[t])[None, None], torch.from_numpy(frames[t - 1])[None, None],
                             torch.from_numpy(flow)[None], ITERS, lg)
        assert np.allclose(lg, lr, rtol=2e-2)
        diffs.append(psnr(clean[t], den[0, 0].cpu().numpy()) - psnr(clean[t], den_ref))
    assert abs(np.mean(diffs)) <= 0.25, diffs
    assert np.abs(diffs).max() <= 0.5, diffs


def test_run_sequence_equals_frame_by_frame(cuda):
    """OnlineF2F.run_sequence (backward flows of `lookahead` consecutive pairs in one batched TV-L1 call) gives
    bit-identical denoised frames and parameters to calling step_frame pair by pair (one TV-L1 call per pair)."""
    from frame2frame_b200 import synth
    from frame2frame_b200.blind_denoising import OnlineF2F
    H, W, NF = 72, 104, 4
    frames = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255))[None, None] for t in range(NF + 1)]
    a, _ = _models(cuda, False)
    b, _ = _models(cuda, False)
    ea, eb = OnlineF2F(a), OnlineF2F(b)
    kw = dict(nscales=3)
    seq = {i: den.clone() for i, den in ea.run_sequence(frames, iters=2, lookahead=3, flow_kw=kw)}
    assert sorted(seq) == list(range(1, NF + 1))
    for t in range(1, NF + 1):
        den = eb.step_frame(frames[t].to(cuda), frames[t - 1].to(cuda), None, 2, flow_kw=kw)
        assert torch.equal(den, seq[t]), t
    assert torch.equal(a.module.flat_params, b.module.flat_params)


def test_blind_denoising_cli(cuda, tmp_path):
    """The reference's command line (blind_denoising.py:262-298) on PNG frames: denoised PNGs, one PSNR per
    line, [model, optimizer] checkpoint; with --flow files (.flo, readFlowFile.py layout) and without (TV-L1 on
    the device) the outputs exist; the .flo files hold the TV-L1 flows of the same frames (what tvl1flow.sh would
    have written), so both routes see the same flows and their PSNRs agree.  (Analytic flows in the files are no
    yardstick here: on 64x96 frames at sigma = 25 the estimated flow differs from the true one by enough to move a
    2-iteration fine-tune by 1-2 dB.)"""
    from PIL import Image
    from frame2frame_b200 import blind_denoising as bd, synth
    from frame2frame_b200.tvl1 import tvl1flow, write_flo
    from frame2frame_b200.tvl1flow_cli import SCRIPT_PARAMS
    H, W, NF = 64, 96, 3
    kw = dict(SCRIPT_PARAMS, fscale=0, nscales=3)
    up = lambda t: torch.from_numpy(synth.noisy_frame(H, W, t).astype(np.uint8).astype(np.float32) / np.float32(255.0)).to(cuda) * 255.0
    sd = {k: torch.from_numpy(v) for k, v in synth.pretrained_state_dict().items()}
    torch.save(sd, str(tmp_path / "net.pth"))
    for t in range(NF + 1):
        Image.fromarray(synth.noisy_frame(H, W, t).astype(np.uint8)).save(str(tmp_path / ("in_%03d.png" % t)))
        Image.fromarray(synth.clean_frame_u8(H, W, t).astype(np.uint8)).save(str(tmp_path / ("ref_%03d.png" % t)))
        if t:
            write_flo(str(tmp_path / ("flow_%03d.flo" % t)), tvl1flow(up(t), up(t - 1), **kw).cpu().numpy())
    common = ["--input", str(tmp_path / "in_%03d.png"), "--ref", str(tmp_path / "ref_%03d.png"), "--first", "0",
              "--last", str(NF), "--iter", "2", "--network", str(tmp_path / "net.pth")]
    res = {}
    for name, extra in (("files", ["--flow", str(tmp_path / "flow_%03d.flo")]), ("tvl1", ["--lookahead", "2", "--flow_fscale", "0", "--flow_nscales", "3"])):
        out = tmp_path / name
        out.mkdir()
        bd.main(common + extra + ["--output", str(out / "%03d.png"), "--output_psnr", str(out / "psnr.txt"),
                                  "--output_network", str(out / "final.pth")])
        res[name] = [float(l) for l in open(out / "psnr.txt").read().split()]
        assert len(res[name]) == NF and all(20.0 < p < 50.0 for p in res[name])
        for t in range(1, NF + 1):
            im = np.array(Image.open(out / ("%03d.png" % t)))
            assert im.shape == (H, W) and im.dtype == np.uint8
        model, opt = torch.load(str(out / "final.pth"), weights_only=False)
        assert any(k.startswith("module.dncnn.0.") for k in model.state_dict())
    assert np.allclose(res["files"], res["tvl1"], atol=0.02), res


def test_cli_on_second_gpu_equals_first(cuda, tmp_path):
    """`--gpu 1` must give what `--gpu 0` gives (the result copy and its `ready` event used to go to
    different devices' streams, so the writer thread could encode a pinned buffer the copy had not reached yet).
    960x540 frames so that the D2H copy is long enough to lose that race; BatchNorm sums are FP64 atomics (order
    varies run to run), hence a tolerance of one grey level / 1e-3 dB instead of bit equality."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from PIL import Image
    from frame2frame_b200 import blind_denoising as bd, synth
    H, W, NF = 540, 960, 3
    torch.save({k: torch.from_numpy(v) for k, v in synth.pretrained_state_dict().items()}, str(tmp_path / "net.pth"))
    for t in range(NF + 1):
        Image.fromarray(synth.noisy_frame(H, W, t).astype(np.uint8)).save(str(tmp_path / ("in_%03d.png" % t)))
        Image.fromarray(synth.clean_frame_u8(H, W, t).astype(np.uint8)).save(str(tmp_path / ("ref_%03d.png" % t)))
    res = {}
    for gpu in (0, 1):
        out = tmp_path / ("gpu%d" % gpu)
        out.mkdir()
        bd.main(["--input", str(tmp_path / "in_%03d.png"), "--ref", str(tmp_path / "ref_%03d.png"), "--first", "0", "--last", str(NF),
                 "--iter", "3", "--network", str(tmp_path / "net.pth"), "--gpu", str(gpu), "--flow_nscales", "4",
                 "--output", str(out / "%03d.png"), "--output_psnr", str(out / "psnr.txt"), "--output_network", str(out / "final.pth")])
        res[gpu] = ([float(l) for l in open(out / "psnr.txt").read().split()],
                    [np.array(Image.open(out / ("%03d.png" % t))).astype(np.int32) for t in range(1, NF + 1)])
    torch.cuda.set_device(0)
    assert np.allclose(res[0][0], res[1][0], atol=1e-3), (res[0][0], res[1][0])
    for a, b in zip(res[0][1], res[1][1]):
        assert np.abs(a - b).max() <= 1


def test_checkpoint_resume_round_trip(cuda, tmp_path):
    """torch.save([model, optimizer]) (blind_denoising.py:258) -> torch.load -> keep training: the reloaded pair must
    continue exactly like the pair that was never saved (flat Adam moments and step count rebuilt from the pickled
    per-parameter state), and load_network() must accept such a file as --network (blind_denoising.py:148-151)."""
    from frame2frame_b200 import synth
    from frame2frame_b200.blind_denoising import OnlineF2F, load_network
    H, W = 48, 64
    fr = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255.0))[None, None] for t in range(3)]
    fl = [None] + [torch.from_numpy(synth.backward_flow(H, W, t))[None] for t in (1, 2)]
    net, _ = _models(cuda, False)
    eng = OnlineF2F(net)
    eng.finetune(fr[1], fr[0], fl[1], 3)
    fn = str(tmp_path / "final.pth")
    torch.save([eng.model, eng.optimizer], fn)
    model2, opt2 = torch.load(fn, weights_only=False)
    eng2 = OnlineF2F(model2, optimizer=opt2)
    assert torch.equal(eng2.net.flat_params, eng.net.flat_params) and torch.equal(eng2.net._running, eng.net._running)
    la, lb = [], []
    eng.finetune(fr[2], fr[1], fl[2], 3, la)
    eng2.finetune(fr[2], fr[1], fl[2], 3, lb)
    assert eng2.optimizer._flat_state[0]["step"] == eng.optimizer._flat_state[0]["step"] == 6
    assert np.allclose(la, lb, rtol=1e-6)
    assert _rel(eng2.net.flat_params, eng.net.flat_params) <= 1e-6
    m3 = load_network(fn, cuda)                 # a saved [model, optimizer] file as --network
    assert _rel(m3.module.flat_params, model2.module.flat_params) <= 3e-3   # model2 trained 3 more iterations since the save


def test_config0_sequence_against_golden_oracle(cuda):
    """BASELINE.json configs[0] (the reference's own CPU-runnable case): 960x540 gray sigma = 25 synthetic sequence with its
    analytic flows, DnCNN-25 layout, iter = 20, in the DEFAULT mode (16-bit tensor-core operands) against the recorded
    FP32 CPU oracle loop (tests/golden/f2f_config0_ref.json, tools/make_f2f_golden.py).  Bars -- what the default build
    meets, r02: first / last loss of every frame within 5e-4; |PSNR difference| <= 0.05 dB (north_star) on the first four
    frames, i.e. as long as two runs of the reference itself agree to 0.025 dB (tests/golden/f2f_config0_floor.json);
    afterwards within 0.2 dB per frame and 0.06 dB on the sequence's mean |difference| (measured 0.133 / 0.035; the FP32
    modes and the reference against itself reach 0.10 / 0.09 on the same frames: tests/test_precise_gpu.py)."""
    from test_precise_gpu import _run_config0
    diffs, ldev = _run_config0(cuda)
    print("default mode, PSNR - oracle per frame (dB):", " ".join("%+.4f" % d for d in diffs))
    assert max(ldev) <= 5e-4, ldev
    assert all(abs(d) <= 0.05 for d in diffs[:4]), diffs
    assert np.abs(diffs).max() <= 0.2 and np.mean(np.abs(diffs)) <= 0.06, diffs


def test_backward_is_linear_in_dy_down_to_tiny_gradients(cuda):
    """Size-independent property of the backward pass: parameter gradients are linear in the incoming gradient.
    With 16-bit (FP16-range) intermediates this holds only because dy is rescaled by a power of two on entry
    (dncnn_net.cu grad_scale): a mean-reduced loss at 1080p hands in |dy| ~ 5e-7, far below FP16's normal range.
    c = 2^-24 must reproduce c times the gradients to FP32 rounding (the rescaled dy is the same tensor; not bit for
    bit, because the FP64 atomics of the BatchNorm sums arrive in a different order from run to run) and c = 3e-7 to
    16-bit rounding (3e-3: its rescaled dy has other mantissas); without the rescaling the tiny gradients flush to
    zero in FP16 and the relative error is 1."""
    from frame2frame_b200 import synth
    H, W = 40, 56
    net, _ = _models(cuda, True)
    net.train()
    x = torch.from_numpy(synth.noisy_frame(H, W, 0) / np.float32(255.0))[None, None].to(cuda)
    dy = torch.from_numpy(np.random.default_rng(1).standard_normal((1, 1, H, W)).astype(np.float32)).to(cuda)
    grads = []
    for c in (1.0, 2.0 ** -24, 3e-7):
        net.zero_grad()
        y = net(x)
        (y * (dy * c)).sum().backward()
        grads.append(net.module.flat_grads.detach().clone())
    g1, g2, g3 = grads
    assert float(g1.abs().max()) > 0
    assert _rel(g2, g1 * 2.0 ** -24) <= 1e-5
    assert _rel(g3, g1 * 3e-7) <= 3e-3
