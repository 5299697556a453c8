"""GPU parity tests: CUDA TV-L1 (through the C ABI) against the oracle.

Bar (BASELINE.json north_star): flow within 1e-3 px max-abs of the OpenMP
reference and -- because one extra/missing inner iteration already moves the
flow by ~0.01 px -- identical per-(scale, warp) iteration counts.
"""
import ast
import os

import numpy as np
import pytest
import torch

from frame2frame_b200 import _lib, synth
from frame2frame_b200.tvl1 import tvl1flow, tvl1flow_host
from oracle.tvl1 import OracleTvl1

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "tvl1_ref.npz"))
FLOW_TOL = 1e-3  # px, north_star


def dev(a, cuda):
    return torch.from_numpy(np.ascontiguousarray(a)).to(cuda)


def _stage(cuda, fn, *arrs):
    return [dev(a, cuda) for a in arrs]


def test_stage_kernels_bit_exact(cuda):
    lib = _lib.lib()
    o = OracleTvl1()
    for (H, W) in [(33, 47), (135, 240)]:
        I0, I1 = synth.noisy_frame(H, W, 3), synth.noisy_frame(H, W, 2)
        # gaussian, mask.c:223-339
        d = dev(I0, cuda)
        tmp = torch.empty_like(d)
        _lib.check(lib.f2f_tvl1_gaussian(d.data_ptr(), W, H, 0.8, tmp.data_ptr(), None), "gaussian")
        assert np.array_equal(d.cpu().numpy(), o.gaussian(I0, 0.8))
        # zoom_out, zoom.c:43-81
        for f in (0.5, 0.7):
            ref = o.zoom_out(I0, f)
            out = torch.empty(ref.shape, device=cuda)
            tmp2 = torch.empty(2 * H * W, device=cuda)
            src = dev(I0, cuda)   # keep a reference: a temporary would be freed before the kernel runs
            _lib.check(lib.f2f_tvl1_zoom_out(src.data_ptr(), out.data_ptr(), W, H, f, tmp2.data_ptr(), None), "zoom_out")
            assert np.array_equal(out.cpu().numpy(), ref)
        # zoom_in + rescale
        z = o.zoom_out(I0, 0.5)
        out = torch.empty((H, W), device=cuda)
        dz = dev(z, cuda)
        _lib.check(lib.f2f_tvl1_zoom_in_scaled(dz.data_ptr(), out.data_ptr(), z.shape[1], z.shape[0], W, H, 0.5, None), "zoom_in")
        assert np.array_equal(out.cpu().numpy(), o.zoom_in_scaled(z, W, H, 0.5))
        # centered gradient
        dx, dy = torch.empty((H, W), device=cuda), torch.empty((H, W), device=cuda)
        src = dev(I0, cuda)
        _lib.check(lib.f2f_tvl1_centered_gradient(src.data_ptr(), dx.data_ptr(), dy.data_ptr(), W, H, None), "cgrad")
        rx, ry = o.centered_gradient(I0)
        assert np.array_equal(dx.cpu().numpy(), rx) and np.array_equal(dy.cpu().numpy(), ry)
        # normalisation (arbitrary value range, negative minimum)
        a, b = I0 * 0.3 - 7, I1 * 0.3 - 7
        oa, ob = torch.empty((H, W), device=cuda), torch.empty((H, W), device=cuda)
        scratch = torch.empty(8, dtype=torch.uint8, device=cuda)
        da, db = dev(a, cuda), dev(b, cuda)
        _lib.check(lib.f2f_tvl1_normalize(da.data_ptr(), db.data_ptr(), oa.data_ptr(), ob.data_ptr(), W, H, scratch.data_ptr(), None), "normalize")
        ra, rb = o.normalize(a, b)
        assert np.array_equal(oa.cpu().numpy(), ra) and np.array_equal(ob.cpu().numpy(), rb)


@pytest.mark.parametrize("case", ["a", "b", "c"])
def test_flow_matches_reference_golden(cuda, case):
    H, W = GOLD[case + "_hw"]
    kw = ast.literal_eval(str(GOLD[case + "_kw"]))
    I0, I1 = synth.noisy_frame(H, W, 1), synth.noisy_frame(H, W, 0)
    flow, iters = tvl1flow(dev(I0, cuda), dev(I1, cuda), return_iters=True, **kw)
    assert np.array_equal(iters, GOLD[case + "_iters"]), (iters, GOLD[case + "_iters"])
    diff = np.abs(flow.cpu().numpy() - GOLD[case + "_flow"]).max()
    assert diff <= FLOW_TOL, diff


def test_batch_of_pairs_vs_oracle_and_host_api(cuda):
    # ragged batch (7 pairs, 3 slots): every pair must equal its own single-pair oracle run
    H, W, n = 90, 160, 7
    I0 = np.stack([synth.noisy_frame(H, W, t + 1, seq=t % 3) for t in range(n)])
    I1 = np.stack([synth.noisy_frame(H, W, t, seq=t % 3) for t in range(n)])
    I1[5] = I0[5]          # identical frames: zero flow
    I0[6] = 17.0           # constant pair: den == 0 branch
    I1[6] = 17.0
    flow, iters = tvl1flow(dev(I0, cuda), dev(I1, cuda), nslots=3, return_iters=True, nscales=4)
    fh, ih = tvl1flow_host(I0, I1, max_slots=4, return_iters=True, nscales=4)
    o = OracleTvl1()
    nexact = 0
    for k in range(n):
        u, it, _ = o.flow(I0[k], I1[k], nscales=4)
        assert np.array_equal(iters[k], it), (k, iters[k], it)
        assert np.array_equal(ih[k], it)
        d = np.abs(flow[k].cpu().numpy() - u).max()
        assert d <= FLOW_TOL, (k, d)
        assert np.abs(fh[k] - u).max() <= FLOW_TOL
        nexact += int(d == 0)
    assert np.all(flow[6].cpu().numpy() == 0)
    print("bit-exact pairs: %d / %d" % (nexact, n))


def test_empty_batch_and_errors(cuda):
    lib = _lib.lib()
    e = torch.empty((0, 64, 64), device=cuda)
    assert tvl1flow(e, e).shape == (0, 2, 64, 64)
    with pytest.raises(ValueError):
        tvl1flow(torch.zeros((1, 64, 64), device=cuda), torch.zeros((1, 64, 65), device=cuda))
    with pytest.raises(RuntimeError):
        tvl1flow(torch.zeros((1, 64, 64)), torch.zeros((1, 64, 64)))   # CPU tensor: no fallback


def test_1080p_properties(cuda):
    # full-size pair: symmetric inputs give mirrored flows (size-independent property) and
    # the flow of (I, I) is exactly zero; iteration counts are reported for the bench.
    H, W = 1080, 1920
    I0, I1 = synth.noisy_frame(H, W, 1), synth.noisy_frame(H, W, 0)
    pair0 = np.stack([I0, I0[:, ::-1].copy(), I0])
    pair1 = np.stack([I1, I1[:, ::-1].copy(), I0])
    flow, iters = tvl1flow(dev(pair0, cuda), dev(pair1, cuda), nslots=3, return_iters=True, nscales=5)
    f = flow.cpu().numpy()
    assert np.all(f[2] == 0)
    assert np.isfinite(f).all()
    # background moves by (1.5,-0.75): median flow from frame 1 to frame 0
    assert abs(np.median(f[0, 0]) - 1.5) < 0.5 and abs(np.median(f[0, 1]) + 0.75) < 0.5
    print("1080p iters (scale x warp):", iters[0].tolist())


def test_1080p_pair_equals_oracle(cuda):
    """One pair at the metric's frame size against the CPU oracle inside the suite (about 2-6 s of host time):
    identical per-(scale, warp) iteration counts, flow within the 1e-3 px bar (in practice bit-equal)."""
    H, W = 1080, 1920
    I0, I1 = synth.noisy_frame(H, W, 1), synth.noisy_frame(H, W, 0)
    flow, iters = tvl1flow(dev(I0, cuda), dev(I1, cuda), return_iters=True, nscales=5)
    u, it, _ = OracleTvl1().flow(I0, I1, nscales=5)
    assert np.array_equal(iters, it), (iters.tolist(), it.tolist())
    d = float(np.abs(flow.cpu().numpy() - u).max())
    assert d <= FLOW_TOL, d
    print("1080p pair: max|flow - oracle| = %g px, %d inner iterations at the finest scale" % (d, int(it[0].sum())))


def test_reference_symbol_drop_in(cuda):
    """The exported drop-in `Dual_TVL1_optic_flow_multiscale` (tvl1flow_lib.c:356-372: same name, same 15 arguments,
    host buffers, caller-allocated u1/u2, no return value) called the way main.c:176-181 calls the reference's, against
    the reference's own library when it is built (oracle/_ref) and the C restatement otherwise."""
    import ctypes as C
    from oracle.tvl1 import RefTvl1, clamp_scales
    lib = _lib.lib()
    H, W = 120, 200
    I0, I1 = synth.noisy_frame(H, W, 1), synth.noisy_frame(H, W, 0)
    fp = C.POINTER(C.c_float)
    for kw in (dict(lam=0.15, nscales=100, fscale=0), dict(lam=0.2, nscales=100, fscale=2)):
        ns, fs = clamp_scales(W, H, 0.5, kw["nscales"], kw["fscale"])          # main.c:159-163 runs before the call
        a0, a1 = I0.copy(), I1.copy()
        u = np.full((2, H, W), np.nan, np.float32)
        lib.Dual_TVL1_optic_flow_multiscale.restype = None
        lib.Dual_TVL1_optic_flow_multiscale(a0.ctypes.data_as(fp), a1.ctypes.data_as(fp), u[0].ctypes.data_as(fp),
                                            u[1].ctypes.data_as(fp), C.c_int(W), C.c_int(H), C.c_float(0.25),
                                            C.c_float(kw["lam"]), C.c_float(0.3), C.c_int(ns), C.c_int(fs), C.c_float(0.5),
                                            C.c_int(5), C.c_float(0.01), C.c_bool(False))
        assert np.array_equal(a0, I0) and np.array_equal(a1, I1)                # inputs are not modified
        ref = RefTvl1().flow(I0, I1, **kw) if RefTvl1.available() else OracleTvl1().flow(I0, I1, **kw)[0]
        assert np.abs(u - ref).max() <= FLOW_TOL, kw


def test_inline_division_is_ieee(cuda):
    # the iteration kernel divides with a refined reciprocal + two residual corrections;
    # it must agree bit-for-bit with IEEE division on the operand ranges of the solver
    import ctypes as C
    bad = C.c_ulonglong(123)
    _lib.check(_lib.lib().f2f_selftest_division(1 << 30, C.byref(bad)), "selftest")
    assert bad.value == 0


def test_inline_hypot_is_the_library_square_root(cuda):
    # the dual update's (float)hypot runs an in-line correctly rounded double square root (no slow-path call);
    # float and double results must agree bit-for-bit with sqrt() on gradient pairs of every magnitude, incl. zeros
    import ctypes as C
    bad = C.c_ulonglong(123)
    _lib.check(_lib.lib().f2f_selftest_hypot(1 << 30, C.byref(bad)), "selftest")
    assert bad.value == 0


def test_tvl1flow_command_lines(cuda, tmp_path):
    """The reference's command lines (main.c:86-99 positionals, tvl1flow.sh:9-19 sequence script) through
    frame2frame_b200.tvl1flow_cli: .flo files byte-compatible with readFlowFile.py, flows equal to the oracle's."""
    from PIL import Image
    from frame2frame_b200 import tvl1flow_cli
    from frame2frame_b200.tvl1 import read_flo
    H, W = 72, 104
    fr = [synth.noisy_frame(H, W, t) for t in range(3)]
    for t, f in enumerate(fr):
        Image.fromarray(f.astype(np.uint8)).save(str(tmp_path / ("f_%03d.png" % t)))
    o = OracleTvl1()
    # single pair, explicit positionals (nproc is accepted and ignored)
    out = str(tmp_path / "a.flo")
    assert tvl1flow_cli.main([str(tmp_path / "f_001.png"), str(tmp_path / "f_000.png"), out, "4", "0.25", "0.15", "0.3",
                              "3", "0", "0.5", "5", "0.01", "0"]) == 0
    u, _, _ = o.flow(fr[1], fr[0], nscales=3)
    got = read_flo(out)
    assert got.shape == (H, W, 2) and np.abs(got.transpose(2, 0, 1) - u).max() <= FLOW_TOL
    # the sequence script: bflow / fflow of frames 0..1 with lambda 0.2, fscale 2
    assert tvl1flow_cli.main(["--sequence", str(tmp_path / "f_%03d.png"), "0", "1", str(tmp_path / "o_%03d")]) == 0
    kw = dict(lam=0.2, nscales=100, fscale=2)
    for name, (a, b) in {"o_001_bflow.flo": (1, 0), "o_000_fflow.flo": (0, 1), "o_002_bflow.flo": (2, 1),
                         "o_001_fflow.flo": (1, 2)}.items():
        u, _, _ = o.flow(fr[a], fr[b], **kw)
        assert np.abs(read_flo(str(tmp_path / name)).transpose(2, 0, 1) - u).max() <= FLOW_TOL, name
    # size mismatch -> error return like main.c:189-193
    Image.fromarray(fr[0][:40].astype(np.uint8)).save(str(tmp_path / "small.png"))
    assert tvl1flow_cli.main([str(tmp_path / "small.png"), str(tmp_path / "f_000.png"), out]) == 1


def test_both_directions_share_their_pyramids(cuda):
    """the forward and backward flow of a frame pair (tvl1flow.sh:11-18 runs two processes) in ONE pass with
    one set of pyramids (f2f_tvl1_flow_bidir) must equal the two separate batches bit for bit, iteration counts
    included -- an odd number of pairs, fewer slots than fields, the script's parameters (lambda 0.2, fscale 2)."""
    from frame2frame_b200.tvl1 import tvl1flow_bidir
    H, W, n = 90, 160, 3
    A = dev(np.stack([synth.noisy_frame(H, W, t + 1, seq=t) for t in range(n)]), cuda)
    B = dev(np.stack([synth.noisy_frame(H, W, t, seq=t) for t in range(n)]), cuda)
    for kw in (dict(nscales=4), dict(lam=0.2, nscales=100, fscale=2)):
        fab, fba, it = tvl1flow_bidir(A, B, nslots=4, return_iters=True, **kw)
        ref_ab, it_ab = tvl1flow(A, B, return_iters=True, **kw)
        ref_ba, it_ba = tvl1flow(B, A, return_iters=True, **kw)
        assert torch.equal(fab, ref_ab) and torch.equal(fba, ref_ba)
        assert np.array_equal(it[:, 0], it_ab) and np.array_equal(it[:, 1], it_ba)
    f1, f2 = tvl1flow_bidir(A[0], B[0], nscales=3)
    assert f1.shape == (2, H, W) and torch.equal(f1, tvl1flow(A[0], B[0], nscales=3))
