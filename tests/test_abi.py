"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every
symbol include/f2f_b200.h declares; host-side parameter logic (no compute)."""
import os
import ctypes as C

import numpy as np

from frame2frame_b200 import _lib
from frame2frame_b200.tvl1 import make_params, read_flo, write_flo
from oracle.tvl1 import clamp_scales


def test_library_exports_every_declared_symbol():
    lib = _lib.lib()
    names = _lib.exported_symbols()
    assert "f2f_tvl1_flow_batch" in names and "Dual_TVL1_optic_flow_multiscale" in names
    for n in names:
        assert hasattr(lib, n), "missing export: " + n
    assert lib.f2f_built_sm() == 100


def test_default_params_are_main_c_defaults():
    p = _lib.Tvl1Params()
    _lib.lib().f2f_tvl1_default_params(C.byref(p))
    assert (round(p.tau, 6), round(p.lam, 6), round(p.theta, 6), p.nscales, p.fscale, round(p.zfactor, 6), p.nwarps,
            round(p.epsilon, 6), p.verbose) == (0.25, 0.15, 0.3, 100, 0, 0.5, 5, 0.01, 0)


def test_validate_resets_and_clamps_like_main_c():
    p = make_params(1920, 1080, tau=0.5, lam=-1, theta=0, nscales=0, zfactor=1.5, nwarps=0, epsilon=-3)
    assert (round(p.tau, 6), round(p.lam, 6), round(p.theta, 6), round(p.zfactor, 6), p.nwarps) == (0.25, 0.15, 0.3, 0.5, 5)
    for (nx, ny, z, ns, fs) in [(1920, 1080, 0.5, 100, 0), (3840, 2160, 0.5, 100, 2), (960, 540, 0.5, 5, 0),
                                (120, 68, 0.5, 100, 9), (640, 480, 0.75, 100, 0)]:
        p = make_params(nx, ny, zfactor=z, nscales=ns, fscale=fs)
        assert (p.nscales, p.fscale) == clamp_scales(nx, ny, z, ns, fs)


def test_workspace_bytes_and_bad_args():
    lib = _lib.lib()
    p = make_params(1920, 1080, nscales=5)
    b1 = lib.f2f_tvl1_workspace_bytes(1920, 1080, C.byref(p), 1)
    b4 = lib.f2f_tvl1_workspace_bytes(1920, 1080, C.byref(p), 4)
    assert 0 < b1 < b4 < 4 * b1 + 4096
    assert lib.f2f_tvl1_workspace_bytes(2, 2, C.byref(p), 1) == 0
    assert lib.f2f_tvl1_workspace_bytes(1920, 1080, C.byref(p), 0) == 0


def test_flo_roundtrip_matches_reference_layout(tmp_path):
    # readFlowFile.py:18-34: float32 tag 202021.25, int32 w, int32 h, h*w*2 float32 interleaved
    flow = np.arange(2 * 3 * 5, dtype=np.float32).reshape(2, 3, 5)
    path = str(tmp_path / "x.flo")
    write_flo(path, flow)
    raw = open(path, "rb").read()
    assert len(raw) == 12 + 2 * 3 * 5 * 4
    assert np.frombuffer(raw[:4], np.float32)[0] == np.float32(202021.25)
    assert tuple(np.frombuffer(raw[4:12], np.int32)) == (5, 3)
    back = read_flo(path)
    assert back.shape == (3, 5, 2)
    assert np.array_equal(back.transpose(2, 0, 1), flow)


def test_product_package_never_imports_the_oracle():
    """oracle/ is the checker: nothing under frame2frame_b200/ may import or execute it (no CPU fallback)."""
    import os
    import re
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "frame2frame_b200")
    bad = []
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(root, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M) or "oracle/_ref" in text or "liboracle" in text:
                    bad.append(os.path.join(root, f))
    assert not bad, bad


def test_tvl1flow_cli_positionals_follow_main_c():
    """main.c:25-35,86-99: 13 positionals, the same defaults; a missing image pair is a usage error."""
    import pytest
    from frame2frame_b200 import tvl1flow_cli
    v = tvl1flow_cli.parse_positionals(["a.png", "b.png"])
    assert (v["out"], v["nproc"], v["tau"], v["lam"], v["theta"], v["nscales"], v["fscale"], v["zfactor"], v["nwarps"],
            v["epsilon"], v["verbose"]) == ("flow.flo", 4, 0.25, 0.15, 0.3, 100, 0, 0.5, 5, 0.01, 0)
    v = tvl1flow_cli.parse_positionals("a b out.flo 8 0.2 0.1 0.4 5 1 0.6 3 0.02 1".split())
    assert (v["out"], v["nproc"], v["tau"], v["lam"], v["theta"], v["nscales"], v["fscale"], v["zfactor"], v["nwarps"],
            v["epsilon"], v["verbose"]) == ("out.flo", 8, 0.2, 0.1, 0.4, 5, 1, 0.6, 3, 0.02, 1)
    with pytest.raises(SystemExit):
        tvl1flow_cli.parse_positionals(["only_one.png"])


def test_tvl1flow_cli_sequence_follows_the_script(tmp_path, monkeypatch):
    """tvl1flow.sh:9-19 -- pair order, file names and parameters of the sequence mode; the solver is replaced by a
    recorder, so this runs without a GPU."""
    import numpy as np
    import torch
    from frame2frame_b200 import tvl1flow_cli as cli
    assert cli.plan_sequence(3, 4, "o/%03d") == [(4, 3, "o/004_bflow.flo"), (3, 4, "o/003_fflow.flo"),
                                                 (5, 4, "o/005_bflow.flo"), (4, 5, "o/004_fflow.flo")]
    for i in range(1, 4):
        np.save(tmp_path / ("f%03d.npy" % i), np.full((6, 8), float(i), np.float32))
    calls, written = [], {}

    def fake_bidir(A, B, **kw):
        # both directions of every pair in one pass (shared pyramids): flow A -> B tagged 10k, B -> A tagged 10k + 1
        calls.append((A[:, 0, 0].tolist(), B[:, 0, 0].tolist(), kw))
        n = A.shape[0]
        return (torch.stack([torch.full((2, 6, 8), 10.0 * k) for k in range(n)]),
                torch.stack([torch.full((2, 6, 8), 10.0 * k + 1) for k in range(n)]))

    monkeypatch.setattr(cli, "tvl1flow_bidir", fake_bidir)
    monkeypatch.setattr(cli, "tvl1flow", lambda *a, **k: (_ for _ in ()).throw(AssertionError("whole pairs go through bidir")))
    monkeypatch.setattr(cli, "write_flo", lambda name, f: written.__setitem__(name, float(f[0, 0, 0])))
    out = str(tmp_path / "o%03d")
    assert cli.main(["--sequence", str(tmp_path / "f%03d.npy"), "1", "2", out], device="cpu") == 0
    (a, b, kw), = calls
    assert a == [2.0, 3.0] and b == [1.0, 2.0]          # pair i: A = frame i+1, B = frame i; A -> B is the bflow, B -> A the fflow
    assert kw == dict(tau=0.25, lam=0.2, theta=0.3, nscales=100, fscale=2, zfactor=0.5, nwarps=5, epsilon=0.01)
    assert written == {out % 2 + "_bflow.flo": 0.0, out % 1 + "_fflow.flo": 1.0, out % 3 + "_bflow.flo": 10.0,
                       out % 2 + "_fflow.flo": 11.0}


def test_missing_library_fails_loudly_no_cpu_fallback():
    """The product path has no CPU fallback: without libf2f_b200.so the first call raises."""
    import subprocess
    import sys
    code = ("from frame2frame_b200 import _lib\n"
            "try:\n    _lib.lib()\nexcept RuntimeError as e:\n    print('RAISED', e)\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, F2F_LIB="/nonexistent/libf2f_b200.so"),
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert "RAISED" in r.stdout and "no CPU fallback" in r.stdout, r.stdout + r.stderr


def test_cuda_ops_refuse_cpu_tensors():
    """warp / DnCNN on CPU tensors raise instead of computing on the host."""
    import pytest
    import torch
    from frame2frame_b200.warped_loss import warp_mask
    with pytest.raises((RuntimeError, ValueError)):
        warp_mask(torch.zeros(1, 1, 8, 8), torch.zeros(1, 2, 8, 8))


def test_tvl1flow_cli_sequence_shards_pairs_over_ranks(tmp_path, monkeypatch):
    """Under torchrun (RANK / WORLD_SIZE) the sequence mode splits the pair list into contiguous blocks, one per rank;
    together the ranks write every file of the single-process run exactly once."""
    import numpy as np
    import torch
    from frame2frame_b200 import tvl1flow_cli as cli
    jobs = cli.plan_sequence(1, 5, "o%03d")
    assert sum((cli.shard_jobs(jobs, r, 4) for r in range(4)), []) == jobs
    assert [len(cli.shard_jobs(jobs, r, 4)) for r in range(4)] == [3, 3, 2, 2]
    assert cli.shard_jobs(jobs[:1], 1, 2) == []
    for i in range(1, 5):
        np.save(tmp_path / ("f%03d.npy" % i), np.full((4, 4), float(i), np.float32))
    written = []
    monkeypatch.setattr(cli, "tvl1flow", lambda I0, I1, **kw: torch.zeros(I0.shape[0], 2, 4, 4))
    monkeypatch.setattr(cli, "tvl1flow_bidir", lambda A, B, **kw: (torch.zeros(A.shape[0], 2, 4, 4), torch.zeros(A.shape[0], 2, 4, 4)))
    monkeypatch.setattr(cli, "write_flo", lambda name, f: written.append(os.path.basename(name)))
    for rank in range(3):
        monkeypatch.setenv("RANK", str(rank)); monkeypatch.setenv("WORLD_SIZE", "3")
        assert cli.main(["--sequence", str(tmp_path / "f%03d.npy"), "1", "3", str(tmp_path / "o%03d")], device="cpu") == 0
    # every file exactly once; within a rank whole pairs (shared pyramids) are written before a block's odd field
    assert sorted(written) == sorted(["o002_bflow.flo", "o001_fflow.flo", "o003_bflow.flo", "o002_fflow.flo", "o004_bflow.flo",
                                      "o003_fflow.flo"])
    # a 4-rank split of 5 frames cuts pairs in the middle: the halves still come out, through the single-direction path
    written.clear()
    for rank in range(4):
        monkeypatch.setenv("RANK", str(rank)); monkeypatch.setenv("WORLD_SIZE", "4")
        assert cli.main(["--sequence", str(tmp_path / "f%03d.npy"), "1", "3", str(tmp_path / "o%03d")], device="cpu") == 0
    assert len(written) == 6 and len(set(written)) == 6
