"""Sharding of the online loop over GPUs: independent --first/--last chunks (blind_denoising.py:187,279-282).
Host logic only -- the per-chunk run is replaced by a recorder."""
import os

import pytest

from frame2frame_b200.blind_denoising import chunk_bounds, run_chunked


@pytest.mark.parametrize("first,last,world", [(1, 120, 8), (1, 10, 3), (5, 9, 8), (0, 1, 2), (1, 300, 1), (7, 7, 4)])
def test_chunks_cover_every_denoised_frame_once(first, last, world):
    covered = []
    for r in range(world):
        b = chunk_bounds(first, last, r, world)
        if b is None:
            continue
        f, l = b
        assert first <= f < l <= last
        covered += list(range(f + 1, l + 1))            # a chunk denoises f+1..l; f is only its first `prev`
    assert covered == list(range(first + 1, last + 1))  # in rank order = in frame order, no gap, no overlap
    sizes = [(lambda b: 0 if b is None else b[1] - b[0])(chunk_bounds(first, last, r, world)) for r in range(world)]
    assert max(sizes) - min(sizes) <= 1


def test_configs4_is_eight_chunks_of_fifteen():
    # BASELINE.json configs[4]: 120 denoised frames over 8 GPUs -> 15 each, + 1 overlap frame as the first prev
    assert [chunk_bounds(0, 120, r, 8) for r in range(8)] == [(15 * r, 15 * r + 15) for r in range(8)]


def test_run_chunked_assembles_outputs_in_frame_order(tmp_path):
    args = {"first": 1, "last": 11, "output_psnr": str(tmp_path / "plot_psnr.txt"), "output_network": str(tmp_path / "final.pth"),
            "input": "in/%03d.png", "iter": 20, "gpu": 0}
    calls = []

    def fake_run(**a):
        calls.append((a["first"], a["last"], a["gpu"], os.path.basename(a["output_network"])))
        with open(a["output_psnr"], "w") as f:
            for i in range(a["first"] + 1, a["last"] + 1):
                f.write("%d\n" % i)                        # one line per denoised frame, like plot_psnr.txt

    world = 4
    for rank in (3, 1, 2, 0):                              # rank 0 assembles after the barrier, i.e. last
        run_chunked(args, rank, world, run=fake_run, local_rank=rank)
    assert sorted(calls) == [(1, 4, 0, "final.chunk000.pth"), (4, 7, 1, "final.chunk001.pth"), (7, 9, 2, "final.chunk002.pth"),
                             (9, 11, 3, "final.chunk003.pth")]
    assert open(args["output_psnr"]).read().split() == [str(i) for i in range(2, 12)]
    assert sorted(os.listdir(tmp_path)) == ["plot_psnr.txt"]   # part files are gone
    assert all(a not in ("first", "last") or True for a in args) and args["first"] == 1   # caller's args untouched


def test_run_chunked_with_more_ranks_than_frames(tmp_path):
    args = {"first": 3, "last": 5, "output_psnr": str(tmp_path / "p.txt"), "output_network": str(tmp_path / "n.pth")}
    ran = []

    def fake_run(**a):
        ran.append((a["first"], a["last"]))
        open(a["output_psnr"], "w").write("x\n" * (a["last"] - a["first"]))

    for rank in (3, 2, 1, 0):
        run_chunked(args, rank, 4, run=fake_run)
    assert sorted(ran) == [(3, 4), (4, 5)] and open(args["output_psnr"]).read() == "x\nx\n"


def test_cli_under_two_ranks_gloo(tmp_path):
    """The CLI as torchrun would start it (world size 2, gloo barrier): every rank runs its chunk, rank 0 assembles the
    PSNR file.  The per-chunk run is replaced inside the child processes, so no GPU is needed."""
    import socket
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    child = (
        "import sys; sys.path.insert(0, %r)\n"
        "from frame2frame_b200 import blind_denoising as bd\n"
        "def fake(**a):\n"
        "    open(a['output_psnr'], 'w').write(''.join('%%d %%d\\n' %% (i, a['gpu']) for i in range(a['first'] + 1, a['last'] + 1)))\n"
        "bd.blind_denoising = fake\n"
        "bd.main(['--first', '1', '--last', '8', '--output_psnr', %r, '--output_network', %r])\n"
    ) % (root, str(tmp_path / "psnr.txt"), str(tmp_path / "final.pth"))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, "-c", child], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    for p in procs:
        out, err = p.communicate(timeout=300)
        assert p.returncode == 0, err[-2000:]
    lines = open(tmp_path / "psnr.txt").read().split("\n")[:-1]
    assert lines == ["2 0", "3 0", "4 0", "5 0", "6 1", "7 1", "8 1"]      # frames 2..5 on GPU 0, 6..8 on GPU 1
