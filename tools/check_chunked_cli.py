"""The blind_denoising command line under torchrun on 2 GPUs (GPU box): every rank denoises its chunk of a small
synthetic sequence; checks the assembled PSNR file, the output frames and the per-chunk networks.

    python tools/check_chunked_cli.py            # prepares inputs, starts torchrun itself, checks the outputs
"""
import os, subprocess, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from PIL import Image
from frame2frame_b200 import synth

H, W, NF, WORLD = 64, 96, 5, 2
tmp = tempfile.mkdtemp()
torch.save({k: torch.from_numpy(v) for k, v in synth.pretrained_state_dict().items()}, os.path.join(tmp, "net.pth"))
for t in range(NF + 1):
    Image.fromarray(synth.noisy_frame(H, W, t).astype(np.uint8)).save(os.path.join(tmp, "in_%03d.png" % t))
    Image.fromarray(synth.clean_frame_u8(H, W, t).astype(np.uint8)).save(os.path.join(tmp, "ref_%03d.png" % t))
cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(WORLD), "--master-addr", "127.0.0.1",
       "--master-port", "29533", "-m", "frame2frame_b200.blind_denoising", "--input", os.path.join(tmp, "in_%03d.png"),
       "--ref", os.path.join(tmp, "ref_%03d.png"), "--first", "0", "--last", str(NF), "--iter", "2", "--lookahead", "2",
       "--network", os.path.join(tmp, "net.pth"), "--output", os.path.join(tmp, "out_%03d.png"),
       "--output_psnr", os.path.join(tmp, "psnr.txt"), "--output_network", os.path.join(tmp, "final.pth")]
r = subprocess.run(cmd, capture_output=True, text=True, timeout=200, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
print(r.stdout[-600:], r.stderr[-1500:] if r.returncode else "")
assert r.returncode == 0
ps = [float(v) for v in open(os.path.join(tmp, "psnr.txt")).read().split()]
print("PSNR per frame:", ps)
assert len(ps) == NF and all(20.0 < p < 50.0 for p in ps)
for t in range(1, NF + 1):
    assert np.array(Image.open(os.path.join(tmp, "out_%03d.png" % t))).shape == (H, W)
assert sorted(f for f in os.listdir(tmp) if f.startswith("final")) == ["final.chunk000.pth", "final.chunk001.pth"]
assert not [f for f in os.listdir(tmp) if "psnr.chunk" in f]
print("CHUNKED CLI OK")
