"""Fixture generator (GPU box, stock PyTorch -- NOT the product path): the reference's default
weights dncnn/logs/DnCNN-S-25/net.pth live in an un-vendored submodule, so a
DnCNN-S-25-layout denoiser is trained here for a few thousand iterations on synthetic textures
(sigma = 25 Gaussian noise, residual learning, MSE) and saved, `module.`-prefixed like the
reference checkpoint (blind_denoising.py:144-147), to tests/golden/dncnn_s25_synth.npz (float16
storage: 1.1 MB).  Both the oracle and the CUDA path load this same file.

    python tools/pretrain_dncnn.py [iters] [out.npz]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from frame2frame_b200 import synth
from oracle.dncnn_oracle import OracleDnCNN, load_numpy_state

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                        "tests", "golden", "dncnn_s25_synth.npz")
dev = torch.device("cuda:0")
torch.manual_seed(0)
rng = np.random.default_rng(0)

# training images: the synthetic sequence's clean frames plus random sinusoid/edge mixtures
H, W = 360, 640
imgs = [synth.clean_frame(H, W, t) for t in range(0, 40, 4)]
yy, xx = np.mgrid[0:H, 0:W].astype(np.float64)
for _ in range(30):
    im = np.full((H, W), 127.5)
    for _ in range(6):
        fx, fy, ph = rng.uniform(-0.6, 0.6), rng.uniform(-0.6, 0.6), rng.uniform(0, 6.28)
        im += rng.uniform(5, 45) * np.sin(fx * xx + fy * yy + ph)
    for _ in range(4):      # a few discs / half planes: edges
        cx, cy, r = rng.uniform(0, W), rng.uniform(0, H), rng.uniform(20, 150)
        im += rng.uniform(-60, 60) * ((xx - cx) ** 2 + (yy - cy) ** 2 < r * r)
    imgs.append(np.clip(im, 0, 255))
clean = torch.from_numpy(np.stack(imgs).astype(np.float32) / 255.0).to(dev)

net = load_numpy_state(OracleDnCNN(residual=False), synth.dncnn_state_dict(seed=0)).to(dev)
opt = torch.optim.Adam(net.parameters(), lr=1e-3)
sched = torch.optim.lr_scheduler.MultiStepLR(opt, [int(iters * 0.6), int(iters * 0.85)], 0.2)
B, PS = 64, 50
net.train()
for it in range(iters):
    idx = torch.randint(0, clean.shape[0], (B,), device=dev)
    ys = torch.randint(0, H - PS, (B,))
    xs = torch.randint(0, W - PS, (B,))
    patch = torch.stack([clean[i, y:y + PS, x:x + PS] for i, y, x in zip(idx.tolist(), ys.tolist(), xs.tolist())])[:, None]
    if it % 2:
        patch = patch.flip(-1)
    noise = torch.randn_like(patch) * (25.0 / 255.0)
    loss = ((net(patch + noise) - noise) ** 2).sum() / (2 * B)
    opt.zero_grad()
    loss.backward()
    opt.step()
    sched.step()
    if it % 500 == 0 or it == iters - 1:
        print("iter %d loss %.4f" % (it, float(loss)), flush=True)

# evaluate on a frame of the synthetic sequence (eval mode, running stats)
net.eval()
with torch.no_grad():
    Hh, Ww = 270, 480
    c = torch.from_numpy(synth.clean_frame_u8(Hh, Ww, 3) / 255.0).float().to(dev)[None, None]
    n = torch.from_numpy(synth.noisy_frame(Hh, Ww, 3) / 255.0).float().to(dev)[None, None]
    den = n - net(n)
    ps = lambda a, b: float(10 * torch.log10(1.0 / ((a - b) ** 2).mean()))
    print("PSNR noisy %.2f dB -> denoised %.2f dB" % (ps(n, c), ps(den, c)))
sd = {"module." + k: v.detach().cpu().numpy() for k, v in net.state_dict().items()}
np.savez_compressed(out, **{k: (v.astype(np.float16) if v.dtype == np.float32 else v) for k, v in sd.items()})
print("saved", out, os.path.getsize(out))
