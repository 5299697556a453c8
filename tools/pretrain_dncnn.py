"""Fixture generator (GPU box, stock PyTorch -- NOT the product path): the reference's default
weights dncnn/logs/DnCNN-S-25/net.pth live in an un-vendored submodule, so a
DnCNN-S-25-layout denoiser is trained here on synthetic textures (sigma = 25 Gaussian noise on
uint8-quantised frames, residual learning, MSE) and saved, `module.`-prefixed like the reference
checkpoint (blind_denoising.py:144-147), to tests/golden/dncnn_s25_synth.npz (float32).  Both the
oracle and the CUDA path load this same file.

    python tools/pretrain_dncnn.py [iters] [out.npz]

Round 2: the round-1 fixture was trained on 64-patch batches drawn from 40 images of very different
contrast and kept the running statistics of that MIXTURE.  The reference fine-tunes in train mode
(blind_denoising.py:212), i.e. with the batch statistics of ONE frame; on that fixture the train-mode
forward of a frame was 8 dB below its eval-mode forward (31.99 vs 23.81 dB at 384x216), so the first
fine-tuning iterations had to repair the network and the eval output lost 2-6 dB -- nothing like the
operating point of a real pretrained DnCNN, whose patch batches, dataset and single images share their
statistics.  This generator therefore (1) trains on images of one family (the synthetic sequence
generator at random phases plus contrast-matched sinusoid/edge mixtures), (2) ends with a phase whose
batches are ONE whole image each (N = 1, full frame), the regime the F2F loop runs in, and (3) re-estimates
the running statistics as the plain average of per-image batch statistics over the generator's frames.  The script prints train- vs
eval-mode PSNR and a 3-frame F2F fine-tuning trajectory so that the consistency is on record.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from frame2frame_b200 import synth
from oracle import f2f_oracle as fo
from oracle.dncnn_oracle import OracleDnCNN, load_numpy_state, make_adam

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                        "tests", "golden", "dncnn_s25_synth.npz")
dev = torch.device("cuda:0" if torch.cuda.is_available() else "cpu")
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
torch.manual_seed(0)
rng = np.random.default_rng(0)

# training images: the synthetic generator at many phases, plus contrast-matched random mixtures
H, W = 360, 640
imgs = [synth.clean_frame(H, W, t) for t in range(0, 120, 3)]
yy, xx = np.mgrid[0:H, 0:W].astype(np.float64)
for _ in range(40):
    im = np.zeros((H, W))
    for _ in range(6):
        fx, fy, ph = rng.uniform(-0.5, 0.5), rng.uniform(-0.5, 0.5), rng.uniform(0, 6.28)
        im += rng.uniform(10, 40) * np.sin(fx * xx + fy * yy + ph)
    for _ in range(3):      # a few discs: edges
        cx, cy, r = rng.uniform(0, W), rng.uniform(0, H), rng.uniform(40, 150)
        im += rng.uniform(-50, 50) * ((xx - cx) ** 2 + (yy - cy) ** 2 < r * r)
    im = 127.5 + (im - im.mean()) * (52.0 / im.std())          # the generator's frames have std ~ 52
    imgs.append(np.clip(im, 0, 255))
clean = torch.from_numpy(np.round(np.stack(imgs)).astype(np.float32)).to(dev)      # uint8-quantised, 0..255
NI = clean.shape[0]


def corrupt(p):
    """clip(round(clean + N(0, 25^2))) like the frames the loop reads (synth.noisy_frame) -> noisy, clean in [0, 1]"""
    return torch.clamp(torch.round(p + torch.randn_like(p) * 25.0), 0, 255) / 255.0, p / 255.0


net = load_numpy_state(OracleDnCNN(residual=False), synth.dncnn_state_dict(seed=0)).to(dev)
opt = torch.optim.Adam(net.parameters(), lr=1e-3)
n1 = int(iters * 0.6)
sched = torch.optim.lr_scheduler.MultiStepLR(opt, [int(n1 * 0.6), int(n1 * 0.85)], 0.2)
net.train()
B, PS = 64, 50
for it in range(n1):                                   # phase 1: mixed patch batches
    idx = torch.randint(0, NI, (B,)).tolist()
    ys = torch.randint(0, H - PS, (B,)).tolist()
    xs = torch.randint(0, W - PS, (B,)).tolist()
    patch = torch.stack([clean[i, y:y + PS, x:x + PS] for i, y, x in zip(idx, ys, xs)])[:, None]
    if it % 2:
        patch = patch.flip(-1)
    noisy, cl = corrupt(patch)
    loss = ((net(noisy) - (noisy - cl)) ** 2).sum() / (2 * B)
    opt.zero_grad()
    loss.backward()
    opt.step()
    sched.step()
    if it % 500 == 0 or it == n1 - 1:
        print("phase 1 iter %d loss %.4f" % (it, float(loss)), flush=True)

# phase 2: every batch is ONE image (N = 1): per-image batch statistics, the regime the F2F loop runs in.  The image is a
# whole 640x360 training frame, a whole 960x540 frame of the generator, or a 640x360 window of one of its 1920x1080
# frames, so that the statistics of small frames, large frames and parts of large frames are all seen.
n2 = iters - n1
NGEN = 40                                              # images 0..39 are frames of the synthetic generator
mk = lambda h, w, ts: torch.from_numpy(np.stack([synth.clean_frame_u8(h, w, t) for t in ts])).to(dev)
clean540, clean1080 = mk(540, 960, range(1, 120, 8)), mk(1080, 1920, range(2, 120, 10))
opt = torch.optim.Adam(net.parameters(), lr=2e-4)
sched = torch.optim.lr_scheduler.MultiStepLR(opt, [int(n2 * 0.5), int(n2 * 0.8)], 0.25)
for it in range(n2):
    r = it % 8
    if r in (3, 7):
        img = clean[int(torch.randint(0, NI, (1,)))]
    elif r in (1, 5):
        img = clean540[int(torch.randint(0, clean540.shape[0], (1,)))]
    elif r in (2, 6):
        big = clean1080[int(torch.randint(0, clean1080.shape[0], (1,)))]
        y, x = int(torch.randint(0, 1080 - H, (1,))), int(torch.randint(0, 1920 - W, (1,)))
        img = big[y:y + H, x:x + W]
    else:
        img = clean[int(torch.randint(0, NGEN, (1,)))]
    img = img[None, None]
    if (it // 8) % 2:
        img = img.flip(-1)
    noisy, cl = corrupt(img)
    loss = ((net(noisy) - (noisy - cl)) ** 2).sum() / 2 * (2500.0 / img.numel())     # per 50x50 px like phase 1
    opt.zero_grad()
    loss.backward()
    opt.step()
    sched.step()
    if it % 500 == 0 or it == n2 - 1:
        print("phase 2 iter %d loss %.4f" % (it, float(loss)), flush=True)
# phase 3: a few whole 1920x1080 frames (the metric's size) at the final learning rate
for it in range(int(n2 * 0.05)):
    noisy, cl = corrupt(clean1080[it % clean1080.shape[0]][None, None])
    loss = ((net(noisy) - (noisy - cl)) ** 2).sum() / 2 * (2500.0 / noisy.numel())
    opt.zero_grad()
    loss.backward()
    opt.step()
print("phase 3 last loss %.4f" % float(loss), flush=True)

# running statistics = plain average of per-image batch statistics over the generator's frames at the three sizes
# (momentum=None: cumulative average) -- the pretrained network is "in distribution" for the sequences it will denoise,
# as DnCNN-S-25 is for natural video
for m in net.modules():
    if isinstance(m, torch.nn.BatchNorm2d):
        m.reset_running_stats()
        m.momentum = None
with torch.no_grad():
    for i in range(NGEN):
        net(corrupt(clean[i][None, None])[0])
    for i in range(clean540.shape[0]):
        net(corrupt(clean540[i][None, None])[0])
    for rep in range(2):
        for i in range(clean1080.shape[0]):
            net(corrupt(clean1080[i][None, None])[0])
for m in net.modules():
    if isinstance(m, torch.nn.BatchNorm2d):
        m.momentum = 0.1
        m.num_batches_tracked.zero_()

sd = {"module." + k: v.detach().cpu().numpy() for k, v in net.state_dict().items()}
np.savez_compressed(out, **sd)
print("saved", out, os.path.getsize(out))

# ---- on record: train- vs eval-mode forward, and the F2F fine-tuning trajectory (stock PyTorch FP32) ----
ps = lambda a, b: float(10 * torch.log10(1.0 / ((a - b) ** 2).mean()))
for Hh, Ww in ((216, 384), (540, 960), (1080, 1920)):
    c = torch.from_numpy(synth.clean_frame_u8(Hh, Ww, 3) / 255.0).float().to(dev)[None, None]
    n = torch.from_numpy(synth.noisy_frame(Hh, Ww, 3) / 255.0).float().to(dev)[None, None]
    with torch.no_grad():
        net.eval()
        pe = ps(n - net(n), c)
        saved = {k: v.clone() for k, v in net.state_dict().items()}
        net.train()
        pt = ps(n - net(n), c)
        net.load_state_dict(saved)
    print("%dx%d: noisy %.2f dB, eval-mode %.2f dB, train-mode %.2f dB" % (Ww, Hh, ps(n, c), pe, pt))

Hh, Ww = 270, 480
opt = make_adam(net)
for t in range(1, 6):
    curr, prev = (synth.noisy_frame(Hh, Ww, s) / np.float32(255.0) for s in (t, t - 1))
    cl = torch.from_numpy(synth.clean_frame_u8(Hh, Ww, t) / 255.0).float().to(dev)[None, None]
    warped, _, mask = fo.warp_and_mask(prev, synth.backward_flow(Hh, Ww, t), f64=True)
    tc, tw, tm = (torch.from_numpy(v).to(dev)[None, None] for v in (curr, warped, mask))
    with torch.no_grad():
        net.eval()
        p0 = ps(tc - net(tc), cl)
    net.train()
    for _ in range(20):
        loss = (tm * (tc - net(tc)) - tm * tw).abs().sum()
        opt.zero_grad()
        loss.backward()
        opt.step()
    with torch.no_grad():
        net.eval()
        p1 = ps(tc - net(tc), cl)
    print("F2F frame %d (%dx%d): eval PSNR before %.2f dB, after 20 iterations %.2f dB, last loss %.1f" % (t, Ww, Hh, p0, p1, float(loss)))
