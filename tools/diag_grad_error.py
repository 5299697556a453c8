"""Diagnostic: the default (16-bit) mode's parameter gradient of ONE F2F iteration against the precise (FP32) mode's, same
weights and frame: relative L2 error and sign agreement per layer.   python tools/diag_grad_error.py H W"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import dncnn, synth
from frame2frame_b200.blind_denoising import load_network
from frame2frame_b200.warped_loss import WarpedLoss

H, W = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda:0")
fix = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "dncnn_s25_synth.npz")
fr = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255.0))[None, None].to(dev) for t in range(2)]
flow = torch.from_numpy(synth.backward_flow(H, W, 1))[None].to(dev)
g = {}
for mode in ("fp16", "precise"):
    dncnn.set_precise(mode == "precise")
    net = load_network(fix, dev)
    net.train()
    crit = WarpedLoss().cuda(dev)
    out = fr[1] - net(fr[1])
    loss = crit(out, fr[0], flow)
    loss.backward()
    g[mode] = net.module.flat_grads.detach().double().cpu().numpy().copy()
    print(mode, "loss %.3f" % float(loss))
dncnn.set_precise(False)
a, b = g["fp16"], g["precise"]
rel = lambda x, y: float(np.linalg.norm(x - y) / np.linalg.norm(y))
print("all parameters: rel L2 error %.4f, sign agreement %.4f" % (rel(a, b), float(np.mean(np.sign(a) == np.sign(b)))))
o = 576
print("conv0  : rel %.4f" % rel(a[:576], b[:576]))
for k in range(1, 16):
    w = slice(o, o + 36864); ga = slice(o + 36864, o + 36928); be = slice(o + 36928, o + 36992)
    print("conv%-2d : rel %.4f  sign agree %.4f | gamma rel %.4f  beta rel %.4f | |g| median %.3g" %
          (k, rel(a[w], b[w]), float(np.mean(np.sign(a[w]) == np.sign(b[w]))), rel(a[ga], b[ga]), rel(a[be], b[be]), float(np.median(np.abs(b[w])))))
    o += 36992
print("conv16 : rel %.4f" % rel(a[o:], b[o:]))
