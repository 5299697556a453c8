"""Whole network on one odd frame shape against the FP32 oracle: relative error of the output and of every
parameter group's gradient (GPU box).   python tools/check_shape.py H W [H W ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import synth
from frame2frame_b200.dncnn import DnCNN
from oracle import dncnn_oracle as do

dev = torch.device("cuda:0")
sd = synth.pretrained_state_dict()
rel = lambda a, b: float((a.double().cpu() - b.double().cpu()).norm() / (b.double().norm() + 1e-30))
args = [int(v) for v in sys.argv[1:]]
for H, W in zip(args[0::2], args[1::2]):
    ref = do.load_numpy_state(do.OracleDnCNN(residual=True), sd)
    net = torch.nn.DataParallel(DnCNN(residual=True), device_ids=[0]).cuda()
    net.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    x = torch.from_numpy((synth.noisy_frame(H, W, 0) / 255.0)[None, None].astype(np.float32))
    dy = torch.from_numpy(np.random.default_rng(1).standard_normal(x.shape).astype(np.float32))
    net.train(); ref.train()
    y, yr = net(x.to(dev)), ref(x)
    (y * dy.to(dev)).sum().backward()
    (yr * dy).sum().backward()
    g, gr = net.module.flat_grads.detach().cpu(), do.flat_grads(ref)
    print("%dx%d: output rel %.3g, all grads rel %.3g" % (H, W, rel(y.detach(), yr.detach()), rel(g, gr)))
    o = 0
    for name, p in ref.named_parameters():
        n = p.numel()
        r = rel(g[o:o + n], gr[o:o + n])
        if r > 0.05:
            print("   %-16s rel %.3g  |ref| %.3g |got| %.3g" % (name, r, float(gr[o:o + n].norm()), float(g[o:o + n].norm())))
        o += n
