"""One 1080p F2F train iteration (+ eval forward) for ncu launch lists / captures (development aid)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import synth
from frame2frame_b200.blind_denoising import OnlineF2F
from frame2frame_b200.dncnn import DnCNN

H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1080, 1920)
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda:0")
net = torch.nn.DataParallel(DnCNN(residual=False), device_ids=[0]).cuda()
net.load_state_dict({k: torch.from_numpy(v) for k, v in synth.pretrained_state_dict().items()})
eng = OnlineF2F(net)
f = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255.0))[None, None].to(dev) for t in range(2)]
flow = torch.from_numpy(synth.backward_flow(H, W, 1))[None].to(dev)
eng.finetune(f[1], f[0], flow, 1)
torch.cuda.synchronize()
e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
e0.record()
eng.finetune(f[1], f[0], flow, iters)
e1.record()
den = eng.denoise(f[1])
e2.record()
torch.cuda.synchronize()
P = H * W
ms = e0.elapsed_time(e1) / iters
print("%dx%d: train iteration %.2f ms (%.0f TFLOP/s), eval forward %.2f ms" % (W, H, ms, 3323520 * P / ms / 1e9, e1.elapsed_time(e2)))
