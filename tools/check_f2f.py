"""End-to-end check of the online F2F loop against the FP32 CPU oracle (GPU box):
same frames, flows and pretrained weights; compares the loss trajectory and per-frame PSNR."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from frame2frame_b200 import synth
from frame2frame_b200.blind_denoising import OnlineF2F, psnr
from frame2frame_b200.dncnn import DnCNN
from oracle import dncnn_oracle as do

H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (135, 240)
NF = int(sys.argv[3]) if len(sys.argv) > 3 else 4
ITERS = int(sys.argv[4]) if len(sys.argv) > 4 else 20
dev = torch.device("cuda:0")
sd = synth.pretrained_state_dict()

ref = do.load_numpy_state(do.OracleDnCNN(residual=False), sd)
opt = do.make_adam(ref)
net = torch.nn.DataParallel(DnCNN(residual=False), device_ids=[0]).cuda()
net.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
eng = OnlineF2F(net)

frames = [synth.noisy_frame(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
clean = [synth.clean_frame_u8(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
ok = True
for t in range(1, NF + 1):
    flow = synth.backward_flow(H, W, t)
    l_ref, l_gpu = [], []
    t0 = time.time()
    den_ref = do.finetune_frame(ref, opt, frames[t], frames[t - 1], flow, ITERS, l_ref)
    t1 = time.time()
    den = eng.step_frame(torch.from_numpy(frames[t])[None, None], torch.from_numpy(frames[t - 1])[None, None],
                         torch.from_numpy(flow)[None], ITERS, l_gpu)
    torch.cuda.synchronize()
    t2 = time.time()
    den = den[0, 0].cpu().numpy()
    p_ref, p_gpu, p_in = psnr(clean[t], den_ref), psnr(clean[t], den), psnr(clean[t], frames[t])
    lr = np.array(l_ref); lg = np.array(l_gpu)
    dl = np.abs(lg - lr) / lr
    good = abs(p_ref - p_gpu) <= 0.05
    ok &= good
    print("frame %d: PSNR noisy %.2f | oracle %.3f dB, B200 %.3f dB (diff %+.4f) | loss first %.1f/%.1f last %.1f/%.1f max rel diff %.2e | "
          "cpu %.1fs gpu %.2fs %s" % (t, p_in, p_ref, p_gpu, p_gpu - p_ref, lr[0], lg[0], lr[-1], lg[-1], dl.max(), t1 - t0, t2 - t1,
                                    "OK" if good else "FAIL"))
print("ALL OK" if ok else "SOME FAILED")
sys.exit(0 if ok else 1)
