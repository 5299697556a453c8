"""Diagnostic: running BatchNorm statistics after one fine-tuned frame in the default (16-bit) and the precise (FP32) mode,
per layer and channel -- where do they differ, and are those the channels whose |mean| / std is large?
    python tools/diag_bn_stats.py H W [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import dncnn, synth
from frame2frame_b200.blind_denoising import OnlineF2F, load_network, psnr

H, W = int(sys.argv[1]), int(sys.argv[2])
ITERS = int(sys.argv[3]) if len(sys.argv) > 3 else 20
dev = torch.device("cuda:0")
fix = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "dncnn_s25_synth.npz")
fr = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255.0))[None, None] for t in range(2)]
flow = torch.from_numpy(synth.backward_flow(H, W, 1))[None]
clean = synth.clean_frame_u8(H, W, 1) / np.float32(255.0)
res = {}
for mode in ("fp16", "precise"):
    dncnn.set_precise(mode == "precise")
    eng = OnlineF2F(load_network(fix, dev))
    r0 = eng.net._running.clone()
    den = eng.step_frame(fr[1], fr[0], flow, ITERS)
    res[mode] = (eng.net._running.clone().cpu().numpy(), float(psnr(clean, den[0, 0].cpu().numpy())), eng)
    print(mode, "eval PSNR after %d iterations: %.3f dB" % (ITERS, res[mode][1]))
dncnn.set_precise(False)
a, b = res["fp16"][0], res["precise"][0]          # [15][2][64]
r0 = r0.cpu().numpy()
for k in range(15):
    dm = np.abs(a[k, 0] - b[k, 0]) / np.sqrt(b[k, 1] + 1e-5)           # mean difference in units of the channel's std
    dv = a[k, 1] / b[k, 1]
    ratio = np.abs(b[k, 0]) / np.sqrt(b[k, 1] + 1e-5)
    j = int(np.argmax(np.abs(np.log(dv))))
    print("layer %2d: max |d mean|/std %.3g  var ratio fp16/fp32 in [%.3f, %.3f]  worst ch %2d: mean %.4g var %.3g (fp16 var %.3g) |mean|/std %.1f  min var %.3g"
          % (k + 1, dm.max(), dv.min(), dv.max(), j, b[k, 0, j], b[k, 1, j], a[k, 1, j], ratio[j], b[k, 1].min()))
# cross evaluation: the fp16-trained weights with the precise run's running statistics, evaluated in the default mode
eng = res["fp16"][2]
eng.net._running.copy_(torch.from_numpy(b).to(dev))
print("fp16 weights + precise running stats, default-mode eval: %.3f dB" % float(psnr(clean, eng.denoise(fr[1])[0, 0].cpu().numpy())))
dncnn.set_precise(True)
print("fp16 weights + precise running stats, precise eval:       %.3f dB" % float(psnr(clean, eng.denoise(fr[1])[0, 0].cpu().numpy())))
eng.net._running.copy_(torch.from_numpy(a).to(dev))
print("fp16 weights + fp16 running stats, precise eval:          %.3f dB" % float(psnr(clean, eng.denoise(fr[1])[0, 0].cpu().numpy())))
dncnn.set_precise(False)
