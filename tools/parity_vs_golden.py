"""The CUDA online loop on BASELINE.json configs[0] against the recorded FP32 oracle run (tests/golden/f2f_config0_ref.json):
per-frame PSNR difference and loss deviation, no CPU work.
    python tools/parity_vs_golden.py [out.json] [golden.json] [--precise]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import _lib, synth
from frame2frame_b200.blind_denoising import OnlineF2F, psnr
from frame2frame_b200.dncnn import DnCNN

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
argv = [a for a in sys.argv[1:] if not a.startswith("--")]
for a in sys.argv:
    if a.startswith("--precise"):
        from frame2frame_b200.dncnn import set_precise
        set_precise(int(a.split("=")[1]) if "=" in a else 1)
gold = json.load(open(argv[1] if len(argv) > 1 else os.path.join(root, "tests", "golden", "f2f_config0_ref.json")))
W, H = (int(v) for v in gold["frame"].split("x"))
ITERS, NF = gold["iters"], len(gold["frames"])
net = torch.nn.DataParallel(DnCNN(residual=False), device_ids=[0]).cuda()
net.load_state_dict({k: torch.from_numpy(v) for k, v in synth.pretrained_state_dict().items()})
eng = OnlineF2F(net)
frames = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255.0))[None, None] for t in range(NF + 1)]
flows = {t: torch.from_numpy(synth.backward_flow(H, W, t))[None] for t in range(1, NF + 1)}
losses, rec = [], []
for t, den in eng.run_sequence(frames, ITERS, flows=flows, losses=losses):
    g = gold["frames"][t - 1]
    clean = synth.clean_frame_u8(H, W, t) / np.float32(255.0)
    p = float(psnr(clean, den[0, 0].cpu().numpy()))
    rec.append({"t": t, "psnr_cuda": p, "diff_db": p - g["psnr_oracle"],
                "loss_first_dev": losses[(t - 1) * ITERS] / g["loss_first"][0] - 1, "loss_last_dev": losses[t * ITERS - 1] / g["loss_last"][0] - 1})
    print("t=%d  PSNR %.3f  oracle %.3f  diff %+.3f dB   loss dev first %+.1e last %+.1e" % (t, p, g["psnr_oracle"], rec[-1]["diff_db"], rec[-1]["loss_first_dev"], rec[-1]["loss_last_dev"]))
d = [r["diff_db"] for r in rec]
dt = ["", "fp32 CUDA cores (precise 1)", "split fp16 pairs on tensor cores (precise 2)"][_lib.lib().f2f_dncnn_get_precise()] if _lib.lib().f2f_dncnn_get_precise() else ("fp16" if _lib.lib().f2f_dncnn_storage_dtype() else "bf16")
print("storage %s: mean diff %+.4f dB, max |diff| %.4f dB" % (dt, np.mean(d), np.max(np.abs(d))))
if argv:
    json.dump({"storage": dt, "frames": rec, "mean_diff_db": float(np.mean(d)), "max_abs_diff_db": float(np.max(np.abs(d)))}, open(argv[0], "w"), indent=1)
