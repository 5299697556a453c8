"""PSNR parity of the online F2F loop at a given frame size: the CUDA path against the FP32 CPU oracle
(oracle/dncnn_oracle.py: stock PyTorch modules, the loop of blind_denoising.py:211-230) on the same synthetic
frames, analytic flows and pretrained weights.  Run on the GPU box; writes a JSON record.

    python tools/parity_f2f.py H W NFRAMES ITERS out.json [--resync]

--resync: before every frame the CUDA engine is re-loaded with the oracle's state (weights, Adam moments, BatchNorm
running statistics), so each frame measures what 20 iterations of BF16/FP16 storage do on their own; without it the
two runs evolve independently like two runs of the reference on different hardware would.
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import synth
from frame2frame_b200.blind_denoising import OnlineF2F, psnr
from frame2frame_b200.dncnn import DnCNN
from oracle import dncnn_oracle as do

H, W, NF, ITERS = (int(v) for v in sys.argv[1:5])
out_fn = sys.argv[5]
resync = "--resync" in sys.argv
dev = torch.device("cuda:0")
torch.set_num_threads(os.cpu_count() or 1)
sd = synth.pretrained_state_dict()
ref = do.load_numpy_state(do.OracleDnCNN(residual=False), sd)
opt = do.make_adam(ref)
net = torch.nn.DataParallel(DnCNN(residual=False), device_ids=[0]).cuda()
net.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
eng = OnlineF2F(net)
frames = [synth.noisy_frame(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
clean = [synth.clean_frame_u8(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
rec = {"frame": "%dx%d" % (W, H), "iters": ITERS, "resync": resync, "frames": []}
for t in range(1, NF + 1):
    flow = synth.backward_flow(H, W, t)
    lr, lg = [], []
    with torch.no_grad():
        ref.eval()
        p0 = psnr(clean[t], frames[t] - ref(torch.from_numpy(frames[t])[None, None])[0, 0].numpy())
    t0 = time.time()
    den_ref = do.finetune_frame(ref, opt, frames[t], frames[t - 1], flow, ITERS, lr)
    t_cpu = time.time() - t0
    torch.cuda.synchronize()
    t0 = time.time()
    den = eng.step_frame(torch.from_numpy(frames[t])[None, None], torch.from_numpy(frames[t - 1])[None, None],
                         torch.from_numpy(flow)[None], ITERS, lg)
    torch.cuda.synchronize()
    t_gpu = time.time() - t0
    pg, pr = float(psnr(clean[t], den[0, 0].cpu().numpy())), float(psnr(clean[t], den_ref))
    r = {"t": t, "psnr_before_finetune": float(p0), "psnr_oracle": pr, "psnr_cuda": pg, "diff_db": pg - pr,
         "loss_first": [lr[0], lg[0]], "loss_last": [lr[-1], lg[-1]],
         "loss_max_rel_dev": float(np.max(np.abs(np.array(lg) / np.array(lr) - 1))), "cpu_s": t_cpu, "gpu_s": t_gpu}
    rec["frames"].append(r)
    print(json.dumps(r), flush=True)
    if resync and t < NF:
        net.load_state_dict({"module." + k: v.clone() for k, v in ref.state_dict().items()})
        st, _ = eng.optimizer.fused_state(eng.net)
        o = 0
        for p in ref.parameters():
            n = p.numel()
            s = opt.state[p]
            st["exp_avg"][o:o + n].copy_(s["exp_avg"].reshape(-1))
            st["exp_avg_sq"][o:o + n].copy_(s["exp_avg_sq"].reshape(-1))
            o += n
d = [f["diff_db"] for f in rec["frames"]]
rec["mean_diff_db"], rec["max_abs_diff_db"] = float(np.mean(d)), float(np.max(np.abs(d)))
print("mean diff %.4f dB, max |diff| %.4f dB" % (rec["mean_diff_db"], rec["max_abs_diff_db"]))
json.dump(rec, open(out_fn, "w"), indent=1)
