"""BASELINE.json configs[4] at full size (3840x2160 gray): TV-L1 flow and the F2F step against the CPU oracles on one
frame pair, then the end-to-end online loop (flows on the device + 20 iterations per frame) timed over a few frames.
Run on the GPU box; writes a JSON record.   python tools/check_4k.py out.json [nframes]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import synth
from frame2frame_b200.blind_denoising import OnlineF2F, load_network, psnr
from frame2frame_b200.tvl1 import tvl1flow
from oracle import dncnn_oracle as do
from oracle.tvl1 import OracleTvl1, RefTvl1

H, W = 2160, 3840
out_fn = sys.argv[1]
NF = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
cores = os.cpu_count() or 1
torch.set_num_threads(cores)
rec = {"frame": "%dx%d" % (W, H), "cores": cores}
frames = [synth.noisy_frame(H, W, t) for t in range(NF + 1)]

# ---- TV-L1, one pair, 5 scales: flow and iteration counts against the oracle ----
RefTvl1.set_threads(cores)
t0 = time.time()
u, it_ref, _ = OracleTvl1().flow(frames[1], frames[0], nscales=5)
rec["tvl1_cpu_s"] = time.time() - t0
flow, it = tvl1flow(torch.from_numpy(frames[1]).to(dev), torch.from_numpy(frames[0]).to(dev), return_iters=True, nscales=5)
d = float(np.abs(flow.cpu().numpy() - u).max())
rec["tvl1_max_abs_px"] = d
rec["tvl1_iters_equal"] = bool(np.array_equal(it, it_ref))
print("TV-L1 4K: max|flow - oracle| = %.3g px, iteration counts equal: %s (CPU %.1f s)" % (d, rec["tvl1_iters_equal"], rec["tvl1_cpu_s"]), flush=True)

# ---- F2F: two inner iterations against the FP32 oracle loop (same flow) ----
gold = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "dncnn_s25_synth.npz")
eng = OnlineF2F(load_network(gold, dev))
ref = do.load_numpy_state(do.OracleDnCNN(residual=False), synth.pretrained_state_dict())
opt = do.make_adam(ref)
f01 = [f / np.float32(255.0) for f in frames]
lr, lg = [], []
t0 = time.time()
do.finetune_frame(ref, opt, f01[1], f01[0], u, 2, lr)
rec["f2f_cpu_s_2iters_plus_eval"] = time.time() - t0
eng.finetune(torch.from_numpy(f01[1])[None, None], torch.from_numpy(f01[0])[None, None], torch.from_numpy(u)[None], 2, lg)
rec["f2f_losses_oracle"], rec["f2f_losses_cuda"] = lr, lg
rec["f2f_loss_max_rel_dev"] = float(np.max(np.abs(np.array(lg) / np.array(lr) - 1)))
print("F2F 4K: losses oracle %s cuda %s (max rel dev %.2e)" % (lr, lg, rec["f2f_loss_max_rel_dev"]), flush=True)

# ---- end to end: flows on the device + 20 iterations per frame ----
eng = OnlineF2F(load_network(gold, dev))
host = [torch.from_numpy(f)[None, None].pin_memory() for f in f01]
outs = [torch.empty((1, 1, H, W)).pin_memory() for _ in f01]
for _ in eng.run_sequence(host[:2], 20, out=outs):     # warm-up: one frame
    pass
torch.cuda.synchronize()
t0 = time.time()
for _ in eng.run_sequence(host, 20, out=outs):
    pass
torch.cuda.synchronize()
dt = time.time() - t0
rec["e2e_frames"] = NF
rec["e2e_frames_per_s"] = NF / dt
rec["e2e_psnr_last"] = float(psnr(synth.clean_frame_u8(H, W, NF) / np.float32(255.0), outs[NF][0, 0].numpy()))
rec["gpu_mem_gb"] = torch.cuda.max_memory_allocated() / 1e9
print("end to end 4K: %.2f frames/s over %d frames, PSNR last %.2f dB, peak torch memory %.1f GB"
      % (rec["e2e_frames_per_s"], NF, rec["e2e_psnr_last"], rec["gpu_mem_gb"]), flush=True)
json.dump(rec, open(out_fn, "w"), indent=1)
