"""How reproducible is the reference loop itself?  Runs the FP32 CPU oracle of the online F2F loop
(oracle/dncnn_oracle.py: stock PyTorch modules, blind_denoising.py:211-230) twice on the same frames, flows and
weights, changing ONLY the number of CPU threads (= the summation order inside the FP32 convolution / reduction
kernels), and reports the per-frame PSNR of both runs.  The difference is the floor below which no implementation
-- including the reference on another machine -- can agree with a given reference run.

    python tools/chaos_floor.py H W NFRAMES ITERS threadsA threadsB out.json
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import synth
from oracle import dncnn_oracle as do

H, W, NF, ITERS, TA, TB = (int(v) for v in sys.argv[1:7])
out_fn = sys.argv[7]
psnr = lambda a, b: float(10 * np.log10(1.0 / np.mean((a - b) ** 2)))
frames = [synth.noisy_frame(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
clean = [synth.clean_frame_u8(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
sd = synth.pretrained_state_dict()


def run(threads):
    torch.set_num_threads(threads)
    ref = do.load_numpy_state(do.OracleDnCNN(residual=False), sd)
    opt = do.make_adam(ref)
    out = []
    for t in range(1, NF + 1):
        losses = []
        den = do.finetune_frame(ref, opt, frames[t], frames[t - 1], synth.backward_flow(H, W, t), ITERS, losses)
        out.append({"psnr": psnr(clean[t], den), "losses": losses})
        print(threads, t, out[-1]["psnr"], flush=True)
    return out


a, b = run(TA), run(TB)
rec = {"frame": "%dx%d" % (W, H), "iters": ITERS, "threads": [TA, TB], "torch": torch.__version__,
       "psnr_a": [f["psnr"] for f in a], "psnr_b": [f["psnr"] for f in b],
       "diff_db": [fb["psnr"] - fa["psnr"] for fa, fb in zip(a, b)],
       "loss_max_rel_dev": [float(np.max(np.abs(np.array(fb["losses"]) / np.array(fa["losses"]) - 1))) for fa, fb in zip(a, b)],
       "losses_a": [f["losses"] for f in a]}
print("PSNR difference between two FP32 reference runs (threads %d vs %d):" % (TA, TB), rec["diff_db"])
json.dump(rec, open(out_fn, "w"), indent=1)
