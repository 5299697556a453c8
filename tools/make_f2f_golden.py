"""Golden record of the FP32 CPU oracle loop (oracle/dncnn_oracle.py: the stock torch modules of the reference's loop,
blind_denoising.py:211-253) on a synthetic sequence with its analytic flows and the pretrained fixture -- host only, no
CUDA.  The -m gpu parity tests, tools/parity_vs_golden.py and bench.py's `psnr_delta_db` compare the CUDA loop with it.

    python tools/make_f2f_golden.py H W NFRAMES ITERS out.json [first_t ...]

`first_t` (optional, several allowed): record chunks -- for each value f the loop restarts from the pretrained weights
with frame f as the first `prev` and denoises NFRAMES frames f+1 .. f+NFRAMES (the `--first/--last` chunking of
blind_denoising.py:279-282, SURVEY 8e: chunk-matched oracle for configs[4]).  Default: one chunk starting at 0.
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import synth
from frame2frame_b200.blind_denoising import psnr
from oracle import dncnn_oracle as do

H, W, NF, ITERS = (int(v) for v in sys.argv[1:5])
out_fn = sys.argv[5]
firsts = [int(v) for v in sys.argv[6:]] or [0]
threads = int(os.environ.get("F2F_THREADS", "0")) or os.cpu_count() or 1
torch.set_num_threads(threads)
sd = synth.pretrained_state_dict()
rec = {"frame": "%dx%d" % (W, H), "iters": ITERS, "threads": threads, "torch": torch.__version__,
       "fixture": "tests/golden/dncnn_s25_synth.npz", "flows": "synth.backward_flow (analytic)", "frames": []}
for f0 in firsts:
    ref = do.load_numpy_state(do.OracleDnCNN(residual=False), sd)
    opt = do.make_adam(ref)
    prev = synth.noisy_frame(H, W, f0) / np.float32(255.0)
    for t in range(f0 + 1, f0 + NF + 1):
        curr = synth.noisy_frame(H, W, t) / np.float32(255.0)
        clean = synth.clean_frame_u8(H, W, t) / np.float32(255.0)
        with torch.no_grad():
            ref.eval()
            p0 = float(psnr(clean, curr - ref(torch.from_numpy(curr)[None, None])[0, 0].numpy()))
        losses = []
        t0 = time.time()
        den = do.finetune_frame(ref, opt, curr, prev, synth.backward_flow(H, W, t), ITERS, losses)
        r = {"t": t, "chunk_first": f0, "psnr_before_finetune": p0, "psnr_oracle": float(psnr(clean, den)),
             "loss_first": [losses[0]], "loss_last": [losses[-1]], "losses": losses, "cpu_s": time.time() - t0}
        rec["frames"].append(r)
        print(json.dumps({k: v for k, v in r.items() if k != "losses"}), flush=True)
        json.dump(rec, open(out_fn, "w"), indent=1)
        prev = curr
