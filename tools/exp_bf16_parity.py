"""CPU experiment: which BF16 rounding (weights / activations / gradients) explains the PSNR difference of
the online F2F loop against the FP32 oracle?  Runs the FP32 oracle loop and the BF16-flow emulation
(oracle/dncnn_bf16.py) with selected roundings on the same frames.   usage: exp_bf16_parity.py H W NF ITERS variants..."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import synth
from oracle import dncnn_oracle as do, dncnn_bf16 as db, f2f_oracle as fo

if os.environ.get("F2F_EMU_STORAGE", "bf16") == "fp16":      # emulate the FP16 (default) build instead of the BF16 one
    db.STORAGE = torch.float16
H, W, NF, ITERS = (int(v) for v in sys.argv[1:5])
variants = sys.argv[5:] or ["w+act+grad", "w", "act", "grad", "none"]
sd = synth.pretrained_state_dict()
frames = [synth.noisy_frame(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
clean = [synth.clean_frame_u8(H, W, t) / np.float32(255.0) for t in range(NF + 1)]
psnr = lambda a, b: 10 * np.log10(1.0 / np.mean((a - b) ** 2))

ref = do.load_numpy_state(do.OracleDnCNN(residual=False), sd)
opt = do.make_adam(ref)
p_ref = []
for t in range(1, NF + 1):
    den = do.finetune_frame(ref, opt, frames[t], frames[t - 1], synth.backward_flow(H, W, t), ITERS)
    p_ref.append(psnr(clean[t], den))
print("fp32 oracle     :", " ".join("%.3f" % p for p in p_ref), flush=True)

def emu(variant):
    for k in db.ROUND:
        db.ROUND[k] = k in variant.split("+")
    m0 = do.load_numpy_state(do.OracleDnCNN(residual=False), sd)
    flat = do.flat_params(m0).clone().requires_grad_(True)
    running = torch.stack([torch.stack([m.running_mean, m.running_var]) for m in m0.dncnn if isinstance(m, torch.nn.BatchNorm2d)]).clone()
    opt = torch.optim.Adam([flat], lr=5e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-5)
    out = []
    for t in range(1, NF + 1):
        warped, _, mask = fo.warp_and_mask(frames[t - 1], synth.backward_flow(H, W, t), f64=True)
        tc, tw, tm = (torch.from_numpy(a)[None, None] for a in (frames[t], warped, mask))
        for _ in range(ITERS):
            with torch.no_grad():
                r = db.forward_backward(flat, tc, None, running, True, True)
                dy = tm * torch.sign(tm * r["y"] - tm * tw)
                r = db.forward_backward(flat, tc, dy, running, True, True)
                running = r["running"]
            flat.grad = r["grads"]
            opt.step()
        with torch.no_grad():
            den = db.forward_backward(flat, tc, None, running, False, True)["y"][0, 0].numpy()
        out.append(psnr(clean[t], den))
    return out

for v in variants:
    t0 = time.time()
    p = emu(v)
    print("%-16s:" % v, " ".join("%.3f" % x for x in p), "| diff", " ".join("%+.3f" % (a - b) for a, b in zip(p, p_ref)), "(%.0fs)" % (time.time() - t0), flush=True)
