"""Experiment: "mixed" precision of the online loop -- train-mode forward in FP32 (precise kernels: the output that the
L1 loss takes the SIGN of), backward in the default 16-bit tensor-core mode on the activations of a 16-bit forward --
against a committed FP32-oracle record.   python tools/exp_mixed_mode.py golden.json [nframes]"""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import dncnn, synth
from frame2frame_b200._lib import check, lib
from frame2frame_b200.blind_denoising import OnlineF2F, load_network, psnr

gold = json.load(open(sys.argv[1]))
W, H = (int(v) for v in gold["frame"].split("x"))
NF = min(int(sys.argv[2]) if len(sys.argv) > 2 else 99, len(gold["frames"]))
ITERS = gold["iters"]
dev = torch.device("cuda:0")
fix = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "dncnn_s25_synth.npz")
eng = OnlineF2F(load_network(fix, dev))
net, L = eng.net, lib()
st_, group = eng.optimizer.fused_state(net)
b1, b2 = group["betas"]
stream = torch.cuda.current_stream().cuda_stream
ws16 = torch.empty(L.f2f_dncnn_workspace_bytes(1, H, W, 1), dtype=torch.uint8, device=dev)
dncnn.set_precise(True)
ws32 = torch.empty(L.f2f_dncnn_workspace_bytes(1, H, W, 1), dtype=torch.uint8, device=dev)
dncnn.set_precise(False)
scr_run, scr_nbt = torch.empty_like(net._running), torch.empty_like(net._nbt)
out, out16, dout = (torch.empty((1, 1, H, W), device=dev) for _ in range(3))
loss = torch.zeros((), dtype=torch.float64, device=dev)
frames = [torch.from_numpy(synth.noisy_frame(H, W, t) / np.float32(255.0))[None, None].to(dev) for t in range(NF + 1)]
diffs = []
for t in range(1, NF + 1):
    flow = torch.from_numpy(synth.backward_flow(H, W, t))[None].to(dev)
    warped, mask = eng.criterion.precompute(frames[t - 1], flow)
    curr = frames[t]
    losses = []
    for _ in range(ITERS):
        scr_run.copy_(net._running); scr_nbt.copy_(net._nbt)
        P = net.flat_params.data_ptr()
        dncnn.set_precise(False)       # 16-bit forward: only to leave its activations for the 16-bit backward
        check(L.f2f_dncnn_forward(P, scr_run.data_ptr(), scr_nbt.data_ptr(), curr.data_ptr(), out16.data_ptr(), ws16.data_ptr(), ws16.numel(), 1, H, W, 1, 1, stream), "fwd16")
        dncnn.set_precise(True)        # FP32 forward: the output, the loss, the running statistics
        check(L.f2f_dncnn_forward(P, net._running.data_ptr(), net._nbt.data_ptr(), curr.data_ptr(), out.data_ptr(), ws32.data_ptr(), ws32.numel(), 1, H, W, 1, 1, stream), "fwd32")
        dncnn.set_precise(False)
        check(L.f2f_masked_loss(out.data_ptr(), warped.data_ptr(), mask.data_ptr(), dout.data_ptr(), loss.data_ptr(), H * W, 0, stream), "loss")
        check(L.f2f_dncnn_backward(P, net.flat_grads.data_ptr(), dout.data_ptr(), ws16.data_ptr(), ws16.numel(), 1, H, W, 1, 0, stream), "bwd16")
        check(L.f2f_adam_step(P, net.flat_grads.data_ptr(), st_["exp_avg"].data_ptr(), st_["exp_avg_sq"].data_ptr(), net.flat_params.numel(), st_["step"] + 1, group["lr"], b1, b2, group["eps"], group["weight_decay"], stream), "adam")
        eng.optimizer.advance()
        losses.append(float(loss.item()))
    dncnn.set_precise(True)
    den = eng.denoise(curr)
    dncnn.set_precise(False)
    g = gold["frames"][t - 1]
    p = float(psnr(synth.clean_frame_u8(H, W, t) / np.float32(255.0), den[0, 0].cpu().numpy()))
    diffs.append(p - g["psnr_oracle"])
    print("t=%d  PSNR %.3f  oracle %.3f  diff %+.4f dB   loss dev first %+.1e last %+.1e" % (t, p, g["psnr_oracle"], diffs[-1], losses[0] / g["loss_first"][0] - 1, losses[-1] / g["loss_last"][0] - 1), flush=True)
print("mixed mode: max |diff| %.4f dB" % np.max(np.abs(diffs)))
