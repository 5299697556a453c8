"""smoke(): one small online-F2F frame (warp + masks, 3 fine-tuning iterations, eval forward) on
cuda:0 through the C ABI, checked against the CPU oracle (oracle/ is the checker only)."""
import numpy as np
import torch

from . import synth
from .blind_denoising import OnlineF2F, psnr
from .dncnn import DnCNN


def run(dev):
    from oracle import dncnn_oracle as do
    H, W, iters = 72, 104, 3
    sd = synth.pretrained_state_dict()
    net = torch.nn.DataParallel(DnCNN(residual=False), device_ids=[dev.index or 0]).to(dev)
    net.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    ref = do.load_numpy_state(do.OracleDnCNN(residual=False), sd)
    opt = do.make_adam(ref)
    curr, prev = (synth.noisy_frame(H, W, t) / np.float32(255.0) for t in (1, 0))
    flow = synth.backward_flow(H, W, 1)
    l_ref, l_gpu = [], []
    den_ref = do.finetune_frame(ref, opt, curr, prev, flow, iters, l_ref)
    den = OnlineF2F(net).step_frame(torch.from_numpy(curr)[None, None], torch.from_numpy(prev)[None, None],
                                    torch.from_numpy(flow)[None], iters, l_gpu)[0, 0].cpu().numpy()
    clean = synth.clean_frame_u8(H, W, 1) / np.float32(255.0)
    p_ref, p_gpu = psnr(clean, den_ref), psnr(clean, den)
    assert np.allclose(l_gpu, l_ref, rtol=2e-2), (l_gpu, l_ref)
    assert abs(p_ref - p_gpu) <= 0.25, (p_ref, p_gpu)
    print("smoke: F2F %dx%d, %d iterations: loss %.1f vs oracle %.1f, PSNR %.2f dB vs oracle %.2f dB"
          % (W, H, iters, l_gpu[-1], l_ref[-1], p_gpu, p_ref))
