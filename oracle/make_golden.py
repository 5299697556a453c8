"""oracle/make_golden.py -- TEST INFRASTRUCTURE.  Regenerates tests/golden/*.npz
by executing the REFERENCE itself in this container:

  * TV-L1: oracle/_ref/libtvl1_ref.so (tvl1flow_lib.c compiled in place);
  * F2F:   /root/reference/blind_denoising.py imported with stub modules
           (see oracle/f2f_ref.py).

The fixtures are small (a few hundred KB) and committed; the GPU box has no
/root/reference, so the -m gpu tests read only these files.
    python oracle/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from frame2frame_b200 import synth  # noqa: E402
from oracle.tvl1 import OracleTvl1, RefTvl1  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def tvl1():
    ref, orc = RefTvl1(), OracleTvl1()
    out = {}
    cases = [("a", 68, 120, dict()), ("b", 97, 131, dict(nscales=3, nwarps=3)),
             ("c", 135, 240, dict(lam=0.2, nscales=100, fscale=2))]   # c = tvl1flow.sh:14 settings
    for name, H, W, kw in cases:
        I0 = synth.noisy_frame(H, W, 1)
        I1 = synth.noisy_frame(H, W, 0)
        u = ref.flow(I0, I1, **kw)
        _, iters, _ = orc.flow(I0, I1, **kw)
        out["%s_hw" % name] = np.array([H, W], np.int32)
        out["%s_flow" % name] = u
        out["%s_iters" % name] = iters
        out["%s_kw" % name] = np.array(repr(kw))
    np.savez_compressed(os.path.join(GOLD, "tvl1_ref.npz"), **out)
    print("tvl1_ref.npz:", {k: getattr(v, "shape", None) for k, v in out.items()})


def dncnn():
    """tests/golden/dncnn_ref.npz: the REFERENCE's DnCNN (lib/frame2frame/dncnn/net.py) and its
    fine-tuning loop body (blind_denoising.py:211-230 with the reference's WarpedLoss and
    torch.optim.Adam) on seeded inputs."""
    import torch
    from oracle import f2f_ref
    ref = f2f_ref.load()
    sd = {k[len("module."):]: torch.from_numpy(v) for k, v in synth.dncnn_state_dict(seed=0).items()}
    out = {}
    H, W = 24, 40
    x = torch.from_numpy(np.stack([synth.noisy_frame(H, W, t) / 255.0 for t in range(2)])[:, None].astype(np.float32))
    dy = torch.from_numpy(np.random.default_rng(11).standard_normal((2, 1, H, W)).astype(np.float32))
    net = ref.LibDnCNN(1, 17)                       # forward = x - dncnn(x), net.py:31
    net.load_state_dict(sd)
    net.train()
    y = net(x)
    (y * dy).sum().backward()
    g = torch.cat([p.grad.reshape(-1) for p in net.parameters()])
    out["net_x"], out["net_dy"], out["net_y_train"] = x.numpy(), dy.numpy(), y.detach().numpy()
    out["net_grad_head"], out["net_grad_tail"] = g[:576].numpy(), g[-576:].numpy()
    out["net_grad_bn1"] = g[576 + 36864:576 + 36864 + 128].numpy()
    out["net_grad_norm"] = np.array(float(g.double().norm()))
    out["net_running_var_1"] = net.dncnn[3].running_var.numpy().copy()
    net.eval()
    with torch.no_grad():
        out["net_y_eval"] = net(x).numpy()
    # fine-tuning loop body, 3 iterations, the paper script's noise-estimate forward
    H, W = 32, 48
    curr, prev, flow = f2f_ref.step_inputs(H, W)
    model = ref.RefDnCNN(1, 17)
    model.load_state_dict(sd)
    crit = ref.WarpedLoss()
    opt = torch.optim.Adam(model.parameters(), lr=5e-5, betas=(0.9, 0.999), eps=1e-08, weight_decay=0.00001, amsgrad=False)
    tc, tp, tf = torch.from_numpy(curr)[None, None], torch.from_numpy(prev)[None, None], torch.from_numpy(flow)[None]
    model.train()
    opt.zero_grad()
    losses = []
    for it in range(3):
        out_train = tc - model(tc)
        loss = crit(out_train, tp, tf)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    model.eval()
    with torch.no_grad():
        den = tc - model(tc)
    out["ft_hw"] = np.array([H, W], np.int32)
    out["ft_losses"] = np.array(losses, np.float64)
    out["ft_den"] = den[0, 0].numpy()
    out["ft_w16"] = model.dncnn[47].weight.detach().numpy().copy()
    np.savez_compressed(os.path.join(GOLD, "dncnn_ref.npz"), **out)
    print("dncnn_ref.npz:", {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    tvl1()
    try:
        from oracle import f2f_ref
        f2f_ref.make_golden(GOLD)
        dncnn()
    except ImportError as e:
        print("f2f golden skipped:", e)
