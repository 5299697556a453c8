"""Development check of the whole-network DnCNN path against the torch FP32 oracle (GPU box).
Prints per-tensor errors; exits non-zero on failure."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from frame2frame_b200 import _lib, synth
from frame2frame_b200.dncnn import DnCNN, Adam
from oracle import dncnn_oracle as do
from oracle import dncnn_bf16 as emu

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
lib = _lib.lib()
H16 = _lib.storage_dtype()      # BF16 or FP16, whatever the library was built with
emu.STORAGE = H16
ok = True


def rel(a, b):
    a, b = a.double().flatten().cpu(), b.double().flatten().cpu()
    return float((a - b).norm() / max(b.norm(), 1e-30))


def names(model):
    return [n for n, _ in model.named_parameters()]


def check_wgrad(H, W):
    """k_wgrad64 alone vs torch: dW = conv2d_weight(a, dz)"""
    global ok
    torch.manual_seed(1)
    z_prev = torch.randn(1, 64, H, W, device=dev).to(H16)
    g = (torch.randn(1, 64, H, W, device=dev) * 0.1).to(H16)
    z = torch.randn(1, 64, H, W, device=dev).to(H16)
    coef_in = torch.zeros(448, device=dev)
    coef_in[0:64] = torch.rand(64, device=dev) + 0.5
    coef_in[64:128] = torch.randn(64, device=dev) * 0.3
    coef = torch.zeros(448, device=dev)
    coef[256:320] = torch.rand(64, device=dev) + 0.5
    coef[320:384] = torch.randn(64, device=dev) * 0.01
    coef[384:448] = torch.randn(64, device=dev) * 0.01
    for bn_in in (1, 0):
        g_nhwc = g.permute(0, 2, 3, 1).contiguous()
        zp_nhwc = z_prev.permute(0, 2, 3, 1).contiguous()
        z_nhwc = z.permute(0, 2, 3, 1).contiguous()
        partial = torch.zeros(torch.cuda.get_device_properties(dev).multi_processor_count * 49152, device=dev)
        dw = torch.full((64, 64, 3, 3), float("nan"), device=dev)
        _lib.check(lib.f2f_dncnn_wgrad64(zp_nhwc.data_ptr(),
                                         coef_in.data_ptr() if bn_in else None, g_nhwc.data_ptr(),
                                         z_nhwc.data_ptr(), coef.data_ptr(),
                                         partial.data_ptr(), dw.data_ptr(), H, W, 0, None), "wgrad64")
        torch.cuda.synchronize()
        a = z_prev.float()
        if bn_in:
            a = torch.relu(a * coef_in[0:64].view(1, -1, 1, 1) + coef_in[64:128].view(1, -1, 1, 1)).to(H16).float()
        dz = (coef[256:320].view(1, -1, 1, 1) * g.float() + coef[320:384].view(1, -1, 1, 1) * z.float()
              + coef[384:448].view(1, -1, 1, 1)).to(H16)
        ref = torch.nn.grad.conv2d_weight(a, (64, 64, 3, 3), dz.float(), padding=1)
        e_dw = rel(dw, ref)
        e_dz = float((g_nhwc.permute(0, 3, 1, 2).float() - dz.float()).abs().max())
        good = e_dw < 2e-3 and e_dz <= 2.0 ** -7 * float(dz.float().abs().max())
        ok &= good
        print("wgrad64 %dx%d bn_in=%d: rel err dW %.3g, dz writeback max err %.3g %s" % (H, W, bn_in, e_dw, e_dz, "OK" if good else "FAIL"))
        if not good:
            d = (dw - ref).abs().cpu()
            print("   per-tap max abs err:", d.amax((0, 1)).numpy(), " ref max", float(ref.abs().max()))


def check_net(H, W, B=1, residual=True, seed=0):
    global ok
    sd = synth.dncnn_state_dict(seed=seed)
    rng = np.random.default_rng(seed + 5)
    # make BN parameters / running stats non-trivial
    for k, v in sd.items():
        if v.shape == (64,):
            if k.endswith("running_var"):
                sd[k] = (0.5 + rng.random(64)).astype(np.float32)
            elif k.endswith("weight"):
                sd[k] = (0.7 + 0.6 * rng.random(64)).astype(np.float32)
            else:
                sd[k] = (0.2 * rng.standard_normal(64)).astype(np.float32)
    ref = do.load_numpy_state(do.OracleDnCNN(residual=residual), sd)
    net = torch.nn.DataParallel(DnCNN(residual=residual), device_ids=[0]).cuda()
    net.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    x = torch.from_numpy(np.stack([synth.noisy_frame(H, W, t) / 255.0 for t in range(B)])[:, None].astype(np.float32))
    dy = torch.from_numpy(rng.standard_normal((B, 1, H, W)).astype(np.float32))
    m = net.module
    flat0 = m.flat_params.detach().cpu().clone()
    running0 = m._running.detach().cpu().clone()
    # eval forward
    ref.eval(); net.eval()
    with torch.no_grad():
        y_ref = ref(x)
        y = net(x.to(dev))
    y_emu = emu.forward_backward(flat0, x, None, running0, train=False, residual=residual)["y"]
    e, e32 = rel(y, y_emu), rel(y, y_ref)
    good = e < 2e-3
    print("net %dx%d B=%d residual=%d eval forward: rel err vs bf16-flow reference %.3g (vs fp32 oracle %.3g) %s"
          % (H, W, B, residual, e, e32, "OK" if good else "FAIL"))
    ok &= good
    # train forward + backward
    ref.train(); net.train()
    y_ref = ref(x)
    (y_ref * dy).sum().backward()
    y = net(x.to(dev))
    (y * dy.to(dev)).sum().backward()
    torch.cuda.synchronize()
    out = emu.forward_backward(flat0, x, dy, running0, train=True, residual=residual)
    e, e32 = rel(y.detach(), out["y"]), rel(y.detach(), y_ref.detach())
    good = e < 2e-3
    print("   train forward: rel err vs bf16-flow %.3g (vs fp32 %.3g) %s" % (e, e32, "OK" if good else "FAIL"))
    ok &= good
    e = rel(m._running, out["running"])
    good = e < 1e-4 and int(m._nbt.sum()) == 15
    print("   running stats: rel err %.3g, nbt sum %d %s" % (e, int(m._nbt.sum()), "OK" if good else "FAIL"))
    ok &= good
    worst = 0.0
    off = 0
    for (n, p), pr in zip(m.named_parameters(), ref.parameters()):
        ge = out["grads"][off:off + p.numel()].view(p.shape)
        off += p.numel()
        e = rel(p.grad, ge)
        worst = max(worst, e)
        if e > 1e-2:
            print("   grad %-22s rel err %.3g vs bf16-flow, %.3g vs fp32 (|ref| %.3g) FAIL" % (n, e, rel(p.grad, pr.grad), float(ge.norm())))
    e, e32 = rel(m.flat_grads, out["grads"]), rel(m.flat_grads, do.flat_grads(ref))
    good = worst < 1e-2
    print("   gradients: flat rel err vs bf16-flow %.3g (vs fp32 %.3g), worst tensor %.3g %s" % (e, e32, worst, "OK" if good else "FAIL"))
    ok &= good
    return net, ref


def check_layers(H, W, seed=0):
    """Layer-by-layer forward check: each kernel against torch on the kernel's OWN previous output,
    so that chaotic amplification through the random 17-layer stack does not blur the comparison."""
    global ok
    import ctypes as C
    import torch.nn.functional as F
    sd = synth.dncnn_state_dict(seed=seed)
    net = DnCNN(residual=True).cuda()
    net.load_state_dict({k[len("module."):]: torch.from_numpy(v) for k, v in sd.items()})
    x = torch.from_numpy((synth.noisy_frame(H, W, 0) / 255.0)[None, None].astype(np.float32)).to(dev)
    net.train()
    y = net(x)
    torch.cuda.synchronize()
    lay = (C.c_size_t * 8)()
    lib.f2f_dncnn_workspace_layout(1, H, W, 1, lay)
    buf = net._ws_train.buf
    act = lambda off: buf[off:off + H * W * 128].view(H16).view(1, H, W, 64).permute(0, 3, 1, 2).float()
    coef = buf[lay[6]:lay[6] + 15 * 448 * 4].view(torch.float32).view(15, 448)
    w0, mid, w16 = emu.unpack(net.flat_params.detach())
    a0 = act(lay[1])
    a0_ref = emu.bf(F.relu(F.conv2d(x, w0, padding=1)))
    worst = 0.0
    bad = int((a0 != a0_ref).sum())
    e = rel(a0, a0_ref)
    worst = max(worst, e)
    print("layers %dx%d: conv0 rel err %.3g (%d of %d elements differ)" % (H, W, e, bad, a0.numel()))
    a_prev = a0
    for k in range(1, 16):
        zk = act(lay[2] + (k - 1) * lay[3])
        z_ref = emu.bf(F.conv2d(a_prev, emu.bf(mid[k - 1][0]), padding=1))
        e = rel(zk, z_ref)
        nd = int((zk != z_ref).sum())
        zd = zk.double()
        mean = zd.mean((0, 2, 3)); var = ((zd * zd).mean((0, 2, 3)) - mean * mean).clamp_min(0)
        invstd = (1 / torch.sqrt(var + 1e-5)).float()
        sc = mid[k - 1][1] * invstd
        sh = mid[k - 1][2] - mean.float() * sc
        ec = max(rel(coef[k - 1, 0:64], sc), rel(coef[k - 1, 64:128], sh))
        worst = max(worst, e, ec)
        print("   conv%-2d z rel err %.3g (%d differ, max abs %.3g), BN scale/shift rel err %.3g"
              % (k, e, nd, float((zk - z_ref).abs().max()), ec))
        a_prev = emu.bf(F.relu(zk * coef[k - 1, 0:64].view(1, -1, 1, 1) + coef[k - 1, 64:128].view(1, -1, 1, 1)))
    y_ref = x - F.conv2d(a_prev, w16, padding=1)
    e = rel(y.detach(), y_ref)
    worst = max(worst, e)
    good = worst < 2e-3
    print("   conv16 + residual rel err %.3g -> %s" % (e, "OK" if good else "FAIL"))
    ok &= good


def check_backward_layers(H, W, seed=0):
    """Layer-by-layer backward check on the kernels' own inputs (f2f_dncnn_set_debug_dump)."""
    global ok
    import ctypes as C
    import torch.nn.functional as F
    sd = synth.dncnn_state_dict(seed=seed)
    net = DnCNN(residual=True).cuda()
    net.load_state_dict({k[len("module."):]: torch.from_numpy(v) for k, v in sd.items()})
    x = torch.from_numpy((synth.noisy_frame(H, W, 0) / 255.0)[None, None].astype(np.float32)).to(dev)
    torch.manual_seed(7)
    dy = torch.randn(1, 1, H, W, device=dev)
    lay = (C.c_size_t * 8)()
    lib.f2f_dncnn_workspace_layout(1, H, W, 1, lay)
    dumpbuf = torch.zeros(31 * lay[3], dtype=torch.uint8, device=dev)
    lib.f2f_dncnn_set_debug_dump(dumpbuf.data_ptr())
    net.train()
    y = net(x)
    y.backward(dy)
    torch.cuda.synchronize()
    lib.f2f_dncnn_set_debug_dump(None)
    buf = net._ws_train.buf
    nhwc = lambda t, off: t[off:off + H * W * 128].view(H16).view(1, H, W, 64).permute(0, 3, 1, 2).float()
    coef = buf[lay[6]:lay[6] + 15 * 448 * 4].view(torch.float32).view(15, 448)
    w0, mid, w16 = emu.unpack(net.flat_params.detach())
    gw0, gmid, gw16 = emu.unpack(net.flat_grads.detach())
    n_el = float(H * W)
    worst = 0.0
    z = [None] + [nhwc(buf, lay[2] + (k - 1) * lay[3]) for k in range(1, 16)]
    a0 = nhwc(buf, lay[1])
    yk = lambda k: z[k] * coef[k - 1, 0:64].view(1, -1, 1, 1) + coef[k - 1, 64:128].view(1, -1, 1, 1)
    act = lambda k: a0 if k == 0 else (emu.bf(F.relu(yk(k))) if k < 15 else F.relu(yk(k)))
    dn = -dy
    # the 16-bit intermediates (g_k, dz_k) live in a power-of-two scaled domain: max|dy| * sc in [1, 2)
    # (dncnn_net.cu grad_scale); parameter gradients come out unscaled
    sc = 2.0 ** -float(torch.floor(torch.log2(dy.abs().max())))
    e = rel(gw16, torch.nn.grad.conv2d_weight(act(15), w16.shape, dn, padding=1))
    g15 = nhwc(dumpbuf, 28 * lay[3])
    e2 = rel(g15, emu.bf(torch.nn.grad.conv2d_input(z[15].shape, w16, dn * sc, padding=1) * (yk(15) > 0)))
    worst = max(worst, e, e2)
    print("backward layers %dx%d: dW16 rel err %.3g, g15 rel err %.3g" % (H, W, e, e2))
    for k in range(15, 0, -1):
        g = nhwc(dumpbuf, 2 * (k - 1) * lay[3])
        dz = nhwc(dumpbuf, (2 * (k - 1) + 1) * lay[3])
        mean, invstd = coef[k - 1, 128:192], coef[k - 1, 192:256]
        gamma = mid[k - 1][1]
        xhat = (z[k] - mean.view(1, -1, 1, 1)) * invstd.view(1, -1, 1, 1)
        dbeta = g.double().sum((0, 2, 3)); dgamma = (g * xhat).double().sum((0, 2, 3))
        e_bn = max(rel(gmid[k - 1][1], dgamma / sc), rel(gmid[k - 1][2], dbeta / sc))
        A = (gamma * invstd).double()
        Bc = -A * invstd.double() * dgamma / n_el
        Cc = A * (-dbeta / n_el + mean.double() * invstd.double() * dgamma / n_el)
        e_cf = max(rel(coef[k - 1, 256:320], A), rel(coef[k - 1, 320:384], Bc), rel(coef[k - 1, 384:448], Cc))
        cA, cB, cC = (coef[k - 1, i:i + 64].view(1, -1, 1, 1) for i in (256, 320, 384))
        e_dz = rel(dz, emu.bf(cA * g + (cB * z[k] + cC)))
        e_dw = rel(gmid[k - 1][0], torch.nn.grad.conv2d_weight(act(k - 1), (64, 64, 3, 3), dz, padding=1) / sc)
        gin = torch.nn.grad.conv2d_input(z[k].shape, emu.bf(mid[k - 1][0]), dz, padding=1)
        mask = (yk(k - 1) > 0) if k >= 2 else (a0 > 0)
        gprev = nhwc(dumpbuf, (2 * (k - 2) if k >= 2 else 30) * lay[3])
        e_g = rel(gprev, emu.bf(gin * mask))
        worst = max(worst, e_bn, e_cf, e_dz, e_dw, e_g)
        print("   layer %-2d dgamma/dbeta %.3g, A/B/C %.3g, dz %.3g, dW %.3g, g_prev %.3g" % (k, e_bn, e_cf, e_dz, e_dw, e_g))
    g0 = nhwc(dumpbuf, 30 * lay[3])
    e = rel(gw0, torch.nn.grad.conv2d_weight(x, w0.shape, g0, padding=1) / sc)
    worst = max(worst, e)
    good = worst < 2e-3
    print("   dW0 rel err %.3g -> %s" % (e, "OK" if good else "FAIL"))
    ok &= good


def check_adam():
    global ok
    torch.manual_seed(3)
    p0 = torch.randn(10007) * 0.1
    pr = p0.clone().requires_grad_(True)
    opt_r = torch.optim.Adam([pr], lr=5e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-5)
    pg = p0.clone().to(dev).requires_grad_(True)
    opt_g = Adam([pg], lr=5e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-5)
    for s in range(25):
        g = torch.randn(10007) * (0.5 if s % 3 else 50.0)
        pr.grad = g.clone(); pg.grad = g.to(dev)
        opt_r.step(); opt_g.step()
    e = float((pg.detach().cpu() - pr.detach()).abs().max())
    good = e < 2e-7
    print("adam 25 steps: max abs diff %.3g (lr 5e-5) %s" % (e, "OK" if good else "FAIL"))
    ok &= good


def timing(H=1080, W=1920):
    net = DnCNN(residual=True).cuda()
    x = torch.rand(1, 1, H, W, device=dev)
    dy = torch.randn(1, 1, H, W, device=dev)
    net.train()
    opt = Adam(net.parameters(), lr=5e-5, weight_decay=1e-5)
    for _ in range(2):
        y = net(x); y.backward(dy); opt.step(); opt.zero_grad()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    n = 5
    tf = tb = ta = 0.0
    for _ in range(n):
        ev[0].record(); y = net(x); ev[1].record(); y.backward(dy); ev[2].record(); opt.step(); opt.zero_grad(); ev[3].record()
        torch.cuda.synchronize()
        tf += ev[0].elapsed_time(ev[1]); tb += ev[1].elapsed_time(ev[2]); ta += ev[2].elapsed_time(ev[3])
    P = H * W
    print("1080p train iteration: fwd %.2f ms (%.0f TF/s), bwd %.2f ms (%.0f TF/s), adam %.3f ms; total %.2f ms -> %.0f TF/s"
          % (tf / n, 1108224 * P / (tf / n) / 1e9, tb / n, (3323520 - 1108224) * P / (tb / n) / 1e9, ta / n,
             (tf + tb + ta) / n, 3323520 * P / ((tf + tb + ta) / n) / 1e9))
    net.eval()
    with torch.no_grad():
        for _ in range(2):
            net(x)
        ev[0].record()
        for _ in range(n):
            net(x)
        ev[1].record(); torch.cuda.synchronize()
    print("1080p eval forward: %.2f ms" % (ev[0].elapsed_time(ev[1]) / n))


if __name__ == "__main__":
    what = sys.argv[1:] or ["wgrad", "layers", "blayers", "net", "adam", "time"]
    if "wgrad" in what:
        for hw in [(16, 16), (32, 48), (40, 72), (135, 240)]:
            check_wgrad(*hw)
    if "layers" in what:
        check_layers(48, 64)
        check_layers(97, 131)
    if "blayers" in what:
        check_backward_layers(48, 64)
        check_backward_layers(97, 131)
    if "net" in what:
        check_net(48, 64)
        check_net(97, 131, residual=False)
        check_net(40, 56, B=2)
    if "adam" in what:
        check_adam()
    if "time" in what:
        timing()
    print("ALL OK" if ok else "SOME FAILED")
    sys.exit(0 if ok else 1)
