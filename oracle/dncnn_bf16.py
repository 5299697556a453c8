"""oracle/dncnn_bf16.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

PyTorch FP32 reference of the DnCNN kernels' *data flow*: the same network as
oracle/dncnn_oracle.py (lib/frame2frame/dncnn/net.py:6-39, train-mode BatchNorm of
blind_denoising.py:212) evaluated with FP32 convolutions but with every tensor the CUDA path
stores in BF16 rounded to BF16 at the same point (64-channel conv weights, pre-BN conv outputs
z_k, post-ReLU activations a_k, masked gradients g_k, BatchNorm-backward outputs dz_k).  A deep
randomly initialised ReLU/BatchNorm stack amplifies rounding differences layer by layer, so the
kernels are checked tightly against THIS reference (does each kernel compute what it should?) and
loosely / end-to-end (PSNR, loss trajectory) against the un-rounded FP32 oracle (is BF16 storage
good enough?).
"""
import torch
import torch.nn.functional as F

EPS = 1e-5


STORAGE = torch.bfloat16     # the CUDA path's 16-bit storage type (tests set it from f2f_dncnn_storage_dtype)


def bf(t):
    return t.to(STORAGE).to(torch.float32)


# Which classes of tensors are rounded to BF16 (all of them = the CUDA path's data flow).  The
# switches exist for tools/exp_bf16_parity.py, which attributes the end-to-end PSNR difference
# against the FP32 oracle to weights / activations / gradients.
ROUND = {"w": True, "act": True, "grad": True, "last": False}
# "last": also round the 64->1 layer's operands (a_15 and its weights) -- NOT what the CUDA path does today (that layer
# runs in FP32 on the CUDA cores); the switch exists to predict the parity cost of moving it to the tensor core.


def bfw(t):
    return bf(t) if ROUND["w"] else t


def bfa(t):
    return bf(t) if ROUND["act"] else t


def bfg(t):
    return bf(t) if ROUND["grad"] else t


def unpack(flat):
    """flat parameter vector (state_dict order) -> (w0, [(w_k, gamma_k, beta_k)]*15, w16)"""
    o = 0
    w0 = flat[o:o + 576].view(64, 1, 3, 3); o += 576
    mid = []
    for _ in range(15):
        w = flat[o:o + 36864].view(64, 64, 3, 3); o += 36864
        g = flat[o:o + 64]; o += 64
        b = flat[o:o + 64]; o += 64
        mid.append((w, g, b))
    w16 = flat[o:o + 576].view(1, 64, 3, 3)
    return w0, mid, w16


def forward_backward(flat, x, dy=None, running=None, train=True, residual=True):
    """x [B,1,H,W] float32.  -> dict(y, grads (flat, if dy given), running (updated copy))."""
    w0, mid, w16 = unpack(flat.detach().float())
    x = x.float()
    n_el = float(x.shape[0] * x.shape[2] * x.shape[3])
    a = [bfa(F.relu(F.conv2d(x, w0, padding=1)))]
    z, coef = [None], [None]
    new_running = None if running is None else running.clone()
    for k in range(1, 16):
        w, gamma, beta = mid[k - 1]
        zk = bfa(F.conv2d(a[k - 1], bfw(w), padding=1))
        if train:
            zd = zk.double()
            mean = zd.mean((0, 2, 3))
            var = ((zd * zd).mean((0, 2, 3)) - mean * mean).clamp_min(0)
            if new_running is not None:
                new_running[k - 1, 0] = 0.9 * running[k - 1, 0] + 0.1 * mean.float()
                new_running[k - 1, 1] = 0.9 * running[k - 1, 1] + 0.1 * (var * n_el / max(n_el - 1, 1)).float()
            invstd = (1.0 / torch.sqrt(var + EPS)).float()
            mean = mean.float()
        else:
            mean = running[k - 1, 0]
            invstd = 1.0 / torch.sqrt(running[k - 1, 1] + EPS)
        scale = gamma * invstd
        shift = beta - mean * scale
        yk = zk * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)
        z.append(zk)
        coef.append((scale, shift, mean, invstd, yk > 0))
        a.append(bfa(F.relu(yk)) if (k < 15 or ROUND["last"]) else F.relu(yk))   # the 64->1 layer consumes BN+ReLU of z_15 in FP32
    if ROUND["last"]:
        w16 = bfw(w16)
    noise = F.conv2d(a[15], w16, padding=1)
    out = {"y": x - noise if residual else noise, "running": new_running}
    if dy is None:
        return out
    dn = -dy.float() if residual else dy.float()
    grads = torch.zeros_like(flat, dtype=torch.float32)
    gw0, gmid, gw16 = unpack(grads)
    gw16.copy_(torch.nn.grad.conv2d_weight(a[15], w16.shape, dn, padding=1))
    g = bfg(torch.nn.grad.conv2d_input(a[15].shape, w16, dn, padding=1) * coef[15][4])
    for k in range(15, 0, -1):
        w, gamma, beta = mid[k - 1]
        scale, shift, mean, invstd, _ = coef[k]
        xhat = (z[k] - mean.view(1, -1, 1, 1)) * invstd.view(1, -1, 1, 1)
        dbeta = g.double().sum((0, 2, 3))
        dgamma = (g * xhat).double().sum((0, 2, 3))
        A = (gamma * invstd).double()
        Bc = -A * invstd.double() * dgamma / n_el
        Cc = A * (-dbeta / n_el + mean.double() * invstd.double() * dgamma / n_el)
        dz = bfg(A.float().view(1, -1, 1, 1) * g + (Bc.float().view(1, -1, 1, 1) * z[k] + Cc.float().view(1, -1, 1, 1)))
        gmid[k - 1][0].copy_(torch.nn.grad.conv2d_weight(a[k - 1], w.shape, dz, padding=1))
        gmid[k - 1][1].copy_(dgamma.float())
        gmid[k - 1][2].copy_(dbeta.float())
        gin = torch.nn.grad.conv2d_input(a[k - 1].shape, bfw(w), dz, padding=1)
        mask = coef[k - 1][4] if k >= 2 else (a[0] > 0)
        g = bfg(gin * mask)
    gw0.copy_(torch.nn.grad.conv2d_weight(x, w0.shape, g, padding=1))
    out["grads"] = grads
    return out
