"""Development check of the tcgen05 conv kernel against torch conv2d (GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
from frame2frame_b200 import _lib
lib = _lib.lib()
H16 = _lib.storage_dtype()
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
torch.manual_seed(0)

def run(H, W, bn, stats):
    x = torch.randn(1, 64, H, W, device=dev)
    w = torch.randn(64, 64, 3, 3, device=dev) * 0.06
    xb = x.to(H16)
    x_nhwc = xb.permute(0, 2, 3, 1).contiguous()           # [1,H,W,64] bf16
    wprep = torch.empty(9 * 64 * 64, dtype=H16, device=dev)
    _lib.check(lib.f2f_dncnn_prep_w64(w.data_ptr(), wprep.data_ptr(), 0, None), "prep")
    out = torch.full((1, H, W, 64), float("nan"), dtype=H16, device=dev)
    scale = (torch.rand(64, device=dev) + 0.5) if bn else None
    shift = (torch.randn(64, device=dev) * 0.3) if bn else None
    st = torch.zeros(128, dtype=torch.float64, device=dev) if stats else None
    _lib.check(lib.f2f_dncnn_conv64(x_nhwc.data_ptr(), out.data_ptr(), wprep.data_ptr(),
                                    scale.data_ptr() if bn else None, shift.data_ptr() if bn else None,
                                    st.data_ptr() if stats else None, H, W, None), "conv64")
    torch.cuda.synchronize()
    a = xb.float()
    if bn:
        a = torch.relu(a * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)).to(H16).float()
    ref = torch.nn.functional.conv2d(a, w.to(H16).float(), padding=1)
    got = out.permute(0, 3, 1, 2).float()
    err = (got - ref).abs().max().item()
    tol = ref.abs().max().item() * 2 ** -7
    msg = "H=%d W=%d bn=%d stats=%d: max err %.4g (tol %.4g, ref max %.3g)" % (H, W, bn, stats, err, tol, ref.abs().max().item())
    if stats:
        gb = got.double()
        s_err = (st[:64] - gb.sum((0, 2, 3))).abs().max().item() / (gb.abs().sum((0, 2, 3)).max().item())
        q_err = ((st[64:] - (gb * gb).sum((0, 2, 3))).abs() / (gb * gb).sum((0, 2, 3))).max().item()
        msg += " stats rel err %.3g %.3g" % (s_err, q_err)
    print(msg, "OK" if err <= tol and not np.isnan(err) else "FAIL")
    return err <= tol

ok = True
for (H, W) in [(16, 16), (32, 48), (40, 72), (270, 480)]:
    for bn in (0, 1):
        for stats in (0, 1):
            ok &= run(H, W, bn, stats)
# timing at 1080p
H, W = 1080, 1920
x = torch.randn(1, H, W, 64, device=dev).to(H16)
w = torch.randn(64, 64, 3, 3, device=dev) * 0.06
wprep = torch.empty(9 * 64 * 64, dtype=H16, device=dev)
lib.f2f_dncnn_prep_w64(w.data_ptr(), wprep.data_ptr(), 0, None)
out = torch.empty_like(x)
scale = torch.ones(64, device=dev); shift = torch.zeros(64, device=dev)
st = torch.zeros(128, dtype=torch.float64, device=dev)
for _ in range(3):
    lib.f2f_dncnn_conv64(x.data_ptr(), out.data_ptr(), wprep.data_ptr(), scale.data_ptr(), shift.data_ptr(), st.data_ptr(), H, W, None)
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(10):
    lib.f2f_dncnn_conv64(x.data_ptr(), out.data_ptr(), wprep.data_ptr(), scale.data_ptr(), shift.data_ptr(), st.data_ptr(), H, W, None)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print("1080p conv64 (BN+ReLU prologue, stats): %.3f ms -> %.1f TFLOP/s" % (ms, 2 * 9 * 64 * 64 * H * W / ms / 1e9))
print("ALL OK" if ok else "SOME FAILED")
