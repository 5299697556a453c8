"""CTA-pair (cta_group::2) variant of the 64->64 convolution kernels against the single-CTA one, on the GPU box:
the four forward variants and the two input-gradient variants, outputs compared BIT FOR BIT (same K order, FP32
accumulation in TMEM), statistics to 1e-12 relative (FP64 atomics, order differs); then 1080p timing of each mode.

    python tools/check_pair.py [quick] [notime]
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from frame2frame_b200 import _lib

lib = _lib.lib()
H16 = _lib.storage_dtype()
dev = torch.device("cuda:0")
torch.manual_seed(0)


def tensors(H, W):
    x = (torch.randn(1, H, W, 64, device=dev) * 0.7).to(H16)
    z = (torch.randn(1, H, W, 64, device=dev) * 0.7).to(H16)
    w = torch.randn(64, 64, 3, 3, device=dev) * 0.06
    wf = torch.empty(9 * 64 * 64, dtype=H16, device=dev)
    wt = torch.empty(9 * 64 * 64, dtype=H16, device=dev)
    _lib.check(lib.f2f_dncnn_prep_w64(w.data_ptr(), wf.data_ptr(), 0, None), "prep")
    _lib.check(lib.f2f_dncnn_prep_w64(w.data_ptr(), wt.data_ptr(), 1, None), "prep")
    c = [(torch.rand(64, device=dev) + 0.5), torch.randn(64, device=dev) * 0.3, torch.randn(64, device=dev) * 0.1,
         (torch.rand(64, device=dev) + 0.5)]
    return x, z, wf, wt, c


def forward(x, wf, c, bn, stats, H, W):
    out = torch.full_like(x, float("nan"))
    st = torch.zeros(128, dtype=torch.float64, device=dev)
    _lib.check(lib.f2f_dncnn_conv64(x.data_ptr(), out.data_ptr(), wf.data_ptr(), c[0].data_ptr() if bn else None,
                                    c[1].data_ptr() if bn else None, st.data_ptr() if stats else None, H, W, None), "conv64")
    return out, st


def dgrad(x, z, wt, c, bn, H, W):
    out = torch.full_like(x, float("nan"))
    st = torch.zeros(128, dtype=torch.float64, device=dev)
    p = [t.data_ptr() for t in c] if bn else [None] * 4
    _lib.check(lib.f2f_dncnn_dgrad64(x.data_ptr(), out.data_ptr(), wt.data_ptr(), z.data_ptr(), p[0], p[1], p[2], p[3],
                                     st.data_ptr() if bn else None, H, W, None), "dgrad64")
    return out, st


def both(fn):
    lib.f2f_set_conv_pair(0)
    a = fn()
    torch.cuda.synchronize()
    lib.f2f_set_conv_pair(1)
    b = fn()
    torch.cuda.synchronize()
    lib.f2f_set_conv_pair(0)
    return a, b


ok = True
sizes = [(16, 8), (5, 7), (48, 64), (97, 131), (270, 480)] + ([] if "quick" in sys.argv else [(1080, 1920)])
for H, W in sizes:
    x, z, wf, wt, c = tensors(H, W)
    cases = [("fwd bn=%d stats=%d" % (bn, st), (lambda bn=bn, st=st: forward(x, wf, c, bn, st, H, W))) for bn in (0, 1) for st in (0, 1)]
    cases += [("dgrad bn=%d" % bn, (lambda bn=bn: dgrad(x, z, wt, c, bn, H, W))) for bn in (0, 1)]
    for name, fn in cases:
        (o0, s0), (o1, s1) = both(fn)
        same = torch.equal(o0.view(torch.int16), o1.view(torch.int16)) and not torch.isnan(o1.float()).any().item()
        srel = float(((s0 - s1).abs() / (s0.abs() + 1e-30)).max())
        good = same and srel < 1e-9
        ok &= good
        print("%4dx%-4d %-18s outputs %s  stats rel diff %.1e  %s" % (H, W, name, "identical" if same else "DIFFER", srel,
                                                                    "OK" if good else "FAIL"), flush=True)

H, W = 1080, 1920
x, z, wf, wt, c = tensors(H, W)
out = torch.empty_like(x)
st = torch.zeros(128, dtype=torch.float64, device=dev)
p = [t.data_ptr() for t in c]
timed = [("forward bn+stats", lambda: lib.f2f_dncnn_conv64(x.data_ptr(), out.data_ptr(), wf.data_ptr(), p[0], p[1], st.data_ptr(), H, W, None)),
         ("dgrad bn", lambda: lib.f2f_dncnn_dgrad64(x.data_ptr(), out.data_ptr(), wt.data_ptr(), z.data_ptr(), p[0], p[1], p[2], p[3],
                                                    st.data_ptr(), H, W, None))]
for name, fn in ([] if "notime" in sys.argv else timed):
    for mode in (0, 1, 0, 1):
        lib.f2f_set_conv_pair(mode)
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print("1080p %-18s pair=%d: %.1f us per launch (20 back to back)" % (name, mode, e0.elapsed_time(e1) * 50), flush=True)
lib.f2f_set_conv_pair(0)
print("ALL OK" if ok else "SOME FAILED")
sys.exit(0 if ok else 1)
