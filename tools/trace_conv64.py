"""Per-role cycle breakdown of the tcgen05 conv kernel variants at 1080p (development aid, GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from frame2frame_b200 import _lib
lib = _lib.lib()
H16 = _lib.storage_dtype()
dev = torch.device("cuda:0")
H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1080, 1920)
torch.manual_seed(0)
x = torch.randn(1, H, W, 64, device=dev).to(H16)
z = torch.randn(1, H, W, 64, device=dev).to(H16)
out = torch.empty_like(x)
w = torch.randn(64, 64, 3, 3, device=dev) * 0.06
wprep = torch.empty(9 * 64 * 64, dtype=H16, device=dev)
lib.f2f_dncnn_prep_w64(w.data_ptr(), wprep.data_ptr(), 0, None)
ones = torch.ones(64, device=dev); zeros = torch.zeros(64, device=dev)
st = torch.zeros(128, dtype=torch.float64, device=dev)
tr = torch.zeros(16, dtype=torch.int64, device=dev)

def conv(bn, stats):
    return lambda: lib.f2f_dncnn_conv64(x.data_ptr(), out.data_ptr(), wprep.data_ptr(), ones.data_ptr() if bn else None,
                                        zeros.data_ptr() if bn else None, st.data_ptr() if stats else None, H, W, None)
def dgrad(bn):
    return lambda: lib.f2f_dncnn_dgrad64(x.data_ptr(), out.data_ptr(), wprep.data_ptr(), z.data_ptr(),
                                         ones.data_ptr() if bn else None, zeros.data_ptr() if bn else None,
                                         zeros.data_ptr() if bn else None, ones.data_ptr() if bn else None,
                                         st.data_ptr() if bn else None, H, W, None)
names = ["prod wait", "prod work", "mma w.acc", "mma w.tile", "mma issue", "epi wait", "epi work", "epi total", "tiles"]
for name, fn in [("conv raw/store", conv(0, 0)), ("conv raw/stats", conv(0, 1)), ("conv bn/store", conv(1, 0)),
                 ("conv bn/stats", conv(1, 1)), ("dgrad relu", dgrad(0)), ("dgrad bn", dgrad(1))]:
    for _ in range(2):
        fn()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(5):
        fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 5 * 1e3
    tr.zero_()
    lib.f2f_dncnn_set_trace(tr.data_ptr())
    fn(); torch.cuda.synchronize()
    lib.f2f_dncnn_set_trace(None)
    t = tr.cpu().tolist()
    tiles = max(t[8], 1)
    if t[9]:                      # TMA-fed kernel: producer = one thread over all tiles, epilogue thread 0 sees every other tile
        per = [t[0] / tiles, 0.0] + [v / tiles for v in t[2:5]] + [t[5] / t[9], t[6] / t[9], t[7] / tiles]
    else:
        ptiles = (tiles + 1) // 2     # the traced producer thread stages every other tile
        per = [t[0] / ptiles, t[1] / ptiles] + [v / tiles for v in t[2:8]]
    exec_cyc = t[12] / max(t[13], 1)          # issue of a tile's first MMA -> its completion seen by the epilogue
    mhz_single = t[10] / max(t[11], 1) * 1e3
    # the same with the trace armed over 20 back-to-back launches: the SM clock the kernel sustains (power cap)
    tr.zero_()
    lib.f2f_dncnn_set_trace(tr.data_ptr())
    e0.record()
    for _ in range(20):
        fn()
    e1.record(); torch.cuda.synchronize()
    lib.f2f_dncnn_set_trace(None)
    t2 = tr.cpu().tolist()
    us20 = e0.elapsed_time(e1) / 20 * 1e3
    mhz_sust = t2[10] / max(t2[11], 1) * 1e3
    print("%-15s %7.1f us %6.0f TFLOP/s | cycles/tile: " % (name, us, 2 * 9 * 64 * 64 * H * W / us / 1e6) +
          "  ".join("%s %.0f" % (n, v) for n, v in zip(names, per)) + "  (%d tiles)" % tiles +
          " | first MMA issued -> tile complete %.0f cycles" % exec_cyc +
          " | SM clock: isolated launch %.0f MHz, 20 back-to-back %.0f MHz (%.1f us/launch, %.0f cycles/tile)"
          % (mhz_single, mhz_sust, us20, (t2[2] + t2[3] + t2[4]) / max(t2[8], 1)))
