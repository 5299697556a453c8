"""wg::k_wgrad64 alone at 1080p: average launch time over back-to-back launches (development aid).  With a
-DF2F_WG_DEBUG build of the library (F2F_LIB=...) the F2F_WGRAD_DBG bit mask knocks out parts of the kernel -- MMAs,
loads, write-back, shared-memory stores, transform arithmetic -- to see what the launch time is made of.

    python tools/perf_wgrad.py [H W [launches]]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from frame2frame_b200 import _lib

H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1080, 1920)
n = int(sys.argv[3]) if len(sys.argv) > 3 else 20
lib = _lib.lib()
dev = torch.device("cuda:0")
H16 = torch.float16 if lib.f2f_dncnn_storage_dtype() == 1 else torch.bfloat16
torch.manual_seed(1)
# several distinct layers' worth of tensors so that consecutive launches do not find their inputs in L2
L = 4
acts = [torch.randn(H, W, 64, device=dev).to(H16) for _ in range(L)]
gs = [(torch.randn(H, W, 64, device=dev) * 0.1).to(H16) for _ in range(L)]
zs = [torch.randn(H, W, 64, device=dev).to(H16) for _ in range(L)]
coef_in = torch.zeros(448, device=dev); coef_in[0:64] = 1.0
coef = torch.zeros(448, device=dev); coef[256:320] = 1.0; coef[320:384] = 0.01
partial = torch.zeros(torch.cuda.get_device_properties(dev).multi_processor_count * 49152, device=dev)
dw = torch.zeros(64, 64, 3, 3, device=dev)


def launch(i):
    k = i % L
    _lib.check(lib.f2f_dncnn_wgrad64(acts[k].data_ptr(), coef_in.data_ptr(), gs[k].data_ptr(), zs[k].data_ptr(), coef.data_ptr(),
                                     partial.data_ptr(), dw.data_ptr(), H, W, 0, None), "wgrad64")


for i in range(4):
    launch(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(n):
    launch(i)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / n * 1e3
print("wgrad64 %dx%d: %.1f us per launch incl. the partial reduction (%d back to back), DBG=%s PF=%s"
      % (W, H, us, n, os.environ.get("F2F_WGRAD_DBG", "0"), os.environ.get("F2F_WGRAD_PF", "default")))

if os.environ.get("F2F_WGRAD_TRACE"):
    tr = torch.zeros(16, dtype=torch.int64, device=dev)
    lib.f2f_dncnn_set_trace(tr.data_ptr())
    launch(0)
    torch.cuda.synchronize()
    lib.f2f_dncnn_set_trace(None)
    t = tr.cpu().tolist()
    names = ["issue loads", "wait free stage", "loads landing", "transform+stores", "fence+arrive"]
    for role, o, n in (("act producer", 0, t[12]), ("dz producer ", 5, t[13])):
        n = max(n, 1)
        print("  %s (group 0, per tile of the group): " % role + ", ".join("%s %d" % (names[i], t[o + i] / n) for i in range(5))
              + "  = %d cycles, %d tiles" % (sum(t[o:o + 5]) / n, n))
    n = max(t[14], 1)
    print("  MMA warp (per tile): wait full stage %d, issue %d = %d cycles, %d tiles" % (t[10] / n, t[11] / n, (t[10] + t[11]) / n, n))
