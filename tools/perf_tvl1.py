"""Quick device-side timing of the TV-L1 batch (development aid, not the bench)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from frame2frame_b200 import synth
from frame2frame_b200.tvl1 import tvl1flow

H, W = 1080, 1920
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
slots = int(sys.argv[2]) if len(sys.argv) > 2 else 16
NS = int(sys.argv[3]) if len(sys.argv) > 3 else 5
dev = torch.device("cuda:0")
base = [synth.noisy_frame(H, W, t) for t in range(5)]
I0 = torch.from_numpy(np.stack([base[(k % 4) + 1] for k in range(n)])).to(dev)
I1 = torch.from_numpy(np.stack([base[(k % 4)] for k in range(n)])).to(dev)
flow, iters = tvl1flow(I0, I1, nslots=slots, return_iters=True, nscales=NS)
torch.cuda.synchronize()
print("iters pair0:", iters[0].tolist(), "sum per scale", iters[0].sum(1).tolist())
P = [1920*1080, 960*540, 480*270, 240*135, 120*68]
pxit = sum(P[s] * iters[:, s].sum() for s in range(NS)) / n
print("px-iter per field: %.1f M" % (pxit / 1e6))
for rep in range(int(os.environ.get('REPS', '3'))):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    flow = tvl1flow(I0, I1, nslots=slots, nscales=NS)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    bytes_alg = pxit * 64 + sum(P) * (12 + 40 * 5)
    print("rep %d: %.2f ms/field, %.1f fields/s, alg %.2f GB/field -> %.0f GB/s" % (rep, ms / n, n / ms * 1e3, bytes_alg / 1e9, bytes_alg * n / ms / 1e6))
