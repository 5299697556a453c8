"""The reference's two TV-L1 command lines on B200.

``python -m frame2frame_b200.tvl1flow_cli I0 I1 [out nproc tau lambda theta nscales fscale zfactor nwarps epsilon
verbose]`` -- the 13 positionals and defaults of tvl1flow/main.c:86-99,25-35 (``nproc`` is accepted and ignored: the
device has no OpenMP thread count); reads two images, writes a Middlebury ``.flo`` (iio.c:2114-2129 layout).

``python -m frame2frame_b200.tvl1flow_cli --sequence I F L O`` -- tvl1flow/tvl1flow.sh:1-19: for i in F..L the
backward flow (I0 = frame i+1, I1 = frame i) to ``O % (i+1) + "_bflow.flo"`` and the forward flow to
``O % i + "_fflow.flo"``, with the script's parameters (lambda 0.2, nscales 100 -> clamp, fscale 2).  All 2*(L-F+1)
pairs go through ONE batched call (frame pairs are independent; slots advance in lock-step on the device) instead
of one process per pair.  Started under torchrun, every rank takes a contiguous block of the pairs on its own GPU.
"""
import os
import sys

import numpy as np
import torch

from .tvl1 import tvl1flow, tvl1flow_bidir, write_flo

DEFAULTS = [("out", str, "flow.flo"), ("nproc", int, 4), ("tau", float, 0.25), ("lam", float, 0.15),
            ("theta", float, 0.3), ("nscales", int, 100), ("fscale", int, 0), ("zfactor", float, 0.5),
            ("nwarps", int, 5), ("epsilon", float, 0.01), ("verbose", int, 0)]
SCRIPT_PARAMS = dict(tau=0.25, lam=0.2, theta=0.3, nscales=100, fscale=2, zfactor=0.5, nwarps=5, epsilon=0.01)


def read_gray(path):
    """iio_read_image_float + the gray conversion of iio.c:808-838 (.299 R + .587 G + .114 B); 8-bit images
    come back as floats in 0..255 like the reference reads them."""
    if path.endswith(".npy"):
        return np.load(path).astype(np.float32)
    from PIL import Image
    a = np.array(Image.open(path)).astype(np.float32)
    if a.ndim == 3:
        a = a[..., 0] * np.float32(0.299) + a[..., 1] * np.float32(0.587) + a[..., 2] * np.float32(0.114) \
            if a.shape[2] >= 3 else a[..., 0]
    return np.ascontiguousarray(a, np.float32)


def parse_positionals(argv):
    if len(argv) < 2:
        raise SystemExit("Usage: tvl1flow I0 I1 [out nproc tau lambda theta nscales fscale zfactor nwarps epsilon verbose]")
    vals = {"I0": argv[0], "I1": argv[1]}
    for k, (name, conv, default) in enumerate(DEFAULTS):
        vals[name] = conv(argv[2 + k]) if len(argv) > 2 + k else default
    return vals


def plan_sequence(first, last, out):
    """tvl1flow.sh:9-19 as a list of (index of I0, index of I1, output file): for every i in first..last the backward
    flow (I0 = frame i+1, I1 = frame i) to out % (i+1) + "_bflow.flo", then the forward flow to out % i + "_fflow.flo"."""
    jobs = []
    for i in range(first, last + 1):
        jobs.append((i + 1, i, (out % (i + 1)) + "_bflow.flo"))
        jobs.append((i, i + 1, (out % i) + "_fflow.flo"))
    return jobs


def shard_jobs(jobs, rank, world):
    """Frame pairs are independent: rank `rank` of `world` takes a contiguous block of the job list."""
    base, extra = divmod(len(jobs), world)
    lo = rank * base + min(rank, extra)
    return jobs[lo:lo + base + (1 if rank < extra else 0)]


def main(argv=None, device=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    # under torchrun every rank computes its own block of pairs on GPU LOCAL_RANK; nothing is exchanged
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))) if device is None else torch.device(device)
    if argv and argv[0] == "--sequence":
        pattern, first, last, out = argv[1], int(argv[2]), int(argv[3]), argv[4]
        jobs = shard_jobs(plan_sequence(first, last, out), rank, world)
        if not jobs:
            return 0
        frames = {i: read_gray(pattern % i) for i in sorted({a for a, _, _ in jobs} | {b for _, b, _ in jobs})}
        # the script's two flows of a frame pair (bflow then fflow, adjacent in the job list) share their pyramids
        k = 0
        pairs, singles = [], []
        while k < len(jobs):
            if k + 1 < len(jobs) and jobs[k][0] == jobs[k + 1][1] and jobs[k][1] == jobs[k + 1][0]:
                pairs.append((jobs[k], jobs[k + 1]))
                k += 2
            else:
                singles.append(jobs[k])
                k += 1
        if pairs:
            A = torch.from_numpy(np.stack([frames[j0[0]] for j0, _ in pairs])).to(dev)
            B = torch.from_numpy(np.stack([frames[j0[1]] for j0, _ in pairs])).to(dev)
            fab, fba = tvl1flow_bidir(A, B, **SCRIPT_PARAMS)
            for (j0, j1), f0, f1 in zip(pairs, fab, fba):
                write_flo(j0[2], f0)
                write_flo(j1[2], f1)
        if singles:                                        # a rank's block may start or end in the middle of a pair
            flow = tvl1flow(torch.from_numpy(np.stack([frames[a] for a, _, _ in singles])).to(dev),
                            torch.from_numpy(np.stack([frames[b] for _, b, _ in singles])).to(dev), **SCRIPT_PARAMS)
            for f, (_, _, name) in zip(flow, singles):
                write_flo(name, f)
        return 0
    v = parse_positionals(argv)
    I0, I1 = read_gray(v["I0"]), read_gray(v["I1"])
    if I0.shape != I1.shape:
        sys.stderr.write("ERROR: input images size mismatch %dx%d != %dx%d\n" % (I0.shape[1], I0.shape[0], I1.shape[1], I1.shape[0]))
        return 1                                                   # main.c:189-193
    kw = {k: v[k] for k in ("tau", "lam", "theta", "nscales", "fscale", "zfactor", "nwarps", "epsilon", "verbose")}
    flow = tvl1flow(torch.from_numpy(I0).to(dev), torch.from_numpy(I1).to(dev), **kw)
    write_flo(v["out"], flow)
    return 0


if __name__ == "__main__":
    sys.exit(main())
