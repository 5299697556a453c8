"""Builds libf2f_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m frame2frame_b200.build [--force] [--verbose]

The library is rebuilt only when a source is newer than the .so.  The .so is
git-ignored but travels with the gpurun snapshot.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libf2f_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
          "-I", INCLUDE]
if os.environ.get("F2F_STORE_BF16", "0") not in ("", "0"):
    COMMON.append("-DF2F_STORE_BF16")     # BF16 instead of FP16 storage of the DnCNN tensors (csrc/tc05.cuh)

# translation unit -> extra flags
#  tvl1.cu: -fmad=false, the reference C is built without FMA contraction (-std=c99)
#  and flow / iteration-count parity needs the same roundings.
UNITS = {
    "tvl1.cu": ["-fmad=false"],
    "f2f_step.cu": ["-fmad=false"],
    "dncnn.cu": [],
    "dncnn_net.cu": [],
}


def nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found")
    return exe


def sources():
    return [u for u in UNITS if os.path.exists(os.path.join(CSRC, u))]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(INCLUDE, "f2f_b200.h"), __file__]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(LIBDIR, "obj")
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    for u in sources():
        obj = os.path.join(objdir, u.replace(".cu", ".o"))
        cmd = [nvcc()] + ARCH + COMMON + UNITS[u] + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", os.path.join(CSRC, u), "-o", obj]
        procs.append((u, cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for u, cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode:
            sys.stderr.write(out)
            raise RuntimeError("nvcc failed on %s: %s" % (u, " ".join(cmd)))
        if verbose:
            print(out)
    cmd = [nvcc()] + ARCH + ["-shared", "-o", LIB] + objs
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
