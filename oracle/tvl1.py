"""oracle/tvl1.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes front-ends for the two CPU checkers of the TV-L1 path:

* ``RefTvl1``    -- the reference's own sources compiled in place
                    (oracle/_ref/libtvl1_ref.so; tvl1flow/tvl1flow_lib.c:356).
* ``OracleTvl1`` -- our C restatement (oracle/tvl1_oracle.c), which also
                    reports per-(scale, warp) iteration counts.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs
import this module.  Nothing under frame2frame_b200/ may.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_FP = C.POINTER(C.c_float)
_IP = C.POINTER(C.c_int)

# main.c:25-35
DEFAULTS = dict(tau=0.25, lam=0.15, theta=0.3, nscales=100, fscale=0,
                zfactor=0.5, nwarps=5, epsilon=0.01)


def build(force=False):
    """Compile the restatement (and _ref when the reference tree is present)."""
    so = os.path.join(HERE, "liboracle_tvl1.so")
    src = os.path.join(HERE, "tvl1_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "--no-print-directory"],
                              stdout=subprocess.DEVNULL)
    return so


def _fp(a):
    return a.ctypes.data_as(_FP)


def clamp_scales(nx, ny, zfactor, nscales, fscale):
    """main.c:159-163 as executed by the restatement (float N, int truncation)."""
    lib = C.CDLL(build())
    ns, fs = C.c_int(nscales), C.c_int(fscale)
    lib.tvo_clamp_scales(C.c_int(nx), C.c_int(ny), C.c_float(zfactor),
                         C.byref(ns), C.byref(fs))
    return ns.value, fs.value


class OracleTvl1:
    def __init__(self):
        self.lib = C.CDLL(build())
        self.lib.tvo_bicubic_at.restype = C.c_float
        self.lib.tvo_multiscale.restype = C.c_int
        self.lib.tvo_gaussian.restype = C.c_int
        self.lib.tvo_zoom_out.restype = C.c_int

    def flow(self, I0, I1, tau=0.25, lam=0.15, theta=0.3, nscales=100, fscale=0,
             zfactor=0.5, nwarps=5, epsilon=0.01, clamp=True):
        """Returns (flow[2,H,W] float32, iters[nscales,nwarps] int32, errs)."""
        I0 = np.ascontiguousarray(I0, np.float32)
        I1 = np.ascontiguousarray(I1, np.float32)
        ny, nx = I0.shape
        if clamp:
            nscales, fscale = clamp_scales(nx, ny, zfactor, nscales, fscale)
        u = np.zeros((2, ny, nx), np.float32)
        iters = np.zeros((nscales, nwarps), np.int32)
        errs = np.zeros((nscales, nwarps), np.float32)
        rc = self.lib.tvo_multiscale(
            _fp(I0), _fp(I1), _fp(u[0]), _fp(u[1]), nx, ny, C.c_float(tau),
            C.c_float(lam), C.c_float(theta), nscales, fscale, C.c_float(zfactor),
            nwarps, C.c_float(epsilon), iters.ctypes.data_as(_IP), _fp(errs))
        if rc:
            raise RuntimeError("tvo_multiscale failed (%d): the reference would abort" % rc)
        return u, iters, errs

    # --- per-stage entry points for unit tests -----------------------------
    def gaussian(self, I, sigma):
        I = np.array(I, np.float32, order="C")
        rc = self.lib.tvo_gaussian(_fp(I), I.shape[1], I.shape[0], C.c_double(sigma))
        if rc:
            raise RuntimeError("sigma too large")
        return I

    def zoom_out(self, I, factor):
        I = np.ascontiguousarray(I, np.float32)
        ny, nx = I.shape
        nxx, nyy = C.c_int(), C.c_int()
        self.lib.tvo_zoom_size(nx, ny, C.byref(nxx), C.byref(nyy), C.c_float(factor))
        out = np.zeros((nyy.value, nxx.value), np.float32)
        if self.lib.tvo_zoom_out(_fp(I), _fp(out), nx, ny, C.c_float(factor)):
            raise RuntimeError("sigma too large")
        return out

    def zoom_in_scaled(self, I, nxx, nyy, zfactor):
        I = np.ascontiguousarray(I, np.float32)
        ny, nx = I.shape
        out = np.zeros((nyy, nxx), np.float32)
        self.lib.tvo_zoom_in_scaled(_fp(I), _fp(out), nx, ny, nxx, nyy, C.c_float(zfactor))
        return out

    def centered_gradient(self, I):
        I = np.ascontiguousarray(I, np.float32)
        dx, dy = np.zeros_like(I), np.zeros_like(I)
        self.lib.tvo_centered_gradient(_fp(I), _fp(dx), _fp(dy), I.shape[1], I.shape[0])
        return dx, dy

    def normalize(self, I0, I1):
        I0 = np.ascontiguousarray(I0, np.float32)
        I1 = np.ascontiguousarray(I1, np.float32)
        a, b = np.zeros_like(I0), np.zeros_like(I1)
        self.lib.tvo_normalize(_fp(I0), _fp(I1), _fp(a), _fp(b), I0.size)
        return a, b

    def solve_scale(self, I0, I1, u1, u2, tau=0.25, lam=0.15, theta=0.3, nwarps=5,
                    epsilon=0.01):
        I0 = np.ascontiguousarray(I0, np.float32)
        I1 = np.ascontiguousarray(I1, np.float32)
        u1 = np.array(u1, np.float32, order="C")
        u2 = np.array(u2, np.float32, order="C")
        ny, nx = I0.shape
        iters = np.zeros(nwarps, np.int32)
        errs = np.zeros(nwarps, np.float32)
        self.lib.tvo_solve_scale(_fp(I0), _fp(I1), _fp(u1), _fp(u2), nx, ny,
                                 C.c_float(tau), C.c_float(lam), C.c_float(theta),
                                 nwarps, C.c_float(epsilon),
                                 iters.ctypes.data_as(_IP), _fp(errs))
        return u1, u2, iters, errs


class RefTvl1:
    """The reference library itself (present only where oracle/_ref was built)."""

    path = os.path.join(HERE, "_ref", "libtvl1_ref.so")

    @classmethod
    def available(cls):
        return os.path.exists(cls.path)

    def __init__(self):
        self.lib = C.CDLL(self.path)
        self.lib.bicubic_interpolation_at.restype = C.c_float

    @staticmethod
    def set_threads(n):
        """main.c:143-146: omp_set_num_threads(nproc).  Explicit, because a launcher (torchrun) may have exported
        OMP_NUM_THREADS=1 before this process started."""
        for name in ("libgomp.so.1", "libgomp.so"):
            try:
                C.CDLL(name).omp_set_num_threads(int(n))
                return True
            except OSError:
                continue
        return False

    def flow(self, I0, I1, tau=0.25, lam=0.15, theta=0.3, nscales=100, fscale=0,
             zfactor=0.5, nwarps=5, epsilon=0.01, clamp=True):
        I0 = np.array(I0, np.float32, order="C")
        I1 = np.array(I1, np.float32, order="C")
        ny, nx = I0.shape
        if clamp:
            nscales, fscale = clamp_scales(nx, ny, zfactor, nscales, fscale)
        u = np.zeros((2, ny, nx), np.float32)
        self.lib.Dual_TVL1_optic_flow_multiscale(
            _fp(I0), _fp(I1), _fp(u[0]), _fp(u[1]), nx, ny, C.c_float(tau),
            C.c_float(lam), C.c_float(theta), nscales, fscale, C.c_float(zfactor),
            nwarps, C.c_float(epsilon), C.c_bool(False))
        return u

    def gaussian(self, I, sigma):
        I = np.array(I, np.float32, order="C")
        self.lib.gaussian(_fp(I), I.shape[1], I.shape[0], C.c_double(sigma))
        return I

    def zoom_out(self, I, factor):
        I = np.ascontiguousarray(I, np.float32)
        ny, nx = I.shape
        nxx, nyy = C.c_int(), C.c_int()
        self.lib.zoom_size(nx, ny, C.byref(nxx), C.byref(nyy), C.c_float(factor))
        out = np.zeros((nyy.value, nxx.value), np.float32)
        self.lib.zoom_out(_fp(I), _fp(out), nx, ny, C.c_float(factor))
        return out

    def zoom_in(self, I, nxx, nyy):
        I = np.ascontiguousarray(I, np.float32)
        ny, nx = I.shape
        out = np.zeros((nyy, nxx), np.float32)
        self.lib.zoom_in(_fp(I), _fp(out), nx, ny, nxx, nyy)
        return out

    def centered_gradient(self, I):
        I = np.ascontiguousarray(I, np.float32)
        dx, dy = np.zeros_like(I), np.zeros_like(I)
        self.lib.centered_gradient(_fp(I), _fp(dx), _fp(dy), I.shape[1], I.shape[0])
        return dx, dy

    def normalize(self, I0, I1):
        I0 = np.ascontiguousarray(I0, np.float32)
        I1 = np.ascontiguousarray(I1, np.float32)
        a, b = np.zeros_like(I0), np.zeros_like(I1)
        self.lib.image_normalization(_fp(I0), _fp(I1), _fp(a), _fp(b), I0.size)
        return a, b

    def solve_scale(self, I0, I1, u1, u2, tau=0.25, lam=0.15, theta=0.3, nwarps=5,
                    epsilon=0.01):
        I0 = np.array(I0, np.float32, order="C")
        I1 = np.array(I1, np.float32, order="C")
        u1 = np.array(u1, np.float32, order="C")
        u2 = np.array(u2, np.float32, order="C")
        ny, nx = I0.shape
        self.lib.Dual_TVL1_optic_flow(_fp(I0), _fp(I1), _fp(u1), _fp(u2), nx, ny,
                                      C.c_float(tau), C.c_float(lam), C.c_float(theta),
                                      nwarps, C.c_float(epsilon), C.c_bool(False))
        return u1, u2

    def bicubic_warp(self, I, u, v, border_out=True):
        I = np.ascontiguousarray(I, np.float32)
        u = np.ascontiguousarray(u, np.float32)
        v = np.ascontiguousarray(v, np.float32)
        out = np.zeros_like(I)
        self.lib.bicubic_interpolation_warp(_fp(I), _fp(u), _fp(v), _fp(out),
                                            I.shape[1], I.shape[0], C.c_bool(border_out))
        return out
