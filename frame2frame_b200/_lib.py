"""ctypes binding of libf2f_b200.so (include/f2f_b200.h).

The product path has no CPU fallback: if the library is missing or a call
fails, a RuntimeError is raised.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# F2F_LIB points at another build of the same library (the storage-format variants of build.py, for A/B runs)
LIB_PATH = os.environ.get("F2F_LIB") or os.path.join(HERE, "lib", "libf2f_b200.so")

_lib = None


class Tvl1Params(C.Structure):
    """struct f2f_tvl1_params (include/f2f_b200.h); defaults of main.c:25-35."""
    _fields_ = [("tau", C.c_float), ("lam", C.c_float), ("theta", C.c_float),
                ("nscales", C.c_int), ("fscale", C.c_int), ("zfactor", C.c_float),
                ("nwarps", C.c_int), ("epsilon", C.c_float), ("verbose", C.c_int)]


def _declare(lib):
    vp, fp, ip, sz = C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t
    lib.f2f_last_error.restype = C.c_char_p
    lib.f2f_tvl1_default_params.argtypes = [C.POINTER(Tvl1Params)]
    lib.f2f_tvl1_default_params.restype = None
    lib.f2f_tvl1_validate_params.argtypes = [C.POINTER(Tvl1Params), C.c_int, C.c_int]
    lib.f2f_tvl1_zoom_size.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_float]
    lib.f2f_tvl1_zoom_size.restype = None
    lib.f2f_tvl1_workspace_bytes.argtypes = [C.c_int, C.c_int, C.POINTER(Tvl1Params), C.c_int]
    lib.f2f_tvl1_workspace_bytes.restype = sz
    lib.f2f_tvl1_flow_batch.argtypes = [fp, fp, fp, C.c_int, C.c_int, C.c_int, C.POINTER(Tvl1Params), ip, vp, sz, vp]
    lib.f2f_tvl1_flow_bidir.argtypes = [fp, fp, fp, C.c_int, C.c_int, C.c_int, C.POINTER(Tvl1Params), ip, vp, sz, vp]
    lib.f2f_tvl1_flow_batch_host.argtypes = [fp, fp, fp, C.c_int, C.c_int, C.c_int, C.POINTER(Tvl1Params), ip, C.c_int]
    lib.f2f_tvl1_gaussian.argtypes = [fp, C.c_int, C.c_int, C.c_double, fp, vp]
    lib.f2f_tvl1_gaussian_noalloc.argtypes = [fp, C.c_int, C.c_int, C.c_double, vp]
    lib.f2f_tvl1_zoom_out.argtypes = [fp, fp, C.c_int, C.c_int, C.c_float, fp, vp]
    lib.f2f_tvl1_zoom_in_scaled.argtypes = [fp, fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp]
    lib.f2f_tvl1_centered_gradient.argtypes = [fp, fp, fp, C.c_int, C.c_int, vp]
    lib.f2f_tvl1_normalize.argtypes = [fp, fp, fp, fp, C.c_int, C.c_int, vp, vp]
    lib.f2f_selftest_division.argtypes = [C.c_ulonglong, C.POINTER(C.c_ulonglong)]
    lib.f2f_selftest_hypot.argtypes = [C.c_ulonglong, C.POINTER(C.c_ulonglong)]
    lib.f2f_warp_mask.argtypes = [fp, fp, fp, fp, fp, fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.f2f_occlusion_mask.argtypes = [fp, fp, fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.f2f_masked_loss.argtypes = [fp, fp, fp, fp, vp, sz, C.c_int, vp]
    lib.f2f_dncnn_prep_w64.argtypes = [fp, vp, C.c_int, vp]
    lib.f2f_dncnn_conv64.argtypes = [vp, vp, vp, fp, fp, vp, C.c_int, C.c_int, vp]
    lib.f2f_dncnn_dgrad64.argtypes = [vp, vp, vp, vp, fp, fp, fp, fp, vp, C.c_int, C.c_int, vp]
    lib.f2f_dncnn_wgrad64.argtypes = [vp, fp, vp, vp, fp, fp, fp, C.c_int, C.c_int, C.c_int, vp]
    lib.f2f_dncnn_param_count.restype = sz
    lib.f2f_dncnn_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
    lib.f2f_dncnn_workspace_bytes.restype = sz
    lib.f2f_dncnn_workspace_layout.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(sz)]
    lib.f2f_dncnn_workspace_layout.restype = None
    lib.f2f_dncnn_set_debug_dump.argtypes = [vp]
    lib.f2f_dncnn_set_debug_dump.restype = None
    lib.f2f_dncnn_set_trace.argtypes = [vp]
    lib.f2f_dncnn_set_trace.restype = None
    lib.f2f_set_pdl.argtypes = [C.c_int]
    lib.f2f_set_pdl.restype = None
    lib.f2f_set_serpentine.argtypes = [C.c_int]
    lib.f2f_set_serpentine.restype = None
    lib.f2f_set_conv_pair.argtypes = [C.c_int]
    lib.f2f_set_conv_pair.restype = None
    lib.f2f_dncnn_storage_dtype.restype = C.c_int
    lib.f2f_dncnn_set_precise.argtypes = [C.c_int]
    lib.f2f_dncnn_set_precise.restype = None
    lib.f2f_dncnn_get_precise.restype = C.c_int
    lib.f2f_dncnn_forward.argtypes = [fp, fp, vp, fp, fp, vp, sz, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.f2f_dncnn_backward.argtypes = [fp, fp, fp, vp, sz, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.f2f_dncnn_param_count_c.argtypes = [C.c_int]
    lib.f2f_dncnn_param_count_c.restype = sz
    lib.f2f_dncnn_workspace_bytes_c.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.f2f_dncnn_workspace_bytes_c.restype = sz
    lib.f2f_dncnn_forward_c.argtypes = [fp, fp, vp, fp, fp, vp, sz, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.f2f_dncnn_backward_c.argtypes = [fp, fp, fp, vp, sz, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.f2f_adam_step.argtypes = [fp, fp, fp, fp, sz, C.c_longlong, C.c_float, C.c_float, C.c_float, C.c_float,
                                  C.c_float, vp]
    lib.f2f_step_train.argtypes = [fp, fp, fp, fp, fp, vp, fp, fp, fp, fp, fp, vp, vp, sz, C.c_int, C.c_int, C.c_int,
                                   C.c_longlong, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, vp]
    lib.f2f_profile_dncnn.argtypes = [C.c_int]
    lib.f2f_profile_dncnn.restype = None
    lib.f2f_profile_dncnn_read.argtypes = [C.POINTER(C.c_double)]
    lib.f2f_profile_dncnn_read.restype = None
    lib.f2f_profile_enable.argtypes = [C.c_int]
    lib.f2f_profile_enable.restype = None
    lib.f2f_profile_read.argtypes = [C.POINTER(C.c_double)]
    lib.f2f_profile_read.restype = None
    lib.f2f_release.restype = None


def lib():
    """The loaded library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libf2f_b200.so is not built (%s); run `python -m frame2frame_b200.build`. "
                "There is no CPU fallback." % LIB_PATH)
        _lib = C.CDLL(LIB_PATH)
        _declare(_lib)
    return _lib


def storage_dtype():
    """torch dtype of the DnCNN activation / operand buffers the library was built with (f2f_dncnn_storage_dtype)."""
    import torch
    return torch.float16 if lib().f2f_dncnn_storage_dtype() else torch.bfloat16


def check(rc, what):
    if rc != 0:
        msg = lib().f2f_last_error()
        raise RuntimeError("%s failed (code %d): %s" % (what, rc, msg.decode() if msg else ""))


def exported_symbols():
    """Names declared in include/f2f_b200.h (parsed), for the ABI test."""
    import re
    hdr = os.path.join(os.path.dirname(HERE), "include", "f2f_b200.h")
    text = open(hdr).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(f2f_\w+|Dual_TVL1_optic_flow_multiscale)\s*\(", text)
    return sorted(set(n for n in names if not n.startswith("f2f_tvl1_params")))
