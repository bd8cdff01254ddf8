"""ctypes binding of ``libmia_b200.so`` (``include/mia_b200.h``).

There is NO CPU fallback: if the shared library is missing this module raises, loudly, at first use.
PyTorch is used for device storage and streams only; every argument that crosses the boundary is a plain
pointer or integer.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MIA_B200_LIB") or os.path.join(_HERE, "libmia_b200.so")   # env override: tuning builds

MIA_I16, MIA_F32 = 0, 1
MIA_MODE_VTK, MIA_MODE_ITK = 0, 1
MIA_MAX_MATS, MIA_MAX_BINS = 32, 128
MIA_STATIC_MAX_MATS, MIA_STATIC_WS_BYTES = 16, 512 * 1024

_c = ctypes
_vp, _i, _d, _f, _sz = _c.c_void_p, _c.c_int, _c.c_double, _c.c_float, _c.c_size_t
_ip, _dp = _c.POINTER(_c.c_int), _c.POINTER(_c.c_double)

# name -> (restype, argtypes): the complete list of symbols include/mia_b200.h declares
SIGNATURES = {
    "mia_version": (_i, []),
    "mia_last_error": (_c.c_char_p, []),
    "mia_debug_last_path": (_i, []),
    "mia_debug_x2_static": (_i, [_i]),
    "mia_resample_rigid": (_i, [_vp, _i, _ip, _vp, _i, _ip, _dp, _d, _i, _vp]),
    "mia_resample_rigid_host": (_i, [_vp, _i, _ip, _vp, _i, _ip, _dp, _d, _i, _i]),
    "mia_autocrop_extent": (_i, [_dp, _dp, _ip, _dp, _i, _ip, _dp]),
    "mia_permute_axes": (_i, [_vp, _i, _ip, _vp, _ip, _ip, _vp]),
    "mia_metric_workspace_bytes": (_sz, [_ip, _i, _i]),
    "mia_resample_ssd": (_i, [_vp, _i, _ip, _vp, _i, _ip, _vp, _i, _i, _dp, _vp, _d, _i, _vp, _vp, _vp, _sz, _vp]),
    "mia_resample_ssd_static": (_i, [_vp, _i, _ip, _vp, _i, _ip, _vp, _i, _i, _dp, _vp, _d, _i, _vp, _vp, _vp]),
    "mia_resample_jointhist": (_i, [_vp, _i, _ip, _vp, _i, _ip, _i, _dp, _vp, _i, _f, _f, _f, _f, _i, _vp, _vp,
                                    _vp, _sz, _vp]),
    "mia_resample_ssd_masked": (_i, [_vp, _i, _ip, _vp, _i, _vp, _vp, _ip, _vp, _i, _i, _dp, _vp, _d, _i, _vp, _vp, _vp,
                                     _sz, _vp]),
    "mia_resample_jointhist_masked": (_i, [_vp, _i, _ip, _vp, _i, _vp, _vp, _ip, _i, _dp, _vp, _i, _f, _f, _f, _f, _i,
                                           _vp, _vp, _vp, _sz, _vp]),
    "mia_resample_mattes": (_i, [_vp, _i, _ip, _vp, _i, _vp, _vp, _ip, _i, _dp, _vp, _i, _f, _f, _f, _f, _i, _vp, _vp, _vp,
                                 _sz, _vp]),
    "mia_mi_from_hist64": (_i, [_vp, _i, _i, _vp, _vp]),
    "mia_mi_from_hist": (_i, [_vp, _i, _i, _vp, _vp]),
    "mia_optimizer_state_bytes": (_sz, []),
    "mia_optimizer_init": (_i, [_vp, _dp, _dp, _dp, _dp, _d, _d, _d, _i, _vp, _vp]),
    "mia_optimizer_step": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "mia_optimizer_read": (_i, [_vp, _dp, _dp, _ip, _ip, _ip, _dp, _dp, _dp, _i, _vp]),
    "mia_warp_dvf": (_i, [_vp, _i, _ip, _vp, _vp, _i, _ip, _dp, _dp, _d, _i, _vp]),
    "mia_sample_field_points": (_i, [_vp, _ip, _i, _vp, _c.c_longlong, _d, _vp, _vp]),
    "mia_mask_workspace_bytes": (_sz, [_ip]),
    "mia_mask_scan": (_i, [_vp, _ip, _vp, _vp, _sz, _vp]),
    "mia_resample_masked": (_i, [_vp, _i, _ip, _vp, _ip, _dp, _d, _i, _vp, _sz, _vp, _vp]),
    "mia_sample_stats_workspace_bytes": (_sz, []),
    "mia_sample_stats": (_i, [_vp, _c.c_ulonglong, _c.POINTER(_c.c_float), _i, _c.POINTER(_c.c_ulonglong), _i, _vp,
                              _vp, _sz, _vp]),
    "mia_read_file_to_device": (_i, [_c.c_char_p, _c.c_ulonglong, _c.c_ulonglong, _vp, _i]),
    "mia_write_file_from_device": (_i, [_c.c_char_p, _c.c_ulonglong, _vp, _c.c_ulonglong, _i, _i]),
    "mia_shrink_mean": (_i, [_vp, _i, _ip, _vp, _ip, _vp]),
    "mia_metric_pack": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "mia_optimizer_step_packed": (_i, [_vp, _vp, _i, _vp, _vp, _i, _vp]),
    "mia_fused_exchange_bytes": (_sz, []),
    "mia_fused_push": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "mia_optimizer_step_fused": (_i, [_vp, _vp, _i, _i, _vp, _vp, _i, _vp]),
    "mia_fused_timeouts": (_i, [_vp, _c.POINTER(_c.c_ulonglong), _vp]),
    "mia_host_register": (_i, [_vp, _sz]),
    "mia_host_unregister": (_i, [_vp]),
    "mia_host_alloc": (_vp, [_sz]),
    "mia_host_free": (_i, [_vp]),
}

_lib = None


class MiaError(RuntimeError):
    """A negative return code from the C ABI, with ``mia_last_error()`` as the message."""

    def __init__(self, code, message):
        super().__init__("mia_b200 error %d: %s" % (code, message))
        self.code = code


def load():
    """Load the library (once).  Raises ``ImportError`` when it has not been built -- no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libmia_b200.so not found at %s: build it with medicalimageanalysis_b200/csrc/build.sh "
            "(or __graft_entry__.build()).  This package has no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise MiaError(rc, (load().mia_last_error() or b"").decode("utf-8", "replace"))


def i3(v):
    return (ctypes.c_int * 3)(*[int(x) for x in v])


def darr(values):
    vals = [float(x) for x in values]
    return (ctypes.c_double * len(vals))(*vals)
