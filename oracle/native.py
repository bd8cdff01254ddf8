"""ctypes binding of the C/OpenMP oracle (``oracle/csrc/oracle_omp.c``).  Test infrastructure and the
timed CPU baseline only -- see ``oracle/__init__.py``."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_omp.so")
_lib = None
_DT = {np.dtype(np.int16): 0, np.dtype(np.float32): 1}
_MODE = {"vtk": 0, "itk": 1}


def build(force=False):
    src = os.path.join(_HERE, "csrc", "oracle_omp.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "_build/liboracle_omp.so"],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.oracle_set_threads.restype = ctypes.c_int
    return _lib


def set_threads(n=0):
    return int(lib().oracle_set_threads(int(n)))


def _i3(v):
    return (ctypes.c_int * 3)(*[int(x) for x in v])


def _A(A):
    A = np.ascontiguousarray(np.asarray(A, dtype=np.float64).reshape(3, 4))
    return A, A.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def reslice(mov, A, odims, background=-3001.0, mode="vtk", out_dtype=None):
    mov = np.ascontiguousarray(mov)
    out_dtype = np.dtype(out_dtype or mov.dtype)
    out = np.empty((odims[2], odims[1], odims[0]), dtype=out_dtype)
    Ak, Ap = _A(A)
    rc = lib().oracle_reslice(mov.ctypes.data_as(ctypes.c_void_p), _DT[mov.dtype],
                              _i3((mov.shape[2], mov.shape[1], mov.shape[0])),
                              out.ctypes.data_as(ctypes.c_void_p), _DT[out_dtype], _i3(odims), Ap,
                              ctypes.c_double(background), _MODE[mode])
    if rc:
        raise RuntimeError("oracle_reslice rc=%d" % rc)
    return out


def reslice_ssd(fixed, mov, A, mode="itk", store=False, background=0.0, out_dtype=None):
    mov = np.ascontiguousarray(mov)
    fixed = np.ascontiguousarray(fixed)
    odims = (fixed.shape[2], fixed.shape[1], fixed.shape[0])
    out_dtype = np.dtype(out_dtype or mov.dtype)
    out = np.empty(fixed.shape, dtype=out_dtype) if store else None
    ssd = ctypes.c_double(0)
    cnt = ctypes.c_ulonglong(0)
    Ak, Ap = _A(A)
    rc = lib().oracle_reslice_ssd(mov.ctypes.data_as(ctypes.c_void_p), _DT[mov.dtype],
                                  _i3((mov.shape[2], mov.shape[1], mov.shape[0])),
                                  fixed.ctypes.data_as(ctypes.c_void_p), _DT[fixed.dtype],
                                  out.ctypes.data_as(ctypes.c_void_p) if store else None, _DT[out_dtype],
                                  _i3(odims), Ap, ctypes.c_double(background), _MODE[mode],
                                  ctypes.byref(ssd), ctypes.byref(cnt))
    if rc:
        raise RuntimeError("oracle_reslice_ssd rc=%d" % rc)
    return (ssd.value, cnt.value, out) if store else (ssd.value, cnt.value)


def joint_hist(fixed, mov, A, bins, frange, mrange, mode="itk"):
    mov = np.ascontiguousarray(mov)
    fixed = np.ascontiguousarray(fixed)
    odims = (fixed.shape[2], fixed.shape[1], fixed.shape[0])
    H = np.zeros((bins, bins), dtype=np.uint64)
    cnt = ctypes.c_ulonglong(0)
    Ak, Ap = _A(A)
    rc = lib().oracle_joint_hist(mov.ctypes.data_as(ctypes.c_void_p), _DT[mov.dtype],
                                 _i3((mov.shape[2], mov.shape[1], mov.shape[0])),
                                 fixed.ctypes.data_as(ctypes.c_void_p), _DT[fixed.dtype], _i3(odims), Ap,
                                 _MODE[mode], ctypes.c_float(frange[0]), ctypes.c_float(frange[1]),
                                 ctypes.c_float(mrange[0]), ctypes.c_float(mrange[1]), int(bins),
                                 H.ctypes.data_as(ctypes.POINTER(ctypes.c_ulonglong)), ctypes.byref(cnt))
    if rc:
        raise RuntimeError("oracle_joint_hist rc=%d" % rc)
    return H, cnt.value
