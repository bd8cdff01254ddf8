"""Thin tensor-level wrappers over the C ABI.  Tensors are storage; all arithmetic is in libmia_b200.so.

Every function takes CUDA ``torch.Tensor`` volumes laid out ``[z, y, x]`` contiguous (the ``Image.array``
layout of the reference, ``read/dicom.py:530-534``) and 3x4 index matrices (see ``geometry.py``), launches
on the current torch stream and returns device tensors without synchronising.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import MIA_F32, MIA_I16, MIA_MODE_ITK, MIA_MODE_VTK, check, darr, i3

_DT = {torch.int16: MIA_I16, torch.float32: MIA_F32}
_NP_DT = {np.dtype(np.int16): MIA_I16, np.dtype(np.float32): MIA_F32}
_MODE = {"vtk": MIA_MODE_VTK, "itk": MIA_MODE_ITK}

FFS_PERMUTATIONS = {
    # case -> (perm, flip) in [z, y, x] axes: out = flip(transpose(in, perm), flip axes)
    "axial_k1": ((0, 2, 1), (0, 1, 0)),   # np.rot90(a, 1, (1, 2))  read/dicom.py:697
    "axial_k3": ((0, 2, 1), (0, 0, 1)),   # np.rot90(a, 3, (1, 2))  read/dicom.py:699
    "axial_k2": ((0, 1, 2), (0, 1, 1)),   # np.rot90(a, 2, (1, 2))  read/dicom.py:701
    "coronal":  ((1, 0, 2), (1, 0, 0)),   # np.rot90(a, 1, (0, 1))  read/dicom.py:709
    "sagittal": ((1, 2, 0), (1, 0, 1)),   # flip(rot90(a,1,(0,1)).transpose(0,2,1), 2)  read/dicom.py:717
}


def _vol(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError("%s must be a CUDA torch.Tensor (there is no CPU path)" % name)
    if t.dim() != 3:
        raise ValueError("%s must be 3-D [z, y, x]" % name)
    if t.dtype not in _DT:
        raise TypeError("%s dtype must be int16 or float32, got %s" % (name, t.dtype))
    if not t.is_contiguous():
        t = t.contiguous()
    return t


def _dims(t):
    return i3((t.shape[2], t.shape[1], t.shape[0]))


def _stream(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _mats(A):
    A = np.ascontiguousarray(np.asarray(A, dtype=np.float64))
    if A.ndim == 2:
        A = A[None]
    if A.shape[1:] != (3, 4):
        raise ValueError("index matrices must be [n, 3, 4] or [3, 4]")
    return A


def resample(mov, A, out_shape, background=-3001.0, mode="vtk", out_dtype=None, out=None):
    """Trilinear resample of ``mov`` onto an ``out_shape`` ([z,y,x]) lattice through the 3x4 ``A``."""
    lib = _lib.load()
    mov = _vol(mov, "mov")
    out_dtype = out_dtype or mov.dtype
    if out is None:
        out = torch.empty(tuple(int(s) for s in out_shape), dtype=out_dtype, device=mov.device)
    else:
        out = _vol(out, "out")
    A = _mats(A)
    if A.shape[0] != 1:
        raise ValueError("resample takes exactly one matrix")
    with torch.cuda.device(mov.device):
        check(lib.mia_resample_rigid(mov.data_ptr(), _DT[mov.dtype], _dims(mov), out.data_ptr(), _DT[out.dtype],
                                     _dims(out), darr(A.ravel()), float(background), _MODE[mode], _stream(mov)))
    return out


def resample_host(mov, A, out_shape, background=-3001.0, mode="vtk", out_dtype=None, out=None, device=0):
    """NumPy in, NumPy out through ``mia_resample_rigid_host`` (H2D / kernel / D2H pipelined by z-slab)."""
    lib = _lib.load()
    mov = np.ascontiguousarray(mov)
    if mov.dtype not in _NP_DT or mov.ndim != 3:
        raise TypeError("mov must be a 3-D int16 or float32 array")
    out_dtype = np.dtype(out_dtype or mov.dtype)
    if out is None:
        out = np.empty(tuple(int(s) for s in out_shape), dtype=out_dtype)
    A = _mats(A)
    md = i3((mov.shape[2], mov.shape[1], mov.shape[0]))
    od = i3((out.shape[2], out.shape[1], out.shape[0]))
    check(lib.mia_resample_rigid_host(mov.ctypes.data, _NP_DT[mov.dtype], md, out.ctypes.data, _NP_DT[out.dtype], od,
                                      darr(A.ravel()), float(background), _MODE[mode], int(device)))
    return out


def pinned_empty(shape, dtype):
    """A NumPy array over page-locked memory from ``mia_host_alloc`` (freed when the array dies)."""
    lib = _lib.load()
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = lib.mia_host_alloc(n)
    if not p:
        raise _lib.MiaError(-4, (lib.mia_last_error() or b"").decode())
    buf = (ctypes.c_char * max(n, 1)).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    import weakref
    weakref.finalize(buf, lib.mia_host_free, p)
    return arr


def host_register(arr):
    """Page-lock the memory of a NumPy array the caller owns (``mia_host_register``), e.g. this rank's planes of a shared
    mapping.  Returns a handle to pass to ``host_unregister``."""
    lib = _lib.load()
    arr = np.asarray(arr)
    if not arr.flags["C_CONTIGUOUS"] or arr.nbytes == 0:
        raise ValueError("host_register needs a non-empty C-contiguous array")
    check(lib.mia_host_register(arr.ctypes.data, arr.nbytes))
    return arr.ctypes.data


def host_unregister(handle):
    check(_lib.load().mia_host_unregister(handle))


def permute_axes(vol, perm, flip, out=None):
    """Exact ``np.flip(np.transpose(vol, perm), flipped axes)`` on the device."""
    lib = _lib.load()
    vol = _vol(vol, "vol")
    oshape = tuple(vol.shape[p] for p in perm)
    if out is None:
        out = torch.empty(oshape, dtype=vol.dtype, device=vol.device)
    with torch.cuda.device(vol.device):
        check(lib.mia_permute_axes(vol.data_ptr(), _DT[vol.dtype], _dims(vol), out.data_ptr(), i3(perm), i3(flip),
                                   _stream(vol)))
    return out


class MetricWorkspace:
    """Caller-owned device scratch + result buffers for the fused metric launches (reused across calls)."""

    def __init__(self, device, n_mats=1, bins=64):
        lib = _lib.load()
        self.device = torch.device(device)
        self.n_mats, self.bins = int(n_mats), int(bins)
        nbytes = int(lib.mia_metric_workspace_bytes(i3((1, 1, 1)), self.n_mats, self.bins))
        self.ws = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
        self.ssd = torch.zeros(self.n_mats, dtype=torch.float64, device=self.device)
        self.count = torch.zeros(self.n_mats, dtype=torch.int64, device=self.device)
        self.hist = torch.zeros((self.n_mats, self.bins, self.bins), dtype=torch.int32, device=self.device)
        self.mi = torch.zeros(self.n_mats, dtype=torch.float64, device=self.device)
        self._hist64 = None

    @property
    def hist64(self):
        """[n_mats, bins, bins] int64 counters of the Mattes (Parzen) histogram, allocated on first use."""
        if self._hist64 is None:
            self._hist64 = torch.zeros((self.n_mats, self.bins, self.bins), dtype=torch.int64, device=self.device)
        return self._hist64


def _A_args(A, d_A, n_expected=None):
    if d_A is not None:
        if d_A.dtype != torch.float64 or not d_A.is_cuda or not d_A.is_contiguous():
            raise TypeError("d_A must be a contiguous CUDA float64 tensor [n, 3, 4]")
        n = d_A.numel() // 12
        return n, None, ctypes.c_void_p(d_A.data_ptr())
    A = _mats(A)
    return A.shape[0], darr(A.ravel()), None


def _fixed_mask(mask, fixed):
    if mask is None:
        return None
    m = _mask_u8(mask, fixed.shape, fixed.device)
    return m


def resample_ssd(fixed, mov, A=None, d_A=None, mode="itk", ws=None, out=None, background=0.0, fixed_mask=None,
                 moving_mask=None):
    """Fused resample + SSD for n candidate matrices in one launch.  Returns (ssd[n] f64, count[n] i64)
    device tensors (views into ``ws``).  ``out`` (n == 1) also stores the resampled volume.  ``fixed_mask``
    (uint8 ``[z, y, x]`` on the fixed lattice): voxels where it is 0 take no part."""
    lib = _lib.load()
    fixed, mov = _vol(fixed, "fixed"), _vol(mov, "mov")
    n, Ah, Ad = _A_args(A, d_A)
    if ws is None or ws.n_mats < n:
        ws = MetricWorkspace(mov.device, n_mats=n)
    if out is not None:
        out = _vol(out, "out")
    fm, mm = _fixed_mask(fixed_mask, fixed), _fixed_mask(moving_mask, mov)
    with torch.cuda.device(mov.device):
        check(lib.mia_resample_ssd_masked(mov.data_ptr(), _DT[mov.dtype], _dims(mov), fixed.data_ptr(), _DT[fixed.dtype],
                                   fm.data_ptr() if fm is not None else None, mm.data_ptr() if mm is not None else None,
                                   _dims(fixed), out.data_ptr() if out is not None else None,
                                   _DT[out.dtype] if out is not None else _DT[mov.dtype], n, Ah, Ad,
                                   float(background), _MODE[mode], ws.ssd.data_ptr(), ws.count.data_ptr(),
                                   ws.ws.data_ptr(), ws.ws.numel(), _stream(mov)))
    return ws.ssd[:n], ws.count[:n]


def resample_jointhist(fixed, mov, frange, mrange, bins=64, A=None, d_A=None, mode="itk", ws=None, fixed_mask=None,
                       moving_mask=None):
    """Fused resample + joint histogram.  Returns (hist[n, bins, bins] int32, count[n] i64) device tensors."""
    lib = _lib.load()
    fixed, mov = _vol(fixed, "fixed"), _vol(mov, "mov")
    n, Ah, Ad = _A_args(A, d_A)
    if ws is None or ws.n_mats < n or ws.bins != bins:
        ws = MetricWorkspace(mov.device, n_mats=n, bins=bins)
    fm, mm = _fixed_mask(fixed_mask, fixed), _fixed_mask(moving_mask, mov)
    with torch.cuda.device(mov.device):
        check(lib.mia_resample_jointhist_masked(mov.data_ptr(), _DT[mov.dtype], _dims(mov), fixed.data_ptr(),
                                         _DT[fixed.dtype], fm.data_ptr() if fm is not None else None,
                                         mm.data_ptr() if mm is not None else None,
                                         _dims(fixed), n, Ah, Ad, _MODE[mode],
                                         float(frange[0]), float(frange[1]), float(mrange[0]), float(mrange[1]),
                                         int(bins), ws.hist.data_ptr(), ws.count.data_ptr(), ws.ws.data_ptr(),
                                         ws.ws.numel(), _stream(mov)))
    return ws.hist[:n], ws.count[:n]


def resample_mattes(fixed, mov, frange, mrange, bins=50, A=None, d_A=None, mode="itk", ws=None, fixed_mask=None,
                    moving_mask=None):
    """Fused resample + Mattes (Parzen-window) joint histogram.  Returns (hist[n, bins, bins] int64 in
    units of 1/1024 sample, count[n] i64) device tensors; ``mi_from_hist`` turns the histogram into MI."""
    lib = _lib.load()
    fixed, mov = _vol(fixed, "fixed"), _vol(mov, "mov")
    n, Ah, Ad = _A_args(A, d_A)
    if ws is None or ws.n_mats < n or ws.bins != bins:
        ws = MetricWorkspace(mov.device, n_mats=n, bins=bins)
    fm, mm = _fixed_mask(fixed_mask, fixed), _fixed_mask(moving_mask, mov)
    with torch.cuda.device(mov.device):
        check(lib.mia_resample_mattes(mov.data_ptr(), _DT[mov.dtype], _dims(mov), fixed.data_ptr(), _DT[fixed.dtype],
                                      fm.data_ptr() if fm is not None else None,
                                      mm.data_ptr() if mm is not None else None, _dims(fixed), n, Ah, Ad, _MODE[mode],
                                      float(frange[0]), float(frange[1]), float(mrange[0]), float(mrange[1]), int(bins),
                                      ws.hist64.data_ptr(), ws.count.data_ptr(), ws.ws.data_ptr(), ws.ws.numel(),
                                      _stream(mov)))
    return ws.hist64[:n], ws.count[:n]


def mi_from_hist(hist, out=None):
    """MI (nats, float64) of each ``[bins, bins]`` integer histogram in ``hist`` ([n, bins, bins] int32, or int64 for
    the Mattes histogram)."""
    lib = _lib.load()
    if hist.dim() == 2:
        hist = hist[None]
    hist = hist.contiguous()
    n, bins = hist.shape[0], hist.shape[1]
    if out is None:
        out = torch.empty(n, dtype=torch.float64, device=hist.device)
    fn = lib.mia_mi_from_hist64 if hist.dtype == torch.int64 else lib.mia_mi_from_hist
    with torch.cuda.device(hist.device):
        check(fn(hist.data_ptr(), bins, n, out.data_ptr(), _stream(hist)))
    return out


def warp_dvf(mov, dvf, A, B, background=-3001.0, mode="itk", out=None):
    """Sample ``mov`` at ``A . (idx, 1) + B . dvf[idx]`` for every voxel of the DVF lattice (second
    stage of ``Deformable.create_image``).  ``dvf``: CUDA float32 ``[z, y, x, 3]`` physical displacements
    (x, y, z); ``A``: 3x4 output index -> image index; ``B``: 3x3 physical -> image index (linear part)."""
    lib = _lib.load()
    mov = _vol(mov, "mov")
    if not (isinstance(dvf, torch.Tensor) and dvf.is_cuda and dvf.dtype == torch.float32 and dvf.dim() == 4
            and dvf.shape[3] == 3 and dvf.is_contiguous()):
        raise TypeError("dvf must be a contiguous CUDA float32 tensor [z, y, x, 3]")
    oshape = tuple(int(v) for v in dvf.shape[:3])
    if out is None:
        out = torch.empty(oshape, dtype=mov.dtype, device=mov.device)
    else:
        out = _vol(out, "out")
    A = _mats(A)
    B = np.ascontiguousarray(np.asarray(B, dtype=np.float64).reshape(3, 3))
    with torch.cuda.device(mov.device):
        check(lib.mia_warp_dvf(mov.data_ptr(), _DT[mov.dtype], _dims(mov), dvf.data_ptr(), out.data_ptr(),
                               _DT[out.dtype], i3((oshape[2], oshape[1], oshape[0])), darr(A.ravel()), darr(B.ravel()),
                               float(background), _MODE[mode], _stream(mov)))
    return out


def sample_field_points(field, points_xyz, scale=1.0, out=None):
    """``scipy.ndimage.map_coordinates(field[..., c], [z, y, x], order=1, mode='nearest')`` for all channels.
    ``field``: CUDA float32 ``[z, y, x, C]``; ``points_xyz``: CUDA float64 ``[n, 3]`` voxel coordinates (x, y, z).
    Returns CUDA float64 ``[n, C]`` (times ``scale``)."""
    lib = _lib.load()
    if not (isinstance(field, torch.Tensor) and field.is_cuda and field.dtype == torch.float32 and field.dim() == 4
            and field.is_contiguous()):
        raise TypeError("field must be a contiguous CUDA float32 tensor [z, y, x, C]")
    pts = points_xyz
    if not (isinstance(pts, torch.Tensor) and pts.is_cuda and pts.dtype == torch.float64 and pts.dim() == 2
            and pts.shape[1] == 3 and pts.is_contiguous()):
        raise TypeError("points_xyz must be a contiguous CUDA float64 tensor [n, 3]")
    n, C = int(pts.shape[0]), int(field.shape[3])
    if out is None:
        out = torch.empty((n, C), dtype=torch.float64, device=field.device)
    with torch.cuda.device(field.device):
        check(lib.mia_sample_field_points(field.data_ptr(), i3((field.shape[2], field.shape[1], field.shape[0])), C,
                                          pts.data_ptr(), n, float(scale), out.data_ptr(), _stream(field)))
    return out


def _mask_u8(mask, shape, device):
    if isinstance(mask, np.ndarray):
        mask = torch.from_numpy(np.ascontiguousarray(mask))
    if not isinstance(mask, torch.Tensor):
        raise TypeError("mask must be a NumPy array or a torch.Tensor [z, y, x]")
    if tuple(mask.shape) != tuple(shape):
        raise ValueError("mask shape %s != lattice shape %s" % (tuple(mask.shape), tuple(shape)))
    mask = mask.to(device)
    if mask.dtype != torch.uint8:
        mask = (mask > 0).to(torch.uint8)          # the reference's test is `mask_arr > 0` (structure/dose.py:767)
    return mask.contiguous()


def resample_masked(mov, A, mask, background=0.0, mode="itk"):
    """``resample(mov, A, mask.shape)[mask > 0]`` without resampling the unmasked voxels: the masked samples as a
    compact float32 device tensor in C order (``np.where`` order; ``structure/dose.py:760-770``)."""
    lib = _lib.load()
    mov = _vol(mov, "mov")
    mask = _mask_u8(mask, mask.shape, mov.device)
    if mask.dim() != 3:
        raise ValueError("mask must be 3-D [z, y, x]")
    od = i3((mask.shape[2], mask.shape[1], mask.shape[0]))
    A = _mats(A)
    nbytes = int(lib.mia_mask_workspace_bytes(od))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=mov.device)
    total = torch.zeros(1, dtype=torch.int64, device=mov.device)
    with torch.cuda.device(mov.device):
        check(lib.mia_mask_scan(mask.data_ptr(), od, total.data_ptr(), ws.data_ptr(), nbytes, _stream(mov)))
        n = int(total.item())                      # the one synchronisation: the output size
        out = torch.empty(n, dtype=torch.float32, device=mov.device)
        if n:
            check(lib.mia_resample_masked(mov.data_ptr(), _DT[mov.dtype], _dims(mov), mask.data_ptr(), od,
                                          darr(A.ravel()), float(background), _MODE[mode], ws.data_ptr(), nbytes,
                                          out.data_ptr(), _stream(mov)))
    return out


def percentile_ranks(n, qs):
    """Order-statistic indices (floor, ceil) and lerp weight of ``np.percentile(x, q)`` (method 'linear') for a
    sample of ``n`` values."""
    lo, hi, t = [], [], []
    for q in qs:
        pos = float(q) / 100.0 * (n - 1)
        f = int(np.floor(pos))
        lo.append(f)
        hi.append(min(f + 1, n - 1))
        t.append(pos - f)
    return lo, hi, t


def sample_statistics(values, percentiles=(), levels=()):
    """min / max / mean / std (population, two-pass like ``np.std``) / ``np.percentile`` values / counts below each
    level of a compact float32 device sample, from ``mia_sample_stats`` (moments + exact radix select)."""
    lib = _lib.load()
    if not (isinstance(values, torch.Tensor) and values.is_cuda and values.dtype == torch.float32 and values.dim() == 1):
        raise TypeError("values must be a 1-D CUDA float32 tensor")
    values = values.contiguous()
    n = int(values.numel())
    if n == 0:
        raise ValueError("empty sample")
    levels = [float(v) for v in levels]
    if levels != sorted(levels):
        raise ValueError("levels must ascend")
    lo, hi, t = percentile_ranks(n, percentiles)
    ranks = lo + hi
    L, R = len(levels), len(ranks)
    nbytes = int(lib.mia_sample_stats_workspace_bytes())
    ws = torch.empty(nbytes, dtype=torch.uint8, device=values.device)
    res = torch.zeros(6 + L + R, dtype=torch.float64, device=values.device)
    lv = (ctypes.c_float * max(L, 1))(*levels)
    rk = (ctypes.c_ulonglong * max(R, 1))(*ranks)
    with torch.cuda.device(values.device):
        check(lib.mia_sample_stats(values.data_ptr(), n, lv, L, rk, R, res.data_ptr(), ws.data_ptr(), nbytes,
                                   _stream(values)))
    r = res.cpu().numpy()
    os_ = r[6 + L:]
    P = len(lo)
    pct = []
    for i in range(P):
        a, b = float(os_[i]), float(os_[P + i])
        # numpy's _lerp: a + (b - a) * t, evaluated from b when t >= 0.5
        pct.append(b - (b - a) * (1.0 - t[i]) if t[i] >= 0.5 else a + (b - a) * t[i])
    return {"count": n, "min": float(r[0]), "max": float(r[1]), "sum": float(r[2]), "mean": float(r[3]),
            "std": float(np.sqrt(r[4] / n)), "percentiles": pct, "below": [float(v) for v in r[6:6 + L]]}


def dose_statistics(mov, A, mask, d_values, levels, background=0.0, mode="itk"):
    """The numbers of ``Dose.compute_roi_dose_statistics`` (``structure/dose.py:774-816``): D-values are
    ``np.percentile(sample, 100 - d)``; the median is the 50th percentile."""
    sample = resample_masked(mov, A, mask, background=background, mode=mode)
    qs = [100.0 - d for d in d_values] + [50.0]
    st = sample_statistics(sample, percentiles=qs, levels=levels)
    st["median"] = st["percentiles"][-1]
    st["percentiles"] = st["percentiles"][:-1]
    st["sample"] = sample
    return st


def shrink_mean(vol, factors, out=None):
    """One pyramid level: block mean over ``factors`` = (fx, fy, fz) voxels (``mia_shrink_mean``)."""
    lib = _lib.load()
    vol = _vol(vol, "vol")
    f = tuple(int(v) for v in factors)
    oshape = tuple((vol.shape[a] + ff - 1) // ff for a, ff in zip((0, 1, 2), (f[2], f[1], f[0])))
    if out is None:
        out = torch.empty(oshape, dtype=vol.dtype, device=vol.device)
    with torch.cuda.device(vol.device):
        check(lib.mia_shrink_mean(vol.data_ptr(), _DT[vol.dtype], _dims(vol), out.data_ptr(), i3(f), _stream(vol)))
    return out


def autocrop_extent(M, p2w_mov, mdims, out_spacing, extent_rule="round"):
    """Host helper mirroring ``geometry.autocrop_extent`` through the C ABI."""
    lib = _lib.load()
    od = (ctypes.c_int * 3)()
    oo = (ctypes.c_double * 3)()
    check(lib.mia_autocrop_extent(darr(np.asarray(M, dtype=np.float64).ravel()),
                                  darr(np.asarray(p2w_mov, dtype=np.float64).ravel()), i3(mdims),
                                  darr(out_spacing), 0 if extent_rule == "round" else 1, od, oo))
    return tuple(oo), tuple(od)
