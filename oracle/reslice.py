"""NumPy float64 restatement of the rigid trilinear resampler (oracle; test infrastructure only).

Two modes, one arithmetic core.  Both map an output index (i, j, k) through the 3x4 matrix ``A``
(see ``medicalimageanalysis_b200/geometry.py::compose_index_matrix``) to a continuous moving index
c = A @ (i, j, k, 1) in float64 and trilinearly interpolate with neighbour indices clamped to the
buffer.  They differ in the inside test, the interpolation formula and the integer conversion:

  mode 'vtk'  -- ``vtkImageReslice`` as configured at ``structure/rigid.py:685-693`` /
                 ``structure/image.py:189-196`` (VTK 9.3.1, ``requirements.txt:82``):
                 Border on with BorderThickness 0.5  => inside iff every c_a in [-0.5, N_a-0.5]
                 (closed; ``vtkAbstractImageInterpolator::CheckBoundsIJK`` with tolerance 0.5);
                 weights-product trilinear (``vtkImageNLCInterpolate::Trilinear``:
                 rx*(ryrz*v000 + ryfz*v001 + fyrz*v010 + fyfz*v011) + fx*(...));
                 integer output = clamp to type range then floor(v + 0.5)
                 (``vtkInterpolationMath::Round``); background -3001 (``rigid.py:692``).
  mode 'itk'  -- ``sitk.Resample(mov, ref, T, sitkLinear, default, pixelID)`` at
                 ``structure/deformable.py:579,788-795`` (SimpleITK 2.4.1 / ITK 5.4):
                 inside iff every c_a in [-0.5, N_a-0.5) (half-open; ``ImageFunction::IsInsideBuffer``);
                 nested-lerp trilinear (``LinearInterpolateImageFunction::EvaluateOptimized``:
                 v0 + (v1 - v0) * d, x then y then z); integer output = clamp to type range then
                 C static_cast, i.e. truncation toward zero (``CastPixelWithBoundsChecking``).

PARITY UNPINNED (see ``oracle/__init__.py``): no VTK/ITK in this image to diff against.
"""
from __future__ import annotations

import numpy as np

ROUND_HALF_UP = 0   # VTK
ROUND_TRUNC = 1     # ITK
INSIDE_CLOSED = 0   # VTK  [-0.5, N-0.5]
INSIDE_HALFOPEN = 1  # ITK  [-0.5, N-0.5)


def _mode_flags(mode):
    if mode == "vtk":
        return ROUND_HALF_UP, INSIDE_CLOSED
    if mode == "itk":
        return ROUND_TRUNC, INSIDE_HALFOPEN
    raise ValueError("mode must be 'vtk' or 'itk'")


def continuous_index(A, odims, k0=0, k1=None):
    """c[a] for a = x, y, z as float64 arrays of shape [k1-k0, ny, nx].

    Association order follows ``vtkImageResliceExecute``: origin + k*zAxis, + j*yAxis, + i*xAxis."""
    A = np.asarray(A, dtype=np.float64).reshape(3, 4)
    nx, ny, nz = odims
    k1 = nz if k1 is None else k1
    i = np.arange(nx, dtype=np.float64)[None, None, :]
    j = np.arange(ny, dtype=np.float64)[None, :, None]
    k = np.arange(k0, k1, dtype=np.float64)[:, None, None]
    return [((A[a, 3] + k * A[a, 2]) + j * A[a, 1]) + i * A[a, 0] for a in range(3)]


def inside_mask(c, mdims, inside):
    ok = None
    for a in range(3):
        hi = mdims[a] - 0.5
        m = (c[a] >= -0.5) & ((c[a] <= hi) if inside == INSIDE_CLOSED else (c[a] < hi))
        ok = m if ok is None else (ok & m)
    return ok


def interpolate(mov, c, formula="vtk"):
    """Trilinear value (float64) at continuous indices ``c`` (x, y, z arrays); neighbour indices
    clamped to the buffer.  ``mov`` is [z, y, x].  Caller masks the outside voxels."""
    nz, ny, nx = mov.shape
    n = (nx, ny, nz)
    i0, i1, f = [], [], []
    for a in range(3):
        fl = np.floor(c[a])
        fa = c[a] - fl
        ia = fl.astype(np.int64)
        ib = ia + (fa != 0)          # vtk: inIdX1 = inIdX0 + (fx != 0)
        i0.append(np.clip(ia, 0, n[a] - 1))
        i1.append(np.clip(ib, 0, n[a] - 1))
        f.append(fa)
    m = mov.astype(np.float64, copy=False)

    def tap(zz, yy, xx):
        return m[zz, yy, xx]

    v000 = tap(i0[2], i0[1], i0[0]); v100 = tap(i0[2], i0[1], i1[0])
    v010 = tap(i0[2], i1[1], i0[0]); v110 = tap(i0[2], i1[1], i1[0])
    v001 = tap(i1[2], i0[1], i0[0]); v101 = tap(i1[2], i0[1], i1[0])
    v011 = tap(i1[2], i1[1], i0[0]); v111 = tap(i1[2], i1[1], i1[0])
    fx, fy, fz = f
    if formula == "vtk":
        rx, ry, rz = 1.0 - fx, 1.0 - fy, 1.0 - fz
        ryrz, fyrz, ryfz, fyfz = ry * rz, fy * rz, ry * fz, fy * fz
        return (rx * (ryrz * v000 + ryfz * v001 + fyrz * v010 + fyfz * v011) +
                fx * (ryrz * v100 + ryfz * v101 + fyrz * v110 + fyfz * v111))
    # itk: nested lerps.  ITK clamps the base index low and lets the distance go <= 0 there; with the
    # index clamp above the two are the same function (distance in the clamped zone multiplies a
    # zero difference).
    vx00 = v000 + (v100 - v000) * fx
    vx10 = v010 + (v110 - v010) * fx
    vx01 = v001 + (v101 - v001) * fx
    vx11 = v011 + (v111 - v011) * fx
    vxx0 = vx00 + (vx10 - vx00) * fy
    vxx1 = vx01 + (vx11 - vx01) * fy
    return vxx0 + (vxx1 - vxx0) * fz


def convert(v, out_dtype, round_mode):
    out_dtype = np.dtype(out_dtype)
    if out_dtype.kind == "f":
        return v.astype(out_dtype)
    info = np.iinfo(out_dtype)
    v = np.clip(v, info.min, info.max)
    if round_mode == ROUND_HALF_UP:
        v = np.floor(v + 0.5)
    else:
        v = np.trunc(v)
    return v.astype(out_dtype)


def oracle_reslice(mov, A, odims, background=-3001.0, mode="vtk", out_dtype=None,
                   slab=16, return_float=False):
    """Resample ``mov`` ([z,y,x]) onto an output lattice of ``odims`` (x,y,z).  Returns [z,y,x].

    ``return_float`` also returns the unrounded float64 samples and the inside mask (for the metrics)."""
    round_mode, inside = _mode_flags(mode)
    mov = np.asarray(mov)
    out_dtype = np.dtype(out_dtype or mov.dtype)
    nx, ny, nz = (int(d) for d in odims)
    mdims = (mov.shape[2], mov.shape[1], mov.shape[0])
    out = np.empty((nz, ny, nx), dtype=out_dtype)
    vals = np.empty((nz, ny, nx), dtype=np.float64) if return_float else None
    mask = np.empty((nz, ny, nx), dtype=bool) if return_float else None
    bg = convert(np.array([float(background)]), out_dtype, round_mode)[0]
    for k0 in range(0, nz, slab):
        k1 = min(nz, k0 + slab)
        c = continuous_index(A, (nx, ny, nz), k0, k1)
        ok = inside_mask(c, mdims, inside)
        v = interpolate(mov, c, "vtk" if mode == "vtk" else "itk")
        o = convert(v, out_dtype, round_mode)
        o[~ok] = bg
        out[k0:k1] = o
        if return_float:
            vals[k0:k1] = v
            mask[k0:k1] = ok
    if return_float:
        return out, vals, mask
    return out
