"""NumPy float64 restatement of the displacement-field stage (oracle = test infrastructure only).

``oracle_warp_dvf``    -- ``sitk.Resample(image, sitk.DisplacementFieldTransform(dvf), sitk.sitkLinear, default,
                          pixel id)`` as called at ``structure/deformable.py:797-806`` when the field lives on the
                          output lattice: output voxel idx at physical p is sampled at p + dvf[idx], i.e. at image
                          index ``A @ (idx, 1) + B @ d`` (A = W2P_image @ P2W_out, B = linear part of W2P_image).
                          Inside rule / interpolation / integer conversion: the 'itk' (or 'vtk') mode of
                          ``oracle/reslice.py``.  PARITY UNPINNED for this function (no SimpleITK in the image); its
                          interpolation core is the one ``oracle_reslice`` uses.
``oracle_sample_field_points`` -- ``scipy.ndimage.map_coordinates(field[..., c], [z, y, x], order=1,
                          mode='nearest')`` (``structure/deformable.py:873-878``, ``:1098-1103``), restated as
                          clamp + trilinear.  PINNED: ``tests/test_oracle_dvf.py`` checks it against
                          ``scipy.ndimage.map_coordinates`` itself, the reference's own dependency, which is installed
                          here.
"""
from __future__ import annotations

import numpy as np

from . import reslice as R


def oracle_warp_dvf(mov, dvf, A, B, background=-3001.0, mode="itk", slab=16):
    """``mov`` [z, y, x]; ``dvf`` [z, y, x, 3] physical displacement (x, y, z) per output voxel.  Returns the warped
    volume on the DVF lattice, dtype of ``mov``."""
    round_mode, inside = R._mode_flags(mode)
    mov = np.asarray(mov)
    dvf = np.asarray(dvf, dtype=np.float64)
    A = np.asarray(A, dtype=np.float64).reshape(3, 4)
    B = np.asarray(B, dtype=np.float64).reshape(3, 3)
    nz, ny, nx = dvf.shape[:3]
    mdims = (mov.shape[2], mov.shape[1], mov.shape[0])
    out = np.empty((nz, ny, nx), dtype=mov.dtype)
    bg = R.convert(np.array([background], dtype=np.float64), mov.dtype, round_mode)[0]
    for k0 in range(0, nz, slab):
        k1 = min(nz, k0 + slab)
        c = R.continuous_index(A, (nx, ny, nz), k0, k1)
        d = dvf[k0:k1]
        c = [c[a] + ((B[a, 0] * d[..., 0] + B[a, 1] * d[..., 1]) + B[a, 2] * d[..., 2]) for a in range(3)]
        ok = R.inside_mask(c, mdims, inside)
        v = R.interpolate(mov, c, "vtk" if mode == "vtk" else "itk")
        o = R.convert(v, mov.dtype, round_mode)
        o[~ok] = bg
        out[k0:k1] = o
    return out


def oracle_sample_field_points(field, points_xyz, scale=1.0):
    """``field`` [z, y, x, C]; ``points_xyz`` [n, 3] continuous voxel coordinates (x, y, z).  Returns [n, C] float64:
    order-1 spline (= trilinear) with the coordinates clamped to the volume (``mode='nearest'``)."""
    f = np.asarray(field, dtype=np.float64)
    p = np.asarray(points_xyz, dtype=np.float64)
    nz, ny, nx, C = f.shape
    n = (nx, ny, nz)
    i0, i1, w = [], [], []
    for a in range(3):
        ca = np.clip(np.where(np.isnan(p[:, a]), 0.0, p[:, a]), 0.0, n[a] - 1.0)
        fl = np.floor(ca)
        i0.append(fl.astype(np.int64))
        i1.append(np.minimum(fl.astype(np.int64) + 1, n[a] - 1))
        w.append((ca - fl)[:, None])
    v000 = f[i0[2], i0[1], i0[0]]; v100 = f[i0[2], i0[1], i1[0]]
    v010 = f[i0[2], i1[1], i0[0]]; v110 = f[i0[2], i1[1], i1[0]]
    v001 = f[i1[2], i0[1], i0[0]]; v101 = f[i1[2], i0[1], i1[0]]
    v011 = f[i1[2], i1[1], i0[0]]; v111 = f[i1[2], i1[1], i1[0]]
    a = v000 + w[0] * (v100 - v000); b = v010 + w[0] * (v110 - v010)
    d = v001 + w[0] * (v101 - v001); e = v011 + w[0] * (v111 - v011)
    lo = a + w[1] * (b - a); hi = d + w[1] * (e - d)
    return scale * (lo + w[2] * (hi - lo))
