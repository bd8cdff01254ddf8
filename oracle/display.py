"""NumPy restatement of the off-axis view reslice (oracle; test infrastructure only).

Follows ``Display.compute_offaxis_array`` (``structure/image.py:160-215``, twin ``structure/dose.py:177-231``):
the volume is handed to VTK with direction = ``image.matrix`` read row-major, the reslice transform is the inverse
of the 3x3 viewing rotation, output spacing = image spacing, AutoCropOutput on, background -3001, linear.  The
geometry below is written out independently of ``medicalimageanalysis_b200.geometry`` (plain 4x4 algebra) so that
the product's composition is checked against a second derivation; the resampler is ``oracle.reslice`` (VTK mode,
parity unpinned -- no VTK in this image)."""
from __future__ import annotations

import numpy as np

from . import reslice


def offaxis_lattice(shape_zyx, spacing, origin, image_matrix, view_matrix):
    """(origin_xyz, dims_xyz, A): the auto-cropped output lattice and output index -> input index matrix."""
    nz, ny, nx = shape_zyx
    sp = np.asarray(spacing, dtype=np.float64)
    D = np.asarray(image_matrix, dtype=np.float64).reshape(3, 3)            # VTK: row-major, columns = axes
    p2w = np.identity(4)
    p2w[:3, :3] = D * sp[None, :]
    p2w[:3, 3] = origin
    R = np.identity(4)
    R[:3, :3] = np.asarray(view_matrix, dtype=np.float64).reshape(3, 3)
    T = np.linalg.inv(R)                                                     # reslice transform: out phys -> in phys
    corners = []
    for c in range(8):
        idx = np.array([(nx - 1) * (c & 1), (ny - 1) * ((c >> 1) & 1), (nz - 1) * ((c >> 2) & 1), 1.0])
        corners.append((R @ (p2w @ idx))[:3])                                # inverse of T
    corners = np.asarray(corners)
    lo, hi = corners.min(axis=0), corners.max(axis=0)
    dims = tuple(int(np.floor(abs(hi[a] - lo[a]) / sp[a] + 0.5)) + 1 for a in range(3))
    o2w = np.identity(4)
    o2w[:3, :3] = np.diag(sp)
    o2w[:3, 3] = lo
    A = (np.linalg.inv(p2w) @ T @ o2w)[:3, :]
    return tuple(lo), dims, A


def oracle_offaxis_array(array, spacing, origin, image_matrix, view_matrix):
    """The resliced volume (``secondary_array``), its VTK output origin and the display origin the reference
    stores afterwards (``transform.TransformPoint(new_origin)`` with the inverted transform, ``image.py:199-200``)."""
    lo, dims, A = offaxis_lattice(array.shape, spacing, origin, image_matrix, view_matrix)
    out = reslice.oracle_reslice(array, A, dims, -3001.0, "vtk")
    R = np.asarray(view_matrix, dtype=np.float64).reshape(3, 3)
    return out, lo, np.linalg.inv(R) @ np.asarray(lo)
