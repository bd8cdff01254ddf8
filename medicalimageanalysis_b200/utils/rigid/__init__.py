"""Operator functions of the rigid resample hot path -- the ``mia.utils.rigid`` slot, which is empty in the
reference (``utils/rigid/__init__.py``).  Each function names the reference call site it stands
in for.  Inputs are ``structure.Image``-like objects (fields ``array, spacing, origin, matrix``) whose
``array`` is a CUDA tensor; there is no CPU path.
"""
from __future__ import annotations

import numpy as np

from ... import geometry as G
from ... import kernels
from ...registration import (RegistrationConfig, RegistrationResult, register_rigid as _register,
                             register_rigid_multires as _register_multires, pyramid_level)
from ...structure.rigid import BACKGROUND, ResampledImage

__all__ = ["matrix_from_dicom_reg", "resample_rigid", "resample_to_reference", "reorient_ffs", "verify_axial_orientation", "metric_ssd",
           "metric_mi", "register_rigid", "register_rigid_multires", "pyramid_level", "RegistrationConfig", "RegistrationResult"]


def matrix_from_dicom_reg(values):
    """DICOM REG ``FrameOfReferenceTransformationMatrix`` (3006,00C6), 16 numbers row-major, to the moving matrix the
    reference stores (``read/dicom.py:1751-1763``): ``moving_matrix = inv(matrix.reshape(4, 4))``."""
    return np.linalg.inv(np.asarray(values, dtype=np.float64).reshape(4, 4))


def resample_rigid(moving, reference, matrix, background=BACKGROUND, inverse=False):
    """``Rigid.create_image`` as a function (``structure/rigid.py:646-695``): VTK-mode reslice of ``moving``
    through ``matrix`` (fixed physical -> moving physical; inverted when ``inverse``) onto an auto-cropped
    lattice with ``reference``'s spacing and identity direction.  Returns a ``ResampledImage``."""
    M = np.asarray(matrix, dtype=np.float64).reshape(4, 4)
    if inverse:
        M = np.linalg.inv(M)
    mgrid = moving.grid
    origin, dims = G.autocrop_extent(mgrid, M, reference.spacing, convention="vtk")
    out_grid = G.Grid(dims, tuple(reference.spacing), origin, np.identity(3))
    A = G.compose_index_matrix(out_grid, mgrid, M, convention="vtk")
    out = kernels.resample(moving.array, A, (dims[2], dims[1], dims[0]), background=background, mode="vtk")
    return ResampledImage(out, origin, tuple(reference.spacing))


def resample_to_reference(moving, reference, matrix=None, default_value=0.0, out_dtype=None):
    """``sitk.Resample(mov, ref, T, sitkLinear, default, mov pixel type)`` (``structure/deformable.py:575-579``
    with default 0.0, ``:788-795`` with -3001; identity-transform variant ``utils/deformable/simpleitk.py:81-94``):
    ITK-mode resample onto ``reference``'s own lattice.  Returns a CUDA tensor shaped like ``reference.array``."""
    M = np.identity(4) if matrix is None else np.asarray(matrix, dtype=np.float64).reshape(4, 4)
    A = G.compose_index_matrix(reference.grid, moving.grid, M, convention="itk")
    return kernels.resample(moving.array, A, tuple(reference.array.shape), background=default_value, mode="itk",
                            out_dtype=out_dtype)


def reorient_ffs(array, case):
    """The array permutation of the Feet-First-Supine normalisation (``read/dicom.py:695-717``), exact.
    ``case`` is one of ``kernels.FFS_PERMUTATIONS``; ``array`` a CUDA tensor ``[z, y, x]``."""
    perm, flip = kernels.FFS_PERMUTATIONS[case]
    return kernels.permute_axes(array, perm, flip)


def verify_axial_orientation(array, origin, orientation, spacing, plane="Axial"):
    """``Read3D._verify_axial_orientation`` (``read/dicom.py:655-740``) for a device-resident volume:
    picks the corner with the smallest x+y+z, permutes the array accordingly (exact copy on the GPU instead
    of a NumPy view), and returns ``(array, origin, orientation, image_matrix)`` rewritten the same way.

    ``spacing`` is ``[x, y, z]`` as on ``Image`` (``read/dicom.py:616-617``)."""
    shape = tuple(int(s) for s in array.shape)
    orientation = np.asarray(orientation, dtype=np.float64).copy()
    origin = np.asarray(origin, dtype=np.float64)
    if plane == "Axial":
        sp = [spacing[0], spacing[1], spacing[2]]
        dimensions = (shape[0], shape[1], shape[2])
    elif plane == "Coronal":
        sp = [spacing[0], spacing[2], spacing[1]]
        dimensions = (shape[1], shape[0], shape[2])
    else:
        sp = [spacing[1], spacing[2], spacing[0]]
        dimensions = (shape[1], shape[2], shape[0])
    row, col = orientation[:3], orientation[3:]
    slc = np.cross(row, col)
    ns, ny, nx = shape[0] - 1, shape[1] - 1, shape[2] - 1
    corners = np.zeros((8, 3))
    for c in range(8):
        corners[c] = (origin + (nx * sp[0] * row if c & 1 else 0) + (ny * sp[1] * col if c & 2 else 0) +
                      (ns * sp[2] * slc if c & 4 else 0))
    corner_idx = int(np.argmin(np.sum(corners, axis=1)))
    if corner_idx == 0:
        return array, origin, orientation, _image_matrix(orientation)
    new_origin = corners[corner_idx]
    if plane == "Axial":
        case = {1: "axial_k1", 2: "axial_k3"}.get(corner_idx, "axial_k2")
        square = corners[:4] if corner_idx < 4 else corners[4:]
    else:
        case = "coronal" if plane == "Coronal" else "sagittal"
        s1 = np.argsort(corners[:4, 2])
        s2 = np.argsort(corners[4:, 2]) + 4
        square = [corners[s1[0]], corners[s1[1]], corners[s2[0]], corners[s2[1]]]
    new_array = reorient_ffs(array, case)
    dist = np.asarray([np.linalg.norm(corners[corner_idx] - s) for s in square])
    order = np.argsort(dist)
    c1 = square[order[1]] - corners[corner_idx]
    c2 = square[order[2]] - corners[corner_idx]
    if abs(c1[0]) > abs(c2[0]):
        first, second = c1, c2
    else:
        first, second = c2, c1
    orientation[:3] = first / (spacing[0] * dimensions[2])
    orientation[3:] = second / (spacing[1] * dimensions[1])
    return new_array, new_origin, orientation, _image_matrix(orientation)


def _image_matrix(orientation):
    """``read/dicom.py:640-653``."""
    mat = np.eye(3, dtype=np.float32)
    mat[0] = orientation[:3]
    mat[1] = orientation[3:]
    mat[2] = np.cross(orientation[:3], orientation[3:])
    return mat


def _index_matrix(fixed, moving, matrix):
    M = np.identity(4) if matrix is None else np.asarray(matrix, dtype=np.float64).reshape(4, 4)
    return G.compose_index_matrix(fixed.grid, moving.grid, M, convention="itk")


def metric_ssd(fixed, moving, matrix=None, mode="itk"):
    """Sum of squared differences of ``fixed`` against ``moving`` resampled through ``matrix`` (SURVEY M1;
    the reference only names it: ``SetMetricAsMeanSquares`` ``utils/deformable/simpleitk.py:115``).
    Returns ``(ssd, n_valid, mse)`` as Python numbers (one device->host read)."""
    ssd, cnt = kernels.resample_ssd(fixed.array, moving.array, A=_index_matrix(fixed, moving, matrix), mode=mode)
    s, n = float(ssd[0].item()), int(cnt[0].item())
    return s, n, (s / n if n else float("nan"))


def metric_mi(fixed, moving, matrix=None, bins=64, frange=None, mrange=None, mode="itk"):
    """Joint-histogram mutual information in nats (named only at
    ``utils/deformable/simpleitk.py:155-156``).  Returns ``(mi, n_valid)``."""
    if frange is None:
        frange = (float(fixed.array.min().item()), float(fixed.array.max().item()))
    if mrange is None:
        mrange = (float(moving.array.min().item()), float(moving.array.max().item()))
    hist, cnt = kernels.resample_jointhist(fixed.array, moving.array, frange, mrange, bins=bins,
                                           A=_index_matrix(fixed, moving, matrix), mode=mode)
    mi = kernels.mi_from_hist(hist)
    return float(mi[0].item()), int(cnt[0].item())


def register_rigid(fixed, moving, params0=None, cfg=None, **kwargs):
    """6-DoF intensity registration of ``moving`` to ``fixed`` on the GPU (see ``registration.py``)."""
    if cfg is None:
        cfg = RegistrationConfig(**kwargs)
    return _register(fixed, moving, params0=params0, cfg=cfg)


def register_rigid_multires(fixed, moving, params0=None, cfg=None, shrink_factors=(4, 2, 1), iterations=None, **kwargs):
    """Coarse-to-fine registration over block-mean pyramids (``registration.register_rigid_multires``)."""
    if cfg is None:
        cfg = RegistrationConfig(**kwargs)
    return _register_multires(fixed, moving, params0=params0, cfg=cfg, shrink_factors=shrink_factors,
                              iterations=iterations)
