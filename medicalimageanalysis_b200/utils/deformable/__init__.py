"""Displacement-vector-field operators next to the rigid path -- the device side of what
``structure/deformable.py`` does with ``self.dvf`` once a deformable registration has produced it.  The
registration itself (SimpleITK / Elastix, ``utils/deformable/simpleitk.py``) and the field inversion
(``sitk.InvertDisplacementFieldImageFilter``, ``structure/deformable.py:802-803``) stay out of scope: these functions
take the field they are given.  ``array`` arguments are CUDA tensors; there is no CPU path.
"""
from __future__ import annotations

import numpy as np
import torch

from ... import geometry as G
from ... import kernels

__all__ = ["warp_image", "resample_and_warp", "displace_points"]


def _field(dvf, device):
    """``Deformable.dvf`` is ``[z, y, x, 3]`` (``structure/deformable.py:797``); float32 on the device."""
    if isinstance(dvf, torch.Tensor):
        t = dvf.to(device=device, dtype=torch.float32)
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(dvf, dtype=np.float32))).to(device)
    if t.dim() != 4 or t.shape[3] != 3:
        raise ValueError("dvf must be [z, y, x, 3]")
    return t.contiguous()


def warp_image(image, dvf, dvf_origin=None, dvf_spacing=None, default_value=-3001.0):
    """``sitk.Resample(image, sitk.DisplacementFieldTransform(dvf), sitk.sitkLinear, default_value, pixel id)``
    (``structure/deformable.py:803-806``) for a field that lives on ``image``'s own lattice (``Deformable.spacing`` /
    ``.origin`` are the reference image's, ``:798-799``): every voxel at physical p is re-sampled at p + dvf.
    ``image`` is ``structure.Image``-like; returns a CUDA tensor shaped like the field."""
    grid = image.grid
    field = _field(dvf, image.array.device)
    nz, ny, nx = (int(v) for v in field.shape[:3])
    out_grid = G.Grid((nx, ny, nz), tuple(dvf_spacing if dvf_spacing is not None else grid.spacing),
                      tuple(dvf_origin if dvf_origin is not None else grid.origin), grid.matrix)
    w2p = G.physical_to_index(grid, "itk")
    A = (w2p @ G.index_to_physical(out_grid, "itk"))[:3, :]
    return kernels.warp_dvf(image.array, field, A, w2p[:3, :3], background=default_value, mode="itk")


def resample_and_warp(moving, reference, rigid_matrix, dvf, default_value=-3001.0):
    """Both stages of ``Deformable.create_image`` (``structure/deformable.py:764-806``): the rigid ITK-mode resample
    of ``moving`` onto ``reference``'s lattice (``:788-795``), then the displacement-field resample of that result
    (``:803-806``).  Two passes with the intermediate rounded to the pixel type, exactly as the reference composes
    them."""
    from ..rigid import resample_to_reference

    class _Stage(object):              # the intermediate image: reference lattice, moving pixel type
        pass
    stage = _Stage()
    stage.array = resample_to_reference(moving, reference, rigid_matrix, default_value=default_value)
    stage.grid = reference.grid
    return warp_image(stage, dvf, default_value=default_value)


def displace_points(points, dvf, origin, spacing, percent=100.0):
    """``points + map_coordinates(percent * dvf[..., i] / 100, [z, y, x], order=1, mode='nearest')`` for i = x, y, z
    (``structure/deformable.py:868-880`` with the inverted field, ``:1095-1104`` with the forward one): moves ROI
    mesh vertices through a displacement field.  ``points``: ``[n, 3]`` physical (x, y, z), NumPy or CUDA tensor;
    returns the same kind."""
    as_numpy = not isinstance(points, torch.Tensor)
    field = _field(dvf, torch.device("cuda") if as_numpy or not points.is_cuda else points.device)
    pts = torch.as_tensor(np.asarray(points, dtype=np.float64) if as_numpy else points, dtype=torch.float64,
                          device=field.device).contiguous()
    o = torch.as_tensor(np.asarray(origin, dtype=np.float64), device=field.device)
    s = torch.as_tensor(np.asarray(spacing, dtype=np.float64), device=field.device)
    voxel = ((pts - o) / s).contiguous()
    moved = pts + kernels.sample_field_points(field, voxel, scale=float(percent) / 100.0)
    return moved.cpu().numpy() if as_numpy else moved
