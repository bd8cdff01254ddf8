"""Host-side float64 geometry for the rigid resample hot path.

Everything here is tiny 4x4 algebra; it runs on the host in float64 and ends in ONE
3x4 matrix ``A`` (row-major, 12 doubles) that maps an output voxel index ``(i, j, k)``
(x, y, z order) to a continuous index in the moving volume.  The CUDA kernels only ever
see ``A`` so neither the VTK nor the ITK direction-matrix convention leaks below this file.

Reference sites restated here (paths relative to the reference checkout):
  * index -> physical / physical -> index: ``structure/image.py:62-108``
  * transform assembly ``M = matrix @ combo_matrix`` (+ Inverse): ``structure/rigid.py:674-683``
  * rotation / translation nudges: ``structure/rigid.py:950-989``
  * Euler parameterisation handed to ITK: ``utils/image/transform.py:15-38``,
    ``structure/deformable.py:575-577,778-783``
  * AutoCrop output grid of vtkImageReslice: ``structure/rigid.py:685-693``
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

__all__ = [
    "Grid", "euler_xyz_matrix", "euler_from_matrix_ZXY", "itk_euler_matrix", "pivot_transform",
    "translation_transform", "index_to_physical", "physical_to_index", "compose_index_matrix",
    "autocrop_extent", "params_to_matrix", "slab_moving_zrange", "is_signed_permutation", "pyramid_grid",
]


@dataclass
class Grid:
    """Geometry of a voxel lattice, field names as on the reference ``Image``
    (``structure/image.py:339-346``): ``dims`` is (x, y, z) voxel counts, ``spacing`` and
    ``origin`` are (x, y, z) in mm, ``matrix`` is 3x3 with ROWS = row/col/slice direction."""
    dims: tuple
    spacing: tuple
    origin: tuple
    matrix: np.ndarray = None

    def __post_init__(self):
        self.dims = tuple(int(d) for d in self.dims)
        self.spacing = tuple(float(s) for s in self.spacing)
        self.origin = tuple(float(o) for o in self.origin)
        if self.matrix is None:
            self.matrix = np.identity(3)
        self.matrix = np.asarray(self.matrix, dtype=np.float64).reshape(3, 3)


def _rot(axis: int, deg: float) -> np.ndarray:
    a = math.radians(float(deg))
    c, s = math.cos(a), math.sin(a)
    r = np.identity(3)
    i, j = [(1, 2), (2, 0), (0, 1)][axis]
    r[i, i] = c
    r[i, j] = -s
    r[j, i] = s
    r[j, j] = c
    return r


def euler_xyz_matrix(r_x: float, r_y: float, r_z: float) -> np.ndarray:
    """``Rotation.from_euler('xyz', [rx, ry, rz], degrees=True).as_matrix()``
    (``structure/rigid.py:957``): extrinsic x, then y, then z  =>  R = Rz @ Ry @ Rx."""
    return _rot(2, r_z) @ _rot(1, r_y) @ _rot(0, r_x)


def euler_from_matrix_ZXY(R: np.ndarray) -> np.ndarray:
    """``Rotation.from_matrix(R).as_euler('ZXY', degrees=True)`` (``structure/rigid.py:756-757``):
    intrinsic Z, X', Y''  =>  R = Rz(a) @ Rx(b) @ Ry(c); returns [a, b, c] in degrees."""
    R = np.asarray(R, dtype=np.float64)[:3, :3]
    # R = Rz(a) Rx(b) Ry(c):  R[2,1] = sin b ; R[0,1] = -sin a cos b ; R[1,1] = cos a cos b
    #                         R[2,0] = -cos b sin c ; R[2,2] = cos b cos c
    sb = min(1.0, max(-1.0, R[2, 1]))
    b = math.asin(sb)
    if abs(sb) < 1.0 - 1e-12:
        a = math.atan2(-R[0, 1], R[1, 1])
        c = math.atan2(-R[2, 0], R[2, 2])
    else:  # gimbal lock: fold everything into the first angle (scipy sets the third to 0)
        a = math.atan2(R[1, 0], R[0, 0])
        c = 0.0
    return np.degrees([a, b, c])


def itk_euler_matrix(angle_x: float, angle_y: float, angle_z: float, zyx: bool = False) -> np.ndarray:
    """Rotation block of ``sitk.Euler3DTransform`` for angles in RADIANS
    (``utils/image/transform.py:23-36``): default R = Rz @ Rx @ Ry, ``ComputeZYX`` R = Rz @ Ry @ Rx."""
    rx = _rot(0, math.degrees(angle_x))
    ry = _rot(1, math.degrees(angle_y))
    rz = _rot(2, math.degrees(angle_z))
    return rz @ ry @ rx if zyx else rz @ rx @ ry


def pivot_transform(center, r_x=0.0, r_y=0.0, r_z=0.0) -> np.ndarray:
    """``T(+c) @ R @ T(-c)`` exactly as ``Rigid.update_rotation`` builds it (``structure/rigid.py:957-965``)."""
    c = np.asarray(center, dtype=np.float64).reshape(3)
    R = np.identity(4)
    R[:3, :3] = euler_xyz_matrix(r_x, r_y, r_z)
    t_neg = np.identity(4)
    t_neg[:3, 3] = -c
    t_pos = np.identity(4)
    t_pos[:3, 3] = c
    return t_pos @ R @ t_neg


def translation_transform(t_x=0.0, t_y=0.0, t_z=0.0) -> np.ndarray:
    """4x4 translation as in ``Rigid.update_translation`` (``structure/rigid.py:977-980``)."""
    T = np.identity(4)
    T[:3, 3] = (t_x, t_y, t_z)
    return T


def _direction(matrix: np.ndarray, convention: str) -> np.ndarray:
    """3x3 whose COLUMNS are the physical directions of the x, y, z index axes.

    ``matrix`` has ROWS = axis directions (``read/dicom.py:640-653``).
      * 'itk' : the reference hands ITK ``matrix.flatten(order='F')`` (``structure/image.py:926-927``),
                i.e. columns = axes: D = matrix.T.  Same as ``structure/image.py:79-83``.
      * 'vtk' : the reference hands VTK ``matrix.reshape(1, 9)[0]`` (``structure/rigid.py:666-669``);
                VTK reads that row-major with columns = axes, so it sees D = matrix (the transpose
                of what ITK sees).  Reproduced as-is, not "fixed" (SURVEY F8).
    For the identity matrix of FFS-normalised axial images the two agree."""
    m = np.asarray(matrix, dtype=np.float64).reshape(3, 3)
    if convention == "itk":
        return m.T.copy()
    if convention == "vtk":
        return m.copy()
    raise ValueError("direction convention must be 'itk' or 'vtk'")


def index_to_physical(grid: Grid, convention: str = "itk") -> np.ndarray:
    """4x4 index(x,y,z,1) -> physical mm.  Restates ``structure/image.py:62-85`` in float64."""
    D = _direction(grid.matrix, convention)
    P = np.identity(4)
    P[:3, :3] = D * np.asarray(grid.spacing, dtype=np.float64)[None, :]
    P[:3, 3] = grid.origin
    return P


def physical_to_index(grid: Grid, convention: str = "itk") -> np.ndarray:
    """4x4 physical mm -> continuous index.  Restates ``structure/image.py:87-108`` in float64."""
    return np.linalg.inv(index_to_physical(grid, convention))


def compose_index_matrix(out_grid: Grid, mov_grid: Grid, M: np.ndarray,
                         convention: str = "itk", out_convention: str | None = None) -> np.ndarray:
    """The single 3x4 ``A``: output index -> moving continuous index.

    ``M`` maps fixed/output physical -> moving physical (the direction shared by VTK's
    ResliceTransform and ITK's resample transform, ``structure/rigid.py:680-686``,
    ``structure/deformable.py:575-579``):   A = W2P_mov @ M @ P2W_out."""
    M = np.asarray(M, dtype=np.float64).reshape(4, 4)
    p2w = index_to_physical(out_grid, out_convention or convention)
    w2p = physical_to_index(mov_grid, convention)
    return np.ascontiguousarray((w2p @ M @ p2w)[:3, :])


def autocrop_extent(mov_grid: Grid, M: np.ndarray, out_spacing, convention: str = "vtk",
                    extent_rule: str = "round"):
    """Output origin (x,y,z) and dims (x,y,z) of ``vtkImageReslice`` with ``AutoCropOutputOn`` and
    identity output direction (``structure/rigid.py:689-691``).

    Restatement of VTK 9.3 ``vtkImageReslice::GetAutoCroppedOutputBounds`` + ``RequestInformation``
    (VTK is a pinned dependency, ``requirements.txt:82``; its source is not vendored): push the 8
    corners of the moving volume through inverse(M), take per-axis min/max; origin = min;
    last index = Round(|max - min| / spacing)  => dims = that + 1.  ``extent_rule='ceil'`` is kept
    for comparison (SURVEY R1 guessed ceil); parity on this arithmetic is UNPINNED (no VTK here)."""
    Minv = np.linalg.inv(np.asarray(M, dtype=np.float64).reshape(4, 4))
    p2w = index_to_physical(mov_grid, convention)
    lo = np.full(3, np.inf)
    hi = np.full(3, -np.inf)
    nx, ny, nz = mov_grid.dims
    for c in range(8):
        idx = np.array([(nx - 1) * (c & 1), (ny - 1) * ((c >> 1) & 1), (nz - 1) * ((c >> 2) & 1), 1.0])
        p = Minv @ (p2w @ idx)
        p = p[:3] / p[3]
        lo = np.minimum(lo, p)
        hi = np.maximum(hi, p)
    sp = np.asarray(out_spacing, dtype=np.float64)
    ext = np.abs((hi - lo) / sp)
    if extent_rule == "round":
        last = np.floor(ext + 0.5)
    elif extent_rule == "ceil":
        last = np.ceil(ext)
    else:
        raise ValueError("extent_rule must be 'round' or 'ceil'")
    dims = tuple(int(v) + 1 for v in last)
    return tuple(float(v) for v in lo), dims


def params_to_matrix(params, center) -> np.ndarray:
    """6-DoF parameterisation of the registration loop: ``[rx, ry, rz]`` degrees,
    'xyz' Euler about ``center`` (the fixed image centre, ``structure/image.py:1041-1069``),
    then ``[tx, ty, tz]`` mm.  Returns M: fixed physical -> moving physical,
    M = T(t) @ T(c) @ R @ T(-c)  (the composition rule of ``structure/rigid.py:957-970``)."""
    p = np.asarray(params, dtype=np.float64).reshape(6)
    return translation_transform(p[3], p[4], p[5]) @ pivot_transform(center, p[0], p[1], p[2])


def slab_moving_zrange(A: np.ndarray, out_dims, k0: int, k1: int, mov_nz: int, halo: int = 1):
    """Moving z-range [z0, z1) touched by output z-slab [k0, k1): AABB of the slab's 8 corners
    under ``A``, +-halo, clipped.  Used by the z-slab sharding."""
    A = np.asarray(A, dtype=np.float64).reshape(3, 4)
    nx, ny, _ = out_dims
    zs = []
    for i in (0, nx - 1):
        for j in (0, ny - 1):
            for k in (k0, k1 - 1):
                zs.append(A[2, 0] * i + A[2, 1] * j + A[2, 2] * k + A[2, 3])
    z0 = int(math.floor(min(zs))) - halo
    z1 = int(math.floor(max(zs))) + 1 + halo + 1
    return max(0, z0), min(mov_nz, max(z1, 0))


def is_signed_permutation(A: np.ndarray, tol: float = 0.0):
    """If the 3x3 part of ``A`` is a signed permutation with integer offsets return (perm, flip):
    output axis a reads moving axis ``perm[a]``... expressed moving-side: moving axis r is driven by
    output axis ``perm[r]`` with sign ``flip[r]``.  Else None.  This is the FFS case."""
    A = np.asarray(A, dtype=np.float64).reshape(3, 4)
    perm, flip = [], []
    for r in range(3):
        row = A[r, :3]
        nz = np.nonzero(np.abs(row) > tol)[0]
        if len(nz) != 1 or abs(abs(row[nz[0]]) - 1.0) > tol or abs(A[r, 3] - round(A[r, 3])) > tol:
            return None
        perm.append(int(nz[0]))
        flip.append(row[nz[0]] < 0)
    if sorted(perm) != [0, 1, 2]:
        return None
    return tuple(perm), tuple(flip)


def pyramid_grid(grid: Grid, factors, convention: str = "itk") -> Grid:
    """Lattice of a block-mean pyramid level: dims = ceil(dims / f), spacing * f, and the origin moved
    to the centre of the first block, i.e. by (f - 1) / 2 voxels along each index axis."""
    f = np.asarray([int(v) for v in factors], dtype=np.float64)
    dims = tuple(int(-(-d // int(ff))) for d, ff in zip(grid.dims, f))
    shift = index_to_physical(grid, convention)[:3, :3] @ ((f - 1.0) / 2.0)
    origin = tuple(np.asarray(grid.origin, dtype=np.float64) + shift)
    return Grid(dims, tuple(np.asarray(grid.spacing) * f), origin, grid.matrix)
