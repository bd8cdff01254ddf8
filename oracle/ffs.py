"""Oracle for the Feet-First-Supine normalisation (test infrastructure only).

This one IS pinned by reference code: ``Read3D._verify_axial_orientation`` (``read/dicom.py:655-740``,
RTDOSE twin ``:2027-2106``) is plain NumPy (``np.rot90`` / ``np.flip`` / ``transpose`` views), restated
here with the same NumPy calls so the CUDA permutation kernel can be checked bit for bit.
"""
from __future__ import annotations

import numpy as np

# case name -> the reference's array expression
FFS_CASES = ("axial_k1", "axial_k2", "axial_k3", "coronal", "sagittal")


def ffs_corner_index(origin, orientation, spacing_xyz, shape_zyx):
    """``read/dicom.py:666-692``: the 8 corners of the volume and argmin of (x + y + z)."""
    origin = np.asarray(origin, dtype=np.float64)
    row = np.asarray(orientation[:3], dtype=np.float64)
    col = np.asarray(orientation[3:], dtype=np.float64)
    slc = np.cross(row, col)
    slices, y, x = shape_zyx[0] - 1, shape_zyx[1] - 1, shape_zyx[2] - 1
    sp = spacing_xyz
    corners = np.zeros((8, 3))
    corners[0] = origin
    corners[1] = origin + x * sp[0] * row
    corners[2] = origin + y * sp[1] * col
    corners[3] = origin + x * sp[0] * row + y * sp[1] * col
    corners[4] = origin + slices * sp[2] * slc
    corners[5] = corners[4] + x * sp[0] * row
    corners[6] = corners[4] + y * sp[1] * col
    corners[7] = corners[4] + x * sp[0] * row + y * sp[1] * col
    return int(np.argmin(np.sum(corners, axis=1))), corners


def axial_case_for_corner(corner_idx):
    """``read/dicom.py:695-701``: corner 1 -> rot90 k=1, corner 2 -> k=3, every other non-zero corner
    (3..7, with no z flip -- reproduced as-is) -> k=2."""
    if corner_idx == 0:
        return None
    return {1: "axial_k1", 2: "axial_k3"}.get(corner_idx, "axial_k2")


def oracle_ffs(array, case):
    """The reference's permutation, materialised (the reference keeps a view)."""
    a = np.asarray(array)
    if case == "axial_k1":
        r = np.rot90(a, 1, (1, 2))                                     # dicom.py:697
    elif case == "axial_k3":
        r = np.rot90(a, 3, (1, 2))                                     # dicom.py:699
    elif case == "axial_k2":
        r = np.rot90(a, 2, (1, 2))                                     # dicom.py:701
    elif case == "coronal":
        r = np.rot90(a, 1, (0, 1))                                     # dicom.py:709
    elif case == "sagittal":
        r = np.flip(np.rot90(a, 1, (0, 1)).transpose(0, 2, 1), axis=2)  # dicom.py:717
    else:
        raise ValueError(case)
    return np.ascontiguousarray(r)
