"""Viewer state of a volume (``Display`` of ``structure/image.py:39-307`` and its twin ``structure/dose.py:53-320``):
slice location per plane, the viewing rotation, and the OFF-AXIS RESLICE of the whole volume that a rotation nudge
triggers (``compute_offaxis_array``, ``image.py:160-215`` / ``dose.py:177-231``) -- a ``vtkImageReslice`` with the
inverse viewing rotation as reslice transform, linear interpolation, output spacing = volume spacing, auto-cropped
extent and background -3001.  Here that reslice is the sm_100a tile kernel (VTK mode) and the resliced volume
(``secondary_array``) stays in HBM; a 2-D slice leaves the device only when ``compute_array`` is asked for it.

The VTK slice / volume objects (``compute_vtk_slice``, ``compute_vtk_volume``) are rendering plumbing and out of
scope (graphics stack: unavailable on this box)."""
from __future__ import annotations

import copy

import numpy as np

from .. import geometry as G
from .. import kernels

BACKGROUND = -3001          # structure/image.py:195, structure/dose.py:211


class VolumeDisplay(object):
    def __init__(self, owner):
        self.owner = owner                                   # an Image or a Dose: array, spacing, origin, matrix
        self.matrix = np.array(owner.matrix, dtype=np.float64)
        self.spacing = copy.deepcopy(owner.spacing)
        self.origin = np.array(owner.origin, dtype=np.float64)
        self.slice_location = owner.compute_center(position=False, zyx=True)
        self.scroll_max = [int(owner.dimensions[0]) - 1, int(owner.dimensions[1]) - 1, int(owner.dimensions[2]) - 1]
        self.secondary_array = None
        self.misc = {}

    # -- index <-> physical with the VIEWING rotation (structure/image.py:62-108) -----------------------------
    def compute_matrix_pixel_to_position(self):
        m = np.identity(4)
        for k in range(3):
            m[:3, k] = self.matrix[k, :] * self.spacing[k]
        m[:3, 3] = self.origin
        return m

    def compute_matrix_position_to_pixel(self):
        hold = np.identity(3)
        for k in range(3):
            hold[k, :] = self.matrix[k, :] / self.spacing[k]
        m = np.identity(4)
        m[:3, :3] = hold
        m[:3, 3] = np.asarray(self.origin).dot(-hold.T)
        return m

    def compute_index_positions(self, xyz):
        """``structure/image.py:141-158``."""
        return np.asarray([xyz[0], xyz[1], xyz[2], 1.0]).dot(self.compute_matrix_pixel_to_position().T)[:3]

    # -- slices -------------------------------------------------------------------------------------------
    def current_array(self):
        return self.owner.array if self.secondary_array is None else self.secondary_array

    def compute_array(self, slice_plane):
        """``structure/image.py:110-139``: one float32 2-D slice (only that slice crosses PCIe)."""
        a = self.current_array()
        if slice_plane == 'Axial':
            s = a[int(self.slice_location[0]), :, :]
        elif slice_plane == 'Coronal':
            s = a[:, int(self.slice_location[1]), :]
        else:
            s = a[:, :, int(self.slice_location[2])]
        return s.float().cpu().numpy()

    def update_slice_location(self, scroll, slice_plane):
        """``structure/image.py:286-306``."""
        self.slice_location[{'Axial': 0, 'Coronal': 1}.get(slice_plane, 2)] = scroll

    def compute_scroll_max(self):
        """``structure/dose.py:233-244`` (the Image twin at ``image.py:217-232`` is a stub)."""
        a = self.current_array()
        self.scroll_max = [int(a.shape[0]) - 1, int(a.shape[1]) - 1, int(a.shape[2]) - 1]

    # -- the off-axis reslice -----------------------------------------------------------------------------
    def offaxis_geometry(self):
        """Output lattice and index matrix of the off-axis reslice: (origin_xyz, dims_xyz, A).

        VTK sees the volume with direction = ``owner.matrix`` read row-major (``image.py:170-176``, the 'vtk'
        convention of ``geometry._direction``), the reslice transform is the INVERSE of the viewing rotation
        (``:178-185``: output physical -> input physical), the output direction is the identity and the output
        spacing the volume's own (``:191``)."""
        R = np.identity(4)
        R[:3, :3] = self.matrix
        M = np.linalg.inv(R)
        mgrid = self.owner.grid
        origin, dims = G.autocrop_extent(mgrid, M, self.owner.spacing, convention="vtk")
        out_grid = G.Grid(dims, tuple(self.owner.spacing), origin, np.identity(3))
        A = G.compose_index_matrix(out_grid, mgrid, M, convention="vtk")
        return origin, dims, A, M

    def compute_offaxis_array(self):
        """``structure/image.py:160-215`` / ``structure/dose.py:177-231``."""
        loc = np.flip(np.asarray(self.slice_location, dtype=np.float64))
        slice_position = np.asarray([loc[0], loc[1], loc[2], 1.0]).dot(self.compute_matrix_pixel_to_position().T)[:3]

        new_origin, dims, A, M = self.offaxis_geometry()
        self.secondary_array = kernels.resample(self.owner.array, A, (dims[2], dims[1], dims[0]),
                                                background=BACKGROUND, mode="vtk")
        # transform.TransformPoint(new_origin) with the inverted transform (image.py:199-200)
        self.origin = (M @ np.asarray([new_origin[0], new_origin[1], new_origin[2], 1.0]))[:3]

        location = np.asarray([slice_position[0], slice_position[1], slice_position[2], 1.0])
        pix = location.dot(self.compute_matrix_position_to_pixel().T)[:3]
        self.slice_location = [int(v) for v in np.flip(np.round(pix)).astype(np.int32)]
        self.scroll_max = [dims[2] - 1, dims[1] - 1, dims[0] - 1]
        for a in range(3):                                   # upper clamp only, as the reference (image.py:207-212)
            if self.slice_location[a] > self.scroll_max[a]:
                self.slice_location[a] = self.scroll_max[a]


class OffAxisMixin(object):
    """``update_rotation`` / ``reset_array`` / ``retrieve_*`` shared by Image (``image.py:1224-1413``) and Dose
    (``dose.py:920-1115``).  The host class provides ``matrix, origin, compute_center`` and ``self.display``."""

    def reset_array(self):
        self.display.secondary_array = None
        self.display.matrix = np.array(self.matrix, dtype=np.float64)
        self.display.origin = np.array(self.origin, dtype=np.float64)
        self.display.slice_location = self.compute_center(position=False, zyx=True)

    def update_rotation(self, r_x=0, r_y=0, r_z=0, base=True):
        """``structure/image.py:1380-1413``: compose the viewing rotation ('xyz' Euler, degrees) on the image
        matrix (``base``) or on the current view, and reslice the volume off-axis on the GPU."""
        if r_x != 0 or r_y != 0 or r_z != 0:
            new_matrix = G.euler_xyz_matrix(r_x, r_y, r_z)
            if base:
                self.display.matrix = new_matrix @ np.array(self.matrix, dtype=np.float64)
            else:
                self.display.matrix = new_matrix @ self.display.matrix
            self.display.compute_offaxis_array()
            self.display.compute_scroll_max()
        else:
            self.display.compute_scroll_max()
            self.reset_array()

    def retrieve_angles(self, order='ZXY'):
        if order != 'ZXY':
            raise NotImplementedError("only the reference's default order 'ZXY' is mirrored")
        return G.euler_from_matrix_ZXY(self.display.matrix[:3, :3])

    def retrieve_array_plane(self, slice_plane):
        return self.display.compute_array(slice_plane=slice_plane)

    def retrieve_slice_location(self, slice_plane):
        return self.display.slice_location[{'Axial': 0, 'Coronal': 1}.get(slice_plane, 2)]

    def retrieve_scroll_max(self, slice_plane):
        return self.display.scroll_max[{'Axial': 0, 'Coronal': 1}.get(slice_plane, 2)]

    def retrieve_slice_position(self, slice_plane=None):
        """``structure/image.py:1294-1323``."""
        sl = self.display.slice_location
        if slice_plane is None:
            location = np.asarray([sl[2], sl[1], sl[0], 1.0])
        elif slice_plane == 'Axial':
            location = np.asarray([0, 0, sl[0], 1.0])
        elif slice_plane == 'Coronal':
            location = np.asarray([0, sl[1], 0, 1.0])
        else:
            location = np.asarray([sl[2], 0, 0, 1.0])
        return location.dot(self.display.compute_matrix_pixel_to_position().T)[:3]
