"""``Rigid`` / ``Display`` with the reference's names and semantics (``structure/rigid.py:36-364,366-1011``)
for the resample path, with ``vtkImageReslice`` replaced by the sm_100a kernel behind the C ABI.

Mirrored: ``matrix, combo_matrix, reference_matrix, inverse``; ``create_image`` (``:646-695``);
``display.compute_reslice`` / ``compute_offset`` / ``compute_array_slice`` (``:180-202, 89-111, 60-87``);
``update_rotation`` / ``update_translation`` (``:950-989``); ``pre_alignment(origin=True)`` (``:718-740``);
``retrieve_angles`` / ``retrieve_translation`` / ``retrieve_center`` / ``retrieve_array_plane``.
Out of scope: ROI mesh transforms, ICP, VTK slice objects, pickle persistence.
"""
from __future__ import annotations

import itertools

import numpy as np

from .. import geometry as G
from .. import kernels
from ..data import Data

BACKGROUND = -3001          # structure/rigid.py:692

# slice plane -> (array axis [z, y, x] that the plane fixes, the two physical axes (x=0, y=1, z=2) spanning the plane)
_PLANES = {'Axial': (0, (0, 1)), 'Coronal': (1, (0, 2)), 'Sagittal': (2, (1, 2))}


class ResampledImage(object):
    """What ``create_image`` returns: the device array plus the output lattice, with the three getters
    ``Display.compute_reslice`` calls on the reference's ``vtkImageData`` (``structure/rigid.py:194-199``)."""

    def __init__(self, array, origin, spacing):
        self.array = array
        self.origin = tuple(origin)
        self.spacing = tuple(spacing)

    def GetOrigin(self):
        return self.origin

    def GetSpacing(self):
        return self.spacing

    def GetDimensions(self):
        z, y, x = self.array.shape
        return (int(x), int(y), int(z))


class Display(object):
    def __init__(self, rigid):
        self.rigid = rigid
        self.origin = None
        self.spacing = None
        self.array = None
        self.matrix = np.identity(4)
        self.slice_location = [0, 0, 0]
        self.scroll_max = [0, 0, 0]
        self.offset = {'Axial': [0, 0], 'Coronal': [0, 0], 'Sagittal': [0, 0]}
        self.misc = {}

    def compute_reslice(self):
        """``structure/rigid.py:180-202``; ``array`` stays a CUDA tensor ``[dz, dy, dx]``."""
        image = self.rigid.create_image()
        self.origin = np.asarray(image.GetOrigin())
        self.spacing = image.GetSpacing()
        self.array = image.array
        self.compute_offset()
        self.compute_scroll_max()

    def compute_offset(self):
        """Where the resliced lattice starts relative to the fixed image, in output voxels, per in-plane axis of each
        slice plane (semantics of ``structure/rigid.py:89-111``; the fixed image is the moving one when inverted)."""
        fixed = Data.image[self.rigid.moving_name if self.rigid.inverse else self.rigid.reference_name]
        shift = (np.asarray(self.origin, dtype=float) - np.asarray(fixed.origin, dtype=float)) / np.asarray(self.spacing, dtype=float)
        for plane, (_, (u, v)) in _PLANES.items():
            self.offset[plane] = [shift[u], shift[v]]

    def compute_scroll_max(self):
        """``structure/rigid.py:237-252``."""
        self.scroll_max = [self.array.shape[0] - 1, self.array.shape[1] - 1, self.array.shape[2] - 1]

    def compute_slice_location(self, position=None):
        """``structure/rigid.py:204-227`` with an explicit position (the viewer state that supplies the
        default is out of scope): nearest voxel of a physical point, [z, y, x]."""
        if position is None:
            position = self.origin
        self.slice_location = np.flip(np.round((np.asarray(position) - self.origin) / self.spacing).astype(np.int32))

    def compute_array_slice(self, slice_plane):
        """One 2-D float64 slice through ``slice_location`` (semantics of ``structure/rigid.py:60-87``; anything that
        is not Axial / Coronal counts as Sagittal, a location outside the array gives None).  Only that slice crosses
        PCIe."""
        axis = _PLANES.get(slice_plane, _PLANES['Sagittal'])[0]
        index = int(self.slice_location[axis])
        if not 0 <= index < self.array.shape[axis]:
            return None
        return self.array.select(axis, index).cpu().numpy().astype(np.double)

    def compute_matrix_pixel_to_position(self):
        """``structure/rigid.py:113-120``."""
        m = np.identity(4)
        m[0, 0], m[1, 1], m[2, 2] = self.spacing
        m[:3, 3] = self.origin
        return m

    def compute_matrix_position_to_pixel(self):
        """``structure/rigid.py:122-132``."""
        return np.linalg.inv(self.compute_matrix_pixel_to_position())


class Rigid(object):
    def __init__(self, reference_name, moving_name, rigid_name=None, roi_names=None, reference_matrix=None,
                 matrix=None, combo_matrix=None, combo_name=None):
        self.reference_name = reference_name
        self.moving_name = moving_name
        self.combo_name = combo_name
        self.roi_names = ['Unknown'] if roi_names is None else roi_names
        self.reference_matrix = np.identity(4) if reference_matrix is None else np.asarray(reference_matrix, dtype=float)
        self.matrix = np.identity(4) if matrix is None else np.asarray(matrix, dtype=float)
        self.combo_matrix = np.identity(4) if combo_matrix is None else np.asarray(combo_matrix, dtype=float)
        self.inverse = False
        self.misc = {}
        self.rotation_center = np.asarray([0, 0, 0])
        self.rigid_name = self.add_rigid(rigid_name)
        self.display = Display(self)

    def add_rigid(self, rigid_name):
        """Register this object in ``Data.rigid`` (semantics of ``structure/rigid.py:434-466``): an unnamed rigid is
        called ``<reference>_<moving>`` (``..._combo`` when a combo matrix is set) and takes the first free ``_<n>``
        suffix, n = 1, 2, ..., if that name exists."""
        if rigid_name is None:
            base = '%s_%s' % (self.reference_name, self.moving_name)
            if not np.array_equal(self.combo_matrix, np.identity(4)):
                base += '_combo'
            candidates = itertools.chain([base], ('%s_%d' % (base, n) for n in itertools.count(1)))
            rigid_name = next(name for name in candidates if name not in Data.rigid_list)
        Data.rigid[rigid_name] = self
        Data.rigid_list += [rigid_name]
        return rigid_name

    # -- the resampler -----------------------------------------------------------------------------
    def _names(self):
        if self.inverse:
            return self.moving_name, self.reference_name      # (ref, mov)  structure/rigid.py:659-664
        return self.reference_name, self.moving_name

    def transform_matrix(self):
        """``structure/rigid.py:674-683``: M = matrix @ combo_matrix, inverted when ``self.inverse``."""
        m = self.matrix @ self.combo_matrix
        return np.linalg.inv(m) if self.inverse else m

    def create_image(self):
        """``structure/rigid.py:646-695`` on the GPU: linear reslice of the moving image, output spacing =
        fixed spacing, identity output direction, auto-cropped extent, background -3001."""
        ref, mov = self._names()
        mov_img, ref_img = Data.image[mov], Data.image[ref]
        M = self.transform_matrix()
        mgrid = mov_img.grid
        origin, dims = G.autocrop_extent(mgrid, M, ref_img.spacing, convention="vtk")
        out_grid = G.Grid(dims, tuple(ref_img.spacing), origin, np.identity(3))
        A = G.compose_index_matrix(out_grid, mgrid, M, convention="vtk")
        out = kernels.resample(mov_img.array, A, (dims[2], dims[1], dims[0]), background=BACKGROUND, mode="vtk")
        return ResampledImage(out, origin, tuple(ref_img.spacing))

    def export_image(self, path=None):
        """``structure/rigid.py:697-716`` (``vtkMetaImageWriter``): the aligned volume as an .mhd / .raw pair, the
        payload written from the device."""
        if self.moving_name is not None and path is not None:
            from ..read.mhd import write_mhd
            image = self.create_image()
            write_mhd(path, image.array, spacing=image.spacing, origin=image.origin)

    # -- parameter updates -------------------------------------------------------------------------
    def retrieve_center(self, slice_plane=None, location=None):
        """``structure/rigid.py:786-812``: centre of the FIXED image, optionally projected on a slice plane."""
        fixed_name = self.moving_name if self.inverse else self.reference_name
        fixed_center = np.asarray(Data.image[fixed_name].compute_center(), dtype=float)
        if slice_plane is None or location is None:
            return fixed_center
        col = {'Axial': 2, 'Coronal': 1}.get(slice_plane, 0)
        normal = self.display.matrix[:3, col]
        normal_hat = normal / np.linalg.norm(normal)
        offset = np.dot(fixed_center - np.asarray(location, dtype=float), normal_hat)
        return fixed_center - offset * normal_hat

    def _apply_increment(self, transform):
        """Compose an increment given in the FIXED image's frame into ``matrix`` and refresh the display: the inverted
        rigid takes it on the left, the forward one takes its inverse on the right (``structure/rigid.py:966-970, 981-985``)."""
        self.matrix = transform @ self.matrix if self.inverse else self.matrix @ np.linalg.inv(transform)
        self.display.compute_reslice()

    def update_rotation(self, center=None, r_x=0, r_y=0, r_z=0, slice_plane=None, location=None):
        """Rotate by 'xyz' Euler degrees about ``center`` (default: the fixed image centre, optionally projected on a
        slice plane) -- ``structure/rigid.py:950-974``."""
        if center is None:
            center = self.retrieve_center(slice_plane=slice_plane, location=location)
        self._apply_increment(G.pivot_transform(center, r_x, r_y, r_z))

    def update_translation(self, t_x=0, t_y=0, t_z=0):
        """Translate by millimetres -- ``structure/rigid.py:976-989``."""
        self._apply_increment(G.translation_transform(t_x, t_y, t_z))

    def pre_alignment(self, superior=False, center=False, origin=False):
        """``structure/rigid.py:718-740``."""
        if superior or center:
            return
        if origin:
            self.matrix[:3, 3] = Data.image[self.moving_name].origin - Data.image[self.reference_name].origin

    def retrieve_angles(self, order='ZXY'):
        """``structure/rigid.py:742-757``."""
        if order != 'ZXY':
            raise NotImplementedError("only the reference's default order 'ZXY' is mirrored")
        return G.euler_from_matrix_ZXY(self.matrix[:3, :3])

    def retrieve_translation(self):
        """``structure/rigid.py:899-911``."""
        return self.matrix[:3, 3]

    def retrieve_array_plane(self, slice_plane, solo=None, position=None):
        """The slice of the resliced volume a viewer shows (``structure/rigid.py:759-784``): resample on first use; unless
        the caller scrolls on its own (``solo``), move the slice location to ``position`` first."""
        if self.display.array is None:
            self.display.compute_reslice()
        if solo is None:
            self.display.compute_slice_location(position=position)
        return self.display.compute_array_slice(slice_plane=slice_plane)

    def retrieve_offset(self, slice_plane):
        return self.display.offset[slice_plane]

    def retrieve_scroll_max(self, slice_plane):
        return self.display.scroll_max[{'Axial': 0, 'Coronal': 1}.get(slice_plane, 2)]
