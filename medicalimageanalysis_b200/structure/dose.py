"""``Dose`` with the reference's field names (``structure/dose.py:380-410``) whose ``array`` lives in HBM, for the
three things on the resample path:

* the off-axis reslice of the dose grid (``Display.compute_offaxis_array``, ``dose.py:177-231``) -- shared with
  ``Image`` through ``display.py``;
* ``compute_roi_dose_array`` (``dose.py:738-772``): the dose resampled onto an image lattice
  (``sitk.ResampleImageFilter``, linear, default 0.0) and gathered under an ROI mask;
* ``compute_roi_dose_statistics`` (``dose.py:774-816``): the dose-volume histogram numbers of that sample.

DICOM tags, meshes and persistence are out of scope."""
from __future__ import annotations

import numpy as np
import torch

from .. import geometry as G
from .. import kernels
from ..data import Data
from ..geometry import Grid, index_to_physical, physical_to_index
from .display import OffAxisMixin, VolumeDisplay

D_VALUES = [1, 2, 5, 10, 15, 20, 25, 30, 35, 40, 45, 50, 55, 60, 65, 70, 75, 80, 85, 90, 95, 98, 99]   # dose.py:809


class Dose(OffAxisMixin):
    """Duck-typed from any object carrying ``array, dose_name, spacing, origin, image_matrix`` (+ optional
    ``modality, plane, dimensions, orientation``), the construction contract of ``structure/dose.py:380-410``.
    ``array`` is float32 on the device (the reference holds ``pixel_array * DoseGridScaling``, float64,
    ``read/dicom.py:1905-1911``; float32 is what ``sitk.Resample`` of this path would be asked for after
    ``GetImageFromArray`` of a float32 grid -- stated as a precision choice in DESIGN.md)."""

    def __init__(self, dose, device="cuda"):
        array = dose.array
        if isinstance(array, np.ndarray):
            array = torch.from_numpy(np.ascontiguousarray(array, dtype=np.float32)).to(device)
        self.array = array.to(torch.float32).contiguous()
        self.dose_name = dose.dose_name
        self.modality = getattr(dose, "modality", "RTDOSE")
        self.plane = getattr(dose, "plane", "Axial")
        self.spacing = np.asarray(dose.spacing, dtype=np.float64)
        self.origin = np.asarray(dose.origin, dtype=np.float64)
        self.orientation = np.asarray(getattr(dose, "orientation", [1, 0, 0, 0, 1, 0]), dtype=np.float64)
        dims = getattr(dose, "dimensions", None)
        self.dimensions = np.asarray(dims if dims is not None else self.array.shape)
        self.matrix = np.asarray(dose.image_matrix, dtype=np.float32)
        self.rois = {}
        self.misc = {}
        self.display = VolumeDisplay(self)

    @property
    def grid(self) -> Grid:
        z, y, x = (int(s) for s in self.array.shape)
        return Grid((x, y, z), tuple(self.spacing), tuple(self.origin), self.matrix)

    def compute_matrix_pixel_to_position(self):
        """``structure/dose.py:874-892``."""
        return index_to_physical(self.grid, "itk")

    def compute_matrix_position_to_pixel(self):
        return physical_to_index(self.grid, "itk")

    def compute_center(self, position=True, zyx=False):
        """``structure/dose.py:645-679``."""
        pixel_index = [int(self.dimensions[2] / 2), int(self.dimensions[1] / 2), int(self.dimensions[0] / 2)]
        if position:
            loc = np.asarray([pixel_index[0], pixel_index[1], pixel_index[2], 1.0])
            center = loc.dot(self.compute_matrix_pixel_to_position().T)[:3]
            return np.flip(center) if zyx else center
        return [pixel_index[2], pixel_index[1], pixel_index[0]] if zyx else pixel_index

    # -- dose on an image lattice, under a mask -----------------------------------------------------------------
    def _image_and_mask(self, image_name, roi_name):
        image = Data.image[image_name] if isinstance(image_name, str) else image_name
        mask = image.rois[roi_name] if isinstance(roi_name, str) else roi_name
        mask = getattr(mask, "mask", mask)                     # an ROI-like object with a .mask, or the mask itself
        if isinstance(mask, np.ndarray):
            mask = torch.from_numpy(np.ascontiguousarray(mask)).to(self.array.device)
        return image, mask

    def compute_roi_dose_array(self, image_name, roi_name):
        """``structure/dose.py:738-772``: 1-D float32 device tensor of the dose at every voxel of the image
        lattice where ``mask > 0``, in C order of the mask (the order of ``np.where``).  ``roi_name`` may be a
        key of ``image.rois`` (whose value carries a ``mask`` array ``[z, y, x]``) or the mask itself."""
        image, mask = self._image_and_mask(image_name, roi_name)
        A = G.compose_index_matrix(image.grid, self.grid, np.identity(4), convention="itk")
        return kernels.resample_masked(self.array, A, mask, background=0.0, mode="itk")

    def compute_roi_dose_statistics(self, image_name, roi_name, max_dose=150, increment=5):
        """``structure/dose.py:774-816``: volume, min / max / mean / median / std, D-percent values
        (``np.percentile(dose, 100 - d)``, linear interpolation) and the "VS" volumes below each dose level, from
        ONE fused launch (resample + mask + moments + fine histogram) and an exact device-side selection of the
        order statistics; nothing but the result dictionary crosses PCIe."""
        image, mask = self._image_and_mask(image_name, roi_name)
        A = G.compose_index_matrix(image.grid, self.grid, np.identity(4), convention="itk")
        levels = [float(d) for d in range(0, max_dose + increment, 5)]       # the reference steps by 5 (dose.py:812)
        st = kernels.dose_statistics(self.array, A, mask, D_VALUES, levels, background=0.0, mode="itk")
        spacing = image.spacing
        voxel_vol_cc = float(np.prod(spacing)) / 1000.0
        n = st["count"]
        name = roi_name if isinstance(roi_name, str) else "ROI"
        dvh = {"ROI": name, "Volume (cc)": n * voxel_vol_cc, "Dmin": st["min"], "Dmax": st["max"],
               "Dmean": st["mean"], "Dmedian": st["median"], "Dstd": st["std"]}
        dvh.update({"D%d" % d: v for d, v in zip(D_VALUES, st["percentiles"])})
        for lvl, below in zip(levels, st["below"]):
            d = int(lvl)
            dvh["VS%dGy_percent" % d] = (below / n * 100.0) if n else float("nan")
            dvh["VS%dGy_cc" % d] = below * voxel_vol_cc
        return dvh
