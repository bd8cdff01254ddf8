"""``Image`` with the reference's field names (``structure/image.py:322-346``) whose ``array`` lives in HBM.

Only the fields and methods the rigid resample path reads are mirrored: ``array, spacing, origin, matrix,
dimensions, orientation, plane, modality, image_name`` and ``compute_center`` / pixel<->position matrices.
The viewer state that drives the off-axis reslice (``Display``, ``update_rotation``) lives in
``display.py``.  Tags, ROIs, POIs and persistence are out of scope."""
from __future__ import annotations

import os

import numpy as np
import torch

from ..geometry import Grid, index_to_physical, physical_to_index
from .display import OffAxisMixin, VolumeDisplay


class Image(OffAxisMixin):
    """Duck-typed from any object carrying ``array, image_name, modality, plane, spacing, dimensions,
    orientation, origin, image_matrix`` -- the same construction contract as the reference
    (``structure/image.py:318-350``; DICOM-free producer ``utils/creation.py:257-295``).

    ``array`` may be a NumPy array (uploaded once to ``device``) or a CUDA tensor (adopted)."""

    def __init__(self, image, device="cuda"):
        array = image.array
        if isinstance(array, np.ndarray):
            if array.dtype not in (np.int16, np.float32):
                # other integer types keep their values: int16 when they fit, float32 otherwise (never a silent wrap)
                fits = array.dtype.kind != "f" and array.size and -32768 <= array.min() and array.max() <= 32767
                array = array.astype(np.int16 if fits else np.float32)
            array = torch.from_numpy(np.ascontiguousarray(array)).to(device)
        self.array = array.contiguous()
        self.image_name = image.image_name
        self.modality = getattr(image, "modality", "CT")
        self.plane = getattr(image, "plane", "Axial")
        self.spacing = np.asarray(image.spacing, dtype=np.float64)          # [x, y, z]  read/dicom.py:616-617
        self.origin = np.asarray(image.origin, dtype=np.float64)            # [x, y, z]  read/dicom.py:674
        self.orientation = np.asarray(getattr(image, "orientation", [1, 0, 0, 0, 1, 0]), dtype=np.float64)
        self.dimensions = np.asarray(getattr(image, "dimensions", None) if getattr(image, "dimensions", None)
                                     is not None else self.array.shape)      # [z, y, x]  read/dicom.py:631-632
        self.matrix = np.asarray(image.image_matrix, dtype=np.float32)      # rows = row/col/slice dirs
        self.rois = {}
        self.pois = {}
        self.display = VolumeDisplay(self)                                  # structure/image.py:348

    # -- geometry ------------------------------------------------------------------------------------
    @property
    def grid(self) -> Grid:
        z, y, x = (int(s) for s in self.array.shape)
        return Grid((x, y, z), tuple(self.spacing), tuple(self.origin), self.matrix)

    def compute_matrix_pixel_to_position(self):
        """``structure/image.py:62-85`` (float64 here; the reference uses float32 for viewer math only)."""
        return index_to_physical(self.grid, "itk")

    def compute_matrix_position_to_pixel(self):
        """``structure/image.py:87-108``."""
        return physical_to_index(self.grid, "itk")

    def compute_center(self, position=True, zyx=False):
        """``structure/image.py:1041-1075``: the voxel at int(dim / 2) per axis, optionally in mm."""
        pixel_index = [int(self.dimensions[2] / 2), int(self.dimensions[1] / 2), int(self.dimensions[0] / 2)]
        if position:
            loc = np.asarray([pixel_index[0], pixel_index[1], pixel_index[2], 1.0])
            center = loc.dot(self.compute_matrix_pixel_to_position().T)[:3]
            return np.flip(center) if zyx else center
        return [pixel_index[2], pixel_index[1], pixel_index[0]] if zyx else pixel_index

    def numpy(self):
        return self.array.cpu().numpy()

    # -- persistence -------------------------------------------------------------------------
    _NOT_SAVED = ('rois', 'pois', 'tags', 'array', 'display')           # structure/image.py:726

    def save_image(self, path, rois=True, pois=True):
        """``structure/image.py:709-741``: ``info.p`` (one-row DataFrame of the scalar fields) + ``array.npy``, the
        latter written from the device (same bytes as ``np.save``).  ROI / POI folders are out of scope."""
        import pandas as pd
        from ..read.npy import save_npy_from_device
        column_names = [name for name in self.__dict__.keys() if name not in self._NOT_SAVED]
        df = pd.DataFrame(index=[0], columns=column_names)
        for name in column_names:
            df.at[0, name] = getattr(self, name)
        df.to_pickle(os.path.join(path, 'info.p'))
        if self.array.is_cuda:
            save_npy_from_device(os.path.join(path, 'array.npy'), self.array)
        else:
            np.save(os.path.join(path, 'array.npy'), self.array.numpy(), allow_pickle=True)

    def load_image(self, image_path, rois=True, pois=True, device=None):
        """``structure/image.py:804-836``: ``array.npy`` lands in HBM without a pageable host copy."""
        import pandas as pd
        from ..read.npy import load_npy_to_device
        device = device or self.array.device
        if torch.device(device).type == "cuda":
            self.array = load_npy_to_device(os.path.join(image_path, 'array.npy'), device=device)
        else:
            self.array = torch.from_numpy(np.load(os.path.join(image_path, 'array.npy'), allow_pickle=True))
        info = pd.read_pickle(os.path.join(image_path, 'info.p'))
        for column in list(info.columns):
            setattr(self, column, info.at[0, column])
        self.display = VolumeDisplay(self)
