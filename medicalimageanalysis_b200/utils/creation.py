"""DICOM-free ``Image`` producer mirroring ``utils/creation.py:232-376`` (``CreateImageFromMask``): the only
way synthetic volumes enter the reference API.  The fake DICOM tag list the reference builds with pydicom
(``:314-365``) is out of scope; the geometry fields are identical."""
from __future__ import annotations

import numpy as np

from ..data import Data
from ..structure.image import Image


class CreateImageFromMask(object):
    def __init__(self, array, origin, spacing, image_name, dimensions=None, orientation=None, plane="Axial",
                 description="Mask to Image", modality="CT", device="cuda"):
        self.array = array
        self.spacing = spacing
        self.origin = origin
        self.image_name = image_name
        self.plane = plane
        self.description = description
        self.modality = modality
        self.device = device
        self.dimensions = tuple(array.shape) if dimensions is None else dimensions
        self.orientation = [1, 0, 0, 0, 1, 0] if orientation is None else orientation
        row = np.asarray(self.orientation[:3], dtype=np.float64)
        col = np.asarray(self.orientation[3:6], dtype=np.float64)
        self.image_matrix = np.identity(3, dtype=np.float32)       # utils/creation.py:290-295
        self.image_matrix[0, :3] = row
        self.image_matrix[1, :3] = col
        self.image_matrix[2, :3] = np.cross(row, col)
        self.image = None

    def add_image(self):
        """``utils/creation.py:367-376``."""
        self.image = Image(self, device=self.device)
        Data.image[self.image_name] = self.image
        Data.image_list += [self.image_name]
        return self.image
