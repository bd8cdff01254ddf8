"""MetaImage (.mhd + .raw / .zraw) volumes straight to the device -- ``MhdReader`` with the reference's
constructor and dispatch (``read/mhd.py:60-252``), minus SimpleITK: the text header is parsed here, the payload goes
from the file to HBM through ``mia_read_file_to_device``.  ``write_mhd`` is what ``Rigid.export_image``
(``structure/rigid.py:697-716``, ``vtkMetaImageWriter``) becomes for a device-resident reslice.

MetaImage header keys honoured: NDims, DimSize, ElementSpacing / ElementSize, Offset / Position / Origin,
TransformMatrix / Orientation / Rotation, ElementType, ElementNumberOfChannels, ElementByteOrderMSB /
BinaryDataByteOrderMSB, HeaderSize, CompressedData (zlib), ElementDataFile (a file name or LOCAL)."""
from __future__ import annotations

import os
import zlib

import numpy as np
import torch

from ..data import Data
from ..utils.creation import CreateImageFromMask
from . import npy as _npy

MET_TYPES = {"MET_CHAR": np.int8, "MET_UCHAR": np.uint8, "MET_SHORT": np.int16, "MET_USHORT": np.uint16,
             "MET_INT": np.int32, "MET_UINT": np.uint32, "MET_LONG": np.int32, "MET_ULONG": np.uint32,
             "MET_LONG_LONG": np.int64, "MET_ULONG_LONG": np.uint64, "MET_FLOAT": np.float32, "MET_DOUBLE": np.float64}
_MET_NAMES = {np.dtype(np.int16): "MET_SHORT", np.dtype(np.float32): "MET_FLOAT", np.dtype(np.uint8): "MET_UCHAR",
              np.dtype(np.int32): "MET_INT", np.dtype(np.float64): "MET_DOUBLE", np.dtype(np.int8): "MET_CHAR"}


def read_mhd_header(path):
    """Parse the ``key = value`` text header.  Returns a dict with ``dims`` (x, y, z), ``spacing``, ``origin``,
    ``direction`` (9 numbers, row-major as ``sitk.Image.GetDirection`` reports them), ``dtype``, ``channels``,
    ``msb``, ``compressed``, ``data_file`` (absolute path) and ``data_offset``."""
    fields = {}
    header_end = 0
    with open(path, "rb") as f:
        while True:
            line = f.readline()
            if not line:
                break
            header_end = f.tell()
            text = line.decode("latin-1").strip()
            if not text or "=" not in text:
                continue
            key, value = (s.strip() for s in text.split("=", 1))
            fields[key] = value
            if key == "ElementDataFile":
                break
    ndims = int(fields.get("NDims", 3))
    if ndims != 3:
        raise ValueError("%s: only 3-D MetaImages are on this path (NDims = %d)" % (path, ndims))
    nums = lambda key, default: [float(v) for v in fields[key].split()] if key in fields else default
    first = lambda keys, default: next((nums(k, None) for k in keys if k in fields), default)
    et = fields.get("ElementType", "MET_SHORT")
    if et not in MET_TYPES:
        raise ValueError("%s: unsupported ElementType %s" % (path, et))
    data_file = fields.get("ElementDataFile", "LOCAL")
    local = data_file == "LOCAL"
    hs = int(fields.get("HeaderSize", 0))
    truthy = lambda key: fields.get(key, "False").lower() in ("true", "1", "yes")
    return {
        "dims": tuple(int(v) for v in fields["DimSize"].split()),
        "spacing": tuple(first(["ElementSpacing", "ElementSize"], [1.0, 1.0, 1.0])),
        "origin": tuple(first(["Offset", "Position", "Origin"], [0.0, 0.0, 0.0])),
        # MetaIO stores the matrix with the axis directions as ROWS; ITK's direction has them as COLUMNS and
        # GetDirection() flattens that row-major, i.e. the transpose of the file's numbers
        "direction": tuple(np.asarray(first(["TransformMatrix", "Orientation", "Rotation"],
                                            [1, 0, 0, 0, 1, 0, 0, 0, 1]), dtype=np.float64).reshape(3, 3).T.ravel()),
        "dtype": np.dtype(MET_TYPES[et]),
        "channels": int(fields.get("ElementNumberOfChannels", 1)),
        "msb": truthy("ElementByteOrderMSB") or truthy("BinaryDataByteOrderMSB"),
        "compressed": truthy("CompressedData"),
        "data_file": path if local else os.path.join(os.path.dirname(os.path.abspath(path)), data_file),
        "data_offset": header_end if local else max(hs, 0),
        "fields": fields,
    }


def read_mhd(path, device="cuda"):
    """(device tensor ``[z, y, x]`` or ``[z, y, x, channels]``, header dict).  Uncompressed little-endian payloads go
    file -> pinned staging -> HBM; compressed or big-endian ones are inflated / byte-swapped on the host first."""
    h = read_mhd_header(path)
    nx, ny, nz = h["dims"]
    shape = (nz, ny, nx) if h["channels"] == 1 else (nz, ny, nx, h["channels"])
    dt = h["dtype"]
    direct = not h["compressed"] and not (h["msb"] and dt.itemsize > 1) and dt in _npy._TORCH
    if direct:
        return _npy.read_payload_to_device(h["data_file"], h["data_offset"], shape, dt, device), h
    with open(h["data_file"], "rb") as f:
        f.seek(h["data_offset"])
        raw = f.read()
    if h["compressed"]:
        raw = zlib.decompress(raw)
    a = np.frombuffer(raw, dtype=dt.newbyteorder(">" if h["msb"] else "<"), count=int(np.prod(shape))).reshape(shape)
    a = a.astype(dt.newbyteorder("="))
    if a.dtype not in _npy._TORCH:                                   # uint16 / uint32 / uint64: widen losslessly
        a = a.astype({2: np.int32, 4: np.int64, 8: np.float64}[a.dtype.itemsize])
    return torch.from_numpy(np.ascontiguousarray(a)).to(device), h


def write_mhd(path, array, spacing=(1.0, 1.0, 1.0), origin=(0.0, 0.0, 0.0), direction=None):
    """Write ``path`` (.mhd) and its sibling .raw, uncompressed little-endian.  ``array``: a CUDA tensor (payload
    written from the device) or a NumPy array, ``[z, y, x]`` or ``[z, y, x, channels]``."""
    is_t = isinstance(array, torch.Tensor)
    np_dtype = np.dtype(str(array.dtype).replace("torch.", "")) if is_t else np.asarray(array).dtype
    if np_dtype not in _MET_NAMES:
        raise TypeError("cannot write dtype %s as a MetaImage" % np_dtype)
    shape = tuple(int(s) for s in array.shape)
    channels = shape[3] if len(shape) == 4 else 1
    raw_name = os.path.splitext(os.path.basename(path))[0] + ".raw"
    d = np.identity(3) if direction is None else np.asarray(direction, dtype=np.float64).reshape(3, 3)
    lines = ["ObjectType = Image", "NDims = 3", "BinaryData = True", "BinaryDataByteOrderMSB = False",
             "CompressedData = False",
             "TransformMatrix = " + " ".join(repr(float(v)) for v in d.T.ravel()),
             "Offset = " + " ".join(repr(float(v)) for v in origin),
             "CenterOfRotation = 0 0 0",
             "ElementSpacing = " + " ".join(repr(float(v)) for v in spacing),
             "DimSize = %d %d %d" % (shape[2], shape[1], shape[0])]
    if channels != 1:
        lines.append("ElementNumberOfChannels = %d" % channels)
    lines += ["ElementType = " + _MET_NAMES[np_dtype], "ElementDataFile = " + raw_name]
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    raw_path = os.path.join(os.path.dirname(os.path.abspath(path)), raw_name)
    if is_t and array.is_cuda:
        _npy.write_payload_from_device(raw_path, 0, array, truncate=True)
    else:
        a = array.cpu().numpy() if is_t else np.asarray(array)
        np.ascontiguousarray(a).astype(np_dtype.newbyteorder("<")).tofile(raw_path)


class DvfRecord(object):
    """What ``MhdReader.create_dvf`` hands to ``Deformable(...)`` in the reference (``read/mhd.py:236-252``): the
    field ``[z, y, x, 3]`` on the device plus its lattice.  (The ``Deformable`` registry object is out of scope.)"""

    def __init__(self, dvf, origin, spacing, dimensions, dvf_matrix, registration_name, reference_name, moving_name):
        self.dvf, self.origin, self.spacing, self.dimensions = dvf, origin, spacing, dimensions
        self.dvf_matrix, self.registration_name = dvf_matrix, registration_name
        self.reference_name, self.moving_name = reference_name, moving_name


def narrow_for_image(array):
    """``Image.array`` on this path is int16 or float32.  Other floating types become float32; other integer types
    become int16 only when every value fits (MET_UCHAR, small MET_USHORT / MET_INT data) and float32 otherwise -- the
    reference keeps the source dtype (read/mhd.py:185), so a silent wrap of values above 32767 would change the image."""
    if array.dtype in (torch.int16, torch.float32):
        return array
    if array.dtype.is_floating_point:
        return array.to(torch.float32)
    lo, hi = int(array.min()), int(array.max())
    return array.to(torch.int16) if (-32768 <= lo and hi <= 32767) else array.to(torch.float32)


class MhdReader(object):
    """``read/mhd.py:60-196`` with the same arguments; ``load()`` registers the image in ``Data`` (or returns a
    ``DvfRecord`` when ``reference_name``, ``moving_name`` and ``dvf`` are given)."""

    def __init__(self, file, modality=None, reference_name=None, moving_name=None, roi_name=None, dose=None, dvf=None,
                 device="cuda"):
        self.file = file
        self.modality = modality
        self.reference_name = reference_name
        self.moving_name = moving_name
        self.roi_name = roi_name
        self.dose = dose
        self.dvf = dvf
        self.device = device
        self.mhd = None
        self.array = None

    def load(self):
        self.array, self.mhd = read_mhd(self.file, device=self.device)
        if self.reference_name is not None:
            if self.dvf is not None and self.moving_name is not None:
                return self.create_dvf()
            return None                                                     # dose / ROI: stubs in the reference too
        return self.create_image()

    def create_image(self):
        """``read/mhd.py:157-196``."""
        if self.modality is None:
            image_name = os.path.splitext(os.path.basename(self.file))[0]
            self.modality = 'CT'
        else:
            idx = len(Data.image_list)
            image_name = "%s %02d" % (self.modality, idx + 1) if idx < 9 else "%s %d" % (self.modality, idx + 1)
        dimensions = np.flip(np.asarray(self.mhd["dims"]))
        array = narrow_for_image(self.array)
        creator = CreateImageFromMask(array, np.asarray(self.mhd["origin"]), np.asarray(self.mhd["spacing"]), image_name,
                                      dimensions=dimensions, orientation=np.asarray(self.mhd["direction"]), plane='Axial',
                                      description='Mhd to Image', modality=self.modality, device=self.device)
        return creator.add_image()

    def create_dvf(self):
        """``read/mhd.py:198-252`` (the field and its lattice; no ``Deformable`` registry here)."""
        if self.array.dim() != 4 or self.array.shape[3] != 3:
            raise ValueError("%s is not a 3-channel displacement field" % self.file)
        # the reference's name, "DVF_<reference>_<moving>", with the first free "_<n>" suffix on a collision
        # (read/mhd.py:222-233); the record is entered in Data.deformable_list so that the next one sees it
        base = "DVF_%s_%s" % (self.reference_name, self.moving_name)
        name, n = base, 0
        while name in Data.deformable_list:
            n += 1
            name = "%s_%d" % (base, n)
        Data.deformable_list += [name]
        return DvfRecord(self.array.to(torch.float32), np.asarray(self.mhd["origin"]), np.asarray(self.mhd["spacing"]),
                         np.flip(np.asarray(self.mhd["dims"])), np.asarray(self.mhd["direction"]).reshape(3, 3), name,
                         self.reference_name, self.moving_name)
