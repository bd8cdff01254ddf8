"""``array.npy`` <-> device memory: the volume payload of ``Image.save_image`` / ``load_image``
(``structure/image.py:709-735, 804-836``: ``np.save(.../array.npy, self.array)`` / ``np.load``) without a pageable
host copy.  The .npy header is parsed / written on the host with NumPy's own ``numpy.lib.format`` routines (so the
files are interchangeable with the reference's), the data block goes through ``mia_read_file_to_device`` /
``mia_write_file_from_device``."""
from __future__ import annotations

import numpy as np
import torch
from numpy.lib import format as npformat

from .. import _lib

_TORCH = {np.dtype(np.int16): torch.int16, np.dtype(np.float32): torch.float32, np.dtype(np.uint8): torch.uint8,
          np.dtype(np.int32): torch.int32, np.dtype(np.float64): torch.float64, np.dtype(np.int8): torch.int8,
          np.dtype(np.int64): torch.int64, np.dtype(np.bool_): torch.bool}


def _device_index(device):
    device = torch.device(device)
    if device.type != "cuda":
        raise TypeError("device I/O needs a CUDA device (there is no CPU path)")
    return device, (device.index if device.index is not None else torch.cuda.current_device())


def read_payload_to_device(path, offset, shape, dtype, device):
    """Raw little-endian ``dtype`` payload at ``offset`` of ``path`` -> new device tensor of ``shape``."""
    dtype = np.dtype(dtype)
    device, idx = _device_index(device)
    out = torch.empty(tuple(int(s) for s in shape), dtype=_TORCH[dtype.newbyteorder("=")], device=device)
    nbytes = out.numel() * out.element_size()
    _lib.check(_lib.load().mia_read_file_to_device(str(path).encode(), int(offset), int(nbytes), out.data_ptr(), idx))
    return out


def write_payload_from_device(path, offset, tensor, truncate=True):
    if not (isinstance(tensor, torch.Tensor) and tensor.is_cuda):
        raise TypeError("tensor must be a CUDA torch.Tensor")
    tensor = tensor.contiguous()
    torch.cuda.current_stream(tensor.device).synchronize()          # the I/O stream does not see torch's stream
    idx = tensor.device.index if tensor.device.index is not None else torch.cuda.current_device()
    _lib.check(_lib.load().mia_write_file_from_device(str(path).encode(), int(offset), tensor.data_ptr(),
                                                      tensor.numel() * tensor.element_size(), 1 if truncate else 0, idx))


def load_npy_to_device(path, device="cuda"):
    """``np.load(path)`` for a C-ordered, native-endian array of a device dtype, landing in HBM."""
    with open(path, "rb") as f:
        version = npformat.read_magic(f)
        if version == (1, 0):
            shape, fortran, dtype = npformat.read_array_header_1_0(f)
        else:
            shape, fortran, dtype = npformat.read_array_header_2_0(f)
        offset = f.tell()
    if fortran or dtype.hasobject or dtype.newbyteorder("=") not in _TORCH or not dtype.isnative:
        raise ValueError("%s: only C-ordered native-endian numeric arrays load straight to the device" % path)
    return read_payload_to_device(path, offset, shape, dtype, device)


def save_npy_from_device(path, tensor):
    """``np.save(path, tensor.cpu().numpy())`` byte for byte, with the data block written from the device."""
    np_dtype = np.dtype(str(tensor.dtype).replace("torch.", ""))
    with open(path, "wb") as f:
        npformat.write_array_header_1_0(f, {"descr": npformat.dtype_to_descr(np_dtype), "fortran_order": False,
                                            "shape": tuple(int(s) for s in tensor.shape)})
        offset = f.tell()
    write_payload_from_device(path, offset, tensor, truncate=False)
