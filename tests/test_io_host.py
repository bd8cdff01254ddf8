"""CPU tests of the volume I/O host logic: MetaImage header parsing / writing and .npy framing.
The device payload path is covered by tests/test_gpu_io.py."""
import os
import zlib

import numpy as np
import pytest

from medicalimageanalysis_b200 import Data, _lib
from medicalimageanalysis_b200.read import mhd as M
from medicalimageanalysis_b200.utils import rigid as R


def _write(tmp_path, name, header, payload):
    p = tmp_path / name
    p.write_text(header)
    return str(p)


def test_header_fields_and_direction_transpose(tmp_path):
    (tmp_path / "v.raw").write_bytes(np.arange(24, dtype="<i2").tobytes())
    p = _write(tmp_path, "v.mhd", "ObjectType = Image\nNDims = 3\nDimSize = 4 3 2\nElementSpacing = 0.5 0.75 2.5\n"
               "Offset = -10 -20 30.5\nTransformMatrix = 0 1 0 -1 0 0 0 0 1\nElementType = MET_SHORT\n"
               "BinaryDataByteOrderMSB = False\nElementDataFile = v.raw\n", None)
    h = M.read_mhd_header(p)
    assert h["dims"] == (4, 3, 2) and h["spacing"] == (0.5, 0.75, 2.5) and h["origin"] == (-10.0, -20.0, 30.5)
    assert h["dtype"] == np.int16 and not h["msb"] and not h["compressed"] and h["channels"] == 1
    assert h["data_file"] == str(tmp_path / "v.raw") and h["data_offset"] == 0
    # file rows are axis directions; ITK's GetDirection() is the transpose, flattened row-major
    assert h["direction"] == (0.0, -1.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0)


def test_local_payload_offset_and_rejects(tmp_path):
    head = "NDims = 3\nDimSize = 2 2 2\nElementType = MET_FLOAT\nElementDataFile = LOCAL\n"
    p = tmp_path / "l.mhd"
    p.write_bytes(head.encode() + np.arange(8, dtype="<f4").tobytes())
    h = M.read_mhd_header(str(p))
    assert h["data_file"] == str(p) and h["data_offset"] == len(head)
    with pytest.raises(ValueError):
        M.read_mhd_header(_write(tmp_path, "b.mhd", "NDims = 2\nDimSize = 2 2\nElementDataFile = x.raw\n", None))
    with pytest.raises(ValueError):
        M.read_mhd_header(_write(tmp_path, "c.mhd", "NDims = 3\nDimSize = 2 2 2\nElementType = MET_WAT\nElementDataFile = x.raw\n", None))


def test_write_mhd_numpy_roundtrip_through_the_parser(tmp_path):
    a = np.random.default_rng(0).integers(-1000, 1000, size=(3, 4, 5)).astype(np.int16)
    p = str(tmp_path / "out.mhd")
    d = np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1.0]])
    M.write_mhd(p, a, spacing=(0.9, 0.9, 2.5), origin=(1.5, -2.5, 3.0), direction=d)
    h = M.read_mhd_header(p)
    assert h["dims"] == (5, 4, 3) and h["spacing"] == (0.9, 0.9, 2.5) and h["origin"] == (1.5, -2.5, 3.0)
    assert np.array_equal(np.asarray(h["direction"]).reshape(3, 3), d)
    raw = np.fromfile(h["data_file"], dtype="<i2").reshape(3, 4, 5)
    assert np.array_equal(raw, a)


def test_reader_on_cpu_handles_compressed_and_big_endian(tmp_path):
    """The host-side branch of read_mhd (inflate / byte swap) needs no GPU."""
    Data.clear()
    a = np.random.default_rng(1).integers(0, 60000, size=(2, 3, 4)).astype(np.uint16)
    (tmp_path / "z.zraw").write_bytes(zlib.compress(a.astype(">u2").tobytes()))
    p = _write(tmp_path, "z.mhd", "NDims = 3\nDimSize = 4 3 2\nElementType = MET_USHORT\nCompressedData = True\n"
               "ElementByteOrderMSB = True\nElementDataFile = z.zraw\n", None)
    t, h = M.read_mhd(p, device="cpu")
    assert h["compressed"] and h["msb"]
    assert np.array_equal(t.numpy(), a.astype(np.int32))
    Data.clear()


def test_dicom_reg_matrix_convention():
    m = np.identity(4)
    m[:3, :3] = [[0, -1, 0], [1, 0, 0], [0, 0, 1]]
    m[:3, 3] = [5, -3, 7]
    got = R.matrix_from_dicom_reg(m.ravel().tolist())
    assert np.allclose(got @ m, np.identity(4))


def test_io_abi_argument_errors_without_gpu(tmp_path):
    lib = _lib.load()
    assert lib.mia_read_file_to_device(None, 0, 0, None, 0) == -1
    assert lib.mia_read_file_to_device(str(tmp_path / "missing.raw").encode(), 0, 16, 1, 0) == -1
    assert b"cannot open" in lib.mia_last_error()
