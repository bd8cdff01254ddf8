"""SURVEY N4 on the GPU: file payloads land in / leave HBM bit for bit (MetaImage and .npy), through the reference's
entry points (MhdReader.load, Image.save_image / load_image, Rigid.export_image)."""
import os

import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import Data, Rigid, geometry as G
from medicalimageanalysis_b200.phantom import make_phantom
from medicalimageanalysis_b200.read import MhdReader, load_npy_to_device, read_mhd, save_npy_from_device, write_mhd
from medicalimageanalysis_b200.utils import CreateImageFromMask

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _clean():
    Data.clear()
    yield
    Data.clear()


@pytest.mark.parametrize("dtype", [np.int16, np.float32, np.uint8])
@pytest.mark.parametrize("shape", [(1, 1, 1), (5, 7, 9), (70, 300, 412)])      # the last one spans several 16 MiB chunks for f32
def test_mhd_roundtrip_device_file_device(cuda_device, tmp_path, dtype, shape):
    rng = np.random.default_rng(sum(shape))
    a = (rng.random(shape) * 200 - 50).astype(dtype)
    t = torch.from_numpy(a).to(cuda_device)
    p = str(tmp_path / "vol.mhd")
    write_mhd(p, t, spacing=(0.9, 0.8, 2.5), origin=(-3.0, 4.0, 5.5))
    assert np.array_equal(np.fromfile(str(tmp_path / "vol.raw"), dtype=dtype).reshape(shape), a)
    back, h = read_mhd(p, device=cuda_device)
    assert back.is_cuda and tuple(back.shape) == shape
    assert np.array_equal(back.cpu().numpy(), a)
    assert h["spacing"] == (0.9, 0.8, 2.5) and h["origin"] == (-3.0, 4.0, 5.5)


def test_mhd_reader_registers_an_image_like_the_reference(cuda_device, tmp_path):
    a = make_phantom((20, 32, 40), "CT", seed=1)
    write_mhd(str(tmp_path / "scan.mhd"), a, spacing=(1.0, 1.0, 3.0), origin=(-20.0, -16.0, -30.0))
    img = MhdReader(str(tmp_path / "scan.mhd")).load()
    assert Data.image_list == ["scan"] and Data.image["scan"] is img         # name = file stem, modality CT
    assert img.array.is_cuda and np.array_equal(img.array.cpu().numpy(), a)
    assert np.allclose(img.spacing, (1.0, 1.0, 3.0)) and np.allclose(img.origin, (-20.0, -16.0, -30.0))
    assert list(img.dimensions) == [20, 32, 40]
    img2 = MhdReader(str(tmp_path / "scan.mhd"), modality="MR").load()
    assert img2.image_name == "MR 02"


def test_dvf_mhd_four_dimensional_payload(cuda_device, tmp_path):
    f = np.random.default_rng(2).standard_normal((6, 8, 10, 3)).astype(np.float32)
    write_mhd(str(tmp_path / "dvf.mhd"), torch.from_numpy(f).to(cuda_device), spacing=(2.0, 2.0, 2.0))
    rec = MhdReader(str(tmp_path / "dvf.mhd"), reference_name="CT 01", moving_name="CT 02", dvf=True).load()
    assert rec.registration_name == "DVF_CT 01_CT 02"                        # the reference's name, read/mhd.py:222-224
    assert np.array_equal(rec.dvf.cpu().numpy(), f) and list(rec.dimensions) == [6, 8, 10]
    again = MhdReader(str(tmp_path / "dvf.mhd"), reference_name="CT 01", moving_name="CT 02", dvf=True).load()
    assert again.registration_name == "DVF_CT 01_CT 02_1"                    # collision suffix, read/mhd.py:226-233


def test_unsigned_short_volume_is_not_wrapped(cuda_device, tmp_path):
    """MET_USHORT data above 32767 must keep its values (the reference keeps the source dtype): it becomes float32, and
    data that fits int16 stays int16."""
    from medicalimageanalysis_b200.structure.image import Image
    from types import SimpleNamespace
    big = (np.arange(4 * 5 * 8, dtype=np.uint16).reshape(4, 5, 8) * 400)          # up to 63600
    small = (np.arange(4 * 5 * 8, dtype=np.uint16).reshape(4, 5, 8) * 3)
    for arr, want in ((big, torch.float32), (small, torch.int16)):
        src = SimpleNamespace(array=arr, image_name="u16", spacing=(1, 1, 1), origin=(0, 0, 0), image_matrix=np.identity(3))
        img = Image(src, device=cuda_device)
        assert img.array.dtype == want and np.array_equal(img.array.cpu().numpy().astype(np.int64), arr.astype(np.int64))


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
def test_npy_is_byte_identical_to_numpy(cuda_device, tmp_path, dtype):
    a = (np.random.default_rng(3).random((33, 65, 129)) * 3000 - 1000).astype(dtype)
    p1, p2 = str(tmp_path / "ours.npy"), str(tmp_path / "theirs.npy")
    save_npy_from_device(p1, torch.from_numpy(a).to(cuda_device))
    np.save(p2, a)
    assert open(p1, "rb").read() == open(p2, "rb").read()
    assert np.array_equal(load_npy_to_device(p2, cuda_device).cpu().numpy(), a)


def test_image_save_load_and_rigid_export(cuda_device, tmp_path):
    sp = (0.9765625, 0.9765625, 2.5)
    a = make_phantom((24, 40, 48), "CT", seed=5, spacing=sp)
    img = CreateImageFromMask(a, (-23.0, -19.0, -30.0), sp, "CT 01").add_image()
    os.makedirs(tmp_path / "CT 01")
    img.save_image(str(tmp_path / "CT 01"))
    assert np.array_equal(np.load(str(tmp_path / "CT 01" / "array.npy")), a)
    other = CreateImageFromMask(np.zeros((2, 2, 2), np.int16), (0, 0, 0), (1, 1, 1), "blank").add_image()
    other.load_image(str(tmp_path / "CT 01"))
    assert other.array.is_cuda and np.array_equal(other.array.cpu().numpy(), a)
    assert other.image_name == "CT 01" and np.allclose(other.spacing, sp) and np.allclose(other.origin, img.origin)
    # export_image: the aligned volume as MetaImage, equal to create_image()
    CreateImageFromMask(a, (-23.0, -19.0, -30.0), sp, "CT 02").add_image()
    r = Rigid("CT 01", "CT 02", matrix=G.params_to_matrix([2, -1, 3, 4, -2, 1], img.compute_center()))
    r.export_image(str(tmp_path / "aligned.mhd"))
    back, h = read_mhd(str(tmp_path / "aligned.mhd"), device=cuda_device)
    want = r.create_image()
    assert torch.equal(back, want.array)
    assert np.allclose(h["origin"], want.origin) and np.allclose(h["spacing"], want.spacing)
