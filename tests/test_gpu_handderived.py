"""The hand-derived known answers of tests/test_handderived.py on the GPU, through the C ABI (per-voxel kernel on the tiny
cases, the TMA tile kernels on volumes whose row pitch allows them) and through ``Rigid.create_image``."""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import Data, Rigid, geometry as G, kernels
from medicalimageanalysis_b200.utils import CreateImageFromMask
from tests.test_handderived import half_voxel_case, rz

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dims", [(12, 9, 5), (136, 24, 16)])
@pytest.mark.parametrize("deg", [0, 90, 180, 270])
def test_rigid_create_image_right_angles_and_shift(cuda_device, dims, deg):
    """Rigid.create_image (auto-cropped VTK lattice) for Rz(deg) composed with a translation: output origin, dims and
    every voxel are known by hand (a rot90 of each plane; AutoCrop follows the translation)."""
    S, O = (0.9, 0.9, 2.5), (-17.1, 4.5, -30.0)
    t = np.array([1.3, -0.7, 2.0])
    mov = np.random.default_rng(deg).integers(-1000, 3000, size=(dims[2], dims[1], dims[0])).astype(np.int16)
    Data.clear()
    CreateImageFromMask(mov, O, S, "moving", device=cuda_device).add_image()
    CreateImageFromMask(mov, O, S, "fixed", device=cuda_device).add_image()
    M = rz(deg)
    M[:3, 3] = t                                        # q = R p + t  (output physical -> moving physical)
    r = Rigid("fixed", "moving", matrix=M)
    img = r.create_image()
    ex, ey = (dims[0] - 1) * S[0], (dims[1] - 1) * S[1]
    # moving box [O, O + e] pulled back through p = R^-1 (q - t)
    lo = {0: (O[0] - t[0], O[1] - t[1]), 90: (O[1] - t[1], -(O[0] - t[0]) - ex),
          180: (-(O[0] - t[0]) - ex, -(O[1] - t[1]) - ey), 270: (-(O[1] - t[1]) - ey, O[0] - t[0])}[deg]
    assert np.allclose(img.GetOrigin(), (lo[0], lo[1], O[2] - t[2]), atol=1e-9)
    assert img.GetDimensions() == ((dims[1], dims[0], dims[2]) if deg in (90, 270) else dims)
    assert np.array_equal(img.array.cpu().numpy(), np.rot90(mov, deg // 90, axes=(1, 2)))
    Data.clear()


@pytest.mark.parametrize("nx", [8, 136])
def test_half_voxel_shift_rounding_and_border_tie_on_the_device(cuda_device, nx):
    """Weights exactly 1/2 and samples exactly on the +-0.5 border, both rules (see test_handderived): per-voxel kernel
    (nx = 8 rows are too short for a box) and tile kernels (nx = 136, the same 8 values repeated)."""
    mov, A, mids = half_voxel_case()
    reps = nx // 8
    row = np.tile(mov.ravel(), reps)
    vol = np.ascontiguousarray(np.broadcast_to(row, (16, 24, nx))).astype(np.int16)
    t = torch.from_numpy(vol).to(cuda_device)
    seam = (int(row[7]) + int(row[0])) / 2                                  # 100 | 3 where the pattern repeats
    want_mid = (mids + [seam]) * reps
    vtk = kernels.resample(t, A, vol.shape, background=-3001.0, mode="vtk").cpu().numpy()
    itk = kernels.resample(t, A, vol.shape, background=0.0, mode="itk").cpu().numpy()
    exp_vtk = [int(np.floor(v + 0.5)) for v in want_mid[:-1]] + [100]       # c = N - 0.5: inside, clamped (closed)
    exp_itk = [int(np.trunc(v)) for v in want_mid[:-1]] + [0]               # c = N - 0.5: outside (half-open)
    assert np.array_equal(vtk, np.broadcast_to(np.array(exp_vtk, dtype=np.int16), vol.shape))
    assert np.array_equal(itk, np.broadcast_to(np.array(exp_itk, dtype=np.int16), vol.shape))
    A2 = np.hstack([np.identity(3), np.array([[-0.5], [0.0], [0.0]])])      # c = -0.5: inside for both, clamps to voxel 0
    for mode in ("vtk", "itk"):
        got = kernels.resample(t, A2, vol.shape, background=-3001.0, mode=mode).cpu().numpy()
        assert (got[:, :, 0] == 3).all()
