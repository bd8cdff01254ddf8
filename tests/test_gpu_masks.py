"""fixed-image masks in the fused metrics (SetMetricFixedMask) on every kernel path -- the one-matrix tile
kernel, the multi-matrix tile kernel, the per-voxel fall-backs (pitch the TMA cannot take; mixed dtypes) -- against
the oracle, and a masked registration."""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import Data, RegistrationConfig, _lib, geometry as G, kernels
from medicalimageanalysis_b200.phantom import make_phantom
from medicalimageanalysis_b200.utils import CreateImageFromMask, rigid
from oracle import metrics
from tests.test_gpu_metrics import mats, pair

pytestmark = pytest.mark.gpu


def _mask(shape, kind, seed=0):
    z, y, x = np.meshgrid(*(np.arange(n) for n in shape), indexing="ij")
    if kind == "ellipsoid":
        c = [s / 2 for s in shape]
        return ((((x - c[2]) / (0.3 * shape[2])) ** 2 + ((y - c[1]) / (0.35 * shape[1])) ** 2 +
                 ((z - c[0]) / (0.4 * shape[0])) ** 2) < 1).astype(np.uint8)
    if kind == "random":
        return (np.random.default_rng(seed).random(shape) > 0.5).astype(np.uint8) * 3          # any non-zero value counts
    if kind == "empty":
        return np.zeros(shape, np.uint8)
    return np.ones(shape, np.uint8)


@pytest.mark.parametrize("kind", ["ellipsoid", "random", "empty", "full"])
@pytest.mark.parametrize("modality", ["CT", "MR"])
def test_masked_ssd_and_hist_one_matrix(cuda_device, kind, modality):
    g, fixed, moving, _ = pair(fdtype=modality, mdtype=modality)
    A = mats(g, 1, seed=4)[0]
    m = _mask(fixed.shape, kind)
    f, mv = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    ssd, cnt = kernels.resample_ssd(f, mv, A=A, fixed_mask=m)
    assert _lib.load().mia_debug_last_path() == 1                                   # the TMA tile kernel took it
    want, n = metrics.oracle_ssd(fixed, moving, A, "itk", fixed_mask=m)
    assert abs(int(cnt[0]) - n) <= 2
    assert abs(float(ssd[0]) - want) <= 1e-5 * max(want, 1.0)
    fr, mr = (float(fixed.min()), float(fixed.max())), (float(moving.min()), float(moving.max()))
    H, hc = kernels.resample_jointhist(f, mv, fr, mr, bins=32, A=A, fixed_mask=m)
    Hw, nw = metrics.oracle_joint_hist(fixed, moving, A, 32, fr, mr, fixed_mask=m)
    assert abs(int(hc[0]) - nw) <= 2
    assert np.abs(H[0].cpu().numpy().astype(np.int64) - Hw.astype(np.int64)).sum() <= 8    # border ties / bin-edge ties
    if kind == "full":
        s0, c0 = kernels.resample_ssd(f, mv, A=A)
        assert float(s0[0]) == float(ssd[0]) and int(c0[0]) == int(cnt[0])          # all-ones mask == no mask, bit for bit


def test_masked_ssd_thirteen_probes_multi_kernel(cuda_device):
    g, fixed, moving, Mgt = pair(shape=(32, 64, 96), fdtype="MR", mdtype="MR")
    base = np.array([2.5, -1.5, 3.5, 2.0, -1.0, 0.5])
    A = np.stack([G.compose_index_matrix(g, g, G.params_to_matrix(base + 0.2 * d, (0, 0, 0)))
                  for d in np.vstack([np.zeros(6), np.identity(6), -np.identity(6)])])
    m = _mask(fixed.shape, "ellipsoid")
    f, mv = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    s, c = kernels.resample_ssd(f, mv, A=A, fixed_mask=m)
    assert _lib.load().mia_debug_last_path() == 2                                   # multi-matrix tile kernel
    for i in range(13):
        want, n = metrics.oracle_ssd(fixed, moving, A[i], "itk", fixed_mask=m)
        assert abs(int(c[i]) - n) <= 2 and abs(float(s[i]) - want) <= 1e-5 * want
        s1, c1 = kernels.resample_ssd(f, mv, A=A[i], fixed_mask=m)                  # one-matrix masked launch
        assert int(c1[0]) == int(c[i]) and abs(float(s1[0]) - float(s[i])) <= 1e-9 * float(s[i])


def test_masked_metrics_on_the_direct_kernel(cuda_device):
    """Odd row pitch (no TMA descriptor) and mixed dtypes run the per-voxel kernel; the mask is honoured there too."""
    shape = (20, 30, 45)
    sp = (1.0, 1.0, 2.0)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    fixed = make_phantom(shape, "CT", seed=5, spacing=sp)
    moving = make_phantom(shape, "MR", seed=5, spacing=sp)
    A = mats(g, 1, seed=6)[0]
    m = _mask(shape, "random", seed=2)
    s, c = kernels.resample_ssd(torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device), A=A,
                                fixed_mask=m)
    assert _lib.load().mia_debug_last_path() == 0
    want, n = metrics.oracle_ssd(fixed, moving, A, "itk", fixed_mask=m)
    assert abs(int(c[0]) - n) <= 2 and abs(float(s[0]) - want) <= 1e-5 * want
    with pytest.raises(_lib.MiaError):                                               # a masked metric does not store
        kernels.resample_ssd(torch.from_numpy(moving).to(cuda_device), torch.from_numpy(moving).to(cuda_device), A=A,
                             fixed_mask=m, out=torch.empty(shape, dtype=torch.float32, device=cuda_device))


def test_masked_registration_ignores_what_the_mask_hides(cuda_device):
    """Half of the fixed image is replaced by garbage; with a mask over the intact half the loop converges as if the
    garbage were not there, without the mask it does not."""
    Data.clear()
    shape, sp = (48, 80, 96), (2.0, 2.0, 2.0)
    gt = [2.0, -1.5, 3.0, 3.0, -2.0, 1.5]
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    M = G.params_to_matrix(gt, (0, 0, 0))
    fixed = make_phantom(shape, "CT", seed=21, spacing=sp, noise=False).astype(np.float32)
    moving = make_phantom(shape, "CT", seed=21, spacing=sp, noise=False, world_matrix=np.linalg.inv(M)).astype(np.float32)
    clean = fixed.copy()
    fixed[:, :, 48:] = np.random.default_rng(0).random((48, 80, 48)).astype(np.float32) * 3000 - 1000
    mask = np.zeros(shape, np.uint8)
    mask[:, :, :48] = 1
    fi = CreateImageFromMask(fixed, g.origin, sp, "F").add_image()
    mi = CreateImageFromMask(moving, g.origin, sp, "M").add_image()
    ci = CreateImageFromMask(clean, g.origin, sp, "C").add_image()
    cfg = RegistrationConfig(metric="ssd", iterations=120, fixed_mask=torch.from_numpy(mask).to(cuda_device))
    center = np.zeros(3)
    from medicalimageanalysis_b200.registration import register_rigid
    got = register_rigid(fi, mi, cfg=cfg, center=center)
    ref = register_rigid(ci, mi, cfg=RegistrationConfig(metric="ssd", iterations=120, fixed_mask=cfg.fixed_mask), center=center)
    assert np.allclose(got.params, ref.params, atol=1e-9)                            # garbage outside the mask: no effect at all
    assert np.abs(got.params[:3] - gt[:3]).max() < 0.3 and np.abs(got.params[3:] - gt[3:]).max() < 0.6
    # multi-resolution with the mask carried down the pyramid
    mr = rigid.register_rigid_multires(fi, mi, cfg=cfg, shrink_factors=(2, 1), iterations=(60, 60))
    assert np.abs(mr.params[:3] - gt[:3]).max() < 0.3 and np.abs(mr.params[3:] - gt[3:]).max() < 0.6
    Data.clear()


@pytest.mark.parametrize("modality", ["CT", "MR"])
def test_moving_mask_nearest_voxel_semantics(cuda_device, modality):
    """SetMetricMovingMask: a sample counts when the moving-mask voxel nearest to its mapped point is set."""
    g, fixed, moving, _ = pair(fdtype=modality, mdtype=modality)
    A = mats(g, 1, seed=11)[0]
    mm = _mask(moving.shape, "ellipsoid")
    fm = _mask(fixed.shape, "random", seed=3)
    f, mv = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    for fmask in (None, fm):
        ssd, cnt = kernels.resample_ssd(f, mv, A=A, fixed_mask=fmask, moving_mask=mm)
        assert _lib.load().mia_debug_last_path() == 0                                   # per-voxel kernel
        want, n = metrics.oracle_ssd(fixed, moving, A, "itk", fixed_mask=fmask, moving_mask=mm)
        # float32 local coordinates against float64: a mapped point within 1e-6 of a half-voxel plane may pick the
        # neighbouring mask voxel
        assert abs(int(cnt[0]) - n) <= max(4, 1e-5 * n)
        assert abs(float(ssd[0]) - want) <= 1e-4 * want
    fr, mr = (float(fixed.min()), float(fixed.max())), (float(moving.min()), float(moving.max()))
    H, hc = kernels.resample_jointhist(f, mv, fr, mr, bins=32, A=A, moving_mask=mm)
    Hw, nw = metrics.oracle_joint_hist(fixed, moving, A, 32, fr, mr, moving_mask=mm)
    assert abs(int(hc[0]) - nw) <= 4 and np.abs(H[0].cpu().numpy().astype(np.int64) - Hw.astype(np.int64)).sum() <= 16
    Hm, mc = kernels.resample_mattes(f, mv, fr, mr, bins=32, A=A, moving_mask=mm)
    Hmw, nmw = metrics.oracle_mattes_hist(fixed, moving, A, 32, fr, mr, moving_mask=mm)
    assert abs(int(mc[0]) - nmw) <= 4
    assert np.abs(Hm[0].cpu().numpy().astype(np.int64) - Hmw.astype(np.int64)).sum() <= 3e-4 * Hmw.sum()
    # all-ones moving mask == no moving mask on the same (per-voxel) kernel
    ones = np.ones(moving.shape, np.uint8)
    s1, c1 = kernels.resample_ssd(f, mv, A=A, moving_mask=ones)
    s1, c1 = float(s1[0]), int(c1[0])
    s0, c0 = kernels.resample_ssd(f, mv, A=A)
    assert c1 == int(c0[0]) and abs(s1 - float(s0[0])) <= 1e-6 * s1                 # tile vs per-voxel arithmetic
