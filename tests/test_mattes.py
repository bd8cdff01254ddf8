"""Mattes (Parzen-window) mutual information -- table, oracle, kernels, registration."""
import numpy as np
import pytest

from oracle import metrics


def test_parzen_table_is_a_partition_of_unity_and_matches_the_bspline():
    tab = metrics.parzen_table()
    assert tab.shape == (257, 4) and (tab.sum(axis=1) == 1024).all() and (tab >= 0).all()
    t = np.arange(257) / 256.0
    w = np.stack([(1 - t) ** 3 / 6, (4 - 6 * t ** 2 + 3 * t ** 3) / 6, (4 - 6 * (1 - t) ** 2 + 3 * (1 - t) ** 3) / 6, t ** 3 / 6], axis=1)
    assert np.abs(tab[:, [0, 1, 3]] / 1024.0 - w[:, [0, 1, 3]]).max() <= 0.5 / 1024      # rounded to nearest
    assert np.abs(tab[:, 2] / 1024.0 - w[:, 2]).max() <= 1.5 / 1024                      # the remainder bin
    assert tab[0, 3] == 0 and tab[256, 0] == 0                             # t = 0 / t = 1: three-bin support
    assert np.array_equal(tab[::-1, 0], tab[:, 3])                         # symmetry of the outer weights


def test_oracle_mattes_identity_pair_is_concentrated_near_the_diagonal():
    rng = np.random.default_rng(0)
    a = (rng.random((8, 12, 16)) * 1000).astype(np.float32)
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    H, n = metrics.oracle_mattes_hist(a, a, A, 24, (0, 1000), (0, 1000))
    assert n == a.size and int(H.sum()) == 1024 * a.size
    assert H[:2].sum() == 0 and H[-2:].sum() == 0                          # fixed padding bins stay empty
    i, j = np.nonzero(H)
    assert np.abs(i - j).max() <= 2                                        # 4-bin window around the fixed bin
    assert metrics.oracle_mi(H) > 1.0


@pytest.mark.gpu
@pytest.mark.parametrize("modality", ["CT", "MR"])
@pytest.mark.parametrize("bins", [8, 32, 50])
def test_mattes_histogram_matches_oracle(cuda_device, modality, bins):
    import torch
    from medicalimageanalysis_b200 import _lib, kernels
    from tests.test_gpu_metrics import mats, pair
    g, fixed, moving, _ = pair(fdtype=modality, mdtype=modality)
    A = mats(g, 1, seed=7)[0]
    fr, mr = (float(fixed.min()), float(fixed.max())), (float(moving.min()), float(moving.max()))
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    H, c = kernels.resample_mattes(f, m, fr, mr, bins=bins, A=A)
    assert _lib.load().mia_debug_last_path() == 1
    Hw, n = metrics.oracle_mattes_hist(fixed, moving, A, bins, fr, mr)
    Hg = H[0].cpu().numpy()
    assert abs(int(c[0]) - n) <= 2 and int(Hg.sum()) == 1024 * int(c[0])
    # float32 interpolation against the float64 oracle moves a sample by at most one 1/256 sub-bin step here and
    # there: the histograms differ by a few weight units in total, MI to 1e-5
    assert np.abs(Hg.astype(np.int64) - Hw.astype(np.int64)).sum() <= 2e-4 * Hw.sum()
    mi = float(kernels.mi_from_hist(H)[0])
    assert mi == pytest.approx(metrics.oracle_mi(Hw), rel=1e-5)
    # masked, and thirteen matrices in one launch == thirteen launches, bit for bit (integer counters)
    mask = (np.random.default_rng(1).random(fixed.shape) > 0.4).astype(np.uint8)
    Hm, cm = kernels.resample_mattes(f, m, fr, mr, bins=bins, A=A, fixed_mask=mask)
    Hmw, nm = metrics.oracle_mattes_hist(fixed, moving, A, bins, fr, mr, fixed_mask=mask)
    assert abs(int(cm[0]) - nm) <= 2
    assert np.abs(Hm[0].cpu().numpy().astype(np.int64) - Hmw.astype(np.int64)).sum() <= 2e-4 * Hmw.sum()
    A13 = mats(g, 13, seed=8)
    H13, c13 = kernels.resample_mattes(f, m, fr, mr, bins=bins, A=A13)
    H13 = H13.clone()
    for i in (0, 5, 12):
        Hi, ci = kernels.resample_mattes(f, m, fr, mr, bins=bins, A=A13[i])
        assert torch.equal(Hi[0], H13[i]) and int(ci[0]) == int(c13[i])


@pytest.mark.gpu
def test_mattes_on_the_direct_kernel_equals_the_tile_kernel(cuda_device, monkeypatch):
    """Mixed dtypes (int16 fixed, float32 moving) have no tile instantiation; same integer histogram definition."""
    import torch
    from medicalimageanalysis_b200 import _lib, kernels
    from tests.test_gpu_metrics import mats, pair
    g, fixed, moving, _ = pair(fdtype="CT", mdtype="MR")
    A = mats(g, 1, seed=9)[0]
    fr, mr = (float(fixed.min()), float(fixed.max())), (float(moving.min()), float(moving.max()))
    H, c = kernels.resample_mattes(torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device), fr, mr,
                                   bins=40, A=A)
    assert _lib.load().mia_debug_last_path() == 0
    Hw, n = metrics.oracle_mattes_hist(fixed, moving, A, 40, fr, mr)
    assert abs(int(c[0]) - n) <= 2
    assert np.abs(H[0].cpu().numpy().astype(np.int64) - Hw.astype(np.int64)).sum() <= 2e-4 * Hw.sum()


@pytest.mark.gpu
def test_mattes_registration_ct_to_mr(cuda_device):
    from medicalimageanalysis_b200 import Data, RegistrationConfig, geometry as G
    from medicalimageanalysis_b200.phantom import make_phantom
    from medicalimageanalysis_b200.registration import register_rigid
    from medicalimageanalysis_b200.utils import CreateImageFromMask
    Data.clear()
    shape, sp = (64, 96, 96), (2.0, 2.0, 2.0)
    gt = [2.0, 1.5, -2.5, 3.0, -2.0, 1.5]
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    M = G.params_to_matrix(gt, (0, 0, 0))
    fixed = make_phantom(shape, "CT", seed=31, spacing=sp).astype(np.float32)
    moving = make_phantom(shape, "MR", seed=31, spacing=sp, world_matrix=np.linalg.inv(M))
    fi = CreateImageFromMask(fixed, g.origin, sp, "F").add_image()
    mi = CreateImageFromMask(moving, g.origin, sp, "M").add_image()
    # From an identity start on identical lattices a single-level loop sits on the grid-aligned interpolation artefact
    # of MI (both flavours; measured), which is what the pyramid is for: coarse level first, as Elastix is asked to do.
    from medicalimageanalysis_b200.registration import register_rigid_multires
    res = register_rigid_multires(fi, mi, cfg=RegistrationConfig(metric="mattes", bins=50, iterations=60), shrink_factors=(2, 1))
    lv = res.extra["levels"]
    assert all(l["cost_last"] < l["cost_first"] for l in lv)              # the optimiser minimises -MI at every level
    assert np.abs(res.params[:3] - gt[:3]).max() < 0.5 and np.abs(res.params[3:] - gt[3:]).max() < 1.0
    single = register_rigid(fi, mi, cfg=RegistrationConfig(metric="mattes", bins=50, iterations=20), center=np.zeros(3))
    assert single.cost_history[-1] < single.cost_history[0]
    Data.clear()
