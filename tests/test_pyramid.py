"""block-mean pyramid level and the coarse-to-fine registration driver."""
import numpy as np
import pytest

from medicalimageanalysis_b200 import geometry as G
from oracle import pyramid as OP


def test_oracle_block_mean_known_answers():
    a = np.arange(2 * 4 * 6, dtype=np.int16).reshape(2, 4, 6)
    m = OP.oracle_shrink_mean(a, (2, 2, 2))
    assert m.shape == (1, 2, 3)
    assert m[0, 0, 0] == np.floor(a[:2, :2, :2].mean() + 0.5)
    f = np.ones((3, 5, 7), dtype=np.float32)
    assert np.array_equal(OP.oracle_shrink_mean(f, (2, 2, 2)), np.ones((2, 3, 4), dtype=np.float32))   # partial blocks
    assert np.array_equal(OP.oracle_shrink_mean(a, (1, 1, 1)), a)


def test_pyramid_grid_keeps_physical_positions():
    g = G.Grid((10, 12, 7), (0.9, 1.1, 2.5), (-4.0, 3.0, 10.0))
    p = G.pyramid_grid(g, (2, 2, 1))
    assert p.dims == (5, 6, 7) and p.spacing == (1.8, 2.2, 2.5)
    # the first coarse voxel sits at the centre of the fine block (0..1, 0..1, 0)
    assert np.allclose(p.origin, (-4.0 + 0.45, 3.0 + 0.55, 10.0))
    # physical position of coarse voxel (i, j, k) = mean position of the fine voxels it covers
    fine = G.index_to_physical(g)
    coarse = G.index_to_physical(p)
    want = (fine @ np.array([2 * 3 + 0.5, 2 * 4 + 0.5, 5, 1.0]))[:3]
    assert np.allclose((coarse @ np.array([3, 4, 5, 1.0]))[:3], want)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("shape,factors", [((8, 12, 16), (2, 2, 2)), ((7, 9, 11), (2, 2, 2)), ((5, 6, 10), (2, 2, 1)),
                                           ((9, 10, 13), (4, 4, 4)), ((1, 1, 1), (2, 2, 2)), ((6, 8, 9), (1, 2, 3)),
                                           ((40, 64, 96), (2, 2, 2))])
def test_shrink_mean_matches_oracle(cuda_device, dtype, shape, factors):
    import torch
    from medicalimageanalysis_b200 import kernels
    rng = np.random.default_rng(sum(shape))
    a = (rng.random(shape) * 4000 - 1000).astype(dtype)
    got = kernels.shrink_mean(torch.from_numpy(a).to(cuda_device), factors).cpu().numpy()
    want = OP.oracle_shrink_mean(a, factors)
    assert got.shape == want.shape and got.dtype == want.dtype
    if dtype == np.int16:
        # float32 sum * reciprocal against float64 mean: exact except on half-way ties of /3, /6 ... block counts
        assert np.abs(got.astype(np.int32) - want.astype(np.int32)).max() <= (0 if all(f in (1, 2, 4) for f in factors) else 1)
    else:
        assert np.allclose(got, want, rtol=1e-6, atol=1e-3)


@pytest.mark.gpu
def test_multires_registration_recovers_a_large_offset(cuda_device):
    """A 12 mm / 8 degree offset is outside the capture range of the single-level loop at 0.5 mm steps in the same
    number of evaluations; the pyramid gets there, and its last level refines like the plain loop."""
    import torch
    from medicalimageanalysis_b200 import Data, RegistrationConfig
    from medicalimageanalysis_b200.phantom import make_phantom
    from medicalimageanalysis_b200.utils import CreateImageFromMask, rigid
    Data.clear()
    shape, sp = (64, 96, 96), (2.0, 2.0, 2.0)
    gt = [3.0, -2.0, 8.0, 12.0, -9.0, 6.0]
    g = G.Grid((shape[2], shape[1], shape[0]), sp, (-96.0, -96.0, -64.0))
    center = (G.index_to_physical(g) @ np.array([48, 48, 32, 1.0]))[:3]
    M = G.params_to_matrix(gt, center)
    fixed = make_phantom(shape, "CT", seed=11, spacing=sp, noise=False).astype(np.float32)
    moving = make_phantom(shape, "CT", seed=11, spacing=sp, noise=False, world_matrix=np.linalg.inv(M)).astype(np.float32)
    fi = CreateImageFromMask(fixed, g.origin, sp, "F").add_image()
    mi = CreateImageFromMask(moving, g.origin, sp, "M").add_image()
    cfg = RegistrationConfig(metric="ssd", iterations=60, rot_step_deg=0.5, trans_step_mm=0.5)
    res = rigid.register_rigid_multires(fi, mi, cfg=cfg, shrink_factors=(4, 2, 1), iterations=(60, 40, 30))
    lv = res.extra["levels"]
    assert [l["factors"] for l in lv] == [(4, 4, 4), (2, 2, 2), (1, 1, 1)]
    assert lv[0]["dims"] == (24, 24, 16)
    assert np.abs(res.params[:3] - gt[:3]).max() < 0.5 and np.abs(res.params[3:] - gt[3:]).max() < 1.0
    # every level starts where the previous one ended and does not get worse on its own cost
    assert all(l["cost_last"] <= l["cost_first"] for l in lv)
    Data.clear()
