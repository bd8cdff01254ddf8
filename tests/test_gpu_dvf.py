"""Displacement-field kernels against the oracle and, for the point sampler, against
``scipy.ndimage.map_coordinates`` itself (the reference's call, ``structure/deformable.py:873-878``)."""
import numpy as np
import pytest
import torch
from scipy import ndimage

from medicalimageanalysis_b200 import Data, geometry as G, kernels
from medicalimageanalysis_b200.phantom import make_phantom
from medicalimageanalysis_b200.utils import CreateImageFromMask, deformable
from oracle import dvf as D
from tests.test_oracle_dvf import _smooth_field

pytestmark = pytest.mark.gpu
STAGED, PIPE = 3, 9            # mia_debug_last_path: serial box-staged kernel (field by LDG) / pipelined kernel (field by TMA)


def _case(shape, sp, seed):
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    w2p = G.physical_to_index(g, "itk")
    return g, (w2p @ G.index_to_physical(g, "itk"))[:3], w2p[:3, :3]


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("mode", ["itk", "vtk"])
def test_warp_dvf_matches_oracle(cuda_device, dtype, mode):
    shape, sp = (37, 45, 70), (1.0, 1.1, 2.0)
    g, A, B = _case(shape, sp, 0)
    img = make_phantom(shape, "CT", seed=4, spacing=sp).astype(dtype)
    field = _smooth_field(shape, 4.0, 5)
    field[:3] += 60.0                                   # pushes the first planes out of the volume: background
    got = kernels.warp_dvf(torch.from_numpy(img).to(cuda_device), torch.from_numpy(field).to(cuda_device), A, B,
                           background=-3001.0, mode=mode).cpu().numpy()
    want = D.oracle_warp_dvf(img, field, A, B, background=-3001.0, mode=mode)
    assert (want == -3001).mean() > 0.02
    if dtype == np.int16:
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        # float32 taps and weights against the float64 oracle: rounding ties only, and inside/outside flips on the
        # half-voxel border where the float32 displacement meets a float64 coordinate
        edge = (got == -3001) != (want == -3001)
        assert edge.mean() < 1e-4
        assert d[~edge].max() <= 1 and (d[~edge] > 0).mean() < 2e-3
    else:
        edge = (got == -3001) != (want == -3001)
        assert edge.mean() < 1e-4
        assert np.allclose(got[~edge], want[~edge], rtol=1e-4, atol=1e-2)


def test_warp_zero_field_is_identity_and_constant_field_is_a_translation(cuda_device):
    shape, sp = (24, 40, 64), (1.0, 1.0, 2.5)
    g, A, B = _case(shape, sp, 0)
    img = make_phantom(shape, "CT", seed=6, spacing=sp)
    t = torch.from_numpy(img).to(cuda_device)
    zero = torch.zeros(shape + (3,), dtype=torch.float32, device=cuda_device)
    assert torch.equal(kernels.warp_dvf(t, zero, A, B, mode="itk"), t)
    shift = np.array([2.0, -3.0, 5.0], dtype=np.float32)            # whole voxels: (2, -3, 2)
    const = torch.from_numpy(np.broadcast_to(shift, shape + (3,)).copy()).to(cuda_device)
    As = A.copy()
    As[:, 3] += B @ shift.astype(np.float64)
    assert torch.equal(kernels.warp_dvf(t, const, A, B, background=-3001.0, mode="itk"),
                       kernels.resample(t, As, shape, background=-3001.0, mode="itk"))


def test_point_sampler_equals_scipy_map_coordinates(cuda_device):
    shape = (19, 23, 31)
    field = _smooth_field(shape, 3.0, 7)
    rng = np.random.default_rng(8)
    pts = rng.uniform(-3.0, 34.0, (20000, 3))
    pts[:100] = np.round(pts[:100])
    got = kernels.sample_field_points(torch.from_numpy(field).to(cuda_device), torch.from_numpy(pts).to(cuda_device),
                                      scale=0.6).cpu().numpy()
    for c in range(3):
        want = ndimage.map_coordinates(0.6 * field[..., c].astype(np.float64), [pts[:, 2], pts[:, 1], pts[:, 0]],
                                       order=1, mode="nearest")
        assert np.allclose(got[:, c], want, rtol=0, atol=1e-12)
    assert np.allclose(got, D.oracle_sample_field_points(field, pts, 0.6), rtol=0, atol=1e-12)
    empty = kernels.sample_field_points(torch.from_numpy(field).to(cuda_device),
                                        torch.zeros((0, 3), dtype=torch.float64, device=cuda_device))
    assert tuple(empty.shape) == (0, 3)


def test_deformable_operators_follow_the_reference_recipe(cuda_device):
    """``utils.deformable``: displace_points == the reference's map_coordinates loop; resample_and_warp == rigid ITK
    resample followed by the field resample, both checked against the oracle composition."""
    from medicalimageanalysis_b200.utils import rigid as R
    from oracle import reslice
    shape, sp = (24, 36, 48), (1.2, 1.2, 2.0)
    origin = (-28.0, -21.0, -24.0)
    Data.clear()
    ref = make_phantom(shape, "CT", seed=9, spacing=sp)
    M = G.params_to_matrix([2.0, -1.0, 3.0, 2.0, 1.0, -2.0], (0, 0, 0))
    mov = make_phantom(shape, "CT", seed=9, spacing=sp, world_matrix=np.linalg.inv(M))
    f = CreateImageFromMask(ref, origin, sp, "ref", device=cuda_device).add_image()
    m = CreateImageFromMask(mov, origin, sp, "mov", device=cuda_device).add_image()
    field = _smooth_field(shape, 2.5, 10)
    got = deformable.resample_and_warp(m, f, M, field).cpu().numpy()
    A1 = G.compose_index_matrix(f.grid, m.grid, M, convention="itk")
    stage = reslice.oracle_reslice(mov, A1, f.grid.dims, background=-3001.0, mode="itk")
    w2p = G.physical_to_index(f.grid, "itk")
    want = D.oracle_warp_dvf(stage, field, (w2p @ G.index_to_physical(f.grid, "itk"))[:3], w2p[:3, :3], mode="itk")
    d = np.abs(got.astype(np.int32) - want.astype(np.int32))
    edge = (got == -3001) != (want == -3001)
    assert edge.mean() < 2e-3 and (d[~edge] > 1).mean() < 2e-3      # stage-1 rounding ties propagate through stage 2
    rng = np.random.default_rng(12)
    pts = rng.uniform(-30, 30, (5000, 3))
    moved = deformable.displace_points(pts, field, origin, sp, percent=40.0)
    vox = (pts - np.asarray(origin)) / np.asarray(sp)
    want_pts = pts.copy()
    for i in range(3):
        want_pts[:, i] += ndimage.map_coordinates((40.0 * field[..., i].astype(np.float64)) / 100,
                                                  [vox[:, 2], vox[:, 1], vox[:, 0]], order=1, mode="nearest")
    assert np.allclose(moved, want_pts, rtol=0, atol=1e-11)
    Data.clear()


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("mode", ["itk", "vtk"])
@pytest.mark.parametrize("kind", ["smooth", "tear", "far"])
def test_box_staged_warp_matches_oracle(cuda_device, dtype, mode, kind):
    """Row pitches the TMA unit accepts take the box-staged, software-pipelined kernel (path 9: image box AND field tile
    by TMA): smooth fields stay inside the box, a field
    with a tear sends the affected tiles to the per-voxel gather, a field that points out of the volume gives
    background -- all against the oracle, and equal to the direct kernel up to float32 coordinate rounding."""
    from medicalimageanalysis_b200 import _lib
    shape, sp = (40, 56, 96), (1.0, 1.1, 2.0)
    g, A, B = _case(shape, sp, 0)
    img = make_phantom(shape, "CT", seed=14, spacing=sp).astype(dtype)
    field = _smooth_field(shape, 4.0, 15)
    if kind == "tear":
        field[:, :, 40:, 0] += 37.0                                  # a jump of 37 mm across one voxel column
        field[20:, 28:, :, 2] -= 24.0
    if kind == "far":
        field[:6] += 400.0                                           # the first planes leave the volume altogether
        field[:, :5, :, 1] -= 2.0 * sp[1]                            # and a strip straddles the lower y border
    t, f = torch.from_numpy(img).to(cuda_device), torch.from_numpy(field).to(cuda_device)
    got = kernels.warp_dvf(t, f, A, B, background=-3001.0, mode=mode).cpu().numpy()
    assert _lib.load().mia_debug_last_path() == PIPE
    want = D.oracle_warp_dvf(img, field, A, B, background=-3001.0, mode=mode)
    edge = (got == -3001) != (want == -3001)
    assert edge.mean() < 1e-4
    if kind == "far":
        assert (want == -3001).mean() > 0.1
    if dtype == np.int16:
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert d[~edge].max() <= 1 and (d[~edge] > 0).mean() < 2e-3
    else:
        assert np.allclose(got[~edge], want[~edge], rtol=1e-4, atol=1e-2)


def test_box_staged_warp_identity_translation_and_partial_tiles(cuda_device):
    from medicalimageanalysis_b200 import _lib
    shape, sp = (21, 45, 72), (1.0, 1.0, 2.5)                         # partial tiles in y and z, full rows of 72
    g, A, B = _case(shape, sp, 0)
    img = make_phantom(shape, "CT", seed=16, spacing=sp)
    t = torch.from_numpy(img).to(cuda_device)
    zero = torch.zeros(shape + (3,), dtype=torch.float32, device=cuda_device)
    assert torch.equal(kernels.warp_dvf(t, zero, A, B, mode="itk"), t)
    assert _lib.load().mia_debug_last_path() == PIPE
    shift = np.array([2.0, -3.0, 5.0], dtype=np.float32)
    const = torch.from_numpy(np.broadcast_to(shift, shape + (3,)).copy()).to(cuda_device)
    As = A.copy()
    As[:, 3] += B @ shift.astype(np.float64)
    assert torch.equal(kernels.warp_dvf(t, const, A, B, background=-3001.0, mode="itk"),
                       kernels.resample(t, As, shape, background=-3001.0, mode="itk"))


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
def test_box_staged_warp_oblique_field_axes(cuda_device, dtype):
    """A field whose physical axes are not the image's index axes (B not diagonal) takes the general coordinate path."""
    from medicalimageanalysis_b200 import _lib
    shape, sp = (32, 48, 64), (1.0, 1.0, 2.0)
    g, A, B = _case(shape, sp, 0)
    Rz = G.euler_xyz_matrix(4.0, -3.0, 10.0)
    B2 = B @ Rz
    img = make_phantom(shape, "CT", seed=18, spacing=sp).astype(dtype)
    field = _smooth_field(shape, 5.0, 19)
    got = kernels.warp_dvf(torch.from_numpy(img).to(cuda_device), torch.from_numpy(field).to(cuda_device), A, B2,
                           background=-3001.0, mode="itk").cpu().numpy()
    assert _lib.load().mia_debug_last_path() == PIPE
    want = D.oracle_warp_dvf(img, field, A, B2, background=-3001.0, mode="itk")
    edge = (got == -3001) != (want == -3001)
    assert edge.mean() < 1e-4
    if dtype == np.int16:
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert d[~edge].max() <= 1 and (d[~edge] > 0).mean() < 2e-3
    else:
        assert np.allclose(got[~edge], want[~edge], rtol=1e-4, atol=1e-2)


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("ox", [70, 61, 72])
def test_field_lattice_other_than_the_image_lattice(cuda_device, dtype, ox):
    """The field lives on the REFERENCE lattice (Deformable.create_image warps the moving image onto it), narrower and
    shorter than the image here.  A field row of ox * 12 bytes is a TMA stride only when ox % 4 == 0: 72 takes the
    pipelined kernel, 70 and 61 the serial one with the field read by LDG -- same oracle, same tolerance."""
    from medicalimageanalysis_b200 import _lib
    shape, sp = (40, 56, 96), (1.0, 1.1, 2.0)
    oshape = (29, 50, ox)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    og = G.Grid((oshape[2], oshape[1], oshape[0]), sp, (-0.5 * sp[0] * shape[2] + 7.3, -0.5 * sp[1] * shape[1] + 2.2, -0.5 * sp[2] * shape[0] + 5.0))
    w2p = G.physical_to_index(g, "itk")
    A, B = (w2p @ G.index_to_physical(og, "itk"))[:3], w2p[:3, :3]
    img = make_phantom(shape, "CT", seed=21, spacing=sp).astype(dtype)
    field = _smooth_field(oshape, 4.0, 22)
    got = kernels.warp_dvf(torch.from_numpy(img).to(cuda_device), torch.from_numpy(field).to(cuda_device), A, B,
                           background=-3001.0, mode="itk").cpu().numpy()
    assert _lib.load().mia_debug_last_path() == (PIPE if ox % 4 == 0 else STAGED)
    want = D.oracle_warp_dvf(img, field, A, B, background=-3001.0, mode="itk")
    edge = (got == -3001) != (want == -3001)
    assert edge.mean() < 1e-4
    if dtype == np.int16:
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert d[~edge].max() <= 1 and (d[~edge] > 0).mean() < 2e-3
    else:
        assert np.allclose(got[~edge], want[~edge], rtol=1e-4, atol=1e-2)


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("oshape", [(8, 8, 32), (3, 5, 8), (17, 8, 64), (64, 24, 96), (200, 104, 160)])
def test_pipelined_warp_on_one_two_three_and_many_tiles_per_cta(cuda_device, dtype, oshape):
    """The pipeline's prologue requests two field tiles and one box ahead; lattices of 1, 1 (partial), 6, 72 and 3250
    tiles give CTAs with one, two, three and ten tiles (every buffer and barrier phase comes round)."""
    from medicalimageanalysis_b200 import _lib
    shape, sp = (200, 112, 168), (1.0, 1.0, 1.0)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, (0.0, 0.0, 0.0))
    og = G.Grid((oshape[2], oshape[1], oshape[0]), sp, (3.25, 2.5, 0.75))
    w2p = G.physical_to_index(g, "itk")
    A, B = (w2p @ G.index_to_physical(og, "itk"))[:3], w2p[:3, :3]
    z, y, x = np.meshgrid(*(np.arange(n, dtype=np.float32) for n in shape), indexing="ij")
    img = (500 * np.sin(x / 7) * np.cos(y / 9) + 300 * np.sin(z / 5)).astype(dtype)
    field = _smooth_field(oshape, 3.0, 23)
    got = kernels.warp_dvf(torch.from_numpy(img).to(cuda_device), torch.from_numpy(field).to(cuda_device), A, B,
                           background=-3001.0, mode="itk").cpu().numpy()
    assert _lib.load().mia_debug_last_path() == PIPE
    want = D.oracle_warp_dvf(img, field, A, B, background=-3001.0, mode="itk")
    edge = (got == -3001) != (want == -3001)
    assert edge.mean() < 1e-4
    if dtype == np.int16:
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert d[~edge].max() <= 1 and (d[~edge] > 0).mean() < 2e-3
    else:
        assert np.allclose(got[~edge], want[~edge], rtol=1e-4, atol=1e-2)


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("ox", [96, 70])
def test_nan_displacements_give_background(cuda_device, dtype, ox):
    """A NaN in the field (a registration that diverged, an unwritten voxel) must not address anything: the column's
    coordinate range keeps the NaN (min.NaN / max.NaN), the column fails the interior test, the voxel compares as
    outside.  Both box-staged kernels (96: pipelined, 70: serial)."""
    from medicalimageanalysis_b200 import _lib
    shape, sp = (40, 56, 96), (1.0, 1.1, 2.0)
    oshape = (40, 56, ox)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, (0.0, 0.0, 0.0))
    og = G.Grid((oshape[2], oshape[1], oshape[0]), sp, (0.0, 0.0, 0.0))
    w2p = G.physical_to_index(g, "itk")
    A, B = (w2p @ G.index_to_physical(og, "itk"))[:3], w2p[:3, :3]
    img = make_phantom(shape, "CT", seed=31, spacing=sp).astype(dtype)
    field = _smooth_field(oshape, 2.0, 32)
    rng = np.random.default_rng(33)
    holes = rng.random(oshape) < 0.01
    for c in range(3):                                               # single components, so that x, y and z ranges all see one
        field[..., c][holes & (rng.integers(0, 3, oshape) == c)] = np.nan
    field[10:14, 20:30, 40:60] = np.nan                              # and a block: whole columns and warps of NaN
    got = kernels.warp_dvf(torch.from_numpy(img).to(cuda_device), torch.from_numpy(field).to(cuda_device), A, B,
                           background=-3001.0, mode="itk").cpu().numpy()
    assert _lib.load().mia_debug_last_path() == (PIPE if ox % 4 == 0 else STAGED)
    bad = np.isnan(field).any(axis=-1)
    assert bad.mean() > 0.01 and np.all(got[bad] == -3001)
    want = D.oracle_warp_dvf(img, field, A, B, background=-3001.0, mode="itk")
    assert np.all(want[bad] == -3001)
    edge = (got == -3001) != (want == -3001)
    assert edge.mean() < 1e-4
    if dtype == np.int16:
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert d[~edge].max() <= 1
    else:
        assert np.allclose(got[~edge], want[~edge], rtol=1e-4, atol=1e-2)
