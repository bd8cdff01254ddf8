"""Parity of the TMA-staged tile kernel (csrc/mia_tile.cuh) with the oracle, through the C ABI.

The tile path needs a 16-byte aligned moving row pitch, so these cases use x sizes that are multiples of 8;
``mia_debug_last_path()`` proves which kernel ran.  Tolerances as in test_gpu_resample.py.
"""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import _lib, geometry as G
from medicalimageanalysis_b200 import kernels
from medicalimageanalysis_b200.phantom import make_phantom
from oracle import metrics, native, reslice
from tests.test_gpu_resample import BG, border_band, check_parity

pytestmark = pytest.mark.gpu
TILE, DIRECT = 1, 0
X2 = 5             # mia_tile_x2.cuh: int16 resample, 64-wide tiles (takes the call when its box fits the matrix)


def last_path():
    return int(_lib.load().mia_debug_last_path())


MULTI = 2          # mia_tile_multi.cuh
MULTI_X2 = 6       # mia_tile_multi_x2.cuh: int16, every probe fits the 64-wide tile's box


def is_multi(dtype=np.int16):
    return last_path() in ((MULTI, MULTI_X2) if np.dtype(dtype) == np.int16 else (MULTI,))


def grid_of(shape, sp):
    return G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))


def run(dev, mov, A, odims, mode, expect=TILE):
    t = torch.from_numpy(mov).to(dev)
    got = kernels.resample(t, A, (odims[2], odims[1], odims[0]), background=BG[mode], mode=mode)
    assert last_path() == expect or (expect == TILE and mov.dtype == np.int16 and last_path() == X2)
    want = reslice.oracle_reslice(mov, A, odims, BG[mode], mode, mov.dtype)
    near = border_band(A, odims, (mov.shape[2], mov.shape[1], mov.shape[0]))
    check_parity(got.cpu().numpy(), want, near, mov.dtype, BG[mode])
    return got


@pytest.mark.parametrize("modality", ["CT", "MR"])
@pytest.mark.parametrize("seed", range(4))
def test_tile_random_rigid_parity(cuda_device, modality, seed):
    rng = np.random.default_rng(300 + seed)
    shape = (int(rng.integers(9, 60)), int(rng.integers(17, 90)), 8 * int(rng.integers(5, 20)))
    sp = (float(rng.uniform(0.6, 1.2)),) * 2 + (float(rng.uniform(1.0, 3.0)),)
    mov = make_phantom(shape, modality, seed=seed, spacing=sp)
    g = grid_of(shape, sp)
    p = np.concatenate([rng.uniform(-4, 4, 3), rng.uniform(-8, 8, 3)])
    A = G.compose_index_matrix(g, g, G.params_to_matrix(p, (0, 0, 0)))
    for mode in ("vtk", "itk"):
        run(cuda_device, mov, A, g.dims, mode)


def test_tile_autocrop_output_larger_than_moving(cuda_device):
    """Output lattice sticks out of the moving volume on every side: outside, border and interior tiles."""
    shape, sp = (40, 64, 96), (0.9765625, 0.9765625, 2.5)
    mov = make_phantom(shape, "CT", seed=5, spacing=sp)
    g = grid_of(shape, sp)
    M = G.params_to_matrix([3, -2, 4, 5, -3, 7], (0, 0, 0))
    origin, odims = G.autocrop_extent(g, M, sp)
    og = G.Grid(tuple(int(d) for d in odims), sp, tuple(origin))
    A = G.compose_index_matrix(og, g, M, "vtk")
    run(cuda_device, mov, A, og.dims, "vtk")
    # a much larger canvas: most tiles are entirely outside
    A2 = A.copy()
    A2[:, 3] -= A2[:, :3] @ np.array([40.0, 30.0, 10.0])
    run(cuda_device, mov, A2, (og.dims[0] + 80, og.dims[1] + 60, og.dims[2] + 20), "vtk")


@pytest.mark.parametrize("name", ["identity", "shift_int", "shift_half"])
@pytest.mark.parametrize("tag", ["i16", "f32"])
def test_tile_exact_cases_bit_exact(cuda_device, name, tag):
    rng = np.random.default_rng(11)
    shape = (21, 37, 64)
    mov = rng.integers(-1024, 3071, size=shape).astype(np.int16 if tag == "i16" else np.float32)
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    if name == "shift_int":
        A[:, 3] = [3.0, -2.0, 1.0]
    elif name == "shift_half":
        A[:, 3] = [0.5, -0.5, 0.5]
    odims = (64, 37, 21)
    for mode in ("vtk", "itk"):
        t = torch.from_numpy(mov).to(cuda_device)
        got = kernels.resample(t, A, shape, background=BG[mode], mode=mode)
        assert last_path() == (X2 if tag == "i16" else TILE)
        want = reslice.oracle_reslice(mov, A, odims, BG[mode], mode, mov.dtype)
        assert np.array_equal(got.cpu().numpy(), want), (name, tag, mode)


def test_tile_thin_volume_and_ragged_output(cuda_device):
    """Moving volume thinner than the box; output sizes that are not multiples of the 32x8x8 tile."""
    rng = np.random.default_rng(12)
    mov = rng.integers(-500, 1500, size=(3, 5, 16)).astype(np.int16)
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    A[:, 3] = [0.3, -0.2, 0.1]
    for odims in [(16, 5, 3), (33, 9, 9), (1, 1, 1), (70, 3, 17)]:
        for mode in ("vtk", "itk"):
            run(cuda_device, mov, A, odims, mode)


def test_large_rotation_host_matrix_goes_direct_and_device_matrix_falls_back_per_tile(cuda_device):
    shape, sp = (32, 64, 64), (1.0, 1.0, 1.0)
    fixed = make_phantom(shape, "CT", seed=8, spacing=sp)
    moving = make_phantom(shape, "CT", seed=9, spacing=sp)
    g = grid_of(shape, sp)
    A = G.compose_index_matrix(g, g, G.params_to_matrix([35, -20, 40, 2, 1, -3], (0, 0, 0)))
    run(cuda_device, moving, A, g.dims, "vtk", expect=DIRECT)
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    s_host, c_host = kernels.resample_ssd(f, m, A=A)
    assert last_path() == DIRECT
    dA = torch.from_numpy(np.ascontiguousarray(A[None])).to(cuda_device)
    s_dev, c_dev = kernels.resample_ssd(f, m, d_A=dA, ws=kernels.MetricWorkspace(cuda_device, n_mats=1))
    assert last_path() == TILE
    want, n = native.reslice_ssd(fixed, moving, A, "itk")
    for s, c in ((s_host, c_host), (s_dev, c_dev)):
        assert abs(int(c[0]) - n) <= 2
        assert abs(float(s[0]) - want) <= 1e-5 * want


@pytest.mark.parametrize("modality", ["CT", "MR"])
def test_tile_metrics_match_oracle(cuda_device, modality):
    shape, sp = (40, 56, 80), (1.0, 1.0, 2.0)
    g = grid_of(shape, sp)
    Mgt = G.params_to_matrix([3, -2, 4, 3, -2, 1], (0, 0, 0))
    fixed = make_phantom(shape, modality, seed=3, spacing=sp)
    moving = make_phantom(shape, modality, seed=3, spacing=sp, world_matrix=np.linalg.inv(Mgt))
    rng = np.random.default_rng(5)
    A = np.stack([G.compose_index_matrix(g, g, G.params_to_matrix(
        np.concatenate([rng.uniform(-4, 4, 3), rng.uniform(-4, 4, 3)]), (0, 0, 0))) for _ in range(13)])
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    ws = kernels.MetricWorkspace(cuda_device, n_mats=13, bins=64)
    s, c = kernels.resample_ssd(f, m, A=A, ws=ws)
    assert is_multi(fixed.dtype)           # 13 unrelated matrices: one shared box, per-tile gather for the misfits
    fr = (float(fixed.min()), float(fixed.max()))
    mr = (float(moving.min()), float(moving.max()))
    s, c = s.cpu().numpy().copy(), c.cpu().numpy().copy()
    h, hc = kernels.resample_jointhist(f, m, fr, mr, bins=64, A=A, ws=ws)
    assert last_path() == TILE
    mi = kernels.mi_from_hist(h).cpu().numpy()
    h, hc = h.cpu().numpy(), hc.cpu().numpy()
    for i in range(13):
        want, n = native.reslice_ssd(fixed, moving, A[i], "itk")
        assert abs(int(c[i]) - n) <= 2
        assert abs(float(s[i]) - want) <= 1e-5 * want
        H = native.joint_hist(fixed, moving, A[i], 64, fr, mr, "itk")[0]
        assert abs(int(hc[i]) - int(H.sum())) <= 2 and int(h[i].sum()) == int(hc[i])
        # float32 vs float64 interpolation may move a handful of samples across a bin edge
        assert np.abs(h[i].astype(np.int64) - H.astype(np.int64)).sum() <= max(8, 2e-5 * H.sum())
        want_mi = metrics.oracle_mi(H)
        assert abs(mi[i] - want_mi) <= 1e-5 * abs(want_mi)


def test_tile_ssd_store_equals_plain_resample_and_is_deterministic(cuda_device):
    shape, sp = (24, 48, 72), (1.0, 1.0, 2.0)
    g = grid_of(shape, sp)
    fixed = make_phantom(shape, "CT", seed=3, spacing=sp)
    moving = make_phantom(shape, "CT", seed=4, spacing=sp)
    A = G.compose_index_matrix(g, g, G.params_to_matrix([2, 1, -3, 1, -2, 2], (0, 0, 0)))
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    out = torch.empty_like(m)
    s, c = kernels.resample_ssd(f, m, A=A, mode="itk", out=out, background=0.0)
    assert last_path() == TILE
    s0 = float(s[0])
    plain = kernels.resample(m, A, m.shape, background=0.0, mode="itk")
    assert torch.equal(out, plain)
    for _ in range(3):
        s2, _ = kernels.resample_ssd(f, m, A=A, mode="itk")
        assert float(s2[0]) == s0


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("mode", ["itk", "vtk"])
def test_multi_matrix_ssd_equals_one_matrix_launches(cuda_device, dtype, mode):
    """The 13 probes of an optimiser iteration share one box per tile (csrc/mia_tile_multi.cuh).  Every voxel
    value is computed by the same tile_compute as in the one-matrix kernel, so counts are equal and sums agree
    to summation-order rounding; the oracle pins both."""
    from medicalimageanalysis_b200.registration import probe_params
    shape, sp = (56, 72, 96), (1.0, 1.0, 1.5)
    g = grid_of(shape, sp)
    fixed = make_phantom(shape, "CT", seed=7, spacing=sp).astype(dtype)
    moving = make_phantom(shape, "CT", seed=7, spacing=sp,
                          world_matrix=np.linalg.inv(G.params_to_matrix([2, -3, 2.5, 3, -2, 1.5], (0, 0, 0)))).astype(dtype)
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    for scale in (1.0, 4.0):                       # probe offsets 0.5 and 2 (degrees / mm)
        probes = probe_params(np.array([1.5, -2.5, 2.0, 2.0, -1.0, 1.0]), scale, np.array([0.5] * 6))
        A = np.stack([G.compose_index_matrix(g, g, G.params_to_matrix(q, (0, 0, 0))) for q in probes])
        for use_dev in (False, True):
            ws = kernels.MetricWorkspace(cuda_device, n_mats=13)
            if use_dev:
                s, c = kernels.resample_ssd(f, m, d_A=torch.from_numpy(A).to(cuda_device), mode=mode, ws=ws)
            else:
                s, c = kernels.resample_ssd(f, m, A=A, mode=mode, ws=ws)
            assert is_multi(dtype)
            s, c = s.cpu().numpy().copy(), c.cpu().numpy().copy()
            s2, c2 = kernels.resample_ssd(f, m, A=A, mode=mode, ws=ws)        # workspace reuse: ticket self-reset
            assert np.array_equal(s2.cpu().numpy(), s) and np.array_equal(c2.cpu().numpy(), c)
            for i in range(13):
                s1, c1 = kernels.resample_ssd(f, m, A=A[i], mode=mode)
                assert last_path() == TILE
                assert int(c1[0]) == int(c[i])
                # probes half a voxel apart share every box: identical voxel values, summation order only; at 2
                # voxels / degrees some tiles of some probes miss the shared box and take the per-voxel gather
                # (float64 row origin: last-bit differences in float32)
                # (int16 launches the x-pair kernel takes: its second voxel of a pair sits at fx + a00 in float32 where
                # the one-matrix 32-wide kernel has an exact Q31.32 base -- 1e-7 of a voxel, 1e-6 relative in the sum)
                x2 = np.dtype(dtype) == np.int16
                assert abs(float(s1[0]) - float(s[i])) <= (1e-6 if x2 else (1e-12 if scale == 1.0 else 1e-7)) * float(s1[0])
            for i in (0, 5, 12):
                want, n = native.reslice_ssd(fixed, moving, A[i], mode)
                assert abs(int(c[i]) - n) <= 2 and abs(float(s[i]) - want) <= 1e-5 * want


def test_multi_matrix_handles_outside_and_ragged(cuda_device):
    """Matrices that push the tile outside the volume, a ragged output lattice and 2 / 16 matrices per launch."""
    shape, sp = (19, 45, 70), (1.0, 1.0, 2.0)
    g = grid_of(shape, sp)
    fixed = make_phantom(shape, "CT", seed=9, spacing=sp)
    moving = make_phantom(shape, "CT", seed=10, spacing=sp)
    mov = np.ascontiguousarray(np.pad(moving, ((0, 0), (0, 0), (0, 2))))      # nx = 72: 16-byte row pitch
    gm = G.Grid((72, 45, 19), sp, g.origin)
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(mov).to(cuda_device)
    for nm in (2, 16):
        rng = np.random.default_rng(nm)
        prm = [np.concatenate([rng.uniform(-2, 2, 3), rng.uniform(-1, 1, 3) * (40 if i % 3 == 2 else 2)]) for i in range(nm)]
        A = np.stack([G.compose_index_matrix(g, gm, G.params_to_matrix(q, (0, 0, 0))) for q in prm])
        s, c = kernels.resample_ssd(f, m, A=A)
        assert is_multi()
        for i in range(nm):
            want, n = native.reslice_ssd(fixed, mov, A[i], "itk")
            assert abs(int(c[i]) - n) <= 2
            assert abs(float(s[i]) - want) <= 1e-5 * max(want, 1.0)


def test_histogram_sizes_in_any_order(cuda_device):
    """The dynamic shared-memory limit of the histogram kernels is only ever raised: a small launch between two
    large ones must not break the cached configuration of the large one (regression)."""
    shape, sp = (24, 40, 64), (1.0, 1.0, 2.0)
    g = grid_of(shape, sp)
    fixed = make_phantom(shape, "CT", seed=2, spacing=sp)
    moving = make_phantom(shape, "CT", seed=3, spacing=sp)
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    A = G.compose_index_matrix(g, g, G.params_to_matrix([1, -1, 2, 1, 0, -1], (0, 0, 0)))
    fr = (-1024.0, 3071.0)
    ref = {}
    for bins in (128, 16, 128, 64, 16, 128):
        h, c = kernels.resample_jointhist(f, m, fr, fr, bins=bins, A=A, ws=kernels.MetricWorkspace(cuda_device, 1, bins))
        assert last_path() == TILE and int(h.sum()) == int(c[0])
        if bins in ref:
            assert torch.equal(h, ref[bins])
        ref[bins] = h.clone()


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int16, np.float32])
def test_multi_matrix_kernel_groups_probes_into_several_boxes(cuda_device, dtype):
    """A wide isotropic lattice (1024 x 1024 x 24 voxels of 0.5 mm): +-0.5 degree probes move the far tiles by +-4.5
    voxels, so one box cannot serve all 13 matrices there and the producer opens a second / third box per tile
    position (mia_tile_multi.cuh).  Every probe must equal its own one-matrix launch and the C/OpenMP oracle."""
    import torch
    from medicalimageanalysis_b200 import _lib, geometry as G, kernels
    from medicalimageanalysis_b200.registration import probe_params
    from oracle import native
    shape, sp = (24, 1024, 1024), (0.5, 0.5, 0.5)
    rng = np.random.default_rng(5)
    z, y, x = np.meshgrid(np.arange(shape[0]), np.arange(shape[1]), np.arange(shape[2]), indexing="ij", sparse=True)
    base = 600 * np.sin(x / 17.0) * np.cos(y / 23.0) * np.sin(z / 5.0 + 0.3)
    fixed = (base + rng.normal(0, 10, shape)).astype(dtype)
    moving = (600 * np.sin(x / 17.0 + 0.1) * np.cos(y / 23.0 - 0.05) * np.sin(z / 5.0 + 0.35)).astype(dtype)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    p2w, w2p = G.index_to_physical(g), G.physical_to_index(g)
    A = np.stack([(w2p @ G.params_to_matrix(p, np.zeros(3)) @ p2w)[:3]
                  for p in probe_params(np.array([0.2, -0.1, 0.3, 0.4, -0.2, 0.1]), 1.0, np.array([0.5] * 6))])
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    s13, c13 = kernels.resample_ssd(f, m, A=A)
    assert _lib.load().mia_debug_last_path() in (2, 6)
    s13, c13 = s13.clone(), c13.clone()
    for i in range(13):
        s1, c1 = kernels.resample_ssd(f, m, A=A[i])
        assert int(c1[0]) == int(c13[i]), i
        assert abs(float(s1[0]) - float(s13[i])) <= 1e-7 * float(s13[i]), i      # gather-path tiles differ in the last bits
    for i in (0, 1, 4, 6):                                                        # centre, +rx, -ry, -rz
        want, n = native.reslice_ssd(fixed, moving, A[i], "itk")
        assert abs(int(c13[i]) - n) <= max(4, 2e-6 * n) and abs(float(s13[i]) - want) <= 1e-5 * want


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("mode", ["vtk", "itk"])
@pytest.mark.parametrize("angles", [(0, 0, 20), (1.5, -1.0, 45), (0, 0, 90), (-2.0, 1.0, 135), (0, 0, -60)])
def test_large_in_plane_rotations_take_the_tall_box(cuda_device, dtype, mode, angles):
    """Rotations about z beyond the standard box (about 11 degrees) run on the second, 36-row box family of the
    store kernel (path 4) instead of the per-voxel kernel -- the interactive view rotation of SURVEY N2."""
    shape, sp = (28, 88, 96), (1.0, 1.0, 2.5)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    mov = make_phantom(shape, "CT", seed=23, spacing=sp).astype(dtype)
    M = G.params_to_matrix(list(angles) + [3.0, -2.0, 1.0], (0, 0, 0))
    A = G.compose_index_matrix(g, g, M, convention=mode)
    got = kernels.resample(torch.from_numpy(mov).to(cuda_device), A, shape, background=BG[mode], mode=mode).cpu().numpy()
    assert _lib.load().mia_debug_last_path() == 4
    want = reslice.oracle_reslice(mov, A, g.dims, BG[mode], mode, mov.dtype)
    near = border_band(A, g.dims, g.dims)
    d = np.abs(got.astype(np.float64) - want.astype(np.float64))
    if dtype == np.int16:
        assert not ((d > 1) & ~near).any()                        # +-1 LSB
    else:
        # 1e-4 relative; at the zero crossings of a +-1000 HU signal "relative" is taken against 10 HU: the float32
        # lerps of taps of that magnitude carry ~5e-4 of absolute rounding whatever the kernel
        assert not ((d > 1e-4 * np.maximum(np.abs(want.astype(np.float64)), 10.0)) & ~near).any()
    assert near.mean() < 0.02


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int16, np.float32])
@pytest.mark.parametrize("angles", [(15, 0, 0), (30, 0, 0), (45, 0, 0), (90, 0, 0), (0, 15, 0), (0, 30, 0), (0, 45, 0), (0, 90, 0),
                                    (0, 0, 15), (0, 0, 30), (0, 0, 45), (0, 0, 90), (20, -15, 10)])
def test_no_single_axis_rotation_falls_to_the_per_voxel_kernel(cuda_device, dtype, angles):
    """Rigid.update_rotation takes any Euler triple (structure/rigid.py:950-974).  Rotations about z run on the tall box
    family, about y on the z-heavy one, about x and moderate compound rotations on the cube -- never on the per-voxel
    gather (path 0) -- and all of them match the oracle."""
    shape, sp = (64, 72, 96), (1.0, 1.0, 1.0)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    mov = make_phantom(shape, "CT", seed=31, spacing=sp).astype(dtype)
    M = G.params_to_matrix(list(angles) + [2.0, -1.0, 1.5], (0, 0, 0))
    for mode in ("vtk", "itk"):
        A = G.compose_index_matrix(g, g, M, convention=mode)
        got = kernels.resample(torch.from_numpy(mov).to(cuda_device), A, shape, background=BG[mode], mode=mode).cpu().numpy()
        assert _lib.load().mia_debug_last_path() in (1, 4, 5, 7, 8), angles
        want = reslice.oracle_reslice(mov, A, g.dims, BG[mode], mode, mov.dtype)
        near = border_band(A, g.dims, g.dims)
        d = np.abs(got.astype(np.float64) - want.astype(np.float64))
        if dtype == np.int16:
            assert not ((d > 1) & ~near).any()
        else:
            assert not ((d > 1e-4 * np.maximum(np.abs(want.astype(np.float64)), 10.0)) & ~near).any()
