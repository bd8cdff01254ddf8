"""Parity of the x-pair tile kernel (csrc/mia_tile_x2.cuh: int16 resample, 64 x 8 x 8 tiles) with the oracle, through the
C ABI.  ``mia_debug_last_path() == 5`` proves which kernel ran.  Tolerances as in test_gpu_resample.py: +-1 LSB, exact
where the weights are 0 / 1/2 / 1, inside / outside decisions identical away from exact ties on the half-voxel border."""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import _lib, geometry as G
from medicalimageanalysis_b200 import kernels
from medicalimageanalysis_b200.phantom import make_phantom
from oracle import reslice
from tests.test_gpu_resample import BG, border_band, check_parity

pytestmark = pytest.mark.gpu
X2 = 5


def last_path():
    return int(_lib.load().mia_debug_last_path())


def grid_of(shape, sp):
    return G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))


def run(dev, mov, A, odims, mode, out=None):
    t = torch.from_numpy(mov).to(dev)
    got = kernels.resample(t, A, (odims[2], odims[1], odims[0]), background=BG[mode], mode=mode, out=out)
    assert last_path() == X2
    want = reslice.oracle_reslice(mov, A, odims, BG[mode], mode, mov.dtype)
    near = border_band(A, odims, (mov.shape[2], mov.shape[1], mov.shape[0]))
    check_parity(got.cpu().numpy(), want, near, mov.dtype, BG[mode])
    return got, want


@pytest.mark.parametrize("seed", range(6))
def test_x2_random_rigid_parity(cuda_device, seed):
    rng = np.random.default_rng(700 + seed)
    shape = (int(rng.integers(9, 70)), int(rng.integers(17, 120)), 8 * int(rng.integers(9, 40)))
    sp = (float(rng.uniform(0.6, 1.2)),) * 2 + (float(rng.uniform(1.0, 3.0)),)
    mov = make_phantom(shape, "CT", seed=seed, spacing=sp)
    g = grid_of(shape, sp)
    p = np.concatenate([rng.uniform(-2, 2, 3), rng.uniform(-8, 8, 3)])      # 64-wide rows: the 16-row box takes about 4 degrees
    A = G.compose_index_matrix(g, g, G.params_to_matrix(p, (0, 0, 0)))
    for mode in ("vtk", "itk"):
        run(cuda_device, mov, A, g.dims, mode)


def test_x2_autocrop_lattice_with_odd_pitch_and_clipped_last_layer(cuda_device):
    """The config-2 situation in small: the auto-cropped lattice sticks out of the moving volume on every side (outside,
    border and interior tiles), its row length and plane size are odd (pairs straddle 4-byte words on every other row)
    and its depth is not a multiple of 8 (the rolled loop of the last tile layer)."""
    shape, sp = (44, 72, 136), (0.9765625, 0.9765625, 2.5)
    mov = make_phantom(shape, "CT", seed=5, spacing=sp)
    g = grid_of(shape, sp)
    M = G.params_to_matrix([3, -2, 4, 5, -3, 7], (0, 0, 0))
    origin, dims = G.autocrop_extent(g, M, sp, convention="vtk")
    og = G.Grid(dims, sp, origin)
    A = G.compose_index_matrix(og, g, M, convention="vtk")
    for od in (og.dims, (og.dims[0] | 1, og.dims[1] | 1, og.dims[2] + (5 - og.dims[2] % 8) % 8)):
        assert od[2] % 8 != 0 or od is og.dims
        run(cuda_device, mov, A, od, "vtk")
        run(cuda_device, mov, A, od, "itk")


def test_x2_output_base_on_an_odd_element(cuda_device):
    """mia_resample_rigid_host hands the kernel output slabs that start wherever a plane starts: 2-byte alignment only."""
    shape, sp = (24, 40, 136), (1.0, 1.0, 2.0)
    mov = make_phantom(shape, "CT", seed=8, spacing=sp)
    g = grid_of(shape, sp)
    A = G.compose_index_matrix(g, g, G.params_to_matrix([1.5, -1, 2, 3, -2, 1], (0, 0, 0)))
    od = (131, 37, 21)
    n = od[0] * od[1] * od[2]
    buf = torch.full((n + 9,), 12345, dtype=torch.int16, device=cuda_device)
    out = buf[5:5 + n].view(od[2], od[1], od[0])
    run(cuda_device, mov, A, od, "vtk", out=out)
    assert bool((buf[:5] == 12345).all()) and bool((buf[5 + n:] == 12345).all())     # nothing written outside


def test_x2_int16_extremes_need_no_saturation(cuda_device):
    """The int16 -> int16 store rounds through the mantissa without a saturating conversion (x2_to_i16): an interpolant of
    taps at -32768 / 32767 must still land inside the range, on the value the float64 oracle rounds to."""
    rng = np.random.default_rng(11)
    shape, sp = (20, 40, 136), (1.0, 1.0, 1.0)
    mov = rng.choice(np.array([-32768, 32767, -32767, 32766, 0], dtype=np.int16), size=shape)
    g = grid_of(shape, sp)
    A = G.compose_index_matrix(g, g, G.params_to_matrix([1.0, -0.5, 1.5, 0.37, -0.21, 0.43], (0, 0, 0)))
    t = torch.from_numpy(mov).to(cuda_device)
    got = kernels.resample(t, A, shape, background=-3001.0, mode="vtk").cpu().numpy()
    assert last_path() == X2
    want = reslice.oracle_reslice(mov, A, g.dims, -3001.0, "vtk", np.int16)
    near = border_band(A, g.dims, g.dims)
    d = np.abs(got.astype(np.int64) - want.astype(np.int64))
    # float32 lerps of +-32768 carry 0.004 per step: a value within that of a half-integer may round the other way
    assert d[~near].max() <= 1
    assert (d[~near] == 1).mean() < 0.02


def test_x2_exact_cases_bit_exact(cuda_device):
    shape, sp = (24, 40, 128), (1.0, 1.0, 1.0)
    mov = make_phantom(shape, "CT", seed=3, spacing=sp)
    t = torch.from_numpy(mov).to(cuda_device)
    for shift in ((0, 0, 0), (3, -2, 1), (0.5, 0, 0), (0, 0.5, 0.5)):
        A = np.hstack([np.eye(3), np.array(shift, dtype=np.float64)[:, None]])
        for mode in ("vtk", "itk"):
            got = kernels.resample(t, A, shape, background=BG[mode], mode=mode).cpu().numpy()
            assert last_path() == X2
            want = reslice.oracle_reslice(mov, A, (128, 40, 24), BG[mode], mode, np.int16)
            assert np.array_equal(got, want), (shift, mode)


def test_rotation_beyond_the_x2_box_takes_the_older_kernels(cuda_device):
    shape, sp = (24, 48, 136), (1.0, 1.0, 1.0)
    mov = make_phantom(shape, "CT", seed=4, spacing=sp)
    g = grid_of(shape, sp)
    A = G.compose_index_matrix(g, g, G.params_to_matrix([0, 0, 9.0, 1, 2, 0], (0, 0, 0)))
    t = torch.from_numpy(mov).to(cuda_device)
    got = kernels.resample(t, A, shape, background=-3001.0, mode="vtk").cpu().numpy()
    assert last_path() in (1, 4)
    want = reslice.oracle_reslice(mov, A, g.dims, -3001.0, "vtk", np.int16)
    check_parity(got, want, border_band(A, g.dims, g.dims), np.int16, -3001.0)


# ---- multi-matrix SSD on the x-pair tile (csrc/mia_tile_multi_x2.cuh) ------------------------------------------------
MULTI_X2 = 6


def _pair(shape, sp, seed):
    from oracle import native  # noqa: F401  (import check: the oracle is the checker below)
    g = grid_of(shape, sp)
    Mgt = G.params_to_matrix([1.0, -0.8, 1.2, 2.0, -1.0, 1.5], (0, 0, 0))
    fixed = make_phantom(shape, "CT", seed=seed, spacing=sp)
    moving = make_phantom(shape, "CT", seed=seed, spacing=sp, world_matrix=np.linalg.inv(Mgt))
    return g, fixed, moving


def _probes(g, centre, spread=0.5):
    from medicalimageanalysis_b200.registration import probe_params
    p2w, w2p = G.index_to_physical(g), G.physical_to_index(g)
    return np.ascontiguousarray(np.stack([(w2p @ G.params_to_matrix(q, np.zeros(3)) @ p2w)[:3]
                                          for q in probe_params(np.asarray(centre, dtype=float), 1.0, np.array([spread] * 6))]))


@pytest.mark.parametrize("mode", ["itk", "vtk"])
@pytest.mark.parametrize("shape", [(40, 56, 136), (21, 45, 200)])          # whole tiles / ragged rows, odd depth
def test_multi_x2_equals_one_matrix_launches_and_oracle(cuda_device, shape, mode):
    from oracle import native
    sp = (1.0, 1.0, 2.0)
    g, fixed, moving = _pair(shape, sp, seed=21)
    A = _probes(g, [0.6, -0.5, 0.8, 2.0, -1.0, 1.5])       # within the gate's slack: the x-pair kernel takes the launch
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    for as_device in (False, True):                      # host matrices: fit-tested on the host; device: FitGate
        kw = {"d_A": torch.from_numpy(A).to(cuda_device)} if as_device else {"A": A}
        s13, c13 = kernels.resample_ssd(f, m, mode=mode, ws=kernels.MetricWorkspace(cuda_device, n_mats=13), **kw)
        assert last_path() == MULTI_X2
        s13, c13 = s13.clone(), c13.clone()
        for i in range(13):
            s1, c1 = kernels.resample_ssd(f, m, A=A[i], mode=mode)
            assert int(c1[0]) == int(c13[i]), i
            assert abs(float(s1[0]) - float(s13[i])) <= 1e-6 * float(s13[i]), i
            want, n = native.reslice_ssd(fixed, moving, A[i], mode)
            assert abs(int(c13[i]) - n) <= 2 and abs(float(s13[i]) - want) <= 1e-5 * want, i
    s2, c2 = kernels.resample_ssd(f, m, A=A, mode=mode)
    assert torch.equal(s2, s13) and torch.equal(c2, c13)          # device == host matrices, run to run identical


def test_multi_x2_gate_hands_large_rotations_to_the_32_wide_kernel(cuda_device):
    """Device matrices the host never sees: probes around a 10-degree rotation do not fit the 64-wide tile's box, the
    x-pair kernel's gate fails and the 32-wide kernel enqueued behind it serves the launch -- same numbers as host
    matrices (which the host routes to the 32-wide kernel itself) and as the oracle."""
    from oracle import native
    shape, sp = (32, 64, 136), (1.0, 1.0, 1.0)
    g, fixed, moving = _pair(shape, sp, seed=22)
    A = _probes(g, [0.5, -0.5, 10.0, 1.0, 2.0, -1.0])
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    sh, ch = kernels.resample_ssd(f, m, A=A)
    assert last_path() != MULTI_X2                        # the host fit test sends these to the older kernels
    sh, ch = sh.clone(), ch.clone()
    sd, cd = kernels.resample_ssd(f, m, d_A=torch.from_numpy(A).to(cuda_device), ws=kernels.MetricWorkspace(cuda_device, n_mats=13))
    assert torch.equal(cd, ch) and float(((sd - sh).abs() / sh).max()) <= 1e-6
    want, n = native.reslice_ssd(fixed, moving, A[0], "itk")
    assert abs(int(cd[0]) - n) <= 2 and abs(float(sd[0]) - want) <= 1e-5 * want


def _autocrop_case(shape=(60, 136, 264), sp=(0.9765625, 0.9765625, 2.5), seed=11):
    """A lattice with outside, border and interior tiles (what makes the tile costs uneven)."""
    mov = make_phantom(shape, "CT", seed=seed, spacing=sp)
    g = grid_of(shape, sp)
    M = G.params_to_matrix([3, -2, 4, 5, -3, 7], (0, 0, 0))
    origin, dims = G.autocrop_extent(g, M, sp, convention="vtk")
    A = G.compose_index_matrix(G.Grid(dims, sp, origin), g, M, convention="vtk")
    return mov, A, (dims[2], dims[1], dims[0])


def test_x2_claimed_tile_order_writes_the_bytes_of_the_static_order(cuda_device):
    """Tiles are claimed from a per-launch counter (mia_b200.cu::x2_claim_slot).  Which CTA computes a tile changes no
    bit of it, a launch leaves its counter pair at zero for the next one, launches on two streams own different pairs,
    and a captured launch keeps its pair over the replays of its graph."""
    lib = _lib.load()
    mov, A, oshape = _autocrop_case()
    t = torch.from_numpy(mov).to(cuda_device)
    prev = lib.mia_debug_x2_static(1)
    try:
        want = kernels.resample(t, A, oshape, background=BG["vtk"], mode="vtk")
        assert last_path() == X2
        assert lib.mia_debug_x2_static(0) == 1
        for _ in range(6):                                   # counters come back to zero: every launch writes every voxel
            got = torch.full(oshape, 777, dtype=torch.int16, device=cuda_device)
            kernels.resample(t, A, oshape, background=BG["vtk"], mode="vtk", out=got)
            assert last_path() == X2 and torch.equal(got, want)
        # two streams at once
        s1, s2 = torch.cuda.Stream(cuda_device), torch.cuda.Stream(cuda_device)
        outs = [torch.full(oshape, 777, dtype=torch.int16, device=cuda_device) for _ in range(8)]
        torch.cuda.synchronize()
        for i, o in enumerate(outs):
            with torch.cuda.stream(s1 if i % 2 == 0 else s2):
                kernels.resample(t, A, oshape, background=BG["vtk"], mode="vtk", out=o)
        torch.cuda.synchronize()
        assert all(torch.equal(o, want) for o in outs)
        # captured once, replayed three times, with eager launches in between
        o_g = torch.full(oshape, 777, dtype=torch.int16, device=cuda_device)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            kernels.resample(t, A, oshape, background=BG["vtk"], mode="vtk", out=o_g)
            kernels.resample(t, A, oshape, background=BG["vtk"], mode="vtk", out=outs[0])
        for _ in range(3):
            o_g.fill_(777); outs[0].fill_(777)
            graph.replay()
            kernels.resample(t, A, oshape, background=BG["vtk"], mode="vtk", out=outs[1])
            torch.cuda.synchronize()
            assert torch.equal(o_g, want) and torch.equal(outs[0], want) and torch.equal(outs[1], want)
    finally:
        lib.mia_debug_x2_static(prev)


def test_x2_fewer_chunks_than_ctas(cuda_device):
    """A lattice of a few tiles: the grid shrinks to the number of chunks, and chunks past the last tile hold skip records."""
    shape, sp = (9, 17, 72), (1.0, 1.0, 2.0)
    mov = make_phantom(shape, "CT", seed=3, spacing=sp)
    g = grid_of(shape, sp)
    A = G.compose_index_matrix(g, g, G.params_to_matrix([0.5, -0.5, 1, 2, -1, 1], (0, 0, 0)))
    for od in ((72, 17, 9), (65, 9, 8), (130, 8, 17), (64, 8, 8), (1, 1, 1)):
        run(cuda_device, mov, A, od, "vtk")
        run(cuda_device, mov, A, od, "itk")
