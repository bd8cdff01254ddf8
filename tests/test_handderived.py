"""Known answers derived by hand from the reference's call recipe (``structure/rigid.py:685-693``: linear, AutoCropOutputOn,
output spacing = fixed spacing, identity direction, background -3001; ``structure/deformable.py:579``: sitk.Resample onto
the reference grid) and the published VTK / ITK rules -- NOT from the oracle's code.  The oracle, the
Python geometry, the C helper ``mia_autocrop_extent`` and (test_gpu_handderived below, GPU) the kernels must all land on
them.  This narrows what "parity unpinned" leaves open: output lattice of right-angle rotations and translations,
exact values on 0 / 1/2 / 1 weights, both sides of the +-0.5 border tie, and both integer rounding rules."""
import numpy as np
import pytest

from medicalimageanalysis_b200 import geometry as G
from oracle import reslice

S = (0.9, 0.9, 2.5)
O = (-17.1, 4.5, -30.0)
DIMS = (12, 9, 5)                                      # x, y, z


def volume(seed=0):
    return np.random.default_rng(seed).integers(-1000, 3000, size=(DIMS[2], DIMS[1], DIMS[0])).astype(np.int16)


def autocrop_all(grid, M, spacing):
    """The lattice from the Python geometry and from the C helper behind the ABI (loads on a CPU box)."""
    from medicalimageanalysis_b200 import kernels
    o1, d1 = G.autocrop_extent(grid, M, spacing, convention="vtk")
    o2, d2 = kernels.autocrop_extent(M, G.index_to_physical(grid, "vtk"), grid.dims, spacing)
    assert tuple(d1) == tuple(d2) and np.allclose(o1, o2, atol=1e-9)
    return np.asarray(o1), tuple(d1)


def rz(deg):
    c, s = np.cos(np.radians(deg)), np.sin(np.radians(deg))
    M = np.identity(4)
    M[:2, :2] = [[c, -s], [s, c]]
    return M


def test_identity_and_translations_keep_the_volume():
    """AutoCrop follows the volume: for M = T(t) the output bounds are the moving bounds shifted by -t, so origin =
    o - t, dims = n and every voxel is copied -- whatever t is, integer or not, because the lattice moves with it."""
    g = G.Grid(DIMS, S, O)
    mov = volume()
    for t in ((0, 0, 0), (3 * S[0], -2 * S[1], S[2]), (0.45, 0.0, -1.25), (0.123, 7.7, 0.01)):
        M = np.identity(4)
        M[:3, 3] = t
        origin, dims = autocrop_all(g, M, S)
        assert dims == DIMS and np.allclose(origin, np.asarray(O) - np.asarray(t), atol=1e-12)
        A = G.compose_index_matrix(G.Grid(dims, S, tuple(origin)), g, M, convention="vtk")
        assert np.allclose(A, np.hstack([np.identity(3), np.zeros((3, 1))]), atol=1e-9)
        assert np.array_equal(reslice.oracle_reslice(mov, A, dims, -3001, "vtk", np.int16), mov)


@pytest.mark.parametrize("deg", [90, 180, 270])
def test_right_angle_rotations_about_z(deg):
    """M = Rz(deg) maps output physical -> moving physical.  The moving box [o, o + (n-1)s] pushed through M^-1 swaps /
    negates the in-plane axes, so (isotropic in-plane spacing) the output lattice and every value follow by hand:
    90: origin (oy, -ox-(nx-1)s), dims (ny, nx), plane = rot90(plane, 1); 180: origin (-ox-(nx-1)s, -oy-(ny-1)s),
    plane = rot90(plane, 2); 270: origin (-oy-(ny-1)s, ox), plane = rot90(plane, 3)."""
    g = G.Grid(DIMS, S, O)
    mov = volume(1)
    ex, ey = (DIMS[0] - 1) * S[0], (DIMS[1] - 1) * S[1]
    want_origin = {90: (O[1], -O[0] - ex, O[2]), 180: (-O[0] - ex, -O[1] - ey, O[2]), 270: (-O[1] - ey, O[0], O[2])}[deg]
    want_dims = (DIMS[1], DIMS[0], DIMS[2]) if deg in (90, 270) else DIMS
    M = rz(deg)
    origin, dims = autocrop_all(g, M, S)
    assert dims == want_dims and np.allclose(origin, want_origin, atol=1e-9)
    A = G.compose_index_matrix(G.Grid(dims, S, tuple(origin)), g, M, convention="vtk")
    got = reslice.oracle_reslice(mov, A, dims, -3001, "vtk", np.int16)
    assert np.array_equal(got, np.rot90(mov, deg // 90, axes=(1, 2)))


def test_autocrop_extent_rounding_rule():
    """A 10 mm box on a 0.75 mm output lattice: last index = Round(10 / 0.75) = Round(13.33) = 13 -> 14 voxels (VTK's
    rule as SURVEY R1 states it; ceil would give 15), and 10 / 0.8 = 12.5 rounds half up to 13 -> 14."""
    g = G.Grid((11, 11, 11), (1.0, 1.0, 1.0), (0.0, 0.0, 0.0))
    _, d = autocrop_all(g, np.identity(4), (0.75, 0.8, 1.0))
    assert d == (14, 14, 11)
    assert G.autocrop_extent(g, np.identity(4), (0.75, 0.8, 1.0), extent_rule="ceil")[1] == (15, 14, 11)


def half_voxel_case():
    """Moving row 0..7 sampled at c = i + 0.5 on the SAME lattice (the sitk.Resample situation: no auto-crop)."""
    mov = np.array([3, 4, -3, -4, 10, 11, -7, 100], dtype=np.int16).reshape(1, 1, 8)
    A = np.hstack([np.identity(3), np.array([[0.5], [0.0], [0.0]])])
    mids = [(3 + 4) / 2, (4 - 3) / 2, (-3 - 4) / 2, (-4 + 10) / 2, (10 + 11) / 2, (11 - 7) / 2, (-7 + 100) / 2]   # 3.5 .5 -3.5 3 10.5 2 46.5
    return mov, A, mids


def test_half_voxel_shift_rounding_and_border_tie():
    """Weights exactly 1/2.  VTK: int out = floor(v + 0.5) (half up: 3.5 -> 4, -3.5 -> -3, 0.5 -> 1) and the last sample,
    c = N - 0.5, is INSIDE (closed interval) and clamps to the edge voxel.  ITK: int out = trunc toward zero
    (3.5 -> 3, -3.5 -> -3, 0.5 -> 0) and c = N - 0.5 is OUTSIDE (half-open) -> default pixel."""
    mov, A, mids = half_voxel_case()
    vtk = reslice.oracle_reslice(mov, A, (8, 1, 1), -3001, "vtk", np.int16).ravel()
    itk = reslice.oracle_reslice(mov, A, (8, 1, 1), 0, "itk", np.int16).ravel()
    assert list(vtk) == [int(np.floor(v + 0.5)) for v in mids] + [100]
    assert list(itk) == [int(np.trunc(v)) for v in mids] + [0]
    # the low side: c = -0.5 is inside for both rules (closed on the left) and clamps to voxel 0
    A2 = np.hstack([np.identity(3), np.array([[-0.5], [0.0], [0.0]])])
    for mode in ("vtk", "itk"):
        assert reslice.oracle_reslice(mov, A2, (8, 1, 1), -3001, mode, np.int16).ravel()[0] == 3
    # one ulp beyond either border is outside for both
    A3 = np.hstack([np.identity(3), np.array([[-0.5 - 1e-9], [0.0], [0.0]])])
    for mode in ("vtk", "itk"):
        assert reslice.oracle_reslice(mov, A3, (8, 1, 1), -3001, mode, np.int16).ravel()[0] == -3001
