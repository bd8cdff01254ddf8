"""Host geometry against scipy (the reference's own Rotation calls, structure/rigid.py:756,957) and against
hand-derived known answers of the reference formulas."""
import numpy as np
import pytest
from scipy.spatial.transform import Rotation

from medicalimageanalysis_b200 import geometry as G


@pytest.mark.parametrize("angles", [(3, -2, 4), (0, 0, 90), (45, 30, -60), (-170, 80, 10), (0, 0, 0)])
def test_euler_xyz_matches_scipy(angles):
    want = Rotation.from_euler('xyz', list(angles), degrees=True).as_matrix()
    assert np.allclose(G.euler_xyz_matrix(*angles), want, atol=1e-14)


@pytest.mark.parametrize("angles", [(3, -2, 4), (20, 40, -75), (-100, 10, 33)])
def test_euler_from_matrix_ZXY_matches_scipy(angles):
    R = Rotation.from_euler('xyz', list(angles), degrees=True)
    assert np.allclose(G.euler_from_matrix_ZXY(R.as_matrix()), R.as_euler('ZXY', degrees=True), atol=1e-10)


def test_pivot_transform_keeps_the_pivot_fixed():
    c = np.array([10.0, -20.0, 5.0])
    T = G.pivot_transform(c, 7, -3, 11)
    assert np.allclose(T @ np.append(c, 1.0), np.append(c, 1.0))
    assert np.allclose(T[:3, :3], G.euler_xyz_matrix(7, -3, 11))


def test_itk_euler_orders():
    ax, ay, az = 0.1, -0.2, 0.3
    rx = Rotation.from_euler('x', ax).as_matrix()
    ry = Rotation.from_euler('y', ay).as_matrix()
    rz = Rotation.from_euler('z', az).as_matrix()
    assert np.allclose(G.itk_euler_matrix(ax, ay, az), rz @ rx @ ry)
    assert np.allclose(G.itk_euler_matrix(ax, ay, az, zyx=True), rz @ ry @ rx)


def test_index_physical_roundtrip_and_reference_formula():
    m = Rotation.from_euler('xyz', [5, 10, 15], degrees=True).as_matrix()      # rows = axis directions
    g = G.Grid((10, 12, 14), (0.5, 0.75, 2.0), (1.0, 2.0, 3.0), m)
    P = G.index_to_physical(g, "itk")
    # structure/image.py:79-83: column k = matrix[k, :] * spacing[k]
    for k in range(3):
        assert np.allclose(P[:3, k], m[k, :] * g.spacing[k])
    assert np.allclose(P[:3, 3], g.origin)
    assert np.allclose(G.physical_to_index(g, "itk") @ P, np.identity(4), atol=1e-12)
    # the VTK call site hands the same 9 numbers row-major => transpose of the ITK view
    assert np.allclose(G.index_to_physical(g, "vtk")[:3, :3], (m * np.asarray(g.spacing)[None, :]))


def test_compose_identity_same_grid_is_identity():
    g = G.Grid((10, 12, 14), (0.5, 0.75, 2.0), (1.0, 2.0, 3.0))
    A = G.compose_index_matrix(g, g, np.identity(4))
    assert np.allclose(A, np.hstack([np.identity(3), np.zeros((3, 1))]), atol=1e-12)


def test_compose_translation_in_voxels():
    g = G.Grid((10, 12, 14), (0.5, 0.75, 2.0), (1.0, 2.0, 3.0))
    A = G.compose_index_matrix(g, g, G.translation_transform(1.0, 1.5, -4.0))
    assert np.allclose(A[:, 3], [2.0, 2.0, -2.0])


def test_autocrop_identity_reproduces_the_moving_lattice():
    g = G.Grid((10, 12, 14), (0.5, 0.75, 2.0), (1.0, 2.0, 3.0))
    origin, dims = G.autocrop_extent(g, np.identity(4), g.spacing)
    assert dims == g.dims and np.allclose(origin, g.origin)


def test_autocrop_rotation_bounds_contain_all_corners():
    g = G.Grid((32, 24, 20), (1, 1, 1), (-16, -12, -10))
    M = G.params_to_matrix([0, 0, 30, 3, 0, 0], (0, 0, 0))
    origin, dims = G.autocrop_extent(g, M, (1, 1, 1))
    Minv = np.linalg.inv(M)
    for c in range(8):
        idx = np.array([31 * (c & 1), 23 * ((c >> 1) & 1), 19 * ((c >> 2) & 1), 1.0])
        p = Minv @ (G.index_to_physical(g) @ idx)
        assert np.all(p[:3] >= np.asarray(origin) - 1e-9)
        assert np.all(p[:3] <= np.asarray(origin) + (np.asarray(dims) - 1) + 0.5 + 1e-9)


def test_update_rotation_composition_rule():
    """structure/rigid.py:967-970: forward mode right-multiplies by the inverse pivot transform."""
    c = np.array([1.0, 2.0, 3.0])
    m = np.identity(4)
    t = G.pivot_transform(c, 0, 0, 10)
    m = m @ np.linalg.inv(t)
    assert np.allclose(m @ t, np.identity(4))


def test_slab_moving_zrange_covers_taps():
    g = G.Grid((32, 24, 40), (1, 1, 1), (0, 0, 0))
    A = G.compose_index_matrix(g, g, G.params_to_matrix([4, -3, 2, 1, 2, 3], (16, 12, 20)))
    z0, z1 = G.slab_moving_zrange(A, g.dims, 10, 20, 40)
    i, j, k = np.meshgrid([0, 31], [0, 23], np.arange(10, 20), indexing="ij")
    cz = A[2, 0] * i + A[2, 1] * j + A[2, 2] * k + A[2, 3]
    assert z0 <= max(0, int(np.floor(cz.min()))) and z1 >= min(40, int(np.floor(cz.max())) + 2)


def test_is_signed_permutation():
    A = np.array([[0, -1, 0, 23], [1, 0, 0, 0], [0, 0, 1, 0]], dtype=float)
    assert G.is_signed_permutation(A) == ((1, 0, 2), (True, False, False))
    A[0, 0] = 0.01
    assert G.is_signed_permutation(A) is None
