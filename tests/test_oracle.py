"""Pins the oracle: committed golden vectors, NumPy vs C/OpenMP restatement, scipy interior cross-check,
and known answers derivable from reference code alone (SURVEY 4 i-iii)."""
import numpy as np
import pytest
from scipy.ndimage import map_coordinates

from medicalimageanalysis_b200 import geometry as G
from medicalimageanalysis_b200.phantom import make_phantom
from oracle import ffs, metrics, native, reslice

BG = {"vtk": -3001.0, "itk": 0.0}


def test_golden_vectors_numpy_and_c(golden):
    names = [str(n) for n in golden["names"]]
    assert len(names) >= 10
    for name in names:
        A, odims = golden[name + "/A"], tuple(int(v) for v in golden[name + "/odims"])
        for tag in ("i16", "f32"):
            mov = golden["mov_" + tag]
            for mode in ("vtk", "itk"):
                want = golden["%s/%s/%s" % (name, tag, mode)]
                got_np = reslice.oracle_reslice(mov, A, odims, BG[mode], mode)
                got_c = native.reslice(mov, A, odims, BG[mode], mode)
                assert np.array_equal(got_np, want), (name, tag, mode)
                assert np.array_equal(got_c, want), (name, tag, mode)


def test_golden_ffs_are_the_reference_numpy_expressions(golden):
    a = golden["mov_i16"]
    assert np.array_equal(golden["ffs/axial_k1"], np.rot90(a, 1, (1, 2)))
    assert np.array_equal(golden["ffs/axial_k3"], np.rot90(a, 3, (1, 2)))
    assert np.array_equal(golden["ffs/axial_k2"], np.rot90(a, 2, (1, 2)))
    assert np.array_equal(golden["ffs/coronal"], np.rot90(a, 1, (0, 1)))
    assert np.array_equal(golden["ffs/sagittal"], np.flip(np.rot90(a, 1, (0, 1)).transpose(0, 2, 1), axis=2))


def test_ffs_index_maps_of_survey():
    """SURVEY 4(iii) index maps, and the (perm, flip) pairs the product uses."""
    from medicalimageanalysis_b200.kernels import FFS_PERMUTATIONS
    a = np.arange(4 * 5 * 6, dtype=np.int16).reshape(4, 5, 6)
    Z, Y, X = a.shape
    r = ffs.oracle_ffs(a, "axial_k1")
    assert r[2, 1, 3] == a[2, 3, X - 1 - 1]
    r = ffs.oracle_ffs(a, "axial_k3")
    assert r[2, 1, 3] == a[2, Y - 1 - 3, 1]
    r = ffs.oracle_ffs(a, "sagittal")
    assert r[1, 2, 3] == a[Z - 1 - 3, Y - 1 - 1, 2]
    for case, (perm, flip) in FFS_PERMUTATIONS.items():
        want = ffs.oracle_ffs(a, case)
        got = np.transpose(a, perm)
        for ax in range(3):
            if flip[ax]:
                got = np.flip(got, axis=ax)
        assert np.array_equal(got, want), case


def test_ffs_corner_rule():
    # corner 0 already minimal -> no permutation
    idx, _ = ffs.ffs_corner_index([0, 0, 0], [1, 0, 0, 0, 1, 0], [1, 1, 1], (4, 5, 6))
    assert idx == 0 and ffs.axial_case_for_corner(idx) is None
    # row direction reversed: the far end of the row has the smaller x
    idx, corners = ffs.ffs_corner_index([10, 0, 0], [-1, 0, 0, 0, 1, 0], [1, 1, 1], (4, 5, 6))
    assert idx == 5 or idx == 1       # slice direction is (0,0,-1) here, so corner 5 wins
    assert ffs.axial_case_for_corner(1) == "axial_k1" and ffs.axial_case_for_corner(2) == "axial_k3"
    assert ffs.axial_case_for_corner(5) == "axial_k2"


@pytest.mark.parametrize("mode", ["vtk", "itk"])
def test_identity_and_integer_shift_known_answers(mode):
    mov = make_phantom((12, 14, 16), "CT", seed=3)
    eye = np.hstack([np.identity(3), np.zeros((3, 1))])
    assert np.array_equal(reslice.oracle_reslice(mov, eye, (16, 14, 12), BG[mode], mode), mov)
    A = eye.copy()
    A[:, 3] = [2, -1, 3]
    out = reslice.oracle_reslice(mov, A, (16, 14, 12), BG[mode], mode)
    want = np.full_like(mov, int(BG[mode]))
    want[0:9, 1:14, 0:14] = mov[3:12, 0:13, 2:16]
    assert np.array_equal(out, want)


def test_half_voxel_border_rule_differs_between_modes():
    mov = (np.arange(8, dtype=np.int16) * 10 + 100).reshape(1, 1, 8)
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    A[0, 3] = 0.5           # output i samples moving x = i + 0.5; i = 7 lands exactly on N - 0.5
    v = reslice.oracle_reslice(mov, A, (8, 1, 1), -3001.0, "vtk")
    t = reslice.oracle_reslice(mov, A, (8, 1, 1), 0.0, "itk")
    assert v[0, 0, 7] == 170        # closed interval: clamped edge value
    assert t[0, 0, 7] == 0          # half-open: default pixel
    assert v[0, 0, 0] == 105 and t[0, 0, 0] == 105
    A[0, 3] = -0.5          # i = 0 lands exactly on -0.5: inside in both, clamped to voxel 0
    assert reslice.oracle_reslice(mov, A, (8, 1, 1), -3001.0, "vtk")[0, 0, 0] == 100
    assert reslice.oracle_reslice(mov, A, (8, 1, 1), 0.0, "itk")[0, 0, 0] == 100


def test_rounding_rules():
    mov = np.array([-3, -2], dtype=np.int16).reshape(1, 1, 2)
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    A[0, 3] = 0.5           # value -2.5 at i = 0
    assert reslice.oracle_reslice(mov, A, (1, 1, 1), 0, "vtk")[0, 0, 0] == -2    # floor(-2.5 + 0.5)
    assert reslice.oracle_reslice(mov, A, (1, 1, 1), 0, "itk")[0, 0, 0] == -2    # trunc(-2.5)
    A[0, 3] = 0.25          # value -2.75
    assert reslice.oracle_reslice(mov, A, (1, 1, 1), 0, "vtk")[0, 0, 0] == -3
    assert reslice.oracle_reslice(mov, A, (1, 1, 1), 0, "itk")[0, 0, 0] == -2


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_rigid_numpy_vs_c_vs_scipy(seed):
    rng = np.random.default_rng(seed)
    mov = make_phantom((18, 22, 26), "MR" if seed == 1 else "CT", seed=seed, spacing=(1, 1, 2))
    g = G.Grid((26, 22, 18), (1, 1, 2), (-13, -11, -18))
    p = np.concatenate([rng.uniform(-8, 8, 3), rng.uniform(-4, 4, 3)])
    A = G.compose_index_matrix(g, g, G.params_to_matrix(p, (0, 0, 0)))
    for mode in ("vtk", "itk"):
        o, vals, mask = reslice.oracle_reslice(mov, A, g.dims, BG[mode], mode, return_float=True)
        assert np.array_equal(o, native.reslice(mov, A, g.dims, BG[mode], mode))
        c = reslice.continuous_index(A, g.dims)
        sc = map_coordinates(mov.astype(np.float64), [c[2], c[1], c[0]], order=1, mode="nearest")
        interior = mask.copy()
        for a, n in zip(c, g.dims):
            interior &= (a >= 0) & (a <= n - 1)
        assert interior.mean() > 0.3
        assert np.abs(sc - vals)[interior].max() < 1e-9


def test_metrics_numpy_vs_c():
    fixed = make_phantom((14, 20, 24), "CT", seed=5)
    mov = make_phantom((14, 20, 24), "MR", seed=5)
    g = G.Grid((24, 20, 14), (1, 1, 1), (0, 0, 0))
    A = G.compose_index_matrix(g, g, G.params_to_matrix([2, 1, -3, 1, -1, 0.5], (12, 10, 7)))
    s, n = metrics.oracle_ssd(fixed, mov, A)
    sc, nc = native.reslice_ssd(fixed, mov, A)
    assert n == nc and abs(s - sc) <= 1e-12 * abs(s)
    H, hn = metrics.oracle_joint_hist(fixed, mov, A, 32, (-1024, 3071), (0, 4095))
    Hc, hnc = native.joint_hist(fixed, mov, A, 32, (-1024, 3071), (0, 4095))
    assert hn == hnc == n and np.array_equal(H, Hc)
    assert int(H.sum()) == n


def test_mi_known_answers():
    assert metrics.oracle_mi(np.zeros((4, 4))) == 0.0
    H = np.diag([10, 10, 10, 10])
    assert abs(metrics.oracle_mi(H) - np.log(4)) < 1e-12          # perfectly dependent, uniform
    H = np.full((4, 4), 5)
    assert abs(metrics.oracle_mi(H)) < 1e-12                      # independent


def test_identity_ssd_is_zero():
    v = make_phantom((10, 12, 14), "CT", seed=9)
    eye = np.hstack([np.identity(3), np.zeros((3, 1))])
    s, n = metrics.oracle_ssd(v, v, eye)
    assert s == 0.0 and n == v.size
