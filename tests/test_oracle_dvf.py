"""CPU tests of the displacement-field oracle.

The point sampler is PINNED: it must equal ``scipy.ndimage.map_coordinates(order=1, mode='nearest')`` -- the very
call the reference makes (``structure/deformable.py:873-878``) -- on random, boundary and out-of-range points.
The image warp has no runnable reference here (SimpleITK is absent); it is checked for consistency with the pinned
pieces: zero field == rigid oracle, a constant field == the translated rigid oracle, and interior samples ==
``map_coordinates`` at the displaced coordinates."""
import numpy as np
from scipy import ndimage

from medicalimageanalysis_b200 import geometry as G
from medicalimageanalysis_b200.phantom import make_phantom
from oracle import dvf as D
from oracle import reslice as R


def _smooth_field(shape, amp, seed):
    rng = np.random.default_rng(seed)
    z, y, x = np.meshgrid(*(np.arange(n, dtype=np.float64) for n in shape), indexing="ij")
    f = np.empty(shape + (3,), dtype=np.float32)
    for c in range(3):
        a, b, d = rng.uniform(0.05, 0.2, 3)
        f[..., c] = amp * np.sin(a * x + c) * np.cos(b * y - c) * np.sin(d * z + 0.5 * c)
    return f


def test_point_sampler_equals_scipy_map_coordinates():
    shape = (9, 11, 13)
    field = _smooth_field(shape, 3.0, 1)
    rng = np.random.default_rng(2)
    pts = rng.uniform(-2.0, 15.0, (4000, 3))                       # x, y, z -- inside, on and beyond every face
    pts[:50] = np.round(pts[:50])                                   # exact lattice points
    pts[50:60] = [[0, 0, 0], [12, 10, 8], [12.0, 0, 8], [-1, 5, 4], [13, 5, 4], [6, -3, 4], [6, 11, 4], [6, 5, -0.5],
                  [6, 5, 8.5], [11.999999, 9.999999, 7.999999]]
    got = D.oracle_sample_field_points(field, pts, scale=0.35)
    for c in range(3):
        want = ndimage.map_coordinates(0.35 * field[..., c].astype(np.float64), [pts[:, 2], pts[:, 1], pts[:, 0]],
                                       order=1, mode="nearest")
        assert np.allclose(got[:, c], want, rtol=0, atol=1e-12)


def _pair(shape=(20, 24, 28), sp=(1.0, 1.2, 2.0)):
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    img = make_phantom(shape, "CT", seed=11, spacing=sp)
    w2p = G.physical_to_index(g, "itk")
    A = (w2p @ G.index_to_physical(g, "itk"))[:3]
    return g, img, A, w2p[:3, :3]


def test_zero_and_constant_fields_reduce_to_the_rigid_oracle():
    g, img, A, B = _pair()
    zero = np.zeros(img.shape + (3,), dtype=np.float32)
    assert np.array_equal(D.oracle_warp_dvf(img, zero, A, B, mode="itk"), img)
    shift = np.array([1.3, -0.7, 2.5])                              # physical mm
    const = np.broadcast_to(shift.astype(np.float32), img.shape + (3,)).copy()
    As = A.copy()
    As[:, 3] += B @ shift.astype(np.float32).astype(np.float64)
    want = R.oracle_reslice(img, As, g.dims, background=-3001.0, mode="itk")
    got = D.oracle_warp_dvf(img, const, A, B, background=-3001.0, mode="itk")
    assert np.mean(got != want) < 1e-3 and np.abs(got.astype(int) - want.astype(int)).max() <= 1


def test_interior_of_the_warp_equals_map_coordinates_at_displaced_points():
    g, img, A, B = _pair()
    field = _smooth_field(img.shape, 2.0, 3)
    got = D.oracle_warp_dvf(img.astype(np.float32), field, A, B, background=-3001.0, mode="itk")
    nz, ny, nx = img.shape
    k, j, i = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    c = [A[a, 0] * i + A[a, 1] * j + A[a, 2] * k + A[a, 3] + field.astype(np.float64) @ B[a] for a in range(3)]
    want = ndimage.map_coordinates(img.astype(np.float64), [c[2], c[1], c[0]], order=1, mode="nearest")
    inner = np.ones(img.shape, bool)
    for a, n in zip(range(3), (nx, ny, nz)):
        inner &= (c[a] >= 0) & (c[a] <= n - 1)
    assert inner.mean() > 0.7
    assert np.allclose(got[inner], want[inner].astype(np.float32), rtol=1e-6, atol=1e-3)
