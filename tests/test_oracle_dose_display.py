"""CPU tests for SURVEY N2 / N3: the off-axis view lattice (product geometry against the oracle's independent
derivation and against known answers), the dose oracle, and the percentile rank arithmetic the GPU statistics use."""
import numpy as np
import pytest

from medicalimageanalysis_b200 import Data, geometry as G, kernels, _lib
from medicalimageanalysis_b200.structure import Dose
from medicalimageanalysis_b200.utils import CreateImageFromMask
from oracle import display as OD
from oracle import dose as ODose


class _DoseSrc(object):
    def __init__(self, array, origin, spacing, name="Dose 01"):
        self.array, self.origin, self.spacing, self.dose_name = array, origin, spacing, name
        self.image_matrix = np.identity(3, dtype=np.float32)
        self.dimensions = array.shape


@pytest.fixture(autouse=True)
def _clean():
    Data.clear()
    yield
    Data.clear()


def _image(shape=(20, 28, 36), sp=(1.0, 1.2, 2.0), seed=0):
    rng = np.random.default_rng(seed)
    a = rng.integers(-1000, 2000, size=shape).astype(np.int16)
    origin = (-17.0, -15.0, -20.0)
    return CreateImageFromMask(a, origin, sp, "CT %d" % seed, device="cpu").add_image(), a


@pytest.mark.parametrize("angles", [(0, 0, 30), (10, -5, 3), (0, 90, 0), (-7, 12, 170)])
def test_offaxis_lattice_matches_independent_derivation(angles):
    img, a = _image()
    img.display.matrix = G.euler_xyz_matrix(*angles) @ np.asarray(img.matrix, dtype=np.float64)
    origin, dims, A, _ = img.display.offaxis_geometry()
    o2, d2, A2 = OD.offaxis_lattice(a.shape, img.spacing, img.origin, img.matrix, img.display.matrix)
    assert dims == d2
    assert np.allclose(origin, o2, atol=1e-9)
    assert np.allclose(A, A2, atol=1e-9)


def test_offaxis_90_degrees_about_z_is_a_rot90_of_the_slices():
    """Isotropic volume, viewing rotation 90 deg about z: the reslice is an exact permutation.  The output voxel at
    physical p shows the input at inv(R) p: x_in = y_out, y_in = -x_out."""
    rng = np.random.default_rng(3)
    a = rng.integers(-1000, 2000, size=(6, 10, 14)).astype(np.int16)
    R = G.euler_xyz_matrix(0, 0, 90)
    out, lo, disp_origin = OD.oracle_offaxis_array(a, (1.0, 1.0, 1.0), (0.0, 0.0, 0.0), np.identity(3), R)
    assert out.shape == (6, 14, 10)
    assert np.array_equal(out, np.rot90(a, -1, (1, 2)))
    assert np.allclose(lo, (-9.0, 0.0, 0.0), atol=1e-9)
    assert np.allclose(disp_origin, np.linalg.inv(R) @ np.array(lo))


def test_update_rotation_zero_resets_the_view():
    img, _ = _image()
    img.display.matrix = G.euler_xyz_matrix(1, 2, 3)
    img.display.slice_location = [1, 2, 3]
    img.update_rotation(0, 0, 0)
    assert np.array_equal(img.display.matrix, np.asarray(img.matrix, dtype=np.float64))
    assert img.display.slice_location == img.compute_center(position=False, zyx=True)
    assert img.display.secondary_array is None
    assert np.allclose(img.retrieve_angles(), 0.0)


def test_dose_oracle_identity_and_statistics_keys():
    rng = np.random.default_rng(1)
    dose = rng.random((8, 10, 12)).astype(np.float32) * 70
    mask = rng.random((8, 10, 12)) > 0.7
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    s = ODose.oracle_roi_dose_array(dose, A, mask)
    assert np.array_equal(s, dose[mask])
    dvh = ODose.oracle_dose_statistics(s, (1.0, 1.0, 2.5))
    assert dvh["Volume (cc)"] == pytest.approx(mask.sum() * 2.5 / 1000.0)
    assert dvh["D50"] == pytest.approx(dvh["Dmedian"])
    assert dvh["VS0Gy_percent"] == 0.0 and dvh["VS150Gy_percent"] == 100.0
    assert len([k for k in dvh if k.startswith("VS")]) == 2 * 31


@pytest.mark.parametrize("n", [1, 2, 3, 10, 101, 1000])
def test_percentile_ranks_reproduce_numpy(n):
    rng = np.random.default_rng(n)
    x = (rng.random(n) * 80).astype(np.float32)
    qs = [100.0 - d for d in ODose.D_VALUES] + [50.0, 0.0, 100.0]
    lo, hi, t = kernels.percentile_ranks(n, qs)
    xs = np.sort(x)
    for q, a, b, w in zip(qs, lo, hi, t):
        va, vb = float(xs[a]), float(xs[b])
        got = vb - (vb - va) * (1.0 - w) if w >= 0.5 else va + (vb - va) * w
        assert got == pytest.approx(float(np.percentile(x, q)), rel=1e-6, abs=1e-6)


def test_dose_abi_argument_errors_without_gpu():
    lib = _lib.load()
    assert lib.mia_mask_workspace_bytes(_lib.i3((64, 64, 64))) >= 64 * 8 + 64 * 4
    assert lib.mia_sample_stats_workspace_bytes() > 4096 * 4
    rc = lib.mia_mask_scan(None, _lib.i3((4, 4, 4)), None, None, 0, None)
    assert rc == -1 and b"null pointer" in lib.mia_last_error()
    rc = lib.mia_sample_stats(None, 0, None, 0, None, 0, 1, 1, 1 << 20, None)
    assert rc == -1 and b"empty sample" in lib.mia_last_error()


def test_dose_class_mirrors_reference_fields():
    d = Dose(_DoseSrc(np.zeros((4, 6, 8), dtype=np.float64), (1.0, 2.0, 3.0), (2.0, 2.0, 3.0)), device="cpu")
    assert d.array.dtype.is_floating_point and tuple(d.array.shape) == (4, 6, 8)
    assert d.display.scroll_max == [3, 5, 7]
    assert np.allclose(d.compute_center(), (1.0 + 4 * 2.0, 2.0 + 3 * 2.0, 3.0 + 2 * 3.0))
