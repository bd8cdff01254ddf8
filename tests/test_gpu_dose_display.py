"""SURVEY N2 / N3 on the GPU: the off-axis view reslice of Image / Dose and the masked dose sample with its DVH
statistics, against the oracle and against the reference's own NumPy reductions."""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import Data, geometry as G, kernels
from medicalimageanalysis_b200.phantom import make_phantom
from medicalimageanalysis_b200.structure import Dose
from medicalimageanalysis_b200.utils import CreateImageFromMask
from oracle import display as OD
from oracle import dose as ODose
from tests.test_oracle_dose_display import _DoseSrc

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _clean():
    Data.clear()
    yield
    Data.clear()


def _smooth_dose(shape, peak=70.0):
    z, y, x = np.meshgrid(*(np.linspace(-1, 1, n) for n in shape), indexing="ij")
    return (peak * np.exp(-2.5 * (x * x + 1.4 * y * y + 0.8 * z * z))).astype(np.float32)


@pytest.mark.parametrize("angles", [(0, 0, 25), (8, -6, 4), (0, 90, 0)])
def test_image_update_rotation_matches_oracle(cuda_device, angles):
    shape, sp = (40, 56, 72), (0.9765625, 0.9765625, 2.5)
    a = make_phantom(shape, "CT", seed=7, spacing=sp)
    img = CreateImageFromMask(a, (-35.0, -27.0, -50.0), sp, "CT 07").add_image()
    img.display.slice_location = [12, 30, 40]
    img.update_rotation(*angles)
    want, lo, disp_origin = OD.oracle_offaxis_array(a, sp, img.origin, img.matrix, img.display.matrix)
    got = img.display.secondary_array.cpu().numpy()
    assert got.shape == want.shape and got.dtype == np.int16
    d = np.abs(got.astype(np.int32) - want.astype(np.int32))
    edge = (got == -3001) != (want == -3001)
    assert edge.mean() < 1e-4
    assert d[~edge].max() <= 1                                    # +-1 LSB (north star)
    assert (want == -3001).mean() > 0.01 or angles == (0, 90, 0)
    assert np.allclose(img.display.origin, disp_origin, atol=1e-9)
    assert img.display.scroll_max == [s - 1 for s in want.shape]
    assert all(0 <= v <= m for v, m in zip(img.display.slice_location, img.display.scroll_max))
    assert np.allclose(img.retrieve_angles(), G.euler_from_matrix_ZXY(G.euler_xyz_matrix(*angles)))
    sl = img.retrieve_array_plane('Coronal')
    assert sl.dtype == np.float32
    assert np.array_equal(sl, got[:, img.display.slice_location[1], :].astype(np.float32))
    # the slice position survives the reslice to within one voxel of the new lattice
    img.reset_array()
    assert img.display.secondary_array is None


def test_offaxis_90_degrees_is_bit_exact_permutation(cuda_device):
    rng = np.random.default_rng(3)
    a = rng.integers(-1000, 2000, size=(16, 40, 64)).astype(np.int16)
    img = CreateImageFromMask(a, (0.0, 0.0, 0.0), (1.0, 1.0, 1.0), "CT iso").add_image()
    img.update_rotation(0, 0, 90)
    assert np.array_equal(img.display.secondary_array.cpu().numpy(), np.rot90(a, -1, (1, 2)))
    img.update_rotation(0, 0, 90, base=False)                     # composes on the current view: 180 degrees
    assert np.array_equal(img.display.secondary_array.cpu().numpy(), a[:, ::-1, ::-1])


def test_dose_offaxis_float32(cuda_device):
    shape, sp = (24, 40, 48), (2.0, 2.0, 3.0)
    dose = _smooth_dose(shape)
    d = Dose(_DoseSrc(dose, (-48.0, -40.0, -36.0), sp))
    d.update_rotation(5, 10, -15)
    want, _, disp_origin = OD.oracle_offaxis_array(dose, sp, d.origin, d.matrix, d.display.matrix)
    got = d.display.secondary_array.cpu().numpy()
    assert got.dtype == np.float32 and got.shape == want.shape
    edge = (got == -3001) != (want == -3001)
    assert edge.mean() < 1e-4
    ok = ~edge & (want != -3001)
    assert np.abs(got[ok] - want[ok]).max() <= 1e-4 * np.abs(want[ok]).max()     # 1e-4 relative (north star)
    assert np.allclose(d.display.origin, disp_origin, atol=1e-9)


def _dose_case(seed=0):
    ishape, isp = (40, 64, 80), (1.0, 1.0, 2.5)
    img = CreateImageFromMask(make_phantom(ishape, "CT", seed=seed, spacing=isp), (-40.0, -32.0, -50.0), isp,
                              "CT dose").add_image()
    dshape, dsp = (30, 36, 44), (2.5, 2.5, 3.0)
    dose = _smooth_dose(dshape)
    d = Dose(_DoseSrc(dose, (-52.0, -40.0, -40.0), dsp))             # partly outside the image: default 0.0 appears
    z, y, x = np.meshgrid(*(np.arange(n) for n in ishape), indexing="ij")
    mask = (((x - 40) / 30.0) ** 2 + ((y - 30) / 22.0) ** 2 + ((z - 20) / 16.0) ** 2 < 1.0).astype(np.uint8)
    mask[0, :4, :] = 1                                                # a piece on the lattice edge
    return img, d, dose, mask


def test_roi_dose_array_matches_oracle_in_np_where_order(cuda_device):
    img, d, dose, mask = _dose_case()
    A = G.compose_index_matrix(img.grid, d.grid, np.identity(4), convention="itk")
    want = ODose.oracle_roi_dose_array(dose, A, mask)
    got = d.compute_roi_dose_array(img, mask).cpu().numpy()
    assert got.shape == want.shape and got.dtype == np.float32
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max()
    # the same through the registry, with an int64 mask of other positive values and an ROI-like holder
    class _Roi(object):
        pass
    roi = _Roi()
    roi.mask = mask.astype(np.int64) * 7
    img.rois["PTV"] = roi
    got2 = d.compute_roi_dose_array("CT dose", "PTV").cpu().numpy()
    assert np.array_equal(got, got2)


@pytest.mark.parametrize("shape", [(1, 1, 1), (3, 5, 7), (9, 33, 130)])
def test_masked_sample_edge_shapes(cuda_device, shape):
    rng = np.random.default_rng(sum(shape))
    vol = (rng.random(shape) * 50).astype(np.float32)
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    for m in (np.ones(shape, np.uint8), np.zeros(shape, np.uint8), (rng.random(shape) > 0.5).astype(np.uint8)):
        got = kernels.resample_masked(torch.from_numpy(vol).to(cuda_device), A, m).cpu().numpy()
        assert np.array_equal(got, vol[m > 0])                        # identity: weights 0 / 1, bit exact


def test_dose_statistics_match_the_reference_numpy_reductions(cuda_device):
    img, d, dose, mask = _dose_case()
    dvh = d.compute_roi_dose_statistics(img, mask)
    sample = d.compute_roi_dose_array(img, mask).cpu().numpy()
    want = ODose.oracle_dose_statistics(sample, img.spacing)          # the reference's NumPy calls on OUR sample
    assert set(dvh) == set(want)
    for k, v in want.items():
        if k == "ROI":
            continue
        if k in ("Dmin", "Dmax") or k.endswith("_cc") or k.endswith("_percent") or k == "Volume (cc)":
            assert dvh[k] == pytest.approx(float(v), rel=1e-12, abs=1e-12), k
        else:                                                         # float32 NumPy reductions vs float64 here
            assert dvh[k] == pytest.approx(float(v), rel=2e-6, abs=1e-6), k
    # and against the oracle's own resample (float64 interpolation): 1e-4 relative
    A = G.compose_index_matrix(img.grid, d.grid, np.identity(4), convention="itk")
    full = ODose.oracle_dose_statistics(ODose.oracle_roi_dose_array(dose, A, mask), img.spacing)
    for k in ("Dmean", "Dmedian", "D2", "D50", "D98", "Dmax"):
        assert dvh[k] == pytest.approx(float(full[k]), rel=1e-4, abs=1e-4), k


@pytest.mark.parametrize("n", [1, 2, 31, 4097, 1 << 20])
def test_sample_statistics_exact_order_statistics(cuda_device, n):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) * 20 + 30).astype(np.float32)
    x[rng.integers(0, n, size=max(1, n // 7))] = 0.0                  # ties, as the default pixel produces
    if n > 2:
        x[1] = -0.0
    qs = [0, 1, 2.5, 25, 50, 75, 99, 100]
    levels = [-10.0, 0.0, 5.0, 30.0, 1000.0]
    st = kernels.sample_statistics(torch.from_numpy(x).to(cuda_device), percentiles=qs, levels=levels)
    assert st["count"] == n
    assert st["min"] == float(x.min()) and st["max"] == float(x.max())
    assert st["mean"] == pytest.approx(float(x.astype(np.float64).mean()), rel=1e-12)
    assert st["std"] == pytest.approx(float(x.astype(np.float64).std()), rel=1e-10, abs=1e-12)
    assert st["below"] == [float((x < l).sum()) for l in levels]
    want = np.percentile(x.astype(np.float64), qs)
    assert np.allclose(st["percentiles"], want, rtol=1e-12, atol=1e-12)
