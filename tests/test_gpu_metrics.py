"""Fused resample + SSD / joint-histogram / MI against the oracle (metric values within 1e-5 relative)."""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import geometry as G
from medicalimageanalysis_b200 import kernels
from medicalimageanalysis_b200.phantom import make_phantom
from oracle import metrics, native, reslice

pytestmark = pytest.mark.gpu


def pair(shape=(24, 48, 72), seed=3, fdtype="CT", mdtype="CT"):
    sp = (1.0, 1.0, 2.0)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    Mgt = G.params_to_matrix([3, -2, 4, 3, -2, 1], (0, 0, 0))
    fixed = make_phantom(shape, fdtype, seed=seed, spacing=sp)
    moving = make_phantom(shape, mdtype, seed=seed, spacing=sp, world_matrix=np.linalg.inv(Mgt))
    return g, fixed, moving, Mgt


def mats(g, n, seed=0):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        p = np.concatenate([rng.uniform(-5, 5, 3), rng.uniform(-4, 4, 3)])
        out.append(G.compose_index_matrix(g, g, G.params_to_matrix(p, (0, 0, 0))))
    return np.stack(out)


@pytest.mark.parametrize("fd,md", [("CT", "CT"), ("CT", "MR"), ("MR", "CT"), ("MR", "MR")])
def test_ssd_matches_oracle_all_dtype_pairs(cuda_device, fd, md):
    g, fixed, moving, _ = pair(fdtype=fd, mdtype=md)
    A = mats(g, 1, seed=1)[0]
    ssd, cnt = kernels.resample_ssd(torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device),
                                    A=A, mode="itk")
    want, n = native.reslice_ssd(fixed, moving, A, "itk")
    assert abs(int(cnt[0]) - n) <= 2          # exact-tie border voxels only
    assert abs(float(ssd[0]) - want) <= 1e-5 * want


def test_ssd_batched_equals_single_and_is_deterministic(cuda_device):
    g, fixed, moving, _ = pair()
    A = mats(g, 13, seed=2)
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    ws = kernels.MetricWorkspace(cuda_device, n_mats=13)
    s, c = kernels.resample_ssd(f, m, A=A, ws=ws)
    s1, c1 = s.clone(), c.clone()
    s, c = kernels.resample_ssd(f, m, A=A, ws=ws)
    assert torch.equal(s, s1) and torch.equal(c, c1)              # bit-reproducible run to run
    dA = torch.from_numpy(A).to(cuda_device)
    s2, c2 = kernels.resample_ssd(f, m, d_A=dA, ws=kernels.MetricWorkspace(cuda_device, n_mats=13))
    assert torch.equal(s2, s1) and torch.equal(c2, c1)            # device matrices == host matrices
    for i in range(13):
        si, ci = kernels.resample_ssd(f, m, A=A[i])
        # the batched launch shares one box per tile between the matrices (mia_tile_multi.cuh); these 13 are
        # unrelated, so many tiles of many matrices miss the shared box and take the per-voxel gather, whose
        # float64-row-origin arithmetic differs from the tile arithmetic in the last float32 bits
        assert abs(float(si[0]) - float(s1[i])) <= 1e-7 * float(s1[i]) and int(ci[0]) == int(c1[i])
        want, n = native.reslice_ssd(fixed, moving, A[i], "itk")
        assert abs(float(s1[i]) - want) <= 1e-5 * want and abs(int(c1[i]) - n) <= 2


def test_ssd_with_store_matches_plain_resample(cuda_device):
    g, fixed, moving, _ = pair()
    A = mats(g, 1, seed=3)[0]
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    out = torch.empty_like(m)
    s, c = kernels.resample_ssd(f, m, A=A, mode="itk", out=out, background=0.0)
    plain = kernels.resample(m, A, m.shape, background=0.0, mode="itk")
    assert torch.equal(out, plain)
    s2, _ = kernels.resample_ssd(f, m, A=A, mode="itk")
    assert float(s[0]) == float(s2[0])


def test_identity_ssd_zero_and_full_count(cuda_device):
    g, fixed, _, _ = pair()
    f = torch.from_numpy(fixed).to(cuda_device)
    eye = np.hstack([np.identity(3), np.zeros((3, 1))])
    s, c = kernels.resample_ssd(f, f, A=eye)
    assert float(s[0]) == 0.0 and int(c[0]) == fixed.size


@pytest.mark.parametrize("bins", [16, 64, 128])
def test_joint_hist_and_mi_match_oracle(cuda_device, bins):
    g, fixed, moving, _ = pair(fdtype="CT", mdtype="MR")
    A = mats(g, 3, seed=4)
    fr, mr = (-1024.0, 3071.0), (0.0, 4095.0)
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    hist, cnt = kernels.resample_jointhist(f, m, fr, mr, bins=bins, A=A)
    mi = kernels.mi_from_hist(hist).cpu().numpy()
    hist = hist.cpu().numpy().astype(np.int64)
    for i in range(3):
        H, n = native.joint_hist(fixed, moving, A[i], bins, fr, mr)
        assert abs(int(cnt[i]) - n) <= 2
        assert int(hist[i].sum()) == int(cnt[i])
        # float32 interpolation vs float64 can move a sample across a bin edge: a handful of counts
        assert np.abs(hist[i] - H.astype(np.int64)).sum() <= 2e-4 * n + 4
        want = metrics.oracle_mi(H)
        assert abs(mi[i] - want) <= 1e-5 * abs(want)
        # MI kernel on exactly the oracle histogram: pure float64 arithmetic, tight tolerance
    Hd = torch.from_numpy(H.astype(np.int32)).to(cuda_device)
    assert abs(float(kernels.mi_from_hist(Hd)[0]) - metrics.oracle_mi(H)) <= 1e-12 * abs(want) + 1e-15


def test_hist_is_order_independent_and_batched_equals_single(cuda_device):
    g, fixed, moving, _ = pair(fdtype="CT", mdtype="MR")
    A = mats(g, 5, seed=6)
    fr, mr = (-1024.0, 3071.0), (0.0, 4095.0)
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    h, c = kernels.resample_jointhist(f, m, fr, mr, bins=64, A=A)
    h1 = h.clone()
    for i in range(5):
        hi, ci = kernels.resample_jointhist(f, m, fr, mr, bins=64, A=A[i])
        assert torch.equal(hi[0], h1[i]) and int(ci[0]) == int(c[i])


def test_z_slab_partials_sum_to_the_whole(cuda_device):
    """The sharding rule of SURVEY 8e on one GPU: slabs of the fixed lattice, each with its own shifted A,
    give integer histograms that add up bit for bit and SSD partials that add up to 1e-8 relative: a slab that
    does not start on an 8-plane tile boundary splits the thread-local float32 coordinates differently, so the
    interpolated values differ in their last bits (1000 times below the 1e-5 metric tolerance)."""
    g, fixed, moving, _ = pair(shape=(32, 40, 64), fdtype="CT", mdtype="CT")
    A = mats(g, 1, seed=8)[0]
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    fr = (-1024.0, 3071.0)
    hw, cw = kernels.resample_jointhist(f, m, fr, fr, bins=32, A=A)
    hw, cw = hw.clone(), cw.clone()
    sw, _ = kernels.resample_ssd(f, m, A=A)
    sw = float(sw[0])
    hs, cs, ss = 0, 0, 0.0
    for k0, k1 in ((0, 7), (7, 16), (16, 32)):
        As = A.copy()
        As[:, 3] += As[:, 2] * k0
        h, c = kernels.resample_jointhist(f[k0:k1].contiguous(), m, fr, fr, bins=32, A=As)
        hs = hs + h.clone().long()
        cs += int(c[0])
        s, _ = kernels.resample_ssd(f[k0:k1].contiguous(), m, A=As)
        ss += float(s[0])
    assert torch.equal(hs[0], hw[0].long()) and cs == int(cw[0])
    assert abs(ss - sw) <= 1e-8 * sw


def test_metric_error_paths(cuda_device):
    from medicalimageanalysis_b200 import _lib
    f = torch.zeros((4, 4, 8), dtype=torch.int16, device=cuda_device)
    eye = np.hstack([np.identity(3), np.zeros((3, 1))])
    with pytest.raises(_lib.MiaError, match="bins"):
        kernels.resample_jointhist(f, f, (0, 1), (0, 1), bins=129, A=eye, ws=kernels.MetricWorkspace(cuda_device, 1, 129))
    with pytest.raises(_lib.MiaError, match="range"):
        kernels.resample_jointhist(f, f, (1, 1), (0, 1), bins=16, A=eye)
    with pytest.raises(_lib.MiaError, match="n_mats"):
        kernels.resample_ssd(f, f, A=np.stack([eye] * 33), ws=kernels.MetricWorkspace(cuda_device, 32))


def test_full_size_ssd_properties_c1(cuda_device):
    """BASELINE config 1 size (512x512x100 int16): identity SSD is 0 with a full count; a pure shift has the
    analytically known count."""
    vol = torch.from_numpy(make_phantom((100, 512, 512), "CT", seed=1, spacing=(0.9765625, 0.9765625, 2.5))).to(cuda_device)
    eye = np.hstack([np.identity(3), np.zeros((3, 1))])
    s, c = kernels.resample_ssd(vol, vol, A=eye)
    assert float(s[0]) == 0.0 and int(c[0]) == vol.numel()
    A = eye.copy()
    A[:, 3] = [10, 0, 0]
    s, c = kernels.resample_ssd(vol, vol, A=A)
    assert int(c[0]) == 100 * 512 * 502          # x in [0, 501] are inside under the half-open rule
    want = float(((vol[:, :, :502].double() - vol[:, :, 10:].double()) ** 2).sum())
    assert abs(float(s[0]) - want) <= 1e-9 * want
