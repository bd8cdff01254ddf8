"""Parity of the CUDA resampler (through the C ABI) with the oracle.

Tolerances (BASELINE.json north_star): int16 output within +-1 LSB, float output within 1e-4 relative,
FFS permutations bit exact.  Voxels whose float64 coordinate lies within 1e-4 voxel of the +-0.5 inside/outside
border are compared as "either value is acceptable" (the kernel's local float32 coordinate may fall on the
other side of an exact tie); the tests assert that this set is tiny.
"""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import geometry as G
from medicalimageanalysis_b200 import kernels
from medicalimageanalysis_b200.phantom import make_phantom
from oracle import ffs, reslice

pytestmark = pytest.mark.gpu
BG = {"vtk": -3001.0, "itk": 0.0}


def border_band(A, odims, mdims, eps=1e-4):
    c = reslice.continuous_index(A, odims)
    near = np.zeros(c[0].shape, dtype=bool)
    for a in range(3):
        near |= (np.abs(c[a] + 0.5) < eps) | (np.abs(c[a] - (mdims[a] - 0.5)) < eps)
    return near


def check_parity(got, want, near, out_dtype, bg, max_near=0.02):
    got = np.asarray(got)
    if np.dtype(out_dtype).kind == "i":
        d = np.abs(got.astype(np.int64) - want.astype(np.int64))
        bad = d > 1
    else:
        scale = np.maximum(np.abs(want.astype(np.float64)), 1.0)
        bad = np.abs(got.astype(np.float64) - want.astype(np.float64)) > 1e-4 * scale
    bad &= ~near
    assert not bad.any(), "%d voxels out of tolerance (max |diff| %g)" % (
        bad.sum(), np.abs(got.astype(np.float64) - want.astype(np.float64))[bad].max())
    assert near.mean() < max_near


def run_case(dev, mov, A, odims, mode, out_dtype=None):
    out_dtype = np.dtype(out_dtype or mov.dtype)
    t = torch.from_numpy(mov).to(dev)
    tdt = torch.int16 if out_dtype == np.int16 else torch.float32
    got = kernels.resample(t, A, (odims[2], odims[1], odims[0]), background=BG[mode], mode=mode, out_dtype=tdt)
    want = reslice.oracle_reslice(mov, A, odims, BG[mode], mode, out_dtype)
    near = border_band(A, odims, (mov.shape[2], mov.shape[1], mov.shape[0]))
    check_parity(got.cpu().numpy(), want, near, out_dtype, BG[mode])
    return got


def test_golden_vectors(cuda_device, golden):
    for name in [str(n) for n in golden["names"]]:
        A, odims = golden[name + "/A"], tuple(int(v) for v in golden[name + "/odims"])
        for tag in ("i16", "f32"):
            mov = golden["mov_" + tag]
            for mode in ("vtk", "itk"):
                t = torch.from_numpy(mov).to(cuda_device)
                got = kernels.resample(t, A, (odims[2], odims[1], odims[0]), background=BG[mode], mode=mode)
                want = golden["%s/%s/%s" % (name, tag, mode)]
                near = border_band(A, odims, (mov.shape[2], mov.shape[1], mov.shape[0]))
                # shift_half puts three whole faces exactly on the -0.5 border (bit-exactness of those ties is
                # asserted in test_exact_cases_are_bit_exact); everywhere else the band must be tiny
                check_parity(got.cpu().numpy(), want, near, mov.dtype, BG[mode],
                             max_near=0.2 if name == "shift_half" else 0.02)


@pytest.mark.parametrize("name", ["identity", "shift_int", "shift_half"])
def test_exact_cases_are_bit_exact(cuda_device, golden, name):
    """Weights are all 0, 1 or 0.5 here, so even the float path must agree exactly -- including the +-0.5 ties."""
    A, odims = golden[name + "/A"], tuple(int(v) for v in golden[name + "/odims"])
    mov = golden["mov_i16"]
    for mode in ("vtk", "itk"):
        t = torch.from_numpy(mov).to(cuda_device)
        got = kernels.resample(t, A, (odims[2], odims[1], odims[0]), background=BG[mode], mode=mode)
        assert np.array_equal(got.cpu().numpy(), golden["%s/i16/%s" % (name, mode)]), (name, mode)


@pytest.mark.parametrize("case", list(kernels.FFS_PERMUTATIONS))
@pytest.mark.parametrize("dtype", [np.int16, np.float32])
def test_ffs_permutation_bit_exact(cuda_device, case, dtype):
    rng = np.random.default_rng(4)
    a = rng.integers(-1024, 3071, size=(37, 45, 70)).astype(dtype)
    perm, flip = kernels.FFS_PERMUTATIONS[case]
    got = kernels.permute_axes(torch.from_numpy(a).to(cuda_device), perm, flip)
    assert np.array_equal(got.cpu().numpy(), ffs.oracle_ffs(a, case))


@pytest.mark.parametrize("case", list(kernels.FFS_PERMUTATIONS))
@pytest.mark.parametrize("shape", [(36, 44, 72), (6, 10, 64), (130, 66, 200), (2, 2, 8),
                                   (40, 72, 88), (8, 8, 8), (64, 64, 64), (72, 136, 200), (16, 128, 192), (12, 20, 64), (68, 4, 136)])
def test_ffs_permutation_vector_paths_bit_exact(cuda_device, case, shape):
    """Even sizes take the 16-byte row copy / 64x64 2x2-register transpose kernels (int16 only); sizes that are
    multiples of 8 take the 16-byte swizzled transpose (full and partial 64x64 tiles)."""
    rng = np.random.default_rng(5)
    a = rng.integers(-32768, 32767, size=shape).astype(np.int16)
    perm, flip = kernels.FFS_PERMUTATIONS[case]
    got = kernels.permute_axes(torch.from_numpy(a).to(cuda_device), perm, flip)
    assert np.array_equal(got.cpu().numpy(), ffs.oracle_ffs(a, case))


@pytest.mark.parametrize("case", ["axial_k1", "axial_k2", "axial_k3"])
def test_ffs_as_a_signed_permutation_matrix_through_the_resampler(cuda_device, golden, case):
    """the FFS permutation is a 4x4 signed permutation whose trilinear weights are 0/1, so the
    general resample kernel must reproduce it bit for bit."""
    a = golden["mov_i16"]
    Z, Y, X = a.shape
    if case == "axial_k1":      # out[z,i,j] = a[z, j, X-1-i]
        A = np.array([[0, -1, 0, X - 1], [1, 0, 0, 0], [0, 0, 1, 0]], dtype=float)
        oshape = (Z, X, Y)
    elif case == "axial_k3":    # out[z,i,j] = a[z, Y-1-j, i]
        A = np.array([[0, 1, 0, 0], [-1, 0, 0, Y - 1], [0, 0, 1, 0]], dtype=float)
        oshape = (Z, X, Y)
    else:                       # out[z,i,j] = a[z, Y-1-i, X-1-j]
        A = np.array([[-1, 0, 0, X - 1], [0, -1, 0, Y - 1], [0, 0, 1, 0]], dtype=float)
        oshape = (Z, Y, X)
    assert G.is_signed_permutation(A) is not None
    got = kernels.resample(torch.from_numpy(a).to(cuda_device), A, oshape, mode="vtk")
    assert np.array_equal(got.cpu().numpy(), golden["ffs/" + case])


@pytest.mark.parametrize("seed", range(6))
def test_random_rigid_parity(cuda_device, seed):
    rng = np.random.default_rng(100 + seed)
    modality = "MR" if seed % 2 else "CT"
    shape = (int(rng.integers(9, 40)), int(rng.integers(17, 70)), int(rng.integers(33, 150)))
    sp = (float(rng.uniform(0.6, 1.2)),) * 2 + (float(rng.uniform(1.0, 3.0)),)
    mov = make_phantom(shape, modality, seed=seed, spacing=sp)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    p = np.concatenate([rng.uniform(-10, 10, 3), rng.uniform(-8, 8, 3)])
    A = G.compose_index_matrix(g, g, G.params_to_matrix(p, (0, 0, 0)))
    for mode in ("vtk", "itk"):
        run_case(cuda_device, mov, A, g.dims, mode)


def test_cross_dtype_outputs(cuda_device):
    mov = make_phantom((12, 30, 44), "CT", seed=2)
    g = G.Grid((44, 30, 12), (1, 1, 1), (0, 0, 0))
    A = G.compose_index_matrix(g, g, G.params_to_matrix([2, 3, -4, 1, 2, -1], (22, 15, 6)))
    run_case(cuda_device, mov, A, g.dims, "itk", np.float32)
    movf = make_phantom((12, 30, 44), "MR", seed=2)
    run_case(cuda_device, movf, A, g.dims, "vtk", np.int16)


@pytest.mark.parametrize("shape", [(1, 1, 1), (1, 1, 7), (3, 5, 2), (2, 9, 131), (5, 8, 128), (4, 3, 129)])
def test_ragged_and_tiny_shapes(cuda_device, shape):
    """Row lengths that are not a multiple of the 4-voxel vector, single-voxel axes, one-tile-plus-one."""
    rng = np.random.default_rng(7)
    mov = rng.integers(-500, 1500, size=shape).astype(np.int16)
    odims = (shape[2], shape[1], shape[0])
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    A[:, 3] = [0.3, -0.2, 0.1]
    for mode in ("vtk", "itk"):
        run_case(cuda_device, mov, A, odims, mode)


def test_fully_outside_is_all_background(cuda_device):
    mov = make_phantom((8, 10, 12), "CT", seed=1)
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    A[:, 3] = [1000.0, 0, 0]
    out = kernels.resample(torch.from_numpy(mov).to(cuda_device), A, (8, 10, 12), background=-3001, mode="vtk")
    assert bool((out == -3001).all())
    A[:, 3] = [1e30, -1e30, 1e12]     # far beyond int range: must not overflow into "inside"
    out = kernels.resample(torch.from_numpy(mov).to(cuda_device), A, (8, 10, 12), background=-3001, mode="vtk")
    assert bool((out == -3001).all())


def test_int16_saturation(cuda_device):
    mov = np.full((2, 2, 8), 30000.0, dtype=np.float32)
    mov[1] = 70000.0
    mov[0, 0, 0] = -70000.0
    A = np.hstack([np.identity(3), np.zeros((3, 1))])
    for mode in ("vtk", "itk"):
        run_case(cuda_device, mov, A, (8, 2, 2), mode, np.int16)


def test_error_codes(cuda_device):
    from medicalimageanalysis_b200 import _lib
    mov = torch.zeros((4, 4, 4), dtype=torch.int16, device=cuda_device)
    with pytest.raises(_lib.MiaError, match="non-finite"):
        kernels.resample(mov, np.full((3, 4), np.nan), (4, 4, 4))
    with pytest.raises(TypeError):
        kernels.resample(torch.zeros((4, 4, 4), dtype=torch.float64, device=cuda_device), np.zeros((3, 4)), (4, 4, 4))
    with pytest.raises(TypeError, match="no CPU path"):
        kernels.resample(torch.zeros((4, 4, 4), dtype=torch.int16), np.zeros((3, 4)), (4, 4, 4))


def test_host_entry_point_matches_device_path(cuda_device):
    mov = make_phantom((40, 64, 96), "CT", seed=8, spacing=(1, 1, 2.5))
    g = G.Grid((96, 64, 40), (1, 1, 2.5), (-48, -32, -50))
    A = G.compose_index_matrix(g, g, G.params_to_matrix([3, -2, 4, 5, -3, 7], (0, 0, 0)))
    dev_out = kernels.resample(torch.from_numpy(mov).to(cuda_device), A, mov.shape, mode="vtk").cpu().numpy()
    host_out = kernels.resample_host(mov, A, mov.shape, mode="vtk")
    assert np.array_equal(dev_out, host_out)
    pin_in = kernels.pinned_empty(mov.shape, np.int16)
    pin_in[...] = mov
    pin_out = kernels.pinned_empty(mov.shape, np.int16)
    kernels.resample_host(pin_in, A, mov.shape, mode="vtk", out=pin_out)
    assert np.array_equal(dev_out, pin_out)
    # z-flipping transform: the slab pipeline must still wait for the right input slabs
    Af = A.copy()
    Af[2, :] = [-A[2, 0], -A[2, 1], -A[2, 2], 39 - A[2, 3]]
    d2 = kernels.resample(torch.from_numpy(mov).to(cuda_device), Af, mov.shape, mode="vtk").cpu().numpy()
    assert np.array_equal(d2, kernels.resample_host(mov, Af, mov.shape, mode="vtk"))


def test_full_size_properties_c2(cuda_device):
    """BASELINE config 2 size (512x512x300 int16): size-independent properties instead of an oracle run."""
    gen = torch.Generator(device=cuda_device).manual_seed(5)
    vol = torch.randint(-1024, 3071, (300, 512, 512), generator=gen, device=cuda_device, dtype=torch.int16)
    eye = np.hstack([np.identity(3), np.zeros((3, 1))])
    assert torch.equal(kernels.resample(vol, eye, vol.shape, mode="vtk"), vol)          # identity
    # FFS round trip: rot90 k=1 then k=3 is the identity; checksum preserved by each
    k1 = kernels.permute_axes(vol, *kernels.FFS_PERMUTATIONS["axial_k1"])
    assert int(k1.sum(dtype=torch.int64)) == int(vol.sum(dtype=torch.int64))
    assert torch.equal(kernels.permute_axes(k1, *kernels.FFS_PERMUTATIONS["axial_k3"]), vol)
    assert torch.equal(k1, torch.rot90(vol, 1, (1, 2)))
    # integer shift: shifted copy with background fill
    A = eye.copy()
    A[:, 3] = [5, -3, 2]
    out = kernels.resample(vol, A, vol.shape, mode="vtk")
    assert torch.equal(out[0:298, 3:512, 0:507], vol[2:300, 0:509, 5:512])
    assert bool((out[298:] == -3001).all()) and bool((out[:, :3] == -3001).all())
    # a 4-degree rotation and its inverse: interior voxels return to within interpolation blur
    g = G.Grid((512, 512, 300), (1, 1, 1), (-256, -256, -150))
    smooth = torch.from_numpy(make_phantom((300, 512, 512), "CT", seed=2, noise=False)).to(cuda_device)
    M = G.params_to_matrix([0, 0, 4, 0, 0, 0], (0, 0, 0))
    fwd = kernels.resample(smooth, G.compose_index_matrix(g, g, M), smooth.shape, mode="itk")
    back = kernels.resample(fwd, G.compose_index_matrix(g, g, np.linalg.inv(M)), smooth.shape, mode="itk")
    core = (slice(100, 200), slice(180, 330), slice(180, 330))
    assert float((back[core].float() - smooth[core].float()).abs().mean()) < 12.0
