"""Sharded registration on the GPU.  In-process: world size 1 must reproduce ``register_rigid`` exactly
(no graph vs graph, same kernels).  The multi-rank check is ``scripts/sharded_check.py`` under torchrun
(``gpurun --gpus 2``); its CPU twin with gloo is tests/test_parallel_host.py."""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import Data, RegistrationConfig, geometry as G, parallel as P
from medicalimageanalysis_b200.phantom import make_phantom
from medicalimageanalysis_b200.utils import CreateImageFromMask
from medicalimageanalysis_b200.utils import rigid as R

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("metric", ["ssd", "mi"])
def test_world1_sharded_equals_plain(cuda_device, metric):
    shape, sp = (24, 48, 64), (1.5, 1.5, 2.0)
    Data.clear()
    fixed = make_phantom(shape, "CT", seed=3, spacing=sp).astype(np.float32)
    f = CreateImageFromMask(fixed, (-48.0, -36.0, -24.0), sp, "fixed", device=cuda_device).add_image()
    M = G.params_to_matrix([2.0, -1.0, 1.5, 2.0, 1.0, -1.0], f.compute_center())
    moving = make_phantom(shape, "CT" if metric == "ssd" else "MR", seed=3, spacing=sp,
                          world_matrix=np.linalg.inv(M)).astype(np.float32)
    m = CreateImageFromMask(moving, (-48.0, -36.0, -24.0), sp, "moving", device=cuda_device).add_image()
    cfg = RegistrationConfig(metric=metric, iterations=12, bins=32, graph_iters=0)
    a = R.register_rigid(f, m, cfg=cfg)
    b = P.register_rigid_sharded(f, m, cfg=cfg)
    assert a.iterations == b.iterations == 12
    assert np.array_equal(a.scale_history, b.scale_history)
    assert np.array_equal(a.cost_history, b.cost_history)
    assert np.array_equal(a.params, b.params)
    Data.clear()


@pytest.mark.parametrize("metric", ["ssd", "mi"])
def test_world1_probe_parallel_equals_plain(cuda_device, metric):
    shape, sp = (24, 48, 64), (1.5, 1.5, 2.0)
    Data.clear()
    fixed = make_phantom(shape, "CT", seed=3, spacing=sp).astype(np.float32)
    f = CreateImageFromMask(fixed, (-48.0, -36.0, -24.0), sp, "fixed", device=cuda_device).add_image()
    M = G.params_to_matrix([2.0, -1.0, 1.5, 2.0, 1.0, -1.0], f.compute_center())
    moving = make_phantom(shape, "CT" if metric == "ssd" else "MR", seed=3, spacing=sp,
                          world_matrix=np.linalg.inv(M)).astype(np.float32)
    m = CreateImageFromMask(moving, (-48.0, -36.0, -24.0), sp, "moving", device=cuda_device).add_image()
    cfg = RegistrationConfig(metric=metric, iterations=12, bins=32, graph_iters=0)
    a = R.register_rigid(f, m, cfg=cfg)
    b = P.register_rigid_probe_parallel(f, m, cfg=cfg)
    assert a.iterations == b.iterations == 12 and b.extra["probes"] == (0, 13)
    assert np.array_equal(a.scale_history, b.scale_history)
    assert np.array_equal(a.cost_history, b.cost_history)
    assert np.array_equal(a.params, b.params)
    Data.clear()


def test_resample_slabs_concatenate_to_the_whole(cuda_device):
    from medicalimageanalysis_b200 import kernels
    shape, sp = (40, 48, 64), (1.0, 1.0, 2.0)
    mov = torch.from_numpy(make_phantom(shape, "CT", seed=5, spacing=sp)).to(cuda_device)
    g = G.Grid((64, 48, 40), sp, (-32.0, -24.0, -40.0))
    A = G.compose_index_matrix(g, g, G.params_to_matrix([3, -2, 4, 2, 1, -1], (0, 0, 0)))
    whole = kernels.resample(mov, A, shape, background=-3001.0, mode="vtk")
    for world in (2, 3, 8):
        parts = [P.resample_slab(mov, A, shape, r, world)[0] for r in range(world)]
        cat = torch.cat([p for p in parts if p is not None], dim=0)
        # tile-aligned slabs run the same tiles with the same thread bases: only the float64 matrix shift differs
        d = (cat.int() - whole.int()).abs()
        assert int(d.max()) <= 1 and float((d > 0).float().mean()) < 1e-4


@pytest.mark.parametrize("metric", ["ssd", "mi"])
def test_world1_sharded_graph_replay_equals_eager(cuda_device, metric):
    """The sharded loops captured in a CUDA graph (pack kernel, device step with the slab shift; the collective is a
    no-op at world size 1) walk the trajectory of the eager loop bit for bit."""
    shape, sp = (24, 48, 64), (1.5, 1.5, 2.0)
    Data.clear()
    fixed = make_phantom(shape, "CT", seed=3, spacing=sp).astype(np.float32)
    f = CreateImageFromMask(fixed, (-48.0, -36.0, -24.0), sp, "fixed", device=cuda_device).add_image()
    M = G.params_to_matrix([2.0, -1.0, 1.5, 2.0, 1.0, -1.0], f.compute_center())
    moving = make_phantom(shape, "CT" if metric == "ssd" else "MR", seed=3, spacing=sp,
                          world_matrix=np.linalg.inv(M)).astype(np.float32)
    m = CreateImageFromMask(moving, (-48.0, -36.0, -24.0), sp, "moving", device=cuda_device).add_image()
    a = R.register_rigid(f, m, cfg=RegistrationConfig(metric=metric, iterations=13, bins=32, graph_iters=0))
    for fn in (P.register_rigid_sharded, P.register_rigid_probe_parallel):
        b = fn(f, m, cfg=RegistrationConfig(metric=metric, iterations=13, bins=32, graph_iters=4))
        assert b.iterations == 13
        assert np.array_equal(a.cost_history, b.cost_history) and np.array_equal(a.params, b.params)
    Data.clear()


def test_slab_with_halo_only_is_bit_identical_to_the_slab_of_the_whole(cuda_device):
    """Each simulated rank keeps ONLY the moving planes slab_plan names (bounding box + one plane of halo) and must
    produce the bytes the same rank produces from the whole volume."""
    from medicalimageanalysis_b200 import kernels
    shape, sp = (44, 72, 136), (0.9765625, 0.9765625, 2.5)
    mov_h = make_phantom(shape, "CT", seed=5, spacing=sp)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    M = G.params_to_matrix([3, -2, 4, 5, -3, 7], (0, 0, 0))
    origin, dims = G.autocrop_extent(g, M, sp, convention="vtk")
    A = G.compose_index_matrix(G.Grid(dims, sp, origin), g, M, convention="vtk")
    oshape = (dims[2], dims[1], dims[0])
    mov = torch.from_numpy(mov_h).to(cuda_device)
    for world in (2, 4, 8):
        held = 0
        for r in range(world):
            (k0, k1), (z0, z1), _ = P.slab_plan(A, oshape, shape[0], r, world)
            full, kk = P.resample_slab(mov, A, oshape, r, world)
            if k1 <= k0:
                assert full is None
                continue
            part = mov[z0:z1].clone()                       # what this rank would hold
            held += z1 - z0
            got, kk2 = P.resample_slab(part, A, oshape, r, world, moving_z0=z0, moving_nz=shape[0])
            assert kk == kk2 == (k0, k1) and torch.equal(got, full)
        assert held < 0.75 * world * shape[0] or world == 2


def test_host_slabs_of_one_pinned_volume_equal_the_one_call_result(cuda_device):
    """resample_host_slab: N simulated ranks move only their planes of ONE pinned host volume and fill ONE pinned output;
    the result equals mia_resample_rigid_host on the whole volume byte for byte."""
    from medicalimageanalysis_b200 import kernels
    shape, sp = (44, 72, 136), (0.9765625, 0.9765625, 2.5)
    mov_h = make_phantom(shape, "CT", seed=7, spacing=sp)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    M = G.params_to_matrix([3, -2, 4, 5, -3, 7], (0, 0, 0))
    origin, dims = G.autocrop_extent(g, M, sp, convention="vtk")
    A = G.compose_index_matrix(G.Grid(dims, sp, origin), g, M, convention="vtk")
    oshape = (dims[2], dims[1], dims[0])
    whole = kernels.resample_host(mov_h, A, oshape, mode="vtk", device=cuda_device.index or 0)
    h_in = np.ascontiguousarray(mov_h).copy()
    h_out = np.full(oshape, 7, dtype=np.int16)
    hin, hout = kernels.host_register(h_in), kernels.host_register(h_out)       # caller-owned memory, pinned in place
    try:
        for world in (1, 3, 8):
            h_out[...] = 7
            for r in range(world):
                P.resample_host_slab(h_in, A, h_out, r, world, device=cuda_device.index or 0)
            assert np.array_equal(h_out, whole), world
    finally:
        kernels.host_unregister(hin); kernels.host_unregister(hout)
