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
