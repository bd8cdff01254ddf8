"""Device optimiser loop vs the host statement of the same algorithm run on the CPU oracle
("identical optimiser convergence on the same inputs": same decisions, per-iteration metric within 1e-5)."""
import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import Data, RegistrationConfig, geometry as G
from medicalimageanalysis_b200.registration import host_optimizer
from medicalimageanalysis_b200.utils import CreateImageFromMask
from medicalimageanalysis_b200.utils import rigid as R
from tests.test_registration_host import oracle_evaluate, small_pair

pytestmark = pytest.mark.gpu


def images(g, fixed, moving, dev):
    Data.clear()
    f = CreateImageFromMask(fixed, g.origin, g.spacing, "fixed", device=dev).add_image()
    m = CreateImageFromMask(moving, g.origin, g.spacing, "moving", device=dev).add_image()
    return f, m


@pytest.mark.parametrize("graph_iters", [0, 5])
def test_ssd_trajectory_matches_cpu_oracle(cuda_device, graph_iters):
    g, center, fixed, moving = small_pair()
    f, m = images(g, fixed, moving, cuda_device)
    # the Image centre (voxel int(dim/2)) is the pivot on both sides
    center = f.compute_center()
    cfg = RegistrationConfig(metric="ssd", iterations=30, graph_iters=graph_iters)
    gpu = R.register_rigid(f, m, cfg=cfg)
    cpu = host_optimizer(oracle_evaluate(g, center, fixed, moving), np.zeros(6), cfg)
    assert gpu.iterations == cpu.iterations == 30
    assert np.array_equal(gpu.scale_history, cpu.scale_history)            # same accept / reject decisions
    assert gpu.n_reject == cpu.n_reject
    assert np.allclose(gpu.cost_history, cpu.cost_history, rtol=1e-5, atol=1e-9)
    assert np.allclose(gpu.param_history, cpu.param_history, atol=1e-3)
    assert np.allclose(gpu.params, cpu.params, atol=1e-3)
    assert gpu.cost_history[-1] < 0.4 * gpu.cost_history[0]


def test_mi_registration_recovers_offset(cuda_device):
    from medicalimageanalysis_b200.phantom import make_phantom
    shape, sp = (40, 64, 64), (2.0, 2.0, 2.0)
    g = G.Grid((64, 64, 40), sp, (-64.0, -64.0, -40.0))
    gt = np.array([2.0, 1.5, -2.5, 3.0, -2.0, 1.5])
    Data.clear()
    fixed = make_phantom(shape, "CT", seed=3, spacing=sp).astype(np.float32)
    f = CreateImageFromMask(fixed, g.origin, g.spacing, "fixed", device=cuda_device).add_image()
    Mgt = G.params_to_matrix(gt, f.compute_center())
    moving = make_phantom(shape, "MR", seed=3, spacing=sp, world_matrix=None)
    # same anatomy, MR contrast, displaced: generate in the displaced frame
    c = np.asarray(f.compute_center())
    W = np.identity(4); W[:3, 3] = -c
    Winv = np.identity(4); Winv[:3, 3] = c
    # phantom coordinates are centred on the volume; the Image centre is half a voxel off that, fold it in
    moving = make_phantom(shape, "MR", seed=3, spacing=sp, world_matrix=np.linalg.inv(Mgt))
    m = CreateImageFromMask(moving, g.origin, g.spacing, "moving", device=cuda_device).add_image()
    cfg = RegistrationConfig(metric="mi", iterations=60, bins=32, rot_step_deg=0.5, trans_step_mm=0.5)
    res = R.register_rigid(f, m, cfg=cfg)
    mi0, _ = R.metric_mi(f, m, np.identity(4), bins=32)
    mi1, _ = R.metric_mi(f, m, res.matrix, bins=32)
    assert mi1 > mi0 * 1.1
    assert -res.cost_history[-1] >= -res.cost_history[0]


def test_rigid_api_flow(cuda_device):
    """Rigid.create_image / update_rotation / update_translation / display fields, reference semantics."""
    from medicalimageanalysis_b200 import Rigid
    from oracle import reslice
    g, center, fixed, moving = small_pair()
    f, m = images(g, fixed, moving, cuda_device)
    rig = Rigid("fixed", "moving")
    assert rig.rigid_name == "fixed_moving" and Data.rigid["fixed_moving"] is rig
    rig.display.compute_reslice()
    assert rig.display.array.shape == m.array.shape                      # identity: same lattice
    assert torch.equal(rig.display.array, m.array)
    assert np.allclose(rig.display.origin, m.origin)
    rig.update_rotation(r_x=0, r_y=0, r_z=7.0)
    rig.update_translation(t_x=3.0, t_y=-1.0, t_z=2.0)
    Mexp = np.linalg.inv(G.pivot_transform(f.compute_center(), 0, 0, 7.0)) @ np.linalg.inv(G.translation_transform(3, -1, 2))
    assert np.allclose(rig.matrix, Mexp)
    origin, dims = G.autocrop_extent(m.grid, rig.matrix, f.spacing, "vtk")
    assert tuple(rig.display.array.shape) == (dims[2], dims[1], dims[0])
    og = G.Grid(dims, tuple(f.spacing), origin)
    A = G.compose_index_matrix(og, m.grid, rig.matrix, "vtk")
    want = reslice.oracle_reslice(moving, A, dims, -3001, "vtk")
    got = rig.display.array.cpu().numpy()
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
    assert np.allclose(rig.display.offset['Axial'][0], (origin[0] - f.origin[0]) / f.spacing[0])
    ang = rig.retrieve_angles()
    from scipy.spatial.transform import Rotation
    assert np.allclose(ang, Rotation.from_matrix(rig.matrix[:3, :3]).as_euler('ZXY', degrees=True), atol=1e-9)
    sl = rig.retrieve_array_plane('Axial', position=rig.display.origin + np.array([0, 0, 4 * f.spacing[2]]))
    assert sl.dtype == np.float64 and np.array_equal(sl, got[4].astype(np.float64))
    # ITK-mode operator onto the reference lattice
    t = R.resample_to_reference(m, f, rig.matrix, default_value=0.0)
    A2 = G.compose_index_matrix(f.grid, m.grid, rig.matrix, "itk")
    want2 = reslice.oracle_reslice(moving, A2, g.dims, 0.0, "itk")
    assert np.abs(t.cpu().numpy().astype(int) - want2.astype(int)).max() <= 1
    s, n, mse = R.metric_ssd(f, m, rig.matrix)
    assert n > 0 and mse == s / n


def test_verify_axial_orientation_matches_reference_views(cuda_device):
    from oracle import ffs
    rng = np.random.default_rng(0)
    a = rng.integers(-1000, 1000, size=(6, 10, 14)).astype(np.int16)
    t = torch.from_numpy(a).to(cuda_device)
    for orient, origin in (([-1, 0, 0, 0, 1, 0], [13.0, 0, 0]), ([1, 0, 0, 0, -1, 0], [0, 9.0, 0]),
                           ([-1, 0, 0, 0, -1, 0], [13.0, 9.0, 0]), ([1, 0, 0, 0, 1, 0], [0, 0, 0])):
        idx, corners = ffs.ffs_corner_index(origin, orient, [1, 1, 1], a.shape)
        case = ffs.axial_case_for_corner(idx)
        arr, new_origin, new_orient, mat = R.verify_axial_orientation(t, origin, orient, [1, 1, 1], "Axial")
        if case is None:
            assert arr is t
        else:
            assert np.array_equal(arr.cpu().numpy(), ffs.oracle_ffs(a, case))
            assert np.allclose(new_origin, corners[idx])
