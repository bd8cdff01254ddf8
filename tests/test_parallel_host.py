"""CPU tests of the z-slab sharding logic: slab arithmetic, and a world_size-2 gloo run in which
each rank evaluates its slab with the ORACLE (there is no GPU here) and the all-reduced partials must equal
the whole-volume oracle result -- histograms and counts bit for bit, SSD to 1e-12."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from medicalimageanalysis_b200 import geometry as G
from medicalimageanalysis_b200 import parallel as P
from medicalimageanalysis_b200.phantom import make_phantom
from oracle import native


def test_slab_bounds_cover_and_align():
    for nz in (0, 1, 7, 8, 9, 100, 300, 317, 1200):
        for world in (1, 2, 3, 4, 8):
            b = P.slab_bounds(nz, world)
            assert b[0] == 0 and b[-1] == nz and len(b) == world + 1
            assert all(b[i] <= b[i + 1] for i in range(world))
            assert all(x % P.TILE_Z == 0 for x in b[1:-1] if x != nz)
            sizes = [b[i + 1] - b[i] for i in range(world)]
            assert max(sizes) - min(s for s in sizes) <= 2 * P.TILE_Z or nz < world * P.TILE_Z


def test_shift_matrix_is_the_slab_rule():
    rng = np.random.default_rng(0)
    A = rng.normal(size=(5, 3, 4))
    S = P.shift_matrix(A, 7)
    k = np.array([3.0, 4.0, 2.0, 1.0])
    kk = np.array([3.0, 4.0, 9.0, 1.0])
    assert np.allclose(S @ k, A @ kk) and not np.shares_memory(S, A)
    St = P.shift_matrix(torch.from_numpy(A), 7)
    assert np.allclose(St.numpy(), S)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shape, sp = (20, 24, 40), (1.0, 1.0, 2.0)
        g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
        fixed = make_phantom(shape, "CT", seed=3, spacing=sp)
        moving = make_phantom(shape, "CT", seed=4, spacing=sp)
        A = np.stack([G.compose_index_matrix(g, g, G.params_to_matrix(p, (0, 0, 0)))
                      for p in ([2, -1, 3, 1, 2, -1], [0, 0, 0, 0.5, 0, 0])])
        fr = (-1024.0, 3071.0)

        def ev_ssd(k0, k1, As):
            r = [native.reslice_ssd(fixed[k0:k1], moving, a, "itk") for a in As]
            return torch.tensor([x[0] for x in r], dtype=torch.float64), torch.tensor([x[1] for x in r], dtype=torch.int64)

        def ev_hist(k0, k1, As):
            r = [native.joint_hist(fixed[k0:k1], moving, a, 16, fr, fr, "itk") for a in As]
            return (torch.from_numpy(np.stack([x[0].astype(np.int64) for x in r])),
                    torch.tensor([x[1] for x in r], dtype=torch.int64))

        ssd, cnt = P.sharded_ssd(ev_ssd, shape[0], A, rank, world)
        h, hc = P.sharded_hist(ev_hist, shape[0], A, 16, rank, world)
        if rank == 0:
            q.put((ssd.numpy(), cnt.numpy(), h.numpy(), hc.numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_sharded_metrics_equal_the_whole(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ssd, cnt, h, hc = got
    shape, sp = (20, 24, 40), (1.0, 1.0, 2.0)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    fixed = make_phantom(shape, "CT", seed=3, spacing=sp)
    moving = make_phantom(shape, "CT", seed=4, spacing=sp)
    fr = (-1024.0, 3071.0)
    for i, p in enumerate(([2, -1, 3, 1, 2, -1], [0, 0, 0, 0.5, 0, 0])):
        A = G.compose_index_matrix(g, g, G.params_to_matrix(p, (0, 0, 0)))
        want, n = native.reslice_ssd(fixed, moving, A, "itk")
        assert cnt[i] == n and abs(ssd[i] - want) <= 1e-12 * want
        H, hn = native.joint_hist(fixed, moving, A, 16, fr, fr, "itk")
        assert hc[i] == hn and np.array_equal(h[i], H.astype(np.int64))


# ---------------------------------------------------------------------------------------------------------
# probe-parallel evaluation (small volumes): the 13 probes of an iteration dealt out over the ranks

def test_probe_bounds_partition():
    for n in (0, 1, 5, 13):
        for world in (1, 2, 3, 4, 8, 16):
            b = P.probe_bounds(n, world)
            assert b[0] == 0 and b[-1] == n and len(b) == world + 1
            sizes = [b[i + 1] - b[i] for i in range(world)]
            assert all(s >= 0 for s in sizes) and max(sizes) - min(sizes) <= 1


def _probe_case():
    from medicalimageanalysis_b200.registration import RegistrationConfig, params_to_index_matrix
    shape, sp = (16, 20, 24), (1.0, 1.0, 1.5)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    center = tuple(np.asarray(g.origin) + 0.5 * (np.asarray(g.dims) - 1) * np.asarray(g.spacing))
    fixed = make_phantom(shape, "CT", seed=5, spacing=sp)
    moving = make_phantom(shape, "CT", seed=5, spacing=sp,
                          world_matrix=np.linalg.inv(G.params_to_matrix((1.5, -1.0, 2.0, 1.0, -0.5, 0.8), center)))
    p2w, w2p = G.index_to_physical(g, "itk"), G.physical_to_index(g, "itk")

    def evaluate(plist):
        out = []
        for prm in plist:
            ssd, n = native.reslice_ssd(fixed, moving, params_to_index_matrix(prm, center, p2w, w2p), "itk")
            out.append(ssd / max(n, 1))
        return out
    return evaluate, RegistrationConfig(metric="ssd", iterations=6)


def _probe_worker(rank, world, port, q):
    from medicalimageanalysis_b200.registration import host_optimizer
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        evaluate, cfg = _probe_case()
        res = host_optimizer(lambda pl: P.probe_parallel_costs(evaluate, pl, rank, world).numpy(), np.zeros(6), cfg)
        q.put((rank, res.params, res.cost_history))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_probe_parallel_walks_the_single_process_trajectory(world):
    from medicalimageanalysis_b200.registration import host_optimizer
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_probe_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    evaluate, cfg = _probe_case()
    want = host_optimizer(evaluate, np.zeros(6), cfg)
    for _, params, hist in got:                      # every rank, bit for bit (adding zeros is exact)
        assert np.array_equal(params, want.params) and np.array_equal(hist, want.cost_history)
    assert want.cost_history[-1] < want.cost_history[0]


# ---------------------------------------------------------------------------------------------------------
# one-shot resample cut into z-slabs, each rank holding only the moving planes its slab touches (halo)

def _halo_case():
    shape, sp = (41, 28, 40), (0.9765625, 0.9765625, 2.5)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    mov = make_phantom(shape, "CT", seed=6, spacing=sp)
    M = G.params_to_matrix([3, -2, 4, 5, -3, 7], (0, 0, 0))
    origin, dims = G.autocrop_extent(g, M, sp, convention="vtk")
    A = G.compose_index_matrix(G.Grid(dims, sp, origin), g, M, convention="vtk")
    return mov, A, dims


@pytest.mark.parametrize("mode", ["vtk", "itk"])
def test_slab_plan_halo_reproduces_the_whole_volume(mode):
    """Resampling slab by slab from the halo sub-volume only gives the bytes of the one-piece resample, for any
    number of ranks (empty trailing slabs included)."""
    mov, A, dims = _halo_case()
    oshape = (dims[2], dims[1], dims[0])
    whole = native.reslice(mov, A, dims, -3001.0, mode)
    for world in (1, 2, 3, 5, 8):
        out = np.full(oshape, 7, dtype=np.int16)
        touched = 0
        for rank in range(world):
            (k0, k1), (z0, z1), As = P.slab_plan(A, oshape, mov.shape[0], rank, world)
            if k1 <= k0:
                continue
            assert 0 <= z0 < z1 <= mov.shape[0]
            touched += z1 - z0
            out[k0:k1] = native.reslice(mov[z0:z1], As, (dims[0], dims[1], k1 - k0), -3001.0, mode)
        assert np.array_equal(out, whole), world
        if world >= 5:
            assert touched < 0.6 * world * mov.shape[0]          # each rank holds a fraction of the volume, not all of it


def _halo_worker(rank, world, port, path_in, path_out, shape_in, shape_out, A):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        h_in = np.memmap(path_in, dtype=np.int16, mode="r", shape=shape_in)       # ONE host volume, shared by the ranks
        h_out = np.memmap(path_out, dtype=np.int16, mode="r+", shape=shape_out)
        (k0, k1), (z0, z1), As = P.slab_plan(A, shape_out, shape_in[0], rank, world)
        if k1 > k0:
            h_out[k0:k1] = native.reslice(np.ascontiguousarray(h_in[z0:z1]), As, (shape_out[2], shape_out[1], k1 - k0),
                                          -3001.0, "vtk")
            h_out.flush()
        planes = torch.tensor([z1 - z0], dtype=torch.int64)
        dist.all_reduce(planes)                                                   # doubles as the completion barrier
        assert int(planes[0]) < world * shape_in[0]
    finally:
        dist.destroy_process_group()


def test_gloo_two_ranks_fill_one_shared_output_from_their_halos(tmp_path):
    mov, A, dims = _halo_case()
    oshape = (dims[2], dims[1], dims[0])
    path_in, path_out = str(tmp_path / "in.raw"), str(tmp_path / "out.raw")
    mov.tofile(path_in)
    np.full(oshape, 7, dtype=np.int16).tofile(path_out)
    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_halo_worker, args=(r, 2, port, path_in, path_out, mov.shape, oshape, A)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    got = np.fromfile(path_out, dtype=np.int16).reshape(oshape)
    assert np.array_equal(got, native.reslice(mov, A, dims, -3001.0, "vtk"))
