"""z-slab sharding of the fixed lattice across the GPUs of one box.

The path shards naturally: output / fixed voxels are independent and the metric is one tiny reduction.
Rank r owns the z-slab ``[k0_r, k1_r)`` of the fixed lattice (boundaries on multiples of the kernel's
8-plane tile so every rank runs whole tiles), evaluates it with the matrix shifted by its first plane,
``A[:, 3] += A[:, 2] * k0`` -- the rule ``mia_resample_rigid_host`` uses for its own z-slab pipeline --
and the partial ``{ssd, count}`` or joint histograms are combined by ``all_reduce(SUM)`` over NCCL / NVLink
(``gloo`` in the CPU tests).  In a registration loop the moving volume is replicated: ``A`` changes every iteration,
and even config 5 (2.5 GB) is small next to 180 GB of HBM.  A one-shot resample (``Rigid.create_image``) needs only the
moving planes its slab touches: ``resample_slab`` / ``resample_host_slab`` take ``geometry.slab_moving_zrange`` -- the
bounding box of the slab under ``A`` plus a one-plane halo -- so that N ranks share ONE volume with 1/N of the
traffic each.

Volumes too small to cut (config 3: 256^3 is 50 us of work per evaluation) shard over the 13 PROBES of an
optimiser iteration instead: rank r evaluates probes ``[q0_r, q1_r)`` on a replicated pair and the
13 costs are combined by one ``all_reduce(SUM)`` of a 13-vector that is zero outside each rank's share
(``probe_bounds``, ``probe_parallel_costs``, ``register_rigid_probe_parallel``).

Pure resampling needs no collective at all: the output stays sharded (``resample_slab``), or whole
volumes are dealt out per rank (config 4, ``bench.py --gpus N``).

The reference has nothing distributed; the call sites this serves are the registration loop of
``registration.py`` and ``Rigid.create_image`` (``structure/rigid.py:646-695``) on volumes
too large or too many for one GPU.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import geometry as G

TILE_Z = 8      # csrc/mia_tile.cuh kTZ: slab boundaries on tile boundaries keep every tile whole


def slab_bounds(nz: int, world: int, align: int = TILE_Z):
    """``world + 1`` increasing plane indices; slab r is ``[b[r], b[r+1])``.  Interior boundaries are
    multiples of ``align``; trailing ranks may get an empty slab when there are fewer tiles than ranks."""
    nz, world = int(nz), int(world)
    if nz < 0 or world < 1:
        raise ValueError("need nz >= 0 and world >= 1")
    ntile = (nz + align - 1) // align
    b = [min(nz, align * ((ntile * r) // world)) for r in range(world)]
    return b + [nz]


def slab_of(nz: int, rank: int, world: int, align: int = TILE_Z):
    b = slab_bounds(nz, world, align)
    return b[rank], b[rank + 1]


def shift_matrix(A, k0):
    """Index matrix of the slab that starts at fixed plane ``k0`` (works on ``[3,4]`` or ``[n,3,4]``,
    NumPy arrays or torch tensors; returns a new object)."""
    out = A.clone() if hasattr(A, "clone") else np.array(A, dtype=np.float64, copy=True)
    out[..., 3] = out[..., 3] + out[..., 2] * float(k0)
    return out


def _dist():
    import torch.distributed as dist
    return dist


def all_reduce_sum(*tensors, group=None):
    """In-place SUM over the process group (no-op when torch.distributed is not initialised)."""
    dist = _dist()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return tensors
    for t in tensors:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return tensors


def sharded_ssd(evaluate_slab, nz, A, rank, world, group=None):
    """Generic driver, also used by the CPU tests with the oracle as ``evaluate_slab``:
    ``evaluate_slab(k0, k1, A_shifted) -> (ssd[n] float64 tensor, count[n] int64 tensor)`` on this rank's
    slab; returns the all-reduced pair."""
    import torch
    k0, k1 = slab_of(nz, rank, world)
    n = 1 if np.ndim(A) == 2 else len(A)
    if k1 > k0:
        ssd, cnt = evaluate_slab(k0, k1, shift_matrix(A, k0))
        ssd, cnt = ssd.clone(), cnt.clone()
    else:
        ssd, cnt = torch.zeros(n, dtype=torch.float64), torch.zeros(n, dtype=torch.int64)
    all_reduce_sum(ssd, cnt, group=group)
    return ssd, cnt


def sharded_hist(evaluate_slab, nz, A, bins, rank, world, group=None):
    """Same for joint histograms: ``evaluate_slab(k0, k1, A_shifted) -> (hist[n,B,B] int, count[n] int64)``.
    Integer sums commute, so the result is bit-identical for every world size."""
    import torch
    k0, k1 = slab_of(nz, rank, world)
    n = 1 if np.ndim(A) == 2 else len(A)
    if k1 > k0:
        h, cnt = evaluate_slab(k0, k1, shift_matrix(A, k0))
        h, cnt = h.to(torch.int64).clone(), cnt.clone()
    else:
        h, cnt = torch.zeros((n, bins, bins), dtype=torch.int64), torch.zeros(n, dtype=torch.int64)
    all_reduce_sum(h, cnt, group=group)
    return h, cnt


def slab_plan(A, out_shape, mov_nz, rank, world):
    """This rank's share of a one-shot resample: output planes ``[k0, k1)``, the moving planes ``[z0, z1)`` they
    touch (``geometry.slab_moving_zrange``: bounding box + one plane of halo, clipped) and the matrix that maps the
    slab's own index space to the sub-volume's: ``A[:, 3] += A[:, 2] * k0`` and ``A[2, 3] -= z0``.  Because the halo
    keeps every sample at least one plane away from an artificial face, inside tests and clamps see what they would
    see in the whole volume: the slab is bit-identical to the same planes of a one-GPU resample."""
    A = np.asarray(A, dtype=np.float64).reshape(3, 4)
    k0, k1 = slab_of(out_shape[0], rank, world)
    if k1 <= k0:
        return (k0, k1), (0, 0), A
    z0, z1 = G.slab_moving_zrange(A, (out_shape[2], out_shape[1], out_shape[0]), k0, k1, mov_nz)
    As = shift_matrix(A, k0)
    As[2, 3] -= float(z0)
    return (k0, k1), (z0, z1), As


def resample_slab(moving, A, out_shape, rank, world, background=-3001.0, mode="vtk", moving_z0=0, moving_nz=None):
    """This rank's z-slab of a resample (no collective).  ``moving`` is a CUDA tensor holding moving planes
    ``[moving_z0, moving_z0 + len)`` of a volume ``moving_nz`` planes deep -- the whole volume (the defaults), or just
    what ``slab_plan`` says this rank needs.  Returns ``(tensor [k1-k0, y, x], (k0, k1))``."""
    from . import kernels
    nz_total = int(moving_nz) if moving_nz is not None else moving_z0 + moving.shape[0]
    (k0, k1), (z0, z1), As = slab_plan(A, out_shape, nz_total, rank, world)
    if k1 <= k0:
        return None, (k0, k1)
    if z1 <= z0:                                  # the slab misses the volume altogether: background only
        z0, z1 = moving_z0, moving_z0 + 1
        As = shift_matrix(np.asarray(A, dtype=np.float64).reshape(3, 4), k0)
        As[2, 3] -= float(z0)
    if z0 < moving_z0 or z1 > moving_z0 + moving.shape[0]:
        raise ValueError("resident moving planes [%d, %d) do not cover the slab's range [%d, %d)" %
                         (moving_z0, moving_z0 + moving.shape[0], z0, z1))
    out = kernels.resample(moving[z0 - moving_z0:z1 - moving_z0], As, (k1 - k0, out_shape[1], out_shape[2]),
                           background=background, mode=mode)
    return out, (k0, k1)


def resample_host_slab(h_moving, A, h_out, rank, world, background=-3001.0, mode="vtk", device=0):
    """This rank's z-slab of ONE host volume (``Rigid.create_image`` on NumPy arrays, ``structure/rigid.py:646-695``, cut
    over the GPUs of a box): ``h_moving`` and ``h_out`` are the WHOLE arrays -- typically a shared mapping that every rank
    opened, pinned with ``kernels.host_register`` -- and only planes ``[z0, z1)`` go up and ``[k0, k1)`` come down,
    through ``mia_resample_rigid_host`` (H2D / kernel / D2H pipelined).  No collective.  Returns ``(k0, k1), (z0, z1)``."""
    from . import kernels
    (k0, k1), (z0, z1), As = slab_plan(A, h_out.shape, h_moving.shape[0], rank, world)
    if k1 <= k0:
        return (k0, k1), (z0, z1)
    if z1 <= z0:
        h_out[k0:k1] = np.asarray(background).astype(h_out.dtype)
        return (k0, k1), (z0, z1)
    kernels.resample_host(h_moving[z0:z1], As, (k1 - k0,) + tuple(h_out.shape[1:]), background=background, mode=mode,
                          out=h_out[k0:k1], device=device)
    return (k0, k1), (z0, z1)


class FusedExchange(object):
    """Symmetric, peer-mapped buffers for the collective-free exchange of the 26 packed doubles of an iteration
    (``mia_fused_push`` / ``mia_optimizer_step_fused``).  One ``FusedBuf`` per rank from
    ``torch.distributed._symmetric_memory``; zeroed, then a barrier, so that no rank can see a stale flag.  NVLink
    traffic per iteration and rank: ``world`` x (208 B payload + 8 B flag) of peer stores."""

    _cache = {}      # (device index, group) -> FusedExchange: allocation + rendezvous cost tens of milliseconds, once

    @classmethod
    def get(cls, lib, device, group=None):
        """The exchange of (device, group), created on first use; every use starts from zeroed flags behind a barrier."""
        import torch
        key = (torch.device(device).index, group)
        ex = cls._cache.get(key)
        if ex is None:
            ex = cls._cache[key] = cls(lib, device, group)
        ex.reset()
        return ex

    def __init__(self, lib, device, group=None):
        import torch
        import torch.distributed._symmetric_memory as symm_mem
        dist = _dist()
        self.lib, self.device, self.group_arg = lib, device, group
        self.group = group if group is not None else dist.group.WORLD
        n = int(lib.mia_fused_exchange_bytes()) // 8
        self.buf = symm_mem.empty(n, dtype=torch.float64, device=device)
        self.handle = symm_mem.rendezvous(self.buf, self.group)
        self.rank, self.world = int(self.handle.rank), int(self.handle.world_size)
        self.peers = torch.tensor([int(p) for p in self.handle.buffer_ptrs], dtype=torch.int64, device=device)

    def reset(self):
        """A new loop counts its iterations from 1 again: clear the flags of the previous one.  Every rank has left its
        previous loop (it ends with a synchronising read-back), zeroes its OWN buffer, and only then meets the others."""
        import torch
        self.buf.zero_()
        torch.cuda.synchronize(self.device)
        _dist().barrier(group=self.group_arg)

    def timeouts(self, stream):
        t = ctypes.c_ulonglong(0)
        from . import _lib
        _lib.check(self.lib.mia_fused_timeouts(self.buf.data_ptr(), ctypes.byref(t), stream))
        return int(t.value)


def _loop_and_read(lib, _lib, torch, dev, cfg, state, one_iteration, center, extra):
    """Run ``cfg.iterations`` of ``one_iteration`` on the current stream -- ``cfg.graph_iters`` at a time as a replayed
    CUDA graph when asked (NCCL collectives are capturable; every rank captures the same sequence) -- and read the
    optimiser state back."""
    from .registration import N_PROBES, RegistrationResult

    def stream():
        return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    from .registration import _nvtx_wrapped
    one_iteration = _nvtx_wrapped(torch, one_iteration, "mia.register_rigid_sharded.iteration[%s]" % cfg.metric)
    done = 0
    gi = int(cfg.graph_iters)
    if gi > 0 and cfg.iterations >= gi + 1:
        one_iteration()                       # warm-up outside capture (lazy module load, NCCL channel set-up)
        done = 1
        torch.cuda.synchronize(dev)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for _ in range(gi):
                one_iteration()
        while cfg.iterations - done >= gi:
            graph.replay()
            done += gi
    while done < cfg.iterations:
        one_iteration()
        done += 1
    n = min(cfg.iterations, 4096)
    params = (ctypes.c_double * 6)()
    scale = ctypes.c_double()
    iters, conv, nrej = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    hc = np.zeros(n); hp = np.zeros((n, 6)); hs = np.zeros(n)
    dp = ctypes.POINTER(ctypes.c_double)
    _lib.check(lib.mia_optimizer_read(state.data_ptr(), params, ctypes.byref(scale), ctypes.byref(iters),
                                      ctypes.byref(conv), ctypes.byref(nrej), hc.ctypes.data_as(dp),
                                      hp.ctypes.data_as(dp), hs.ctypes.data_as(dp), n, stream()))
    p = np.array(list(params))
    return RegistrationResult(params=p, matrix=G.params_to_matrix(p, center), iterations=int(iters.value),
                              converged=bool(conv.value), n_reject=int(nrej.value), cost_history=hc[:iters.value],
                              param_history=hp[:iters.value], scale_history=hs[:iters.value],
                              evaluations=N_PROBES * int(iters.value), extra=extra)


def register_rigid_sharded(fixed, moving, params0=None, cfg=None, group=None):
    """``registration.register_rigid`` with the fixed lattice cut into z-slabs over the ranks of ``group``.

    Every rank passes the same ``fixed`` / ``moving`` images (CUDA arrays on its own device; only this rank's
    slab of ``fixed.array`` is read).  Per iteration, all on the device and on one stream:

    * SSD: the fused 13-probe launch on the slab; ``mia_metric_pack`` puts ``{ssd[13], count[13]}`` into ONE float64
      buffer (counts are exact in float64); ONE ``all_reduce(SUM)`` of 208 bytes; ``mia_optimizer_step_packed``
      updates the parameters and writes the next probes both globally and shifted to this rank's slab.
    * MI: the 13 joint histograms of the slab; ONE ``all_reduce(SUM)`` of ``13 * B * B`` uint32 (integer: bit-identical
      for every world size); MI from the summed histograms; the same device step.

    All ranks see identical sums and walk the same trajectory without exchanging parameters.  With
    ``cfg.graph_iters > 0`` the loop (collective included) is captured once and replayed.  NVLink traffic per
    iteration and rank: 208 B (SSD) or 213 KB (MI, 64 bins).  Returns the same ``RegistrationResult`` on every rank."""
    import torch

    from . import _lib, kernels
    from .registration import N_PROBES, RegistrationConfig, _intensity_range
    dist = _dist()
    lib = _lib.load()
    cfg = cfg or RegistrationConfig()
    on = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if on else 0
    world = dist.get_world_size(group) if on else 1
    params0 = np.zeros(6) if params0 is None else np.asarray(params0, dtype=np.float64)
    dev = fixed.array.device
    center = np.asarray(fixed.compute_center(), dtype=np.float64)
    p2w = G.index_to_physical(fixed.grid, "itk")
    w2p = G.physical_to_index(moving.grid, "itk")
    is_ssd = cfg.metric == "ssd"
    if cfg.metric not in ("ssd", "mi"):
        raise ValueError("metric must be 'ssd' or 'mi'")
    k0, k1 = slab_of(fixed.array.shape[0], rank, world)
    slab = fixed.array[k0:k1] if k1 > k0 else None
    ws = kernels.MetricWorkspace(dev, n_mats=N_PROBES, bins=cfg.bins)
    d_A = torch.zeros((N_PROBES, 3, 4), dtype=torch.float64, device=dev)
    d_Aloc = torch.zeros_like(d_A)
    packed = torch.zeros(2 * N_PROBES, dtype=torch.float64, device=dev)
    state = torch.zeros(int(lib.mia_optimizer_state_bytes()), dtype=torch.uint8, device=dev)
    if not is_ssd:
        ws.hist.zero_()
        if cfg.frange and cfg.mrange:
            frange, mrange = cfg.frange, cfg.mrange
        else:                                   # ranges must be global: reduce min / max over the slabs
            lo_hi = torch.tensor([float("inf"), float("-inf")], dtype=torch.float64, device=dev)
            if slab is not None:
                lo_hi = torch.stack([slab.min().double(), slab.max().double()])
            if world > 1:
                neg = torch.stack([-lo_hi[0], lo_hi[1]])
                dist.all_reduce(neg, op=dist.ReduceOp.MAX, group=group)
                lo_hi = torch.stack([-neg[0], neg[1]])
            frange = cfg.frange or (float(lo_hi[0]), float(lo_hi[1]))
            mrange = cfg.mrange or _intensity_range(moving.array)
        hist_i32 = ws.hist.view(torch.int32)    # same bits; counts stay far below 2^31 per bin

    def stream():
        return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    fused = None
    if cfg.collective == "fused" and world > 1 and is_ssd:      # MI by z-slab reduces 213 KB of histograms: NCCL's job
        fused = FusedExchange.get(lib, dev, group)
    elif cfg.collective not in ("nccl", "fused"):
        raise ValueError("collective must be 'nccl' or 'fused'")

    with torch.cuda.device(dev):
        _lib.check(lib.mia_optimizer_init(state.data_ptr(), _lib.darr(params0), _lib.darr(center),
                                          _lib.darr(p2w.ravel()), _lib.darr(w2p.ravel()), cfg.rot_step_deg,
                                          cfg.trans_step_mm, cfg.min_step_scale, 0 if is_ssd else 1,
                                          d_A.data_ptr(), stream()))
        d_Aloc.copy_(shift_matrix(d_A, k0))      # once; afterwards the device step keeps it up to date

        def one_iteration():
            if is_ssd:
                if slab is not None:
                    kernels.resample_ssd(slab, moving.array, d_A=d_Aloc, mode=cfg.mode, ws=ws)
                    _lib.check(lib.mia_metric_pack(ws.ssd.data_ptr(), ws.count.data_ptr(), N_PROBES, 0, N_PROBES,
                                                   packed.data_ptr(), stream()))
                else:
                    _lib.check(lib.mia_metric_pack(None, None, 0, 0, N_PROBES, packed.data_ptr(), stream()))
                if fused is None:
                    all_reduce_sum(packed, group=group)
            else:
                if slab is not None:
                    kernels.resample_jointhist(slab, moving.array, frange, mrange, bins=cfg.bins, d_A=d_Aloc,
                                               mode=cfg.mode, ws=ws)
                else:
                    ws.hist.zero_()
                all_reduce_sum(hist_i32, group=group)
                kernels.mi_from_hist(ws.hist, out=ws.mi)
                _lib.check(lib.mia_metric_pack(ws.mi.data_ptr(), None, N_PROBES, 0, N_PROBES, packed.data_ptr(), stream()))
            if fused is not None and is_ssd:
                _lib.check(lib.mia_fused_push(state.data_ptr(), packed.data_ptr(), fused.peers.data_ptr(), rank, world, stream()))
                _lib.check(lib.mia_optimizer_step_fused(state.data_ptr(), fused.buf.data_ptr(), world, 1, d_A.data_ptr(),
                                                        d_Aloc.data_ptr(), int(k0), stream()))
            else:
                _lib.check(lib.mia_optimizer_step_packed(state.data_ptr(), packed.data_ptr(), 1 if is_ssd else 0,
                                                         d_A.data_ptr(), d_Aloc.data_ptr(), int(k0), stream()))

        res = _loop_and_read(lib, _lib, torch, dev, cfg, state, one_iteration, center,
                             {"slab": (k0, k1), "world": world, "collective": "fused" if fused is not None else "nccl"})
        if fused is not None and fused.timeouts(stream()):
            raise RuntimeError("fused exchange: a peer never arrived (timeout); the result is invalid")
        return res


# ---------------------------------------------------------------------------------------------------------
# probe-parallel evaluation (small volumes)

def probe_bounds(n_probes: int, world: int):
    """``world + 1`` increasing probe indices; rank r evaluates ``[b[r], b[r+1])`` (sizes differ by at most 1,
    ranks beyond the probe count get nothing)."""
    n_probes, world = int(n_probes), int(world)
    if n_probes < 0 or world < 1:
        raise ValueError("need n_probes >= 0 and world >= 1")
    return [(n_probes * r) // world for r in range(world)] + [n_probes]


def probe_parallel_costs(evaluate_subset, probes, rank, world, group=None):
    """Generic driver, also used by the CPU tests with the oracle as ``evaluate_subset``:
    ``evaluate_subset(list_of_param_vectors) -> costs`` is called with this rank's share of ``probes``; returns
    the full cost vector (float64 tensor) on every rank.  Adding zeros is exact, so the vector is bit-identical
    on all ranks and to the single-process result."""
    import torch
    b = probe_bounds(len(probes), world)
    q0, q1 = b[rank], b[rank + 1]
    costs = torch.zeros(len(probes), dtype=torch.float64)
    if q1 > q0:
        costs[q0:q1] = torch.as_tensor(np.asarray(evaluate_subset(list(probes[q0:q1])), dtype=np.float64))
    all_reduce_sum(costs, group=group)
    return costs


def register_rigid_probe_parallel(fixed, moving, params0=None, cfg=None, group=None):
    """``registration.register_rigid`` with the 13 probes of every iteration dealt out over the ranks of
    ``group``; fixed and moving are replicated (CUDA arrays on each rank's own device).

    Per iteration: one fused launch over this rank's probes, MI from its own histograms, ``mia_metric_pack`` into the
    rank's share of ONE float64 buffer ``{cost[13], count[13]}`` (zero elsewhere, so the sum is exact), ONE
    ``all_reduce(SUM)`` of 208 bytes, then the one-thread optimiser kernel on every rank -- identical inputs, so all
    ranks walk the same trajectory.  ``cfg.graph_iters > 0`` captures the loop, collective included, in a CUDA graph.
    Returns the same ``RegistrationResult`` everywhere."""
    import torch

    from . import _lib, kernels
    from .registration import N_PROBES, RegistrationConfig, _intensity_range
    dist = _dist()
    lib = _lib.load()
    cfg = cfg or RegistrationConfig()
    on = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if on else 0
    world = dist.get_world_size(group) if on else 1
    params0 = np.zeros(6) if params0 is None else np.asarray(params0, dtype=np.float64)
    dev = fixed.array.device
    center = np.asarray(fixed.compute_center(), dtype=np.float64)
    p2w = G.index_to_physical(fixed.grid, "itk")
    w2p = G.physical_to_index(moving.grid, "itk")
    if cfg.metric not in ("ssd", "mi"):
        raise ValueError("metric must be 'ssd' or 'mi'")
    is_ssd = cfg.metric == "ssd"
    b = probe_bounds(N_PROBES, world)
    q0, q1 = b[rank], b[rank + 1]
    nloc = q1 - q0
    ws = kernels.MetricWorkspace(dev, n_mats=max(nloc, 1), bins=cfg.bins)
    d_A = torch.zeros((N_PROBES, 3, 4), dtype=torch.float64, device=dev)
    packed = torch.zeros(2 * N_PROBES, dtype=torch.float64, device=dev)
    state = torch.zeros(int(lib.mia_optimizer_state_bytes()), dtype=torch.uint8, device=dev)
    if not is_ssd:
        frange = cfg.frange or _intensity_range(fixed.array)
        mrange = cfg.mrange or _intensity_range(moving.array)
        ws.mi                                     # allocate outside any graph capture

    def stream():
        return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    if cfg.collective not in ("nccl", "fused"):
        raise ValueError("collective must be 'nccl' or 'fused'")
    fused = FusedExchange.get(lib, dev, group) if (cfg.collective == "fused" and world > 1) else None

    with torch.cuda.device(dev):
        _lib.check(lib.mia_optimizer_init(state.data_ptr(), _lib.darr(params0), _lib.darr(center),
                                          _lib.darr(p2w.ravel()), _lib.darr(w2p.ravel()), cfg.rot_step_deg,
                                          cfg.trans_step_mm, cfg.min_step_scale, 0 if is_ssd else 1,
                                          d_A.data_ptr(), stream()))
        mine = d_A[q0:q1]

        def one_iteration():
            if nloc > 0:
                if is_ssd:
                    v, c = kernels.resample_ssd(fixed.array, moving.array, d_A=mine, mode=cfg.mode, ws=ws)
                else:
                    h, c = kernels.resample_jointhist(fixed.array, moving.array, frange, mrange, bins=cfg.bins,
                                                      d_A=mine, mode=cfg.mode, ws=ws)
                    v = kernels.mi_from_hist(h, out=ws.mi[:nloc])
                _lib.check(lib.mia_metric_pack(v.data_ptr(), c.data_ptr(), nloc, q0, N_PROBES, packed.data_ptr(), stream()))
            else:
                _lib.check(lib.mia_metric_pack(None, None, 0, 0, N_PROBES, packed.data_ptr(), stream()))
            if fused is not None:
                _lib.check(lib.mia_fused_push(state.data_ptr(), packed.data_ptr(), fused.peers.data_ptr(), rank, world, stream()))
                _lib.check(lib.mia_optimizer_step_fused(state.data_ptr(), fused.buf.data_ptr(), world, 1 if is_ssd else 0,
                                                        d_A.data_ptr(), None, 0, stream()))
            else:
                all_reduce_sum(packed, group=group)
                _lib.check(lib.mia_optimizer_step_packed(state.data_ptr(), packed.data_ptr(), 1 if is_ssd else 0,
                                                         d_A.data_ptr(), None, 0, stream()))

        res = _loop_and_read(lib, _lib, torch, dev, cfg, state, one_iteration, center,
                             {"probes": (q0, q1), "world": world, "collective": "fused" if fused is not None else "nccl"})
        if fused is not None and fused.timeouts(stream()):
            raise RuntimeError("fused exchange: a peer never arrived (timeout); the result is invalid")
        return res
