"""Intensity-based rigid registration loop over the fused resample + metric kernels.

The reference has no such loop (rigid = mesh ICP ``utils/rigid/icp.py:62-149`` or manual nudges
``structure/rigid.py:950-989``); what is preserved from it is the parameterisation -- 'xyz' Euler degrees
about the fixed-image centre composed as ``T(c) R T(-c)``, translations in mm -- and the ITK usage pattern
of ``utils/deformable/simpleitk.py:114-124`` (metric + optimiser + linear interpolator, N iterations).

Algorithm (identical on host and device, so CPU-oracle and GPU runs can be compared step for step):
regular-step gradient descent on the 6 parameters with central finite differences -- 13 probes per
iteration (centre, +-h_i per parameter) -- a unit step in scaled-parameter space, and step halving with a
restore of the previous centre whenever the centre cost rises.  ``host_optimizer`` is the plain-Python
statement of it with a pluggable ``evaluate``; ``register_rigid`` runs it on the GPU with the 13 probes
batched into ONE fused launch per iteration and the update done by a one-thread kernel
(``csrc/mia_optimizer.cuh``), so no scalar crosses PCIe inside the loop and the loop replays as a CUDA graph.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field

import numpy as np

from . import geometry as G

N_PROBES = 13


@dataclass
class RegistrationConfig:
    metric: str = "ssd"              # 'ssd' (mean squares, minimised), 'mi' (hard-binned joint-histogram MI) or
                                     # 'mattes' (Parzen-window MI); both MI flavours are maximised
    iterations: int = 200
    rot_step_deg: float = 0.5        # probe offset and step length per unit scale, rotations
    trans_step_mm: float = 0.5       # same, translations
    min_step_scale: float = 1.0 / 1024.0
    bins: int = 64
    mode: str = "itk"                # inside test of the metric (ITK half-open)
    frange: tuple | None = None      # intensity ranges for MI; None -> from the volumes, once
    mrange: tuple | None = None
    graph_iters: int = 10            # iterations captured per CUDA graph (0 = no graph)
    collective: str = "nccl"         # sharded loops: "nccl" (one all_reduce per iteration) or "fused" (partials pushed into
                                     # the peers' symmetric buffers over NVLink, the device step waits on flags: no NCCL call)
    fixed_mask: object = None        # uint8 CUDA tensor on the fixed lattice (SetMetricFixedMask); None = all voxels
    moving_mask: object = None       # same on the moving lattice (SetMetricMovingMask); runs the per-voxel kernel


@dataclass
class RegistrationResult:
    params: np.ndarray
    matrix: np.ndarray               # 4x4 fixed physical -> moving physical
    iterations: int
    converged: bool
    n_reject: int
    cost_history: np.ndarray
    param_history: np.ndarray
    scale_history: np.ndarray
    evaluations: int = 0
    extra: dict = field(default_factory=dict)

    def save(self, path):
        """Persist the optimiser state as a small ``.npz`` (checkpoint / resume): parameters, step scale,
        iteration count and the per-iteration history.  ``resume_from`` turns it back into a start point."""
        scale = float(self.scale_history[-1]) if len(self.scale_history) else 1.0
        np.savez(path, params=self.params, scale=scale, iterations=self.iterations, converged=self.converged,
                 n_reject=self.n_reject, cost_history=self.cost_history, param_history=self.param_history,
                 scale_history=self.scale_history, evaluations=self.evaluations)


def resume_from(path, cfg: RegistrationConfig):
    """``(params0, cfg')`` that continue a saved run: start at its parameters with the probe / step lengths it had
    reached (the saved step scale folded into ``rot_step_deg`` / ``trans_step_mm``; ``min_step_scale`` is rescaled so
    that the absolute stopping length is unchanged)."""
    import dataclasses
    z = np.load(path if str(path).endswith(".npz") else str(path) + ".npz")
    scale = float(z["scale"])
    cfg2 = dataclasses.replace(cfg, rot_step_deg=cfg.rot_step_deg * scale, trans_step_mm=cfg.trans_step_mm * scale,
                               min_step_scale=cfg.min_step_scale / scale)
    return np.asarray(z["params"], dtype=np.float64), cfg2


def probe_params(params, scale, deltas):
    """The 13 parameter vectors of one iteration: centre, then +h_i, -h_i for i = 0..5."""
    p = np.asarray(params, dtype=np.float64)
    out = [p.copy()]
    for i in range(6):
        h = scale * deltas[i]
        a = p.copy(); a[i] += h
        b = p.copy(); b[i] -= h
        out += [a, b]
    return out


def params_to_index_matrix(params, center, p2w_fixed, w2p_mov):
    """params -> 3x4 A (fixed index -> moving continuous index)."""
    M = G.params_to_matrix(params, center)
    return (w2p_mov @ M @ p2w_fixed)[:3, :]


def host_optimizer(evaluate, params0, cfg: RegistrationConfig):
    """Reference statement of the optimiser.  ``evaluate(list_of_13_param_vectors) -> 13 costs`` (lower is
    better).  Mirrors ``optimizer_step_kernel`` line for line."""
    deltas = np.array([cfg.rot_step_deg] * 3 + [cfg.trans_step_mm] * 3, dtype=np.float64)
    params = np.asarray(params0, dtype=np.float64).copy()
    prev_params = params.copy()
    prev_cost, has_prev = 0.0, False
    scale, converged, n_reject = 1.0, False, 0
    hc, hp, hs = [], [], []
    for _ in range(cfg.iterations):
        c = np.asarray(evaluate(probe_params(params, scale, deltas)), dtype=np.float64)
        hc.append(c[0]); hp.append(params.copy()); hs.append(scale)
        if converged:
            continue
        if has_prev and c[0] > prev_cost:
            params = prev_params.copy()
            scale *= 0.5
            n_reject += 1
            if scale < cfg.min_step_scale:
                converged = True
        else:
            prev_params = params.copy()
            prev_cost, has_prev = c[0], True
            sg = np.empty(6)
            for i in range(6):
                h = scale * deltas[i]
                sg[i] = (c[2 * i + 1] - c[2 * i + 2]) / (2.0 * h) * deltas[i]
            norm = float(np.sqrt(np.sum(sg * sg)))
            if 0.0 < norm < 1.0e300:
                params = params - scale * deltas * (sg / norm)
            else:
                converged = True
    return RegistrationResult(params=params, matrix=None, iterations=len(hc), converged=converged,
                              n_reject=n_reject, cost_history=np.array(hc), param_history=np.array(hp),
                              scale_history=np.array(hs), evaluations=N_PROBES * len(hc))


def _nvtx_wrapped(torch, fn, name):
    """NVTX range around one optimiser iteration (13 evaluations + the device step), so that a timeline shows the loop
    structure; host-side markers only, they cost nothing on the device and are legal during graph capture."""
    def wrapped():
        torch.cuda.nvtx.range_push(name)
        try:
            fn()
        finally:
            torch.cuda.nvtx.range_pop()
    return wrapped


def _intensity_range(t):
    return float(t.min().item()), float(t.max().item())


def register_rigid(fixed, moving, params0=None, cfg: RegistrationConfig | None = None, center=None):
    """Register ``moving`` to ``fixed`` (both ``structure.Image``-like with CUDA arrays) entirely on the GPU.
    ``center``: rotation pivot in mm (default: the fixed image centre); the multi-resolution driver pins it to the
    full-resolution centre so that parameters carry over between levels unchanged.

    Returns a ``RegistrationResult``; ``result.matrix`` maps fixed physical -> moving physical, i.e. it is
    what ``Rigid.matrix`` holds (``structure/rigid.py:674-686``)."""
    import torch

    from . import _lib, kernels
    lib = _lib.load()
    cfg = cfg or RegistrationConfig()
    params0 = np.zeros(6) if params0 is None else np.asarray(params0, dtype=np.float64)
    dev = fixed.array.device
    center = np.asarray(fixed.compute_center() if center is None else center, dtype=np.float64)
    p2w = G.index_to_physical(fixed.grid, "itk")
    w2p = G.physical_to_index(moving.grid, "itk")
    is_ssd = cfg.metric == "ssd"
    if cfg.metric not in ("ssd", "mi", "mattes"):
        raise ValueError("metric must be 'ssd', 'mi' or 'mattes'")
    ws = kernels.MetricWorkspace(dev, n_mats=N_PROBES, bins=cfg.bins)
    d_A = torch.zeros((N_PROBES, 3, 4), dtype=torch.float64, device=dev)
    state = torch.zeros(int(lib.mia_optimizer_state_bytes()), dtype=torch.uint8, device=dev)
    fmask = None
    if cfg.fixed_mask is not None:
        fmask = kernels._mask_u8(cfg.fixed_mask, fixed.array.shape, dev)      # converted once, outside the loop
    mmask = None
    if cfg.moving_mask is not None:
        mmask = kernels._mask_u8(cfg.moving_mask, moving.array.shape, dev)
    if not is_ssd:
        frange = cfg.frange or _intensity_range(fixed.array)
        mrange = cfg.mrange or _intensity_range(moving.array)

    def stream():
        return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    with torch.cuda.device(dev):
        _lib.check(lib.mia_optimizer_init(state.data_ptr(), _lib.darr(params0), _lib.darr(center),
                                          _lib.darr(p2w.ravel()), _lib.darr(w2p.ravel()), cfg.rot_step_deg,
                                          cfg.trans_step_mm, cfg.min_step_scale, 0 if is_ssd else 1,
                                          d_A.data_ptr(), stream()))

        def one_iteration():
            if is_ssd:
                kernels.resample_ssd(fixed.array, moving.array, d_A=d_A, mode=cfg.mode, ws=ws, fixed_mask=fmask,
                                     moving_mask=mmask)
                metric = ws.ssd
            else:
                kernels.resample_jointhist(fixed.array, moving.array, frange, mrange, bins=cfg.bins, d_A=d_A,
                                           mode=cfg.mode, ws=ws, fixed_mask=fmask, moving_mask=mmask)
                kernels.mi_from_hist(ws.hist, out=ws.mi)
                metric = ws.mi
            _lib.check(lib.mia_optimizer_step(state.data_ptr(), metric.data_ptr(), ws.count.data_ptr(),
                                              1 if is_ssd else 0, d_A.data_ptr(), stream()))

        def one_iteration_mattes():
            kernels.resample_mattes(fixed.array, moving.array, frange, mrange, bins=cfg.bins, d_A=d_A, mode=cfg.mode,
                                    ws=ws, fixed_mask=fmask, moving_mask=mmask)
            kernels.mi_from_hist(ws.hist64, out=ws.mi)
            metric = ws.mi
            _lib.check(lib.mia_optimizer_step(state.data_ptr(), metric.data_ptr(), ws.count.data_ptr(),
                                              1 if is_ssd else 0, d_A.data_ptr(), stream()))

        if cfg.metric == "mattes":
            ws.hist64                                 # allocate outside any graph capture
            one_iteration = one_iteration_mattes
        one_iteration = _nvtx_wrapped(torch, one_iteration, "mia.register_rigid.iteration[%s]" % cfg.metric)
        done = 0
        gi = int(cfg.graph_iters)
        if gi > 0 and cfg.iterations >= gi:
            one_iteration()                       # warm-up outside capture (lazy module load, attributes)
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
                              evaluations=N_PROBES * int(iters.value))


class PyramidLevel(object):
    """An ``Image``-like view of one pyramid level: ``array`` (device), ``grid``, ``spacing``, ``origin``, ``matrix``."""

    def __init__(self, array, grid):
        self.array, self.grid = array, grid
        self.spacing, self.origin, self.matrix = np.asarray(grid.spacing), np.asarray(grid.origin), grid.matrix
        self.dimensions = np.asarray(array.shape)

    def compute_center(self):
        d = self.grid.dims
        idx = np.array([int(d[0] / 2), int(d[1] / 2), int(d[2] / 2), 1.0])
        return (G.index_to_physical(self.grid, "itk") @ idx)[:3]


def pyramid_level(image, factors):
    """Block-mean level of ``image``.  ``factors`` = (fx, fy, fz); (1, 1, 1) returns a view."""
    from . import kernels
    f = tuple(int(v) for v in factors)
    if f == (1, 1, 1):
        return PyramidLevel(image.array, image.grid)
    return PyramidLevel(kernels.shrink_mean(image.array, f), G.pyramid_grid(image.grid, f))


def register_rigid_multires(fixed, moving, params0=None, cfg: RegistrationConfig | None = None,
                            shrink_factors=(4, 2, 1), iterations=None):
    """Coarse-to-fine ``register_rigid`` over block-mean pyramids of both images -- the "NumberOfResolutions"
    schedule the reference asks Elastix for (``utils/deformable/simpleitk.py:155``).  ``shrink_factors``: one
    entry per level, an int (isotropic) or an (fx, fy, fz) triple; ``iterations``: per level (default: ``cfg.iterations``
    at every level).  The parameters are physical (degrees about the FULL-resolution fixed centre, mm), so each
    level starts from the previous level's result.  Returns the last level's result with ``extra['levels']``."""
    import dataclasses
    cfg = cfg or RegistrationConfig()
    center = np.asarray(fixed.compute_center(), dtype=np.float64)
    levels = [(f, f, f) if np.isscalar(f) else tuple(f) for f in shrink_factors]
    iters = [cfg.iterations] * len(levels) if iterations is None else list(iterations)
    if len(iters) != len(levels):
        raise ValueError("iterations must have one entry per level")
    frange = cfg.frange or (_intensity_range(fixed.array) if cfg.metric != "ssd" else None)
    mrange = cfg.mrange or (_intensity_range(moving.array) if cfg.metric != "ssd" else None)
    params = np.zeros(6) if params0 is None else np.asarray(params0, dtype=np.float64)
    per_level, evals, res = [], 0, None
    for f, n_it in zip(levels, iters):
        fl, ml = pyramid_level(fixed, f), pyramid_level(moving, f)
        gi = cfg.graph_iters if n_it >= cfg.graph_iters else 0
        lmask = None
        if cfg.fixed_mask is not None:               # a coarse voxel is inside when at least half of its block is
            from . import kernels
            m8 = kernels._mask_u8(cfg.fixed_mask, fixed.array.shape, fixed.array.device)
            lmask = m8 if f == (1, 1, 1) else (kernels.shrink_mean((m8 != 0).float(), f) >= 0.5).to(m8.dtype)
        lmov = None
        if cfg.moving_mask is not None:
            from . import kernels
            mm8 = kernels._mask_u8(cfg.moving_mask, moving.array.shape, moving.array.device)
            lmov = mm8 if f == (1, 1, 1) else (kernels.shrink_mean((mm8 != 0).float(), f) >= 0.5).to(mm8.dtype)
        lcfg = dataclasses.replace(cfg, iterations=int(n_it), frange=frange, mrange=mrange, graph_iters=gi,
                                   fixed_mask=lmask, moving_mask=lmov)
        res = register_rigid(fl, ml, params0=params, cfg=lcfg, center=center)
        params = res.params
        evals += res.evaluations
        per_level.append({"factors": f, "dims": fl.grid.dims, "iterations": res.iterations, "params": res.params.copy(),
                          "cost_first": float(res.cost_history[0]) if len(res.cost_history) else None,
                          "cost_last": float(res.cost_history[-1]) if len(res.cost_history) else None})
    res.evaluations = evals
    res.extra["levels"] = per_level
    return res
