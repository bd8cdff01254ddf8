#!/usr/bin/env python
"""Benchmark of the rigid-resample hot path (BASELINE.json metric: resampled Gvoxel/s, % of HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-extra] [--no-cpu]

A "step" is one pass of the hot path over BASELINE config 2 -- synthetic CT 512x512x300 int16, general rigid
(3,-2,4 deg; 5,-3,7 mm), resampled exactly as ``Rigid.create_image`` does (VTK mode, auto-cropped lattice, background
-3001).  N = 1: one launch on the resident volume.  N > 1 (torchrun, one rank per GPU): the SAME volume cut into N
z-slabs of the output lattice, each rank holding only the moving planes its slab touches plus a one-plane halo
(``parallel.slab_plan``); no data-path collective, strong scaling.  ``e2e`` is the same step through host buffers:
``mia_resample_rigid_host`` at N = 1, and at N > 1 every rank moving only its planes of ONE shared, pinned host volume
(``parallel.resample_host_slab``).  Rank 0 prints ONE JSON line; ``extra`` holds the other BASELINE configs with their
own parity figures against the CPU port.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from medicalimageanalysis_b200 import geometry as G                  # noqa: E402
from medicalimageanalysis_b200.phantom import make_phantom            # noqa: E402

SHAPE = (300, 512, 512)                       # [z, y, x]  BASELINE config 2
SPACING = (0.9765625, 0.9765625, 2.5)
GT = (3.0, -2.0, 4.0, 5.0, -3.0, 7.0)
WORKLOAD = "C2: synthetic CT 512x512x300 int16, rigid (3,-2,4 deg; 5,-3,7 mm) resample, VTK-mode auto-crop"
METRIC = "resampled_output_voxels_per_second"
UNIT = "Gvoxel/s"
TRAFFIC_PROFILE = os.path.join("profiles", "r02_x2_store_ncu_raw.csv")
TIMED_KERNEL = "tile_x2_kernel"               # csrc/mia_tile_x2.cuh; the instantiation the C2 launch runs


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def host_threads():
    """Host threads the CPU arm may use: the affinity mask of this process, NOT OMP_NUM_THREADS (torchrun exports
    OMP_NUM_THREADS=1, which would silently turn the all-cores baseline into a one-thread one)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def workload_geometry():
    grid = G.Grid((SHAPE[2], SHAPE[1], SHAPE[0]), SPACING,
                  tuple(-0.5 * np.array(SPACING) * np.array(SHAPE[::-1])))
    M = G.params_to_matrix(GT, (0.0, 0.0, 0.0))
    origin, dims = G.autocrop_extent(grid, M, SPACING, convention="vtk")
    out_grid = G.Grid(dims, SPACING, origin)
    A = G.compose_index_matrix(out_grid, grid, M, convention="vtk")
    return grid, A, dims


def config_dict(world, dims):
    """The ``config`` object of the JSON line -- the SAME for the GPU arm and for ``--impl reference`` at a given N."""
    return {"workload": WORKLOAD, "in_dims_xyz": [SHAPE[2], SHAPE[1], SHAPE[0]], "out_dims_xyz": list(dims),
            "voxels_out_per_step": int(dims[0] * dims[1] * dims[2]),
            "l2": "ring of input/output buffer sets larger than the 126 MB L2 (>= 3 sets, >= 400 MB per rank)",
            "parallelism": "one GPU" if world == 1 else
            "ONE volume, output lattice cut into %d z-slabs; each rank holds its moving planes + 1-plane halo; no collective" % world}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ncu_traffic():
    """DRAM traffic of the timed kernel per launch.  It cannot be measured inside this run (a number taken under a
    profiler is never a bench value), so it is read from the committed ``ncu --set full`` capture of the SAME kernel on
    the SAME workload and labelled as such; ``None`` when that capture is of another kernel."""
    path = os.path.join(ROOT, TRAFFIC_PROFILE)
    try:
        import csv
        with open(path, newline="") as fh:
            rows = list(csv.reader(fh))
        hdr, unit, row = rows[0], rows[1], rows[2]
        if TIMED_KERNEL not in row[hdr.index("Kernel Name")]:
            return None, "committed capture %s is of another kernel" % TRAFFIC_PROFILE
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = 0.0
        for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(name)
            tot += float(row[i]) * scale[unit[i]]
        return tot, "%s: ncu --set full of %s, one launch of this workload, captured in a separate run of this round" % (
            TRAFFIC_PROFILE, row[hdr.index("Kernel Name")][:60])
    except Exception as exc:
        return None, "no committed capture (%s)" % type(exc).__name__


def resample_parity(got, want, A, mdims, tag):
    """Full-size comparison of a GPU-resampled int16 volume with the CPU port's: +-1 LSB everywhere except on exact
    ties of the inside test (a sample within 1e-4 voxel of the half-voxel border plane may fall on either side: one
    arithmetic says background, the other a value)."""
    d = np.abs(got.astype(np.int32) - want.astype(np.int32))
    gt = np.argwhere(d > 1)                                        # [n, 3] as (k, j, i)
    in_band = 0
    if len(gt):
        A = np.asarray(A, dtype=np.float64).reshape(3, 4)
        c = gt[:, ::-1].astype(np.float64) @ A[:, :3].T + A[:, 3]
        near = np.zeros(len(gt), dtype=bool)
        for a in range(3):
            near |= (np.abs(c[:, a] + 0.5) < 1e-4) | (np.abs(c[:, a] - (mdims[a] - 0.5)) < 1e-4)
        in_band = int(near.sum())
    return {"config": tag, "voxels": int(d.size), "max_abs_lsb": int(d.max()), "n_off_by_one": int((d == 1).sum()),
            "n_gt_one": int(len(gt)), "n_gt_one_outside_band": int(len(gt) - in_band), "band_voxels": in_band}


def _sync_max(torch, dist, dev, world, seconds):
    t = torch.tensor([seconds], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def _all_true(torch, dist, dev, world, flag):
    t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(int(t[0]))


def sharded_registration_extra(torch, dist, dev, world, rank, peak):
    """Config 5: CT 1024x1024x1200 int16 at 0.5 mm (seeded phantom generated on the device, ground-truth offset as config
    1), 50 iterations of SSD registration, 13 probes each, the fixed lattice z-slab sharded over the ranks: per
    iteration one fused launch, ONE all-reduce of {ssd, count}[13] as 26 float64, the device step; ten iterations per
    CUDA graph.  Collective call: every rank runs it.  Parity: one 64-plane slab against the CPU port, and at N > 1
    the all-reduced slab sums of the 13 first-iteration probes against the same rank's whole-lattice evaluation."""
    from medicalimageanalysis_b200 import Data, RegistrationConfig, kernels, parallel
    from medicalimageanalysis_b200.phantom import make_phantom_device
    from medicalimageanalysis_b200.registration import params_to_index_matrix, probe_params
    from medicalimageanalysis_b200.utils import CreateImageFromMask
    shape, sp = (1200, 1024, 1024), (0.5, 0.5, 0.5)
    origin = (-256.0, -256.0, -300.0)
    Mgt = G.params_to_matrix(GT, (0.0, 0.0, 0.0))
    fixed = make_phantom_device(shape, "CT", seed=5, spacing=sp, device=dev)
    moving = make_phantom_device(shape, "CT", seed=5, spacing=sp, world_matrix=np.linalg.inv(Mgt), device=dev)
    Data.clear()
    f = CreateImageFromMask(fixed, origin, sp, "fixed", device=dev).add_image()
    m = CreateImageFromMask(moving, origin, sp, "moving", device=dev).add_image()
    out = {"n_gpus": world, "iterations": 50, "graph_iters": 10,
           "data": "make_phantom_device(seed=5), moving under the config-1 ground-truth offset"}

    # ---- parity 1: a 64-plane slab of the centre probe against the CPU port (moving planes: slab + halo only)
    center = np.asarray(f.compute_center(), dtype=np.float64)
    p2w, w2p = G.index_to_physical(f.grid, "itk"), G.physical_to_index(m.grid, "itk")
    deltas = np.array([0.5] * 3 + [0.5] * 3)
    A13 = np.stack([params_to_index_matrix(p, center, p2w, w2p) for p in probe_params(np.zeros(6), 1.0, deltas)])
    if rank == 0:
        from oracle import native
        native.build(); native.set_threads(host_threads())
        k0, k1 = 512, 576
        z0, z1 = G.slab_moving_zrange(A13[0], (shape[2], shape[1], shape[0]), k0, k1, shape[0])
        As = parallel.shift_matrix(A13[0], k0)
        s, c = kernels.resample_ssd(fixed[k0:k1], moving, A=As, mode="itk")
        As_sub = As.copy(); As_sub[2, 3] -= z0
        want, n = native.reslice_ssd(fixed[k0:k1].cpu().numpy(), moving[z0:z1].cpu().numpy(), As_sub, "itk")
        out["parity_slab_vs_cpu_port"] = {"planes": [k0, k1], "ssd_rel_diff": abs(float(s[0]) - want) / want,
                                          "count_diff": int(c[0]) - int(n), "voxels": int((k1 - k0) * shape[1] * shape[2])}

    # ---- parity 2 (N > 1): z-slab sums after the all-reduce == whole lattice on one GPU
    if world > 1:
        def ev_ssd(k0, k1, As):
            s, c = kernels.resample_ssd(fixed[k0:k1], moving, A=np.asarray(As), mode="itk")
            return s.clone(), c.clone()

        def ev_hist(k0, k1, As):
            h, c = kernels.resample_jointhist(fixed[k0:k1], moving, (-1024.0, 3071.0), (-1024.0, 3071.0), bins=64,
                                              A=np.asarray(As), mode="itk")
            return h.clone(), c.clone()
        ssd, cnt = parallel.sharded_ssd(ev_ssd, shape[0], A13, rank, world)
        s1, c1 = ev_ssd(0, shape[0], A13)
        h, hc = parallel.sharded_hist(ev_hist, shape[0], A13[:2], 64, rank, world)
        h1, hc1 = ev_hist(0, shape[0], A13[:2])
        ident = bool(torch.equal(cnt, c1) and torch.equal(h, h1.to(torch.int64)) and torch.equal(hc, hc1))
        rel = float(((ssd - s1).abs() / s1).max())
        out["bit_identical_hist_count"] = _all_true(torch, dist, dev, world, ident)
        out["ssd_rel_diff"] = _sync_max(torch, dist, dev, world, rel)

    # ---- timing
    cfg = RegistrationConfig(metric="ssd", iterations=50, graph_iters=10)
    parallel.register_rigid_sharded(f, m, cfg=RegistrationConfig(metric="ssd", iterations=12, graph_iters=10))
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    e0.record()
    res = parallel.register_rigid_sharded(f, m, cfg=cfg)
    e1.record()
    torch.cuda.synchronize(dev)
    dt = _sync_max(torch, dist, dev, world, e0.elapsed_time(e1) * 1e-3)
    nv = fixed.numel()
    out.update({"iters_per_s": res.iterations / dt, "evals_per_s": res.evaluations / dt,
                "ms_per_evaluation": 1e3 * dt / res.evaluations, "gvoxel_s": nv * res.evaluations / dt / 1e9,
                "hbm_frac_launch_level_per_gpu": (2.0 * nv + 13 * 2.0 * nv) * res.iterations / dt / 1e9 / peak / world,
                "hbm_frac_per_eval_L2_amortised": 4.0 * nv * res.evaluations / dt / 1e9 / peak / world,
                "final_params": [round(float(v), 3) for v in res.params], "gt_params": list(GT),
                "mse_first_last": [float(res.cost_history[0]), float(res.cost_history[-1])],
                "nvlink_bytes_per_iteration_per_rank": 208 if world > 1 else 0,
                "note": "hbm_frac_per_eval counts 4 B per voxel per probe although the 13 probes of a launch share one "
                        "staged box and one read of the fixed image (L2 / shared-memory amortised: not HBM utilisation); "
                        "the launch-level fraction counts N_o*b_f + 13*N_m*b_m per launch"})
    if world > 1:
        try:
            cfgf = RegistrationConfig(metric="ssd", iterations=50, graph_iters=10, collective="fused")
            parallel.register_rigid_sharded(f, m, cfg=RegistrationConfig(metric="ssd", iterations=12, graph_iters=10, collective="fused"))
            torch.cuda.synchronize(dev)
            dist.barrier()
            e0.record()
            resf = parallel.register_rigid_sharded(f, m, cfg=cfgf)
            e1.record()
            torch.cuda.synchronize(dev)
            dtf = _sync_max(torch, dist, dev, world, e0.elapsed_time(e1) * 1e-3)
            out["fused_exchange"] = {"iters_per_s": resf.iterations / dtf,
                                     "same_accept_reject_sequence": bool(np.array_equal(res.scale_history, resf.scale_history)),
                                     "cost_rel_diff_max": float(np.max(np.abs(res.cost_history - resf.cost_history) / np.abs(res.cost_history))),
                                     "nvlink_bytes_per_iteration_per_rank": 216 * world}
        except Exception as exc:
            out["fused_exchange"] = {"error": repr(exc)}
    Data.clear()
    del fixed, moving, f, m
    torch.cuda.empty_cache()
    return out


def batch_extra(torch, dist, dev, world, rank, peak):
    """Config 4: 16 CT volumes 512x512x300 (seeded device phantoms 100..115, each moving under its own random rigid,
    |theta| <= 5 deg, |t| <= 10 mm, default_rng(1000 + i)), each resampled onto its fixed lattice with the SSD fused into
    the same launch (6 algorithmic bytes per voxel); sharded by VOLUME over the ranks, no data-path collective -- only
    the 16 {ssd, count} pairs are gathered.  Parity: volume 0 against the CPU port."""
    from medicalimageanalysis_b200 import kernels
    from medicalimageanalysis_b200.phantom import make_phantom_device
    nvol = 16
    mine = [v for v in range(nvol) if v % world == rank]
    g = G.Grid((SHAPE[2], SHAPE[1], SHAPE[0]), SPACING, tuple(-0.5 * np.array(SPACING) * np.array(SHAPE[::-1])))
    pairs, mats = [], []
    for v in mine:
        rng = np.random.default_rng(1000 + v)
        prm = np.concatenate([rng.uniform(-5, 5, 3), rng.uniform(-10, 10, 3)])
        M = G.params_to_matrix(prm, (0.0, 0.0, 0.0))
        f = make_phantom_device(SHAPE, "CT", seed=100 + v, spacing=SPACING, device=dev)
        m = make_phantom_device(SHAPE, "CT", seed=100 + v, spacing=SPACING, world_matrix=np.linalg.inv(M), device=dev)
        pairs.append((f, m, torch.empty_like(m)))
        mats.append(G.compose_index_matrix(g, g, M))
    ws = kernels.MetricWorkspace(dev, n_mats=1)
    res = torch.zeros((nvol, 2), dtype=torch.float64, device=dev)

    def run_all():
        for v, (f, m, o), A in zip(mine, pairs, mats):
            s, c = kernels.resample_ssd(f, m, A=A, mode="itk", ws=ws, out=o)
            res[v, 0] = s[0]
            res[v, 1] = c[0].to(torch.float64)
    run_all()
    torch.cuda.synchronize(dev)
    parity = None
    if 0 in mine:
        from oracle import native
        native.build(); native.set_threads(host_threads())
        f, m, o = pairs[mine.index(0)]
        want, n, wout = native.reslice_ssd(f.cpu().numpy(), m.cpu().numpy(), mats[mine.index(0)], "itk", store=True)
        parity = resample_parity(o.cpu().numpy(), wout, mats[mine.index(0)], (SHAPE[2], SHAPE[1], SHAPE[0]), "C4 volume 0")
        parity.update(ssd_rel_diff=abs(float(res[0, 0]) - want) / want, count_diff=int(res[0, 1]) - int(n))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    reps = 3
    e0.record()
    for _ in range(reps):
        run_all()
    e1.record()
    torch.cuda.synchronize(dev)
    dt = _sync_max(torch, dist, dev, world, e0.elapsed_time(e1) * 1e-3 / reps)
    if world > 1:
        dist.all_reduce(res, op=dist.ReduceOp.SUM)                    # every rank filled only its own rows
    nv = nvol * SHAPE[0] * SHAPE[1] * SHAPE[2]
    out = {"volumes": nvol, "n_gpus": world, "ms_per_batch": dt * 1e3, "gvoxel_s": nv / dt / 1e9,
           "hbm_frac_per_gpu": 6.0 * nv / dt / 1e9 / peak / world,
           "mse_first_last": [float(res[0, 0] / res[0, 1]), float(res[nvol - 1, 0] / res[nvol - 1, 1])],
           "data": "make_phantom_device(seed=100+i), moving under default_rng(1000+i) rigid",
           "note": "sharded by volume, resample + SSD with store in one launch per volume, 16 {ssd,count} pairs gathered"}
    if parity is not None:
        out["parity_vs_cpu_port"] = parity
    del pairs
    torch.cuda.empty_cache()
    return out


def c3_pair(dev):
    """Config 3 pair: fixed CT-like float32 256^3 1 mm (seed 3), moving MR-like under (4,3,-5 deg; 6,-4,3 mm)."""
    from medicalimageanalysis_b200 import Data
    from medicalimageanalysis_b200.utils import CreateImageFromMask
    shape3, sp3 = (256, 256, 256), (1.0, 1.0, 1.0)
    gt3 = (4.0, 3.0, -5.0, 6.0, -4.0, 3.0)
    Data.clear()
    fixed = make_phantom(shape3, "CT", seed=3, spacing=sp3).astype(np.float32)
    f3 = CreateImageFromMask(fixed, (-128.0, -128.0, -128.0), sp3, "fixed", device=dev).add_image()
    M3 = G.params_to_matrix(gt3, f3.compute_center())
    moving = make_phantom(shape3, "MR", seed=3, spacing=sp3, world_matrix=np.linalg.inv(M3))
    m3 = CreateImageFromMask(moving, (-128.0, -128.0, -128.0), sp3, "moving", device=dev).add_image()
    return f3, m3, gt3, fixed, moving


def probe_parallel_extra(torch, dist, dev, world):
    """Config 3 (256^3 float32 CT -> MR, 64-bin MI, 200 iterations) with the 13 probes of every iteration dealt out
    over the ranks (the volume is too small for z-slabs), ten iterations per CUDA graph, with both
    exchanges: ONE NCCL all-reduce of 26 float64 per iteration, and the fused exchange (partials pushed into the peers'
    symmetric buffers over NVLink, the device step waits on flags: no NCCL call).  Collective: every rank calls it."""
    from medicalimageanalysis_b200 import Data, RegistrationConfig, parallel
    f3, m3, gt3, _, _ = c3_pair(dev)
    out = {"n_gpus": world, "gt_params": list(gt3)}
    ref = None
    for coll in ("nccl", "fused"):
        try:
            parallel.register_rigid_probe_parallel(f3, m3, cfg=RegistrationConfig(metric="mi", iterations=21, bins=64,
                                                                                 graph_iters=10, collective=coll))
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if world > 1:
                dist.barrier()
            e0.record()
            res = parallel.register_rigid_probe_parallel(f3, m3, cfg=RegistrationConfig(metric="mi", iterations=200, bins=64,
                                                                                       graph_iters=10, collective=coll))
            e1.record()
            torch.cuda.synchronize(dev)
            dt = _sync_max(torch, dist, dev, world, e0.elapsed_time(e1) * 1e-3)
            out[coll] = {"iters_per_s": res.iterations / dt, "evals_per_s": res.evaluations / dt,
                         "final_params": [round(float(v), 3) for v in res.params]}
            if ref is None:
                ref = res
            else:
                out["fused_trajectory_identical_to_nccl"] = _all_true(torch, dist, dev, world, bool(
                    np.array_equal(ref.cost_history, res.cost_history) and np.array_equal(ref.params, res.params)))
        except Exception as exc:
            out[coll] = {"error": repr(exc)}
    Data.clear()
    out["note"] = ("13 probes per iteration dealt out over the ranks, volumes replicated, 208 B exchanged per iteration and "
                   "rank, 10 iterations per CUDA graph")
    return out


def cpu_reslice_runs(mov, A, dims, runs, warm=1):
    """The CPU arm: this repo's C/OpenMP restatement of the reference path (oracle/csrc/oracle_omp.c) on ALL host threads
    of this process's affinity mask.  (VTK / SimpleITK, the reference's real native code, are not installable in this
    image.)  It is an untuned float64 statement of the algorithm (per-voxel floors, no row-incremental coordinates): a
    fair baseline for WHAT is computed, a soft one for how fast a tuned CPU library could do it.
    Returns (voxels, times, threads, last output)."""
    from oracle import native
    native.build()
    threads = native.set_threads(host_threads())
    nvox = dims[0] * dims[1] * dims[2]
    out = None
    for _ in range(max(1, warm)):                         # page faults, thread pool
        out = native.reslice(mov, A, dims, -3001.0, "vtk")
    times = []
    for _ in range(runs):
        t0 = time.perf_counter()
        out = native.reslice(mov, A, dims, -3001.0, "vtk")
        times.append(time.perf_counter() - t0)
    return nvox, times, threads, out


def run_reference(args, rank):
    """--impl reference: the CPU implementation of the path on the box's host cores; same config / steps / warmup
    keys as the GPU arm.  Under torchrun rank 0 alone runs it."""
    if rank != 0:
        return
    grid, A, dims = workload_geometry()
    mov = make_phantom(SHAPE, "CT", seed=2, spacing=SPACING)
    W, K = max(3, args.warmup), max(1, args.steps)
    nvox, times, threads, _ = cpu_reslice_runs(mov, A, dims, K, warm=W)
    total = sum(times)
    value = nvox * len(times) / total / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": K, "warmup": W, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f64 interpolation, i16 storage",
        "data": "synthetic", "config": config_dict(args.gpus, dims),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS"),
                         "sample": "full %dx%dx%d output volume per step, %d steps after %d warm-up runs" % (dims + (K, W))},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if threads == 1 and (os.cpu_count() or 1) > 1:
        line["warning"] = "CPU arm ran on ONE thread of a multi-core host: do not form a ratio against it"
    print(json.dumps(line), file=_JSON_OUT, flush=True)


def extras(dev, torch, kernels, peak):
    """Secondary measurements of the same path (not the headline): FFS permutation, fused SSD with store
    (config 1), and the CT->MR MI registration loop (config 3)."""
    from medicalimageanalysis_b200 import Data, RegistrationConfig
    from medicalimageanalysis_b200.utils import CreateImageFromMask
    from medicalimageanalysis_b200.utils import rigid as R
    out = {}

    def time_launches(fn, n=20, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize(dev)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        for a, b in ev:
            a.record(); fn(); b.record()
        torch.cuda.synchronize(dev)
        return statistics.median(a.elapsed_time(b) for a, b in ev) * 1e-3

    # FFS permutation of the C2 volume (exact copy, 4 B / voxel)
    vols = [torch.randint(-1024, 3071, SHAPE, device=dev, dtype=torch.int16) for _ in range(2)]
    outs = [torch.empty((SHAPE[0], SHAPE[2], SHAPE[1]), device=dev, dtype=torch.int16) for _ in range(2)]
    state = {"i": 0}
    for case in ("axial_k1", "axial_k2", "axial_k3"):
        perm, flip = kernels.FFS_PERMUTATIONS[case]

        def fn():
            i = state["i"] = (state["i"] + 1) % 2
            kernels.permute_axes(vols[i], perm, flip, out=outs[i].view(tuple(vols[i].shape[p] for p in perm)))
        t = time_launches(fn)
        nv = vols[0].numel()
        out["ffs_" + case] = {"gvoxel_s": nv / t / 1e9, "hbm_frac": 4.0 * nv / t / 1e9 / peak, "us": t * 1e6}
    del vols, outs

    # config 1: CT 512x512x100, one resample + SSD with store (6 B / voxel)
    shape1 = (100, 512, 512)
    g1 = G.Grid((512, 512, 100), SPACING, tuple(-0.5 * np.array(SPACING) * np.array(shape1[::-1])))
    M = G.params_to_matrix(GT, (0.0, 0.0, 0.0))
    A1 = G.compose_index_matrix(g1, g1, M)
    f1 = torch.from_numpy(make_phantom(shape1, "CT", seed=1, spacing=SPACING)).to(dev)
    m1 = torch.from_numpy(make_phantom(shape1, "CT", seed=1, spacing=SPACING, world_matrix=np.linalg.inv(M))).to(dev)
    o1 = torch.empty_like(m1)
    ws = kernels.MetricWorkspace(dev, n_mats=13)
    t = time_launches(lambda: kernels.resample_ssd(f1, m1, A=A1, mode="itk", ws=ws, out=o1))
    out["c1_resample_ssd_store"] = {"gvoxel_s": f1.numel() / t / 1e9, "hbm_frac": 6.0 * f1.numel() / t / 1e9 / peak,
                                   "us": t * 1e6}
    # parity at full size against the CPU port (the reference's CPU-runnable config)
    from oracle import native
    native.build(); native.set_threads(host_threads())
    s1, c1 = kernels.resample_ssd(f1, m1, A=A1, mode="itk", ws=ws, out=o1)
    want, n, wout = native.reslice_ssd(f1.cpu().numpy(), m1.cpu().numpy(), A1, "itk", store=True)
    par = resample_parity(o1.cpu().numpy(), wout, A1, (512, 512, 100), "C1")
    par.update(ssd_rel_diff=abs(float(s1[0]) - want) / want, count_diff=int(c1[0]) - int(n))
    out["c1_resample_ssd_store"]["parity_vs_cpu_port"] = par
    t = time_launches(lambda: kernels.resample_ssd(f1, m1, A=A1, mode="itk", ws=ws))
    out["c1_ssd_evaluate"] = {"gvoxel_s": f1.numel() / t / 1e9, "hbm_frac": 4.0 * f1.numel() / t / 1e9 / peak,
                              "us": t * 1e6}
    del f1, m1, o1

    # config 3: CT -> MR 256^3 float32, joint-histogram MI registration loop, 200 iterations, 13 probes each
    f3, m3, gt3, fixed, moving = c3_pair(dev)
    # parity of the FIRST iteration: the 13 joint histograms against the CPU port, bin for bin
    from medicalimageanalysis_b200.registration import params_to_index_matrix, probe_params
    c3c = np.asarray(f3.compute_center(), dtype=np.float64)
    A13 = np.stack([params_to_index_matrix(p, c3c, G.index_to_physical(f3.grid, "itk"), G.physical_to_index(m3.grid, "itk"))
                    for p in probe_params(np.zeros(6), 1.0, np.array([0.5] * 6))])
    fr, mr = (float(fixed.min()), float(fixed.max())), (float(moving.min()), float(moving.max()))
    ws13 = kernels.MetricWorkspace(dev, n_mats=13, bins=64)
    h, hc = kernels.resample_jointhist(f3.array, m3.array, fr, mr, bins=64, A=A13, mode="itk", ws=ws13)
    mi = kernels.mi_from_hist(h).cpu().numpy()
    hh, hcc = h.cpu().numpy(), hc.cpu().numpy()
    from oracle import metrics as oracle_metrics
    equal, mi_rel, diff_bins = True, 0.0, 0
    for q in range(13):
        H, n = native.joint_hist(fixed, moving, A13[q], 64, fr, mr, "itk")
        same = np.array_equal(hh[q].astype(np.int64), H.astype(np.int64)) and int(hcc[q]) == int(n)
        equal &= bool(same)
        diff_bins += int((hh[q].astype(np.int64) != H.astype(np.int64)).sum())
        ref_mi = oracle_metrics.oracle_mi(H)
        mi_rel = max(mi_rel, abs(float(mi[q]) - ref_mi) / abs(ref_mi))
    out["c3_first_iteration_parity"] = {"histograms_bit_equal": bool(equal), "bins_that_differ": diff_bins,
                                        "mi_rel_diff_max": mi_rel, "probes": 13, "bins": 64}
    del ws13, h, hc
    for metric in ("mi", "ssd"):
        cfg = RegistrationConfig(metric=metric, iterations=200, bins=64, graph_iters=10)
        R.register_rigid(f3, m3, cfg=RegistrationConfig(metric=metric, iterations=20, bins=64, graph_iters=10))
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        res = R.register_rigid(f3, m3, cfg=cfg)
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        nv = f3.array.numel()
        out["c3_register_" + metric] = {
            "iters_per_s": res.iterations / dt, "evals_per_s": res.evaluations / dt,
            "hbm_frac_launch_level": (4.0 * nv + 13 * 4.0 * nv) * res.iterations / dt / 1e9 / peak,
            "hbm_frac_per_eval_L2_amortised": 8.0 * nv * res.evaluations / dt / 1e9 / peak,
            "role": "registration" if metric == "mi" else "timing exercise only (SSD across CT -> MR has no meaningful optimum)",
            "final_params": [round(float(x), 3) for x in res.params], "gt_params": list(gt3),
            "cost_first": float(res.cost_history[0]), "cost_last": float(res.cost_history[-1])}
    # config 3 with the Mattes (Parzen-window) metric, coarse-to-fine over a 2-level pyramid
    cfgm = RegistrationConfig(metric="mattes", iterations=100, bins=50, graph_iters=10)
    R.register_rigid_multires(f3, m3, cfg=RegistrationConfig(metric="mattes", iterations=20, bins=50), shrink_factors=(2, 1))
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    res = R.register_rigid_multires(f3, m3, cfg=cfgm, shrink_factors=(2, 1))
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    out["c3_register_mattes_multires"] = {
        "iters_per_s": 200 / dt, "evals_per_s": res.evaluations / dt, "levels": "(2,2,2) x 100 + (1,1,1) x 100 iterations",
        "final_params": [round(float(x), 3) for x in res.params], "gt_params": list(gt3)}
    Data.clear()
    del f3, m3

    # the C2 volume through a smooth displacement field (12 + 2 + 2 algorithmic bytes per voxel)
    z, y, x = (torch.arange(n, device=dev, dtype=torch.float32) for n in SHAPE)
    img = (600 * torch.sin(x[None, None, :] / 17) * torch.cos(y[None, :, None] / 23) * torch.sin(z[:, None, None] / 19)).to(torch.int16)
    field = torch.stack([4 * torch.sin(x[None, None, :] / 40 + c) * torch.cos(y[None, :, None] / 50) * torch.sin(z[:, None, None] / 30 + c)
                         for c in range(3)], dim=-1).contiguous()
    gd = G.Grid((SHAPE[2], SHAPE[1], SHAPE[0]), SPACING, (0.0, 0.0, 0.0))
    w2p = G.physical_to_index(gd, "itk")
    Ad, Bd = (w2p @ G.index_to_physical(gd, "itk"))[:3], w2p[:3, :3]
    od = torch.empty_like(img)
    t = time_launches(lambda: kernels.warp_dvf(img, field, Ad, Bd, out=od))
    out["n1_warp_dvf_i16"] = {"gvoxel_s": img.numel() / t / 1e9, "hbm_frac": 16.0 * img.numel() / t / 1e9 / peak, "us": t * 1e6}
    del img, field, od
    return out



_JSON_OUT = sys.stdout


def numa_pin(local_rank):
    """Keep this rank's host threads (and so the first touch of its pinned pages) on the NUMA node of its GPU when
    sysfs says which that is.  Best effort: returns a description for the JSON line."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = torch.cuda.get_device_properties(local_rank).pci_domain_id
        dev_id = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/local_cpulist" % (dom, bus, dev_id)
        cpus = set()
        for part in open(path).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return "%d cpus local to the GPU" % len(cpus)
    except Exception as exc:
        return "not pinned (%s)" % type(exc).__name__
    return "not pinned"


def main():
    # The contract is ONE JSON line on stdout, but NCCL writes its version banner there (neither NCCL_DEBUG nor
    # NCCL_DEBUG_FILE moved it on this image): file descriptor 1 is pointed at stderr for every library, and the JSON
    # line goes out through a private duplicate of the original stdout.
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    from medicalimageanalysis_b200 import kernels, parallel
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peak, peak_kind = measured_peak()
    W, K = max(3, args.warmup), max(1, args.steps)

    grid, A, dims = workload_geometry()
    oshape = (dims[2], dims[1], dims[0])
    nvox_out = dims[0] * dims[1] * dims[2]
    mov_h = make_phantom(SHAPE, "CT", seed=2, spacing=SPACING)          # the SAME volume on every rank
    # this rank's share: output planes [k0, k1), moving planes [z0, z1) = what they touch + a one-plane halo
    (k0, k1), (z0, z1), As = parallel.slab_plan(A, oshape, SHAPE[0], rank, world)
    if k1 <= k0 or z1 <= z0:
        raise SystemExit("rank %d of %d has no work on this lattice" % (rank, world))
    slab_shape = (k1 - k0, oshape[1], oshape[2])
    in_bytes, out_bytes = 2 * (z1 - z0) * SHAPE[1] * SHAPE[2], 2 * slab_shape[0] * slab_shape[1] * slab_shape[2]
    ring = max(3, -(-400_000_000 // (in_bytes + out_bytes)))            # nothing can stay in the 126 MB L2 between steps
    movs = [torch.from_numpy(mov_h[z0:z1]).to(dev) for _ in range(ring)]
    outs = [torch.empty(slab_shape, dtype=torch.int16, device=dev) for _ in range(ring)]

    def step(i):
        kernels.resample(movs[i % ring], As, slab_shape, background=-3001.0, mode="vtk", out=outs[i % ring])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for i in range(W):
        step(i)
    path_taken = int(kernels._lib.load().mia_debug_last_path())
    # The K timed steps are ONE CUDA graph of K kernel launches (step i on ring slot i % ring), replayed once inside the
    # timed region: at N = 8 a slab kernel takes ~30 us, while an eager call costs 56 us of ctypes + tensor-map + launch
    # path and even one graph launch per step leaves ~9 us between kernels -- un-graphed, the timed region would measure
    # the host.  (One graph per ring slot, replayed step by step between CUDA events, gives the per-kernel times below.)
    torch.cuda.synchronize(dev)
    eager_step = step
    graphs = []
    for i in range(ring):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            eager_step(i)
        graphs.append(g)
    whole = torch.cuda.CUDAGraph()
    with torch.cuda.graph(whole):
        for i in range(K):
            eager_step(i)

    def step(i):                                            # noqa: F811  (from here on a step is a graph replay)
        graphs[i % ring].replay()

    for i in range(2 * ring):                               # the first launch of a graph uploads it: not part of a step
        step(i)
    for _ in range(2):
        whole.replay()
    torch.cuda.synchronize(dev)
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    whole.replay()                                          # EXACTLY K steps
    e1.record()
    barrier()
    total_ms = own_total_ms = e0.elapsed_time(e1)
    # per-kernel times: the same K steps, one graph launch each, a CUDA event on both sides of every step
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for i in range(K):
        ev[i][0].record()
        step(i)
        ev[i][1].record()
    torch.cuda.synchronize(dev)
    kernel_ms = [a.elapsed_time(b) for a, b in ev]
    # keep the GPU busy a little longer so the clock sampler sees load even for a very short timed region
    t_end = time.perf_counter() + 0.6
    i = 0
    while time.perf_counter() < t_end:
        step(i); i += 1
        if i % 50 == 0:
            torch.cuda.synchronize(dev)
    torch.cuda.synchronize(dev)
    clocks = sampler.stop()

    # ---- N > 1: every rank's slab must be the bytes of the same planes of a one-GPU resample of the whole volume
    slabs_identical = None
    if world > 1:
        whole = kernels.resample(torch.from_numpy(mov_h).to(dev), A, oshape, background=-3001.0, mode="vtk")
        slabs_identical = _all_true(torch, dist, dev, world, bool(torch.equal(whole[k0:k1], outs[(K - 1) % ring])))
        del whole

    # ---- end to end through host buffers, H2D and D2H inside the timed region, result read on the host every step
    Ke = max(3, min(K, 10))
    e2e_info = {}
    if world == 1:
        pin_in = kernels.pinned_empty(SHAPE, np.int16)
        pin_in[...] = mov_h
        pin_out = kernels.pinned_empty(oshape, np.int16)
        for _ in range(2):
            kernels.resample_host(pin_in, A, oshape, mode="vtk", out=pin_out, device=local_rank)
        barrier()
        t0 = time.perf_counter()
        for _ in range(Ke):
            kernels.resample_host(pin_in, A, oshape, mode="vtk", out=pin_out, device=local_rank)
            _ = int(pin_out[0, 0, 0])
        barrier()
        e2e_s = time.perf_counter() - t0
        e2e_info = {"api": "mia_resample_rigid_host (pinned host in/out, H2D+kernel+D2H pipelined by z-slab)",
                    "h2d_bytes_per_step": 2 * mov_h.size, "d2h_bytes_per_step": 2 * nvox_out,
                    "max_abs_diff_vs_device_path": int(np.abs(pin_out.astype(np.int32) - outs[(K - 1) % ring].cpu().numpy().astype(np.int32)).max())}
        # What the link alone allows: the same bytes as bare copies, both directions at once, no kernel (the pipeline can
        # at best hide everything else behind the longer of the two)
        pin_floor = kernels.pinned_empty(oshape, np.int16)      # not pin_out: its bytes are compared with the device path below
        t_in, t_out = torch.from_numpy(pin_in), torch.from_numpy(pin_floor)
        d_in = torch.empty(SHAPE, dtype=torch.int16, device=dev)
        s_up, s_down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        floor = []
        for _ in range(4):
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            with torch.cuda.stream(s_up):
                d_in.copy_(t_in, non_blocking=True)
            with torch.cuda.stream(s_down):
                t_out.copy_(outs[(K - 1) % ring], non_blocking=True)
            torch.cuda.synchronize(dev)
            floor.append(time.perf_counter() - t0)
        del d_in, t_out, pin_floor
        e2e_info["link_floor_ms"] = 1e3 * min(floor[1:])
        e2e_info["link_floor_note"] = ("bare pinned H2D of the input and D2H of the output, concurrently, best of 3: the "
                                       "time the host link alone needs for one step")
        host_out = pin_out
    else:
        # ONE host volume in a shared mapping; each rank pins and moves only its own planes of it
        tag = os.environ.get("MASTER_PORT", "0")
        p_in, p_out = "/dev/shm/mia_bench_%s_in.raw" % tag, "/dev/shm/mia_bench_%s_out.raw" % tag
        pinned = numa_pin(local_rank)
        if rank == 0:
            for pth, nbytes in ((p_in, 2 * mov_h.size), (p_out, 2 * nvox_out)):
                with open(pth, "wb") as fh:
                    fh.truncate(nbytes)
        barrier()
        h_in = np.memmap(p_in, dtype=np.int16, mode="r+", shape=SHAPE)
        h_out = np.memmap(p_out, dtype=np.int16, mode="r+", shape=oshape)
        # first touch: every rank writes the planes it will read (all ranks hold the same seeded volume, so planes that
        # two halos share get the same bytes twice) and zeroes the planes it will fill -- the pages then sit on the
        # NUMA node of the GPU that moves them
        h_in[z0:z1] = mov_h[z0:z1]
        h_out[k0:k1] = 0
        barrier()
        if rank == 0:
            h_in[...] = mov_h                                          # planes no halo covers (none here) still get their data
        barrier()
        handles = [kernels.host_register(h_in[z0:z1]), kernels.host_register(h_out[k0:k1])]
        for _ in range(2):
            parallel.resample_host_slab(h_in, A, h_out, rank, world, device=local_rank)
        barrier()
        t0 = time.perf_counter()
        for _ in range(Ke):
            parallel.resample_host_slab(h_in, A, h_out, rank, world, device=local_rank)
            _ = int(h_out[k0, 0, 0])
            if world > 1:
                dist.barrier()                                          # the step ends when the WHOLE volume is back
        torch.cuda.synchronize(dev)
        e2e_s = time.perf_counter() - t0
        # rank 0 checks the assembled volume against its own one-GPU result (device path, whole volume), byte for byte
        ok = True
        if rank == 0:
            ref = kernels.resample(torch.from_numpy(mov_h).to(dev), A, oshape, background=-3001.0, mode="vtk").cpu().numpy()
            ok = bool(np.array_equal(np.asarray(h_out), ref))
            del ref
        ok = _all_true(torch, dist, dev, world, ok)
        for hnd in handles:
            kernels.host_unregister(hnd)
        del h_in, h_out
        barrier()
        if rank == 0:
            for pth in (p_in, p_out):
                try:
                    os.unlink(pth)
                except OSError:
                    pass
        hb = torch.tensor([float(in_bytes), float(out_bytes)], dtype=torch.float64, device=dev)
        dist.all_reduce(hb, op=dist.ReduceOp.SUM)
        e2e_info = {"api": "parallel.resample_host_slab: ONE shared pinned host volume, each rank moves its z-slab + halo "
                           "through mia_resample_rigid_host; a barrier closes every step",
                    "h2d_bytes_per_step": int(hb[0]), "d2h_bytes_per_step": int(hb[1]),
                    "identical_to_one_gpu_result": ok, "host_threads": pinned}

    tm = torch.tensor([total_ms, e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    total_ms, e2e_s = float(tm[0]), float(tm[1])

    # ---- CPU port beside it (rank 0, N = 1): the timed baseline AND the full-size parity check of this very output
    cpu_line, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu:
        nv, times, threads, cpu_out = cpu_reslice_runs(mov_h, A, dims, 3)
        cpu_line = {"value": nv / statistics.median(times) / 1e9, "unit": UNIT, "cores": threads, "kind": "port",
                    "sample": "full %dx%dx%d output volume, median of 3 runs (C/OpenMP restatement of vtkImageReslice, "
                              "float64, untuned)" % dims}
        parity = resample_parity(outs[(K - 1) % ring].cpu().numpy(), cpu_out, A, (SHAPE[2], SHAPE[1], SHAPE[0]), "C2")
        parity["e2e_host_path_identical_to_device_path"] = bool(np.array_equal(host_out, outs[(K - 1) % ring].cpu().numpy()))
        del cpu_out

    sharded = probe_par = batch = None
    if not args.no_extra:
        del movs, outs
        movs = outs = None
        torch.cuda.empty_cache()
        try:
            sharded = sharded_registration_extra(torch, dist, dev, world, rank, peak)
        except Exception as exc:               # secondary measurement: never take the headline down
            sharded = {"error": repr(exc)}
        if world > 1:
            try:
                probe_par = probe_parallel_extra(torch, dist, dev, world)
            except Exception as exc:
                probe_par = {"error": repr(exc)}
        try:
            batch = batch_extra(torch, dist, dev, world, rank, peak)
        except Exception as exc:
            batch = {"error": repr(exc)}

    if rank == 0:
        value = nvox_out * K / (total_ms * 1e-3) / 1e9
        # average launch duration over the timed region (this rank's K back-to-back launches, CUDA events on both sides)
        kavg = own_total_ms / K * 1e-3
        alg_bytes = float(in_bytes + out_bytes)             # N_m*b_m + N_o*b_o of this rank's launch, int16 both
        achieved = alg_bytes / kavg / 1e9
        traffic, traffic_src = ncu_traffic() if world == 1 else (None, "captured at N = 1 only")
        cfg = config_dict(world, dims)                      # identical to the reference arm's at this N
        launch = {"out_planes": [k0, k1], "moving_planes_resident": [z0, z1], "ring_sets": ring, "kernel_path": path_taken}
        if slabs_identical is not None:
            launch["slabs_bit_identical_to_one_gpu"] = slabs_identical
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": "f32 interpolation (Q31.32 thread base), i16 storage", "data": "synthetic",
            "config": cfg, "clocks": clocks,
            "e2e": dict({"value": nvox_out * Ke / e2e_s / 1e9, "unit": UNIT, "steps": Ke, "ms_per_step": 1e3 * e2e_s / Ke},
                        **e2e_info),
            "gpu_launches": K,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_kind": peak_kind,
                         "kernel": "tile_x2_kernel<vtk, Box<128,16,12>> (TMA-staged x-pair tile kernel, csrc/mia_tile_x2.cuh)"
                         if path_taken == 5 else "path %d" % path_taken,
                         "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms_avg": kavg * 1e3,
                         "kernel_ms_single_launch_avg": statistics.mean(kernel_ms), "kernel_ms_single_launch_min": min(kernel_ms),
                         "timing": "kernel_ms_avg = the timed region (one graph of K launches) / K; single_launch = the same "
                                   "steps launched one graph at a time with an event on both sides of each",
                         "of_rank": 0, "launch": launch, "parity_c2": parity},
            "cpu_baseline": cpu_line,
        }
        if parity is not None:
            line["parity"] = parity
        if world == 1 and not args.no_extra:
            try:
                line["extra"] = extras(dev, torch, kernels, peak)
            except Exception as exc:           # the headline must survive a failing secondary measurement
                line["extra"] = {"error": repr(exc)}
        if sharded is not None:
            line.setdefault("extra", {})["c5_sharded_register_ssd"] = sharded
        if probe_par is not None:
            line.setdefault("extra", {})["c3_probe_parallel_mi"] = probe_par
        if batch is not None:
            line.setdefault("extra", {})["c4_batch16_resample_ssd"] = batch
        print(json.dumps(line), file=_JSON_OUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
