"""Kernel development probe: times the C2 resample (and C1 SSD) through the C ABI, prints the path taken and a hash of the
output so that two builds / two paths (MIA_NO_X2=1) can be compared bit for bit."""
import hashlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                                   # noqa: E402
from medicalimageanalysis_b200 import _lib, geometry as G, kernels  # noqa: E402
from medicalimageanalysis_b200.phantom import make_phantom      # noqa: E402


def timeit(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for i, (a, b) in enumerate(ev):
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return {"us_med": 1e3 * t[len(t) // 2], "us_min": 1e3 * t[0]}


def main():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    out = {"env_no_x2": os.environ.get("MIA_NO_X2", "")}
    grid, A, dims = bench.workload_geometry()
    oshape = (dims[2], dims[1], dims[0])
    mov_h = make_phantom(bench.SHAPE, "CT", seed=2, spacing=bench.SPACING)
    movs = [torch.from_numpy(mov_h).to(dev) for _ in range(3)]
    outs = [torch.empty(oshape, dtype=torch.int16, device=dev) for _ in range(3)]
    st = {"i": 0}

    def c2():
        i = st["i"] = (st["i"] + 1) % 3
        kernels.resample(movs[i], A, oshape, background=-3001.0, mode="vtk", out=outs[i])
    r = timeit(c2)
    r["path"] = int(lib.mia_debug_last_path())
    r["sha"] = hashlib.sha1(outs[0].cpu().numpy().tobytes()).hexdigest()[:16]
    r["gvoxel_s"] = dims[0] * dims[1] * dims[2] / r["us_med"] / 1e3
    r["frac"] = (2.0 * mov_h.size + 2.0 * dims[0] * dims[1] * dims[2]) / (r["us_med"] * 1e-6) / 1e9 / bench.measured_peak()[0]
    out["c2_store_vtk"] = r
    # one eighth of the lattice (the z-slab one of 8 GPUs computes): 11 tiles per CTA, where the tile order matters most
    nz8 = (dims[2] + 7) // 8 // 8 * 8
    k0 = 3 * nz8
    A8 = np.array(A, dtype=np.float64).reshape(3, 4).copy()
    A8[:, 3] += A8[:, 2] * k0
    o8 = [torch.empty((nz8, dims[1], dims[0]), dtype=torch.int16, device=dev) for _ in range(3)]

    def c2_slab():
        i = st["i"] = (st["i"] + 1) % 3
        kernels.resample(movs[i], A8, (nz8, dims[1], dims[0]), background=-3001.0, mode="vtk", out=o8[i])
    r = timeit(c2_slab)
    r["path"] = int(lib.mia_debug_last_path())
    r["identical_to_whole"] = bool(torch.equal(o8[0], outs[0][k0:k0 + nz8]))
    out["c2_slab_of_8"] = r
    # the same slabs as ONE graph of 20 launches (what bench.py replays at N > 1): no host launch path between kernels
    for parts in (8, 4, 2):
        nzp = (dims[2] + 7) // 8 // parts * 8
        kp = min((parts // 2) * nzp, (dims[2] - nzp) // 8 * 8)
        Ap = np.array(A, dtype=np.float64).reshape(3, 4).copy()
        Ap[:, 3] += Ap[:, 2] * kp
        op = [torch.empty((nzp, dims[1], dims[0]), dtype=torch.int16, device=dev) for _ in range(3)]
        side = torch.cuda.Stream(dev)
        with torch.cuda.stream(side):
            for i in range(3):
                kernels.resample(movs[i], Ap, (nzp, dims[1], dims[0]), background=-3001.0, mode="vtk", out=op[i])
            side.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for i in range(21):
                    kernels.resample(movs[i % 3], Ap, (nzp, dims[1], dims[0]), background=-3001.0, mode="vtk", out=op[i % 3])
        torch.cuda.synchronize()
        r = timeit(graph.replay, n=10, warm=3)
        out["c2_slab_of_%d_graph" % parts] = {"us_per_launch": r["us_med"] / 21, "planes": nzp,
                                               "identical_to_whole": bool(torch.equal(op[0], outs[0][kp:kp + nzp]))}
    if os.environ.get("K2_DEV_ONLY_C2"):
        print(json.dumps(out))
        return
    # ITK mode onto the fixed lattice (C4 flavour) with SSD + store, and SSD evaluate
    g1 = G.Grid((512, 512, 300), bench.SPACING, tuple(-0.5 * np.array(bench.SPACING) * np.array(bench.SHAPE[::-1])))
    M = G.params_to_matrix(bench.GT, (0.0, 0.0, 0.0))
    A1 = G.compose_index_matrix(g1, g1, M)
    fixed = torch.from_numpy(make_phantom(bench.SHAPE, "CT", seed=2, spacing=bench.SPACING, world_matrix=M)).to(dev)
    ws = kernels.MetricWorkspace(dev, n_mats=1)
    o = torch.empty_like(movs[0])
    r = timeit(lambda: kernels.resample_ssd(fixed, movs[0], A=A1, mode="itk", ws=ws, out=o))
    s, c = kernels.resample_ssd(fixed, movs[0], A=A1, mode="itk", ws=ws, out=o)
    r.update(path=int(lib.mia_debug_last_path()), ssd=float(s[0]), count=int(c[0]),
             sha=hashlib.sha1(o.cpu().numpy().tobytes()).hexdigest()[:16])
    out["c2_ssd_store_itk"] = r
    r = timeit(lambda: kernels.resample_ssd(fixed, movs[0], A=A1, mode="itk", ws=ws))
    s, c = kernels.resample_ssd(fixed, movs[0], A=A1, mode="itk", ws=ws)
    r.update(path=int(lib.mia_debug_last_path()), ssd=float(s[0]), count=int(c[0]))
    out["c2_ssd_eval_itk"] = r
    # config 1 (512x512x100): one optimiser iteration = 13 probes in one launch, device matrices (as the loop has them)
    from medicalimageanalysis_b200.registration import params_to_index_matrix, probe_params
    shape1 = (100, 512, 512)
    gq = G.Grid((512, 512, 100), bench.SPACING, tuple(-0.5 * np.array(bench.SPACING) * np.array(shape1[::-1])))
    f1 = torch.from_numpy(make_phantom(shape1, "CT", seed=1, spacing=bench.SPACING)).to(dev)
    m1 = torch.from_numpy(make_phantom(shape1, "CT", seed=1, spacing=bench.SPACING, world_matrix=np.linalg.inv(M))).to(dev)
    p2w, w2p = G.index_to_physical(gq, "itk"), G.physical_to_index(gq, "itk")
    ws13 = kernels.MetricWorkspace(dev, n_mats=13)
    for tag, centre in (("identity", np.zeros(6)), ("gt", np.array(bench.GT))):
        A13 = np.stack([params_to_index_matrix(q, np.zeros(3), p2w, w2p) for q in probe_params(centre, 1.0, np.array([0.5] * 6))])
        dA = torch.from_numpy(np.ascontiguousarray(A13)).to(dev)
        r = timeit(lambda: kernels.resample_ssd(f1, m1, d_A=dA, mode="itk", ws=ws13))
        s, c = kernels.resample_ssd(f1, m1, d_A=dA, mode="itk", ws=ws13)
        r.update(path=int(lib.mia_debug_last_path()), us_per_eval=r["us_med"] / 13, ssd0=float(s[0]), count0=int(c[0]),
                 gvoxel_s=13 * f1.numel() / r["us_med"] / 1e3)
        out["c1_multi13_" + tag] = r
    # config-5-like slab (1024x1024x304 at 0.5 mm): 13 probes per launch around identity and around the ground truth
    from medicalimageanalysis_b200.phantom import make_phantom_device
    del f1, m1, movs, outs, fixed, o
    torch.cuda.empty_cache()
    shape5, sp5 = (304, 1024, 1024), (0.5, 0.5, 0.5)
    g5 = G.Grid((1024, 1024, 304), sp5, (-256.0, -256.0, -76.0))
    f5 = make_phantom_device(shape5, "CT", seed=5, spacing=sp5, device=dev)
    m5 = make_phantom_device(shape5, "CT", seed=5, spacing=sp5, world_matrix=np.linalg.inv(M), device=dev)
    p2w, w2p = G.index_to_physical(g5, "itk"), G.physical_to_index(g5, "itk")
    for tag, centre in (("identity", np.zeros(6)), ("gt", np.array(bench.GT))):
        A13 = np.stack([params_to_index_matrix(q, np.zeros(3), p2w, w2p) for q in probe_params(centre, 1.0, np.array([0.5] * 6))])
        dA = torch.from_numpy(np.ascontiguousarray(A13)).to(dev)
        r = timeit(lambda: kernels.resample_ssd(f5, m5, d_A=dA, mode="itk", ws=ws13), n=6, warm=2)
        s, c = kernels.resample_ssd(f5, m5, d_A=dA, mode="itk", ws=ws13)
        r.update(path=int(lib.mia_debug_last_path()), us_per_eval=r["us_med"] / 13, ssd0=float(s[0]), count0=int(c[0]),
                 gvoxel_s=13 * f5.numel() / r["us_med"] / 1e3)
        out["c5slab_multi13_" + tag] = r
    print(json.dumps(out))


if __name__ == "__main__":
    main()
