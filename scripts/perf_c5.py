"""One optimiser iteration (13 probes, identity start, +-0.5 deg / +-0.5 mm) on a config-5-like lattice: int16
1024 x 1024 x Z at 0.5 mm isotropic, SSD through the multi-matrix tile kernel.  Usage: perf_c5.py [Z]"""
import os
import statistics
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medicalimageanalysis_b200 import _lib, geometry as G, kernels           # noqa: E402
from medicalimageanalysis_b200.registration import probe_params              # noqa: E402

dev = torch.device("cuda")
Z = int(sys.argv[1]) if len(sys.argv) > 1 else 304
shape, sp = (Z, 1024, 1024), (0.5, 0.5, 0.5)
z, y, x = (torch.arange(n, device=dev, dtype=torch.float32) for n in shape)
f = (600 * torch.sin(x[None, None, :] / 17) * torch.cos(y[None, :, None] / 23) * torch.sin(z[:, None, None] / 19)).to(torch.int16)
m = (600 * torch.sin(x[None, None, :] / 17 + 0.1) * torch.cos(y[None, :, None] / 23) * torch.sin(z[:, None, None] / 19 + 0.2)).to(torch.int16)
g = G.Grid((1024, 1024, Z), sp, tuple(-0.5 * np.array(sp) * np.array([1024, 1024, Z])))
p2w, w2p = G.index_to_physical(g), G.physical_to_index(g)
deltas = np.array([0.5] * 3 + [0.5] * 3)


def timed(fn, n=10, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in ev) * 1e3


ws = kernels.MetricWorkspace(dev, n_mats=13)
for name, base in (("identity", np.zeros(6)), ("gt", np.array([3.0, -2.0, 4.0, 5.0, -3.0, 7.0]))):
    A = np.stack([(w2p @ G.params_to_matrix(p, np.zeros(3)) @ p2w)[:3] for p in probe_params(base, 1.0, deltas)])
    dA = torch.from_numpy(A).to(dev)
    us = timed(lambda: kernels.resample_ssd(f, m, d_A=dA, ws=ws)) / 13
    s13, c13 = kernels.resample_ssd(f, m, d_A=dA, ws=ws)
    s13, c13 = s13.clone(), c13.clone()
    s1, c1 = kernels.resample_ssd(f, m, A=A[3])
    ok = int(c1[0]) == int(c13[3]) and abs(float(s1[0]) - float(s13[3])) <= 1e-7 * float(s13[3])
    print("%s: %.1f us per evaluation, %.1f Gvoxel/s, path %d, probe 3 equals its one-matrix launch: %s"
          % (name, us, f.numel() / us / 1e3, _lib.load().mia_debug_last_path(), ok))
