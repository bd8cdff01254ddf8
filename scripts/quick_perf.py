"""Kernel timings for fast iteration (a few seconds on the GPU; not the bench contract):
C2 int16 resample, config-1 int16 SSD / joint histogram, config-3 float32 SSD / joint histogram with 13 probes.
Usage: [MIA_B200_LIB=build/libX.so] python scripts/quick_perf.py"""
import os
import statistics
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                            # noqa: E402
from medicalimageanalysis_b200 import _lib, geometry as G, kernels      # noqa: E402

dev = torch.device("cuda")


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in ev) * 1e3


def synth(shape, dtype, k):
    z, y, x = (torch.arange(n, device=dev, dtype=torch.float32) for n in shape)
    v = 600 * torch.sin(x[None, None, :] / (17 + k)) * torch.cos(y[None, :, None] / (23 - k)) * torch.sin(z[:, None, None] / 19) + 200 * k
    return v.to(dtype).contiguous()


res = {}
# C2: the bench workload (auto-cropped output, VTK mode)
grid, A, dims = bench.workload_geometry()
mov = synth(bench.SHAPE, torch.int16, 0)
out = torch.empty(dims[::-1], dtype=torch.int16, device=dev)
res["c2_resample_i16"] = timed(lambda: kernels.resample(mov, A, dims[::-1], background=-3001.0, mode="vtk", out=out))
res["c2_path"] = _lib.load().mia_debug_last_path()
del mov, out

# config 1 shape, int16: SSD and joint histogram, one host matrix
sh1 = (100, 512, 512)
g1 = G.Grid((512, 512, 100), bench.SPACING, tuple(-0.5 * np.array(bench.SPACING) * np.array(sh1[::-1])))
A1 = G.compose_index_matrix(g1, g1, G.params_to_matrix(bench.GT, (0.0, 0.0, 0.0)))
f1, m1 = synth(sh1, torch.int16, 1), synth(sh1, torch.int16, 2)
ws = kernels.MetricWorkspace(dev, n_mats=13, bins=64)
fr, mr = (float(f1.min()), float(f1.max())), (float(m1.min()), float(m1.max()))
res["c1_ssd_i16"] = timed(lambda: kernels.resample_ssd(f1, m1, A=A1, mode="itk", ws=ws))
res["c1_hist_i16"] = timed(lambda: kernels.resample_jointhist(f1, m1, fr, mr, bins=64, A=A1, mode="itk", ws=ws))
# 13 probes around one pose, device matrices (one registration iteration on int16 data)
rng = np.random.default_rng(1)
base1 = np.array(bench.GT, dtype=np.float64)
A13 = np.stack([G.compose_index_matrix(g1, g1, G.params_to_matrix(base1 + rng.uniform(-0.5, 0.5, 6), (0.0, 0.0, 0.0))) for _ in range(13)])
dA1 = torch.from_numpy(A13).to(dev)
res["c1_ssd_i16_13_per_eval"] = timed(lambda: kernels.resample_ssd(f1, m1, d_A=dA1, mode="itk", ws=ws)) / 13
res["path"] = _lib.load().mia_debug_last_path()
del f1, m1

# config 3 shape, float32, 13 device matrices per launch
n = 256
g3 = G.Grid((n, n, n), (1.0, 1.0, 1.0), (-128.0, -128.0, -128.0))
f3, m3 = synth((n, n, n), torch.float32, 1), synth((n, n, n), torch.float32, 3)
rng = np.random.default_rng(0)
base = np.array([4.0, 3.0, -5.0, 6.0, -4.0, 3.0])
A3 = np.stack([G.compose_index_matrix(g3, g3, G.params_to_matrix(base + rng.uniform(-0.5, 0.5, 6), (0, 0, 0))) for _ in range(13)])
dA = torch.from_numpy(A3).to(dev)
fr, mr = (float(f3.min()), float(f3.max())), (float(m3.min()), float(m3.max()))
res["c3_ssd_f32_per_eval"] = timed(lambda: kernels.resample_ssd(f3, m3, d_A=dA, ws=ws)) / 13
res["c3_hist_f32_per_eval"] = timed(lambda: kernels.resample_jointhist(f3, m3, fr, mr, bins=64, d_A=dA, ws=ws)) / 13
print(" ".join("%s=%.1f" % kv for kv in res.items()))
