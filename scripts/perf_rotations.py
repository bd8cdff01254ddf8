"""Resample time against rotation angle and axis (no rotation may fall to the per-voxel kernel).
An isotropic 384^3 int16 volume and the anisotropic config-2 volume (0.98 x 0.98 x 2.5 mm), resampled onto their own
lattice about the volume centre.  Prints one JSON line: microseconds and the kernel path per (axis, angle)."""
import json
import os
import statistics
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medicalimageanalysis_b200 import _lib, geometry as G, kernels      # noqa: E402

dev = torch.device("cuda:0")


def timed(fn, n=10, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in ev) * 1e3


out = {}
for tag, shape, sp in (("iso384", (384, 384, 384), (1.0, 1.0, 1.0)), ("c2", (300, 512, 512), (0.9765625, 0.9765625, 2.5))):
    z, y, x = (torch.arange(n, device=dev, dtype=torch.float32) for n in shape)
    vol = (600 * torch.sin(x[None, None, :] / 17) * torch.cos(y[None, :, None] / 23) * torch.sin(z[:, None, None] / 19)).to(torch.int16)
    res = torch.empty_like(vol)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    rows = {}
    for axis in range(3):
        for ang in (3, 15, 30, 45, 90):
            e = [0.0, 0.0, 0.0]
            e[axis] = float(ang)
            A = G.compose_index_matrix(g, g, G.params_to_matrix(e + [1.0, -2.0, 1.5], (0, 0, 0)), convention="itk")
            us = timed(lambda: kernels.resample(vol, A, shape, background=0.0, mode="itk", out=res))
            rows["r%s%d" % ("xyz"[axis], ang)] = [round(us, 1), int(_lib.load().mia_debug_last_path())]
    out[tag] = rows
    del vol, res
print(json.dumps(out))
