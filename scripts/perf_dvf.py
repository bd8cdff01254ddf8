"""Timing of the displacement-field kernels on the C2 shape: int16 CT 512x512x300 warped through a smooth
field (12 + 2 + 2 algorithmic bytes per voxel), and 1 M mesh points through the same field."""
import os
import statistics
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                            # noqa: E402
from medicalimageanalysis_b200 import geometry as G, kernels           # noqa: E402

dev = torch.device("cuda")
peak, _ = bench.measured_peak()
shape, sp = bench.SHAPE, bench.SPACING
z, y, x = (torch.arange(n, device=dev, dtype=torch.float32) for n in shape)
img = (600 * torch.sin(x[None, None, :] / 17) * torch.cos(y[None, :, None] / 23) * torch.sin(z[:, None, None] / 19)).to(torch.int16)
field = torch.stack([4 * torch.sin(x[None, None, :] / 40 + c) * torch.cos(y[None, :, None] / 50) * torch.sin(z[:, None, None] / 30 + c)
                     for c in range(3)], dim=-1).contiguous()
g = G.Grid((shape[2], shape[1], shape[0]), sp, (0.0, 0.0, 0.0))
w2p = G.physical_to_index(g, "itk")
A, B = (w2p @ G.index_to_physical(g, "itk"))[:3], w2p[:3, :3]
out = torch.empty_like(img)


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in ev) * 1e-3


t = timed(lambda: kernels.warp_dvf(img, field, A, B, out=out))
nv = img.numel()
print("warp_dvf i16 %dx%dx%d: %.1f us, %.1f Gvoxel/s, %.0f GB/s algorithmic (16 B/voxel) = %.1f %% of %.0f GB/s"
      % (shape[2], shape[1], shape[0], t * 1e6, nv / t / 1e9, 16.0 * nv / t / 1e9, 100 * 16.0 * nv / t / 1e9 / peak, peak))
imgf = img.float()
outf = torch.empty_like(imgf)
t = timed(lambda: kernels.warp_dvf(imgf, field, A, B, out=outf))
print("warp_dvf f32: %.1f us, %.1f Gvoxel/s, %.0f GB/s algorithmic (20 B/voxel) = %.1f %%"
      % (t * 1e6, nv / t / 1e9, 20.0 * nv / t / 1e9, 100 * 20.0 * nv / t / 1e9 / peak))
pts = torch.rand((1_000_000, 3), dtype=torch.float64, device=dev) * torch.tensor([511.0, 511.0, 299.0], dtype=torch.float64, device=dev)
t = timed(lambda: kernels.sample_field_points(field, pts))
print("sample_field_points 1M points x 3 channels: %.1f us (%.2f Gpoint/s)" % (t * 1e6, 1e6 / t / 1e9))
