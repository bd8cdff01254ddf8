"""C2 resample launches for ncu (the bench workload without the bench harness)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                    # noqa: E402
from medicalimageanalysis_b200 import kernels                    # noqa: E402
from medicalimageanalysis_b200.phantom import make_phantom       # noqa: E402

dev = torch.device("cuda")
grid, A, dims = bench.workload_geometry()
mov = torch.from_numpy(make_phantom(bench.SHAPE, "CT", seed=2, spacing=bench.SPACING)).to(dev)
out = torch.empty(dims[::-1], dtype=torch.int16, device=dev)
for _ in range(6):
    kernels.resample(mov, A, dims[::-1], background=-3001.0, mode="vtk", out=out)
torch.cuda.synchronize()
print("ok", int(out.sum()))
