"""Smallest possible call into the tile kernel (debug aid; run under compute-sanitizer on the GPU box)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from medicalimageanalysis_b200 import _lib, kernels

rng = np.random.default_rng(0)
mov = rng.integers(-1000, 3000, size=(24, 40, 64)).astype(np.int16)
A = np.hstack([np.identity(3), np.zeros((3, 1))])
A[:, 3] = [0.25, -0.5, 0.75]
t = torch.from_numpy(mov).cuda()
out = kernels.resample(t, A, mov.shape, background=-3001, mode="vtk")
torch.cuda.synchronize()
print("path", _lib.load().mia_debug_last_path(), "sum", int(out.long().sum()))
