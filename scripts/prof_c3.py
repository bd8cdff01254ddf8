"""Config-3-shaped launches for ncu: 256^3 float32 pair, 13 candidate matrices, fused SSD and joint histogram."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from medicalimageanalysis_b200 import geometry as G, kernels

dev = torch.device("cuda")
n = 256
g = G.Grid((n, n, n), (1.0, 1.0, 1.0), (-128.0, -128.0, -128.0))
x = torch.arange(n, device=dev, dtype=torch.float32)
f = (600 * torch.sin(x[None, None, :] / 17) * torch.cos(x[None, :, None] / 23) * torch.sin(x[:, None, None] / 19) + 1000).contiguous()
m = (500 * torch.cos(x[None, None, :] / 15) * torch.sin(x[None, :, None] / 21) * torch.cos(x[:, None, None] / 18) + 1200).contiguous()
rng = np.random.default_rng(0)
base = np.array([4.0, 3.0, -5.0, 6.0, -4.0, 3.0])
A = np.stack([G.compose_index_matrix(g, g, G.params_to_matrix(base + rng.uniform(-0.5, 0.5, 6), (0, 0, 0))) for _ in range(13)])
dA = torch.from_numpy(A).to(dev)
ws = kernels.MetricWorkspace(dev, n_mats=13, bins=64)
fr, mr = (float(f.min()), float(f.max())), (float(m.min()), float(m.max()))
for _ in range(3):
    kernels.resample_ssd(f, m, d_A=dA, ws=ws)
    kernels.resample_jointhist(f, m, fr, mr, bins=64, d_A=dA, ws=ws)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
ev[0].record()
for _ in range(10):
    kernels.resample_ssd(f, m, d_A=dA, ws=ws)
ev[1].record()
for _ in range(10):
    kernels.resample_jointhist(f, m, fr, mr, bins=64, d_A=dA, ws=ws)
ev[2].record()
torch.cuda.synchronize()
print("ssd  13 probes: %.1f us per launch, %.1f us per evaluation" % (ev[0].elapsed_time(ev[1]) * 100, ev[0].elapsed_time(ev[1]) * 100 / 13))
print("hist 13 probes: %.1f us per launch, %.1f us per evaluation" % (ev[1].elapsed_time(ev[2]) * 100, ev[1].elapsed_time(ev[2]) * 100 / 13))
from medicalimageanalysis_b200 import _lib
print("last path:", _lib.load().mia_debug_last_path())
