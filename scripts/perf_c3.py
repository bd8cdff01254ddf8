"""Config 3 on the seeded phantoms (CT -> MR 256^3 float32; also the int16 config-1 pair): time of the 13 joint histograms
of one iteration and of the MI registration loop.  Usage: [MIA_B200_LIB=build/libX.so] python scripts/perf_c3.py"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                                                  # noqa: E402
from medicalimageanalysis_b200 import Data, RegistrationConfig, geometry as G, kernels        # noqa: E402
from medicalimageanalysis_b200.registration import params_to_index_matrix, probe_params      # noqa: E402
from medicalimageanalysis_b200.utils import rigid as R                                        # noqa: E402

dev = torch.device("cuda:0")
f3, m3, gt3, fixed, moving = bench.c3_pair(dev)
c = np.asarray(f3.compute_center(), dtype=np.float64)
p2w, w2p = G.index_to_physical(f3.grid, "itk"), G.physical_to_index(m3.grid, "itk")
out = {}
ws = kernels.MetricWorkspace(dev, n_mats=13, bins=64)
fr, mr = (float(fixed.min()), float(fixed.max())), (float(moving.min()), float(moving.max()))
for tag, centre in (("identity", np.zeros(6)), ("gt", np.array(gt3))):
    A13 = np.stack([params_to_index_matrix(q, c, p2w, w2p) for q in probe_params(centre, 1.0, np.array([0.5] * 6))])
    dA = torch.from_numpy(np.ascontiguousarray(A13)).to(dev)
    for _ in range(3):
        kernels.resample_jointhist(f3.array, m3.array, fr, mr, bins=64, d_A=dA, ws=ws)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        h, hc = kernels.resample_jointhist(f3.array, m3.array, fr, mr, bins=64, d_A=dA, ws=ws)
    e1.record()
    torch.cuda.synchronize()
    out["hist13_us_per_eval_" + tag] = e0.elapsed_time(e1) * 1e3 / 130
    out["hist_sum_" + tag] = int(h.sum())
    out["hist_sha_" + tag] = hash(h.cpu().numpy().tobytes()) & 0xffffffff
for metric in ("mi", "ssd"):
    cfg = RegistrationConfig(metric=metric, iterations=200, bins=64, graph_iters=10)
    R.register_rigid(f3, m3, cfg=RegistrationConfig(metric=metric, iterations=20, bins=64, graph_iters=10))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = R.register_rigid(f3, m3, cfg=cfg)
    torch.cuda.synchronize()
    out["iters_per_s_" + metric] = res.iterations / (time.perf_counter() - t0)
    out["params_" + metric] = [round(float(v), 3) for v in res.params]
Data.clear()
print(json.dumps(out))
