"""Multi-rank check of the z-slab sharded metric path (run under torchrun on N GPUs of one box):
histograms / counts of the all-reduced slabs must equal rank 0's single-GPU result bit for bit, SSD to 1e-8,
and the sharded registration must walk the single-GPU trajectory.  Prints one JSON line on rank 0."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medicalimageanalysis_b200 import Data, RegistrationConfig, geometry as G, kernels, parallel as P   # noqa: E402
from medicalimageanalysis_b200.phantom import make_phantom                                               # noqa: E402
from medicalimageanalysis_b200.utils import CreateImageFromMask                                         # noqa: E402
from medicalimageanalysis_b200.utils import rigid as R                                                  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    shape, sp = (96, 128, 128), (1.0, 1.0, 1.5)
    fixed = make_phantom(shape, "CT", seed=3, spacing=sp)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    Mgt = G.params_to_matrix([2.0, -1.5, 2.5, 2.0, 1.0, -1.0], (0, 0, 0))
    moving = make_phantom(shape, "CT", seed=3, spacing=sp, world_matrix=np.linalg.inv(Mgt))
    f, m = torch.from_numpy(fixed).to(dev), torch.from_numpy(moving).to(dev)
    A = np.stack([G.compose_index_matrix(g, g, G.params_to_matrix(p, (0, 0, 0)))
                  for p in ([1, -1, 2, 1, 0, -1], [0, 0, 0, 0, 0, 0], [3, 2, -3, -2, 1, 2])])
    fr = (-1024.0, 3071.0)

    def ev_ssd(k0, k1, As):
        s, c = kernels.resample_ssd(f[k0:k1], m, A=np.asarray(As))
        return s.clone(), c.clone()

    def ev_hist(k0, k1, As):
        h, c = kernels.resample_jointhist(f[k0:k1], m, fr, fr, bins=64, A=np.asarray(As))
        return h.clone(), c.clone()

    ssd, cnt = P.sharded_ssd(lambda a, b, c: tuple(t.to(dev) for t in ev_ssd(a, b, c)), shape[0], A, rank, world)
    h, hc = P.sharded_hist(lambda a, b, c: tuple(t.to(dev) for t in ev_hist(a, b, c)), shape[0], A, 64, rank, world)
    s1, c1 = ev_ssd(0, shape[0], A)
    h1, hc1 = ev_hist(0, shape[0], A)
    ok_hist = bool(torch.equal(h, h1.to(torch.int64)) and torch.equal(hc, hc1) and torch.equal(cnt, c1))
    ssd_rel = float(((ssd - s1).abs() / s1).max())

    Data.clear()
    fi = CreateImageFromMask(fixed.astype(np.float32), g.origin, sp, "fixed", device=dev).add_image()
    mi = CreateImageFromMask(moving.astype(np.float32), g.origin, sp, "moving", device=dev).add_image()
    cfg = RegistrationConfig(metric="ssd", iterations=30, graph_iters=0)
    single = R.register_rigid(fi, mi, cfg=cfg)
    torch.cuda.synchronize(dev)
    dist.barrier()
    t0 = time.perf_counter()
    shard = P.register_rigid_sharded(fi, mi, cfg=cfg)
    torch.cuda.synchronize(dev)
    dist.barrier()
    dt = time.perf_counter() - t0
    same_path = bool(np.array_equal(single.scale_history, shard.scale_history))
    cost_rel = float(np.max(np.abs(single.cost_history - shard.cost_history) / np.abs(single.cost_history)))
    par_abs = float(np.max(np.abs(single.params - shard.params)))
    # probe-parallel (13 probes dealt out over the ranks, volumes replicated): bit-identical to one GPU
    pp = {}
    for metric in ("ssd", "mi"):
        cfgp = RegistrationConfig(metric=metric, iterations=30, bins=64, graph_iters=0)
        one = R.register_rigid(fi, mi, cfg=cfgp)
        torch.cuda.synchronize(dev)
        dist.barrier()
        t0 = time.perf_counter()
        par = P.register_rigid_probe_parallel(fi, mi, cfg=cfgp)
        torch.cuda.synchronize(dev)
        dist.barrier()
        pp[metric] = {"identical": bool(np.array_equal(one.cost_history, par.cost_history)
                                        and np.array_equal(one.params, par.params)),
                      "iters_per_s": par.iterations / (time.perf_counter() - t0)}
    flags = torch.tensor([int(ok_hist), int(same_path), int(ssd_rel < 1e-8), int(cost_rel < 1e-7),
                          int(pp["ssd"]["identical"]), int(pp["mi"]["identical"])], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"world": world, "hist_count_bit_identical": bool(flags[0]), "same_trajectory": bool(flags[1]),
                          "ssd_rel_err": ssd_rel, "cost_rel_err": cost_rel, "param_abs_err": par_abs,
                          "sharded_iters_per_s": shard.iterations / dt, "slab": shard.extra["slab"],
                          "probe_parallel": pp}), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if not bool(flags.min()):
        raise SystemExit(1)


if __name__ == "__main__":
    main()
