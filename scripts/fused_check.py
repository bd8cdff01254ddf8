"""Multi-rank check of the fused exchange (run under torchrun on N GPUs of one box): the probe-parallel loop with
collective="fused" must walk the NCCL loop's trajectory bit for bit (its sums are exact), the z-slab SSD loop must
follow the same accept / reject sequence with costs within 1e-12, graphed and un-graphed.  One JSON line on rank 0."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medicalimageanalysis_b200 import Data, RegistrationConfig, geometry as G, parallel as P   # noqa: E402
from medicalimageanalysis_b200.phantom import make_phantom                                       # noqa: E402
from medicalimageanalysis_b200.utils import CreateImageFromMask                                  # noqa: E402


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
    Data.clear()
    fi = CreateImageFromMask(fixed.astype(np.float32), g.origin, sp, "fixed", device=dev).add_image()
    mi = CreateImageFromMask(moving.astype(np.float32), g.origin, sp, "moving", device=dev).add_image()
    out, ok = {"world": world}, True
    for gi in (0, 5):
        for metric in ("ssd", "mi"):
            a = P.register_rigid_probe_parallel(fi, mi, cfg=RegistrationConfig(metric=metric, iterations=31, bins=64, graph_iters=gi))
            b = P.register_rigid_probe_parallel(fi, mi, cfg=RegistrationConfig(metric=metric, iterations=31, bins=64, graph_iters=gi,
                                                                               collective="fused"))
            same = bool(np.array_equal(a.cost_history, b.cost_history) and np.array_equal(a.params, b.params))
            out["probe_parallel_%s_graph%d_identical" % (metric, gi)] = same
            ok &= same
        a = P.register_rigid_sharded(fi, mi, cfg=RegistrationConfig(metric="ssd", iterations=31, graph_iters=gi))
        b = P.register_rigid_sharded(fi, mi, cfg=RegistrationConfig(metric="ssd", iterations=31, graph_iters=gi, collective="fused"))
        rel = float(np.max(np.abs(a.cost_history - b.cost_history) / np.abs(a.cost_history)))
        same = bool(np.array_equal(a.scale_history, b.scale_history)) and rel < 1e-12
        out["sharded_ssd_graph%d" % gi] = {"same_accept_reject": bool(np.array_equal(a.scale_history, b.scale_history)), "cost_rel": rel,
                                          "collective": b.extra["collective"]}
        ok &= same
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        out["all_ranks_ok"] = bool(flag[0])
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if not bool(flag[0]):
        raise SystemExit(1)


if __name__ == "__main__":
    main()
