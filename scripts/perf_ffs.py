"""FFS permutation timings on the C2 volume (int16 512x512x300) plus a bit-exactness check against torch.
Tuning: MIA_TR_DEPTH (tiles of loads in flight per block), MIA_TR_BLOCKS (persistent blocks per SM).
Usage: python scripts/perf_ffs.py [case ...]"""
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medicalimageanalysis_b200 import kernels      # noqa: E402

dev = torch.device("cuda")
PEAK = 6547.5


def timed(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in ev) * 1e3


cases = sys.argv[1:] or list(kernels.FFS_PERMUTATIONS)
vols = [torch.randint(-1024, 3071, (300, 512, 512), dtype=torch.int16, device=dev) for _ in range(3)]   # 3 x 157 MB > L2
res = {}
for case in cases:
    perm, flip = kernels.FFS_PERMUTATIONS[case]
    ref = vols[0].permute(*perm)
    fl = [a for a in range(3) if flip[a]]
    if fl:
        ref = ref.flip(fl)
    got = kernels.permute_axes(vols[0], perm, flip)
    ok = bool(torch.equal(got, ref.contiguous()))
    outs = [torch.empty_like(got) for _ in range(3)]
    it = [0]

    def run():
        i = it[0] % 3
        it[0] += 1
        kernels.permute_axes(vols[i], perm, flip, out=outs[i])
    us = timed(run)
    res[case] = (round(us, 1), round(2 * vols[0].numel() * 2 / us / 1e3 / PEAK, 3), ok)
print("depth=%s blocks=%s" % (os.environ.get("MIA_TR_DEPTH", "-"), os.environ.get("MIA_TR_BLOCKS", "-")),
      " ".join("%s=%s" % kv for kv in res.items()))
