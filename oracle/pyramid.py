"""NumPy restatement of the block-mean pyramid level (oracle; test infrastructure only).  The reference has no
pyramid of its own (it asks Elastix for "NumberOfResolutions", ``utils/deformable/simpleitk.py:155``); the spec is
this repository's: mean over the fx * fy * fz block, partial blocks on the high edges average what they
cover, int16 rounded half-up.  PARITY UNPINNED by construction (no reference implementation to pin against)."""
from __future__ import annotations

import numpy as np


def oracle_shrink_mean(a, factors):
    fx, fy, fz = (int(v) for v in factors)
    nz, ny, nx = a.shape
    oz, oy, ox = -(-nz // fz), -(-ny // fy), -(-nx // fx)
    s = np.zeros((oz, oy, ox), dtype=np.float64)
    c = np.zeros((oz, oy, ox), dtype=np.float64)
    af = a.astype(np.float64)
    for dz in range(fz):
        for dy in range(fy):
            for dx in range(fx):
                blk = af[dz::fz, dy::fy, dx::fx]
                s[:blk.shape[0], :blk.shape[1], :blk.shape[2]] += blk
                c[:blk.shape[0], :blk.shape[1], :blk.shape[2]] += 1.0
    m = s / c
    if a.dtype.kind == "f":
        return m.astype(a.dtype)
    return np.floor(np.clip(m, -32768, 32767) + 0.5).astype(a.dtype)
