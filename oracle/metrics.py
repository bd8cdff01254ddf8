"""Oracle similarity metrics (test infrastructure only; see ``oracle/__init__.py``).

The reference's own code contains NO similarity metric: SSD and mutual information
appear only as option strings handed to SimpleITK / Elastix for deformable registration
(``utils/deformable/simpleitk.py:115`` ``SetMetricAsMeanSquares``, ``:153`` "AdvancedMeanSquares",
``:155-156`` "AdvancedMattesMutualInformation" + ``NumberOfHistogramBins``).  The definitions below
are therefore this repository's written spec, shared word for word by the
CUDA kernels:

  M1  ssd = sum over fixed voxels whose sample is inside (ITK half-open test) of (F - m)^2, m the
      UNROUNDED interpolated moving value; n_valid = number of such voxels; mse = ssd / n_valid.
  M2  joint histogram H[bf][bm], B bins per axis, bf = clamp(floor((F - fmin) * B / (fmax - fmin)), 0, B-1),
      bm likewise from m; hard binning over the same valid voxels; MI = sum p_ij ln(p_ij / (p_i. p_.j))
      in float64 from the integer counts (natural log).
PARITY UNPINNED by the reference: there is nothing in it to pin these against.
"""
from __future__ import annotations

import numpy as np

from .reslice import continuous_index, oracle_reslice


def _moving_mask_hits(moving_mask, A, odims):
    """SetMetricMovingMask: the mask voxel nearest to the mapped point (floor(c + 0.5), clamped like the taps)."""
    mm = np.asarray(moving_mask)
    c = continuous_index(A, odims)
    idx = [np.clip(np.floor(c[a] + 0.5).astype(np.int64), 0, mm.shape[2 - a] - 1) for a in range(3)]
    return mm[idx[2], idx[1], idx[0]] > 0


def bin_index(v, vmin, vmax, bins):
    """The binning arithmetic is done in float32 exactly as the kernel does it
    (scale = B / (max - min) rounded to f32; (v - min) * scale in f32; floor; clamp)."""
    v = np.asarray(v, dtype=np.float32)
    scale = np.float32(bins) / (np.float32(vmax) - np.float32(vmin))
    b = np.floor((v - np.float32(vmin)) * scale)
    return np.clip(b, 0, bins - 1).astype(np.int64)


def oracle_ssd(fixed, mov, A, mode="itk", fixed_mask=None, moving_mask=None):
    """(ssd float64, n_valid int).  ``fixed`` defines the output lattice ([z,y,x]); ``fixed_mask`` (same lattice):
    voxels where it is 0 take no part (SetMetricFixedMask, ``utils/deformable/simpleitk.py:116-117``)."""
    fixed = np.asarray(fixed)
    odims = (fixed.shape[2], fixed.shape[1], fixed.shape[0])
    _, vals, mask = oracle_reslice(mov, A, odims, 0.0, mode, np.float32, return_float=True)
    if fixed_mask is not None:
        mask = mask & (np.asarray(fixed_mask) > 0)
    if moving_mask is not None:
        mask = mask & _moving_mask_hits(moving_mask, A, odims)
    d = fixed.astype(np.float64) - vals
    d = d[mask]
    return float(np.sum(d * d)), int(mask.sum())


def oracle_joint_hist(fixed, mov, A, bins, frange, mrange, mode="itk", fixed_mask=None, moving_mask=None):
    """(H uint64 [bins, bins] indexed [fixed_bin, moving_bin], n_valid)."""
    fixed = np.asarray(fixed)
    odims = (fixed.shape[2], fixed.shape[1], fixed.shape[0])
    _, vals, mask = oracle_reslice(mov, A, odims, 0.0, mode, np.float32, return_float=True)
    if fixed_mask is not None:
        mask = mask & (np.asarray(fixed_mask) > 0)
    if moving_mask is not None:
        mask = mask & _moving_mask_hits(moving_mask, A, odims)
    bf = bin_index(fixed[mask], frange[0], frange[1], bins)
    bm = bin_index(vals[mask].astype(np.float32), mrange[0], mrange[1], bins)
    H = np.bincount(bf * bins + bm, minlength=bins * bins).reshape(bins, bins).astype(np.uint64)
    return H, int(mask.sum())


def oracle_mi(H):
    """Mutual information (nats) of an integer joint histogram, float64."""
    H = np.asarray(H, dtype=np.float64)
    n = H.sum()
    if n <= 0:
        return 0.0
    pf = H.sum(axis=1)
    pm = H.sum(axis=0)
    nzf, nzm = np.nonzero(H)
    h = H[nzf, nzm]
    return float(np.sum((h / n) * np.log(h * n / (pf[nzf] * pm[nzm]))))


# ---- Mattes (Parzen-window) mutual information, SURVEY N5 ---------------------------------------------------------
PARZEN_ONE, PARZEN_STEPS = 1024, 256


def parzen_table():
    """[257, 4] integer weights of the cubic B-spline window on a 1/256 sub-bin lattice, each row summing to 1024.
    Integer arithmetic only (round half up of 1024 * w): w0 = (1-t)^3/6, w1 = (4 - 6t^2 + 3t^3)/6, w3 = t^3/6."""
    T = PARZEN_STEPS
    den = 6 * T ** 3
    tab = np.zeros((T + 1, 4), dtype=np.int64)
    for k in range(T + 1):
        n0, n1, n3 = (T - k) ** 3, 4 * T ** 3 - 6 * k * k * T + 3 * k ** 3, k ** 3
        q0, q1, q3 = ((2 * PARZEN_ONE * n + den) // (2 * den) for n in (n0, n1, n3))
        tab[k] = (q0, q1, PARZEN_ONE - q0 - q1 - q3, q3)
    return tab


def oracle_mattes_hist(fixed, mov, A, bins, frange, mrange, mode="itk", fixed_mask=None, moving_mask=None):
    """(H uint64 [bins, bins] in units of 1/1024 sample, n_valid).  Structure of
    itk::MattesMutualInformationImageToImageMetricv4 (named at ``utils/deformable/simpleitk.py:155-156``): 2 padding
    bins per side, box window on the fixed image, cubic B-spline over 4 bins on the moving image; the spec of the
    integer quantisation is this repository's (``csrc/mia_kernels.cuh::mattes_add``).  PARITY UNPINNED."""
    fixed = np.asarray(fixed)
    odims = (fixed.shape[2], fixed.shape[1], fixed.shape[0])
    _, vals, mask = oracle_reslice(mov, A, odims, 0.0, mode, np.float32, return_float=True)
    if fixed_mask is not None:
        mask = mask & (np.asarray(fixed_mask) > 0)
    if moving_mask is not None:
        mask = mask & _moving_mask_hits(moving_mask, A, odims)
    fb = bin_index(fixed[mask], frange[0], frange[1], bins - 4) + 2
    v = vals[mask].astype(np.float32)
    mscale = np.float32(bins - 4) / (np.float32(mrange[1]) - np.float32(mrange[0]))
    b = np.maximum((v - np.float32(mrange[0])) * mscale, np.float32(0))
    i0 = np.minimum(np.floor(b), np.float32(bins - 5))
    t = np.minimum(b - i0, np.float32(1))
    tq = np.floor(t * np.float32(PARZEN_STEPS) + np.float32(0.5)).astype(np.int64)
    tab = parzen_table()
    H = np.zeros(bins * bins, dtype=np.int64)
    base = fb * bins + i0.astype(np.int64) + 1
    for k in range(4):
        H += np.bincount(base + k, weights=tab[tq, k], minlength=bins * bins).astype(np.int64)
    return H.reshape(bins, bins).astype(np.uint64), int(mask.sum())
