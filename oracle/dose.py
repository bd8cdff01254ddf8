"""NumPy restatement of the dose-on-image-lattice sample and its DVH statistics (oracle; test infrastructure only).

Follows ``structure/dose.py:738-772`` (``compute_roi_dose_array``: ``sitk.ResampleImageFilter`` with the image as
reference, ``sitkLinear``, default pixel 0.0; then ``dose_arr[np.where(mask_arr > 0)]``) and ``:774-816``
(``compute_roi_dose_statistics``).  The resampler is ``oracle.reslice`` in ITK mode (parity unpinned, see
``oracle/__init__.py``); the statistics are the reference's own NumPy calls, restated verbatim, so that half IS
pinned: ``np.min / max / mean / median / std``, ``np.percentile(x, 100 - d)``, ``np.mean(x < d) * 100``."""
from __future__ import annotations

import numpy as np

from . import reslice

D_VALUES = [1, 2, 5, 10, 15, 20, 25, 30, 35, 40, 45, 50, 55, 60, 65, 70, 75, 80, 85, 90, 95, 98, 99]   # dose.py:809


def oracle_roi_dose_array(dose, A, mask, default=0.0):
    """``dose_arr[np.where(mask_arr > 0)]`` with ``dose_arr`` = dose resampled onto the mask's lattice
    (``structure/dose.py:760-770``).  ``A``: image index -> dose continuous index (3x4)."""
    mask = np.asarray(mask)
    odims = (mask.shape[2], mask.shape[1], mask.shape[0])
    dose_arr = reslice.oracle_reslice(np.asarray(dose, dtype=np.float32), A, odims, default, "itk", out_dtype=np.float32)
    return dose_arr[np.where(mask > 0)]


def oracle_dose_statistics(dose_in_roi, spacing, roi_name="ROI", max_dose=150, increment=5):
    """``structure/dose.py:790-816``, line for line (including the fixed step of 5 in the VS loop)."""
    voxel_vol_cc = np.prod(spacing) / 1000.0
    dvh = {"ROI": roi_name,
           "Volume (cc)": dose_in_roi.size * voxel_vol_cc,
           "Dmin": np.min(dose_in_roi),
           "Dmax": np.max(dose_in_roi),
           "Dmean": np.mean(dose_in_roi),
           "Dmedian": np.median(dose_in_roi),
           "Dstd": np.std(dose_in_roi)}
    dvh.update({"D%d" % d: np.percentile(dose_in_roi, 100 - d) for d in D_VALUES})
    for d in range(0, max_dose + increment, 5):
        m = dose_in_roi < d
        dvh["VS%dGy_percent" % d] = np.mean(m) * 100
        dvh["VS%dGy_cc" % d] = np.sum(m) * voxel_vol_cc
    return dvh
