"""Generates tests/golden/reslice_golden.npz.

The reference has NO golden vectors, tests or fixtures and cannot be imported in this image
(needs gdcm / pydicom / vtk / SimpleITK), so these vectors are produced by this repository's NumPy float64
restatement (``oracle/reslice.py``) and, for every case, cross-checked here against
``scipy.ndimage.map_coordinates(order=1)`` -- a function the reference itself calls
(``structure/deformable.py:873-878``) -- on the voxels whose 8 taps are all inside the buffer.  The FFS cases
are produced with the reference's own NumPy expressions (``read/dicom.py:695-717``).

Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
from scipy.ndimage import map_coordinates

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from medicalimageanalysis_b200 import geometry as G          # noqa: E402
from medicalimageanalysis_b200.phantom import make_phantom    # noqa: E402
from oracle import ffs, reslice                               # noqa: E402

SHAPE = (20, 24, 32)          # [z, y, x]


def cases():
    g = G.Grid((32, 24, 20), (1.0, 1.0, 1.0), (-16.0, -12.0, -10.0))
    c = np.array([0.0, 0.0, 0.0])
    ident = np.identity(4)
    out = {}
    out["identity"] = (G.compose_index_matrix(g, g, ident), g.dims)
    out["shift_int"] = (G.compose_index_matrix(g, g, G.translation_transform(3, -2, 1)), g.dims)
    out["shift_sub"] = (G.compose_index_matrix(g, g, G.translation_transform(0.25, -0.6, 0.4)), g.dims)
    out["shift_half"] = (G.compose_index_matrix(g, g, G.translation_transform(0.5, -0.5, 0.5)), g.dims)
    out["rot_x5"] = (G.compose_index_matrix(g, g, G.pivot_transform(c, 5, 0, 0)), g.dims)
    out["rot_y5"] = (G.compose_index_matrix(g, g, G.pivot_transform(c, 0, 5, 0)), g.dims)
    out["rot_z5"] = (G.compose_index_matrix(g, g, G.pivot_transform(c, 0, 0, 5)), g.dims)
    out["compound"] = (G.compose_index_matrix(g, g, G.params_to_matrix([3, -2, 4, 2.5, -1.5, 1.0], c)), g.dims)
    # anisotropic moving (0.98, 0.98, 2.5) resampled onto a 1 mm isotropic lattice
    gm = G.Grid((32, 24, 20), (0.98, 0.98, 2.5), (-15.68, -11.76, -25.0))
    gf = G.Grid((30, 22, 48), (1.0, 1.0, 1.0), (-15.0, -11.0, -24.0))
    out["aniso"] = (G.compose_index_matrix(gf, gm, G.params_to_matrix([1, 2, -3, 1, 0.5, -2], c)), gf.dims)
    # moving smaller than the fixed lattice: lots of background
    gbig = G.Grid((45, 37, 29), (1.0, 1.0, 1.0), (-22.0, -18.0, -14.0))
    out["bigger_fixed"] = (G.compose_index_matrix(gbig, g, G.pivot_transform(c, 0, 0, 10)), gbig.dims)
    # auto-cropped VTK lattice of a rotated volume
    M = G.params_to_matrix([0, 0, 12, 4, 0, 0], c)
    origin, dims = G.autocrop_extent(g, M, (1.0, 1.0, 1.0))
    gauto = G.Grid(dims, (1.0, 1.0, 1.0), origin)
    out["autocrop"] = (G.compose_index_matrix(gauto, g, M, "vtk"), dims)
    return out


def main():
    mov_i16 = make_phantom(SHAPE, "CT", seed=11)
    mov_f32 = make_phantom(SHAPE, "MR", seed=12)
    store = {"mov_i16": mov_i16, "mov_f32": mov_f32}
    names = []
    for name, (A, odims) in cases().items():
        names.append(name)
        store[name + "/A"] = A
        store[name + "/odims"] = np.asarray(odims, dtype=np.int32)
        for tag, mov in (("i16", mov_i16), ("f32", mov_f32)):
            for mode, bg in (("vtk", -3001.0), ("itk", 0.0)):
                o, vals, mask = reslice.oracle_reslice(mov, A, odims, bg, mode, return_float=True)
                store["%s/%s/%s" % (name, tag, mode)] = o
                # independent cross-check on interior voxels
                c = reslice.continuous_index(A, odims)
                sc = map_coordinates(mov.astype(np.float64), [c[2], c[1], c[0]], order=1, mode="nearest")
                interior = mask.copy()
                for a, n in zip(c, (mov.shape[2], mov.shape[1], mov.shape[0])):
                    interior &= (a >= 0) & (a <= n - 1)
                err = np.abs(sc - vals)[interior].max() if interior.any() else 0.0
                assert err < 1e-9, (name, tag, mode, err)
    store["names"] = np.array(names)
    for case in ffs.FFS_CASES:
        store["ffs/" + case] = ffs.oracle_ffs(mov_i16, case)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reslice_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes,", len(names), "cases")


if __name__ == "__main__":
    main()
