"""The C-ABI library loads on a CPU-only box and exports every symbol include/mia_b200.h declares."""
import ctypes
import os
import re

import pytest

from medicalimageanalysis_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mia_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mia_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_bound_symbols():
    assert declared_symbols() == sorted(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build with __graft_entry__.build()"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_version_and_error_string_without_gpu():
    lib = _lib.load()
    assert lib.mia_version() == 100
    # argument validation happens before any CUDA call, so error paths work on a CPU box
    rc = lib.mia_resample_rigid(None, 0, _lib.i3((1, 1, 1)), None, 0, _lib.i3((1, 1, 1)), None, 0.0, 0, None)
    assert rc == -1
    assert b"null pointer" in lib.mia_last_error()
    with pytest.raises(_lib.MiaError):
        _lib.check(rc)


def test_workspace_and_state_sizes():
    lib = _lib.load()
    assert lib.mia_metric_workspace_bytes(_lib.i3((8, 8, 8)), 13, 64) % 256 == 0
    assert lib.mia_optimizer_state_bytes() > 4096 * 8 * 8


def test_autocrop_matches_python_geometry():
    import numpy as np
    from medicalimageanalysis_b200 import geometry as G, kernels
    g = G.Grid((32, 24, 20), (0.9, 0.9, 2.5), (-16.0, -12.0, -10.0))
    for p in ([0, 0, 12, 4, 0, 0], [3, -2, 4, 5, -3, 7], [0, 0, 0, 0, 0, 0]):
        M = G.params_to_matrix(p, (1.0, 2.0, 3.0))
        o1, d1 = G.autocrop_extent(g, M, (1.0, 1.0, 2.0), convention="vtk")
        o2, d2 = kernels.autocrop_extent(M, G.index_to_physical(g, "vtk"), g.dims, (1.0, 1.0, 2.0))
        assert d1 == tuple(d2)
        assert np.allclose(o1, o2, atol=1e-9)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libmia_b200.so")
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.load()
