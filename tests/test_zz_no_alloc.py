"""mia_resample_ssd_static: the fused resample + SSD launch with its scratch in a file-scope device array.

No counterpart in the reference (added for callers that cannot call
cudaMalloc).  Called through the C ABI and compared with the existing entry point, mia_resample_ssd, which takes a
caller workspace.  This file is named so that pytest collects it last.
"""
import ctypes
import gc

import numpy as np
import pytest
import torch

from medicalimageanalysis_b200 import _lib
from medicalimageanalysis_b200 import geometry as G
from medicalimageanalysis_b200.phantom import make_phantom

I16 = _lib.MIA_I16
ITK = _lib.MIA_MODE_ITK


def _pair(shape=(24, 48, 72), seed=7):
    sp = (1.0, 1.0, 2.0)
    g = G.Grid((shape[2], shape[1], shape[0]), sp, tuple(-0.5 * np.array(sp) * np.array(shape[::-1])))
    Mgt = G.params_to_matrix([3, -2, 4, 3, -2, 1], (0, 0, 0))
    fixed = make_phantom(shape, "CT", seed=seed, spacing=sp)
    moving = make_phantom(shape, "CT", seed=seed, spacing=sp, world_matrix=np.linalg.inv(Mgt))
    return g, fixed, moving


def _probes(g, n, seed):
    """n matrices a fraction of a voxel apart, like the probes of one optimiser iteration."""
    rng = np.random.default_rng(seed)
    base = np.array([3.0, -2.0, 4.0, 3.0, -2.0, 1.0])
    out = [G.compose_index_matrix(g, g, G.params_to_matrix(base + rng.uniform(-0.3, 0.3, 6), (0, 0, 0))) for _ in range(n)]
    return np.ascontiguousarray(np.stack(out), dtype=np.float64)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _call_static(lib, f, m, A, ssd, cnt, out=None):
    n = A.shape[0]
    dims = _lib.i3((m.shape[2], m.shape[1], m.shape[0]))
    s = ctypes.c_void_p(torch.cuda.current_stream(m.device).cuda_stream)
    return lib.mia_resample_ssd_static(_ptr(m), I16, dims, _ptr(f), I16, dims, _ptr(out) if out is not None else None, I16, n,
                                       A.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), None, -3001.0, ITK,
                                       _ptr(ssd), _ptr(cnt), s)


def _call_workspace(lib, f, m, A, ssd, cnt, ws, out=None):
    n = A.shape[0]
    dims = _lib.i3((m.shape[2], m.shape[1], m.shape[0]))
    s = ctypes.c_void_p(torch.cuda.current_stream(m.device).cuda_stream)
    return lib.mia_resample_ssd(_ptr(m), I16, dims, _ptr(f), I16, dims, _ptr(out) if out is not None else None, I16, n,
                                A.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), None, -3001.0, ITK,
                                _ptr(ssd), _ptr(cnt), _ptr(ws), ws.numel(), s)


def test_static_rejects_sizes_beyond_the_maximum():
    """The limit is checked before any CUDA call, so this runs on a CPU box too."""
    lib = _lib.load()
    dims = _lib.i3((8, 8, 8))
    for n in (0, _lib.MIA_STATIC_MAX_MATS + 1, _lib.MIA_MAX_MATS):
        rc = lib.mia_resample_ssd_static(None, I16, dims, None, I16, dims, None, I16, n, None, None, 0.0, ITK, None, None, None)
        assert rc == -1
        msg = lib.mia_last_error().decode()
        assert "n_mats must be 1..%d" % _lib.MIA_STATIC_MAX_MATS in msg and str(_lib.MIA_STATIC_WS_BYTES) in msg
    # the largest admitted launch fits the array
    assert lib.mia_metric_workspace_bytes(dims, _lib.MIA_STATIC_MAX_MATS, 64) <= _lib.MIA_STATIC_WS_BYTES


@pytest.mark.gpu
def test_static_scratch_matches_workspace_entry_point_on_calls_in_a_row(cuda_device):
    lib = _lib.load()
    g, fixed, moving = _pair()
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    ws = torch.zeros(lib.mia_metric_workspace_bytes(_lib.i3(g.dims), 13, 64), dtype=torch.uint8, device=cuda_device)
    # everything the calls write is allocated up front: the entry point itself must allocate nothing
    ssd_s = torch.zeros(16, dtype=torch.float64, device=cuda_device)
    cnt_s = torch.zeros(16, dtype=torch.int64, device=cuda_device)
    ssd_w, cnt_w = torch.zeros_like(ssd_s), torch.zeros_like(cnt_s)
    out_s, out_w = torch.empty_like(m), torch.empty_like(m)

    with torch.cuda.device(cuda_device):
        # call 1: 13 probes; call 2: ONE matrix with the resampled volume stored (the scratch still holds the partials
        # and staged matrices of call 1); call 3: 13 other probes again; call 4 repeats call 3 (run-to-run identical)
        plan = [(_probes(g, 13, 1), False), (_probes(g, 1, 2), True), (_probes(g, 13, 3), False), (_probes(g, 13, 3), False)]
        results = []
        for A, store in plan:
            n = A.shape[0]
            gc.collect()                                 # tensors of earlier tests must not be freed in mid-measurement
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated(cuda_device)
            ssd_s.fill_(-1.0); cnt_s.fill_(-1)
            _lib.check(_call_static(lib, f, m, A, ssd_s, cnt_s, out_s if store else None))
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated(cuda_device) <= before          # the call allocates nothing
            ssd_w.fill_(-1.0); cnt_w.fill_(-1)
            _lib.check(_call_workspace(lib, f, m, A, ssd_w, cnt_w, ws, out_w if store else None))
            torch.cuda.synchronize()
            got, want = ssd_s[:n].cpu().numpy(), ssd_w[:n].cpu().numpy()
            assert np.all(want > 0)
            assert np.all(np.abs(got - want) <= 1e-5 * np.abs(want)), (got, want)      # the tolerance of tests/test_gpu_metrics.py
            assert torch.equal(cnt_s[:n], cnt_w[:n]) and int(cnt_s[0]) > 0
            assert float(ssd_s[n]) == -1.0 and int(cnt_s[n]) == -1                   # nothing written past n_mats
            if store:
                assert torch.equal(out_s, out_w)
            results.append((got.copy(), cnt_s[:n].cpu().numpy().copy()))
        assert np.array_equal(results[2][0], results[3][0]) and np.array_equal(results[2][1], results[3][1])


@pytest.mark.gpu
def test_static_scratch_with_device_matrices(cuda_device):
    """d_A path (what a graph-captured optimiser loop passes): no host matrix is staged into the scratch."""
    lib = _lib.load()
    g, fixed, moving = _pair(seed=9)
    f, m = torch.from_numpy(fixed).to(cuda_device), torch.from_numpy(moving).to(cuda_device)
    A = _probes(g, _lib.MIA_STATIC_MAX_MATS, 4)
    dA = torch.from_numpy(A).to(cuda_device)
    n = A.shape[0]
    dims = _lib.i3(g.dims)
    ssd_s = torch.zeros(n, dtype=torch.float64, device=cuda_device)
    cnt_s = torch.zeros(n, dtype=torch.int64, device=cuda_device)
    ssd_w, cnt_w = torch.zeros_like(ssd_s), torch.zeros_like(cnt_s)
    ws = torch.zeros(lib.mia_metric_workspace_bytes(dims, n, 64), dtype=torch.uint8, device=cuda_device)
    with torch.cuda.device(cuda_device):
        s = ctypes.c_void_p(torch.cuda.current_stream(cuda_device).cuda_stream)
        for _ in range(2):
            _lib.check(lib.mia_resample_ssd_static(_ptr(m), I16, dims, _ptr(f), I16, dims, None, I16, n, None, _ptr(dA), 0.0, ITK,
                                                   _ptr(ssd_s), _ptr(cnt_s), s))
            _lib.check(lib.mia_resample_ssd(_ptr(m), I16, dims, _ptr(f), I16, dims, None, I16, n, None, _ptr(dA), 0.0, ITK,
                                            _ptr(ssd_w), _ptr(cnt_w), _ptr(ws), ws.numel(), s))
            torch.cuda.synchronize()
            got, want = ssd_s.cpu().numpy(), ssd_w.cpu().numpy()
            assert np.all(np.abs(got - want) <= 1e-5 * np.abs(want)) and torch.equal(cnt_s, cnt_w)
