"""GPU: the allocation-free entry point xgpr_b200_cg_matvec_accum_static_f64 (include/xgpr_b200.h) against the existing
xgpr_b200_cg_matvec_accum_f64, both called straight through the C ABI (ctypes), on calls made back to back.

The static variant keeps the (n, r) intermediate Z v in a file-scope device array shared by every call on the device, so
the cases run in an order that would expose stale contents: a wide (r = 26) call first, then narrower and shorter ones,
then the split-K shape whose first kernel ACCUMULATES into the intermediate (it has to be zero-filled by the call itself).
This file is named to be collected last.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-10, 1e-11          # the bounds tests/test_gpu_l1.py::test_downstream_consumers puts on this matvec

# (n, m, r): multi-RHS DMMA path, single-RHS path, odd m (the fallback tiles), few rows x many columns (split-K over m)
CASES = [(1500, 1024, 26), (4000, 2048, 1), (333, 1001, 7), (96, 16384, 26), (1500, 1024, 26)]


def _lib():
    from xgpr_b200 import _native
    return _native.lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _both(lib, z, v):
    n, m = z.shape
    r = v.shape[1]
    w_ref = torch.zeros((m, r), dtype=torch.float64, device="cuda")
    w_new = torch.zeros((m, r), dtype=torch.float64, device="cuda")
    scratch = torch.empty((n, r), dtype=torch.float64, device="cuda")
    rc = lib.xgpr_b200_cg_matvec_accum_f64(z.data_ptr(), n, m, v.data_ptr(), r, w_ref.data_ptr(), scratch.data_ptr(),
                                           _stream())
    assert rc == 0, lib.xgpr_b200_last_error()
    rc = lib.xgpr_b200_cg_matvec_accum_static_f64(z.data_ptr(), n, m, v.data_ptr(), r, w_new.data_ptr(), _stream())
    assert rc == 0, lib.xgpr_b200_last_error()
    torch.cuda.synchronize()
    return w_ref.cpu().numpy(), w_new.cpu().numpy()


def test_static_scratch_matvec_matches_workspace_matvec_on_consecutive_calls():
    lib = _lib()
    assert lib.xgpr_b200_cg_matvec_static_max_elems() == 1 << 20
    rng = np.random.default_rng(7)
    for n, m, r in CASES:
        zh = rng.standard_normal((n, m)) / np.sqrt(m)
        vh = rng.standard_normal((m, r))
        z, v = torch.from_numpy(zh).cuda(), torch.from_numpy(vh).cuda()
        w_ref, w_new = _both(lib, z, v)
        np.testing.assert_allclose(w_new, w_ref, rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(w_new, zh.T @ (zh @ vh), rtol=RTOL, atol=ATOL)
        # the same call again, immediately: the array now holds THIS call's intermediate
        w_ref2, w_new2 = _both(lib, z, v)
        np.testing.assert_allclose(w_new2, w_ref2, rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(w_new2, w_new, rtol=RTOL, atol=ATOL)


def test_static_scratch_matvec_accumulates_like_the_original():
    lib = _lib()
    rng = np.random.default_rng(8)
    zh = rng.standard_normal((700, 512)) / np.sqrt(512)
    vh = rng.standard_normal((512, 3))
    z, v = torch.from_numpy(zh).cuda(), torch.from_numpy(vh).cuda()
    w = torch.ones((512, 3), dtype=torch.float64, device="cuda")
    for _ in range(2):
        rc = lib.xgpr_b200_cg_matvec_accum_static_f64(z.data_ptr(), 700, 512, v.data_ptr(), 3, w.data_ptr(), _stream())
        assert rc == 0, lib.xgpr_b200_last_error()
    np.testing.assert_allclose(w.cpu().numpy(), 1.0 + 2.0 * (zh.T @ (zh @ vh)), rtol=RTOL, atol=ATOL)


def test_static_scratch_matvec_rejects_sizes_beyond_the_array():
    lib = _lib()
    cap = lib.xgpr_b200_cg_matvec_static_max_elems()
    z = torch.zeros((8, 64), dtype=torch.float64, device="cuda")
    v = torch.zeros((64, 2), dtype=torch.float64, device="cuda")
    w = torch.zeros((64, 2), dtype=torch.float64, device="cuda")
    before = torch.cuda.memory_allocated()
    # n * r one past the limit: refused from the sizes alone, before any pointer is touched
    rc = lib.xgpr_b200_cg_matvec_accum_static_f64(z.data_ptr(), cap // 2 + 1, 64, v.data_ptr(), 2, w.data_ptr(), _stream())
    assert rc == -1
    msg = lib.xgpr_b200_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "static scratch" in msg and str(cap) in msg
    # exactly at the limit the size check passes (n * r == cap); a tiny real call afterwards still works
    rc = lib.xgpr_b200_cg_matvec_accum_static_f64(z.data_ptr(), 8, 64, v.data_ptr(), 2, w.data_ptr(), _stream())
    assert rc == 0, lib.xgpr_b200_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert float(w.abs().max()) == 0.0
