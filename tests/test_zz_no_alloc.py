"""qsp_robust_res_noalloc (include/qsp_recon.h): the allocation-free form of get_robust_res, through the C ABI,
against the existing qsp_robust_res on the same residuals - twice in a row (the fixed scratch array keeps the last
call's contents), at a size with many blocks, a ragged one and n = 1, plus the rejection of n beyond the maximum.
robust and w must be bit-identical; the loss differs by the summation order only (rtol 1e-5, the helper tests'
tolerance).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _existing(L, N, res, b):
    rob, w, loss = np.empty_like(res), np.empty_like(res), np.zeros(1, np.float32)
    N.check(L.qsp_robust_res(0, N.ptr(res), res.size, float(b), N.ptr(rob), N.ptr(w), N.ptr(loss)))
    return rob, w, loss[0]


def _noalloc(L, N, res, b, poison):
    d_res = torch.from_numpy(res).cuda()
    d_rob = torch.full((res.size,), poison, device="cuda")
    d_w = torch.full((res.size,), poison, device="cuda")
    d_loss = torch.full((1,), poison, device="cuda")
    torch.cuda.synchronize()
    vp = lambda t: C.c_void_p(t.data_ptr())
    N.check(L.qsp_robust_res_noalloc(0, vp(d_res), res.size, float(b), vp(d_rob), vp(d_w), vp(d_loss), None))
    torch.cuda.synchronize()
    return d_rob.cpu().numpy(), d_w.cpu().numpy(), float(d_loss.cpu()[0])


def test_noalloc_robust_res_matches_the_allocating_entry_point():
    from qsp_slam_b200 import _native as N
    L = N.lib()
    rng = np.random.default_rng(7)
    b = 0.025
    for call, n in enumerate((300_000, 300_000, 1025, 1, 77_777)):       # the first two: same size, two calls in a row
        res = (rng.standard_normal(n) * 0.05).astype(np.float32)
        res[::7] = 0.0                                                   # |res| = 0 rows (weight of a zero residual)
        rob0, w0, loss0 = _existing(L, N, res, b)
        rob1, w1, loss1 = _noalloc(L, N, res, b, poison=float(call + 3))
        np.testing.assert_array_equal(rob1, rob0)
        np.testing.assert_array_equal(w1, w0)
        np.testing.assert_allclose(loss1, loss0, rtol=1e-5, atol=0.0)
        assert np.isclose(loss0, np.mean(rob0.astype(np.float64) ** 2), rtol=1e-5)
    # the same input twice: the block-ordered sum is repeatable bit for bit
    again = _noalloc(L, N, res, b, poison=-1.0)
    assert again[2] == loss1


def test_noalloc_rejects_sizes_beyond_the_scratch_array():
    from qsp_slam_b200 import _native as N
    L = N.lib()
    t = torch.zeros(8, device="cuda")
    vp = C.c_void_p(t.data_ptr())
    rc = L.qsp_robust_res_noalloc(0, vp, N.NOALLOC_MAX_N + 1, 0.02, vp, vp, vp, None)
    assert rc == -1                                                      # QSP_ERR_INVALID
    assert b"QSP_NOALLOC_MAX_N" in L.qsp_last_error()
    assert L.qsp_robust_res_noalloc(0, vp, 0, 0.02, vp, vp, vp, None) == -1
    # the largest accepted size runs
    n = N.NOALLOC_MAX_N
    res = torch.full((n,), 0.01, device="cuda")
    out = torch.empty((2, n), device="cuda")
    loss = torch.zeros(1, device="cuda")
    p = lambda x: C.c_void_p(x.data_ptr())
    N.check(L.qsp_robust_res_noalloc(0, p(res), n, 0.02, p(out[0]), p(out[1]), p(loss), None))
    torch.cuda.synchronize()
    assert abs(float(loss[0]) - 1e-4) < 1e-8 and abs(float(out[1].min()) - 1.0) < 1e-6
