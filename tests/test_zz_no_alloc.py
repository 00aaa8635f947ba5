"""masg_gmm_barycenter_noalloc (include/masg.h): the allocation-free variant of masg_gmm_barycenter, whose scratch is a fixed
file-scope device array instead of the context's growable arena.  Called through the C ABI on two calls in a row and compared
with the allocating entry point on the same inputs.  The two share kernels and the weight sum is done in the same order with
the same rounding, so the comparison is exact; the parity tolerance of the barycentre tests (1e-5, tests/test_gpu_parity.py) is
the stated bar.  Collected last on purpose."""
import ctypes

import pytest
import torch

import masg_b200  # noqa: F401
from masg_b200 import _lib

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
TOL = 1e-5


def _inputs(A, K, D, seed):
    g = torch.Generator().manual_seed(seed)
    pi = torch.softmax(torch.randn(A, K, generator=g), dim=-1)
    mu = torch.randn(A, K, D, generator=g)
    sigma = torch.rand(A, K, D, generator=g) * 0.9 + 0.1
    rate = torch.softmax(torch.randn(A, generator=g), dim=0)
    return pi.to(DEV), mu.to(DEV), sigma.to(DEV), [float(r) for r in rate]


def _call(ctx, fn_name, pi, mu, sigma, rate, A, K, D):
    nb = K ** A
    bmean = torch.full((nb, D), float("nan"), device=DEV)
    bstd = torch.full((nb, D), float("nan"), device=DEV)
    bw = torch.full((nb,), float("nan"), device=DEV)
    assign = torch.full((A * K,), -1, dtype=torch.int32, device=DEV)
    h_rate = (ctypes.c_float * A)(*rate)
    fn = getattr(ctx.lib, fn_name)
    ctx.check(fn(ctx.h, ctx.stream(), _lib.ptr(pi), _lib.ptr(mu), _lib.ptr(sigma), ctypes.cast(h_rate, ctypes.c_void_p), A, K, D,
                 _lib.ptr(bmean), _lib.ptr(bstd), _lib.ptr(bw), _lib.ptr(assign)), fn_name)
    torch.cuda.synchronize()
    return bmean, bstd, bw, assign


def test_barycenter_noalloc_matches_allocating_entry_point_twice():
    ctx = _lib.default_context(DEV, "fp32")
    # two calls in a row with different shapes and values: the second must not see the first one's accumulators (the shared
    # scratch keeps the last call's contents), and a larger problem followed by a smaller one exercises the clear
    for A, K, D, seed in [(4, 3, 256, 11), (2, 3, 256, 12)]:
        pi, mu, sigma, rate = _inputs(A, K, D, seed)
        want = _call(ctx, "masg_gmm_barycenter", pi, mu, sigma, rate, A, K, D)
        got = _call(ctx, "masg_gmm_barycenter_noalloc", pi, mu, sigma, rate, A, K, D)
        for name, w, g in zip(("mean", "std", "weight"), want[:3], got[:3]):
            assert torch.isfinite(g).all(), name
            assert (g - w).abs().max().item() <= TOL, name
            assert torch.equal(g, w), name                                  # same kernels, same summation order
        assert torch.equal(got[3], want[3])
        assert abs(got[2].sum().item() - 1.0) < 1e-5                         # rates and pis each sum to one


def test_barycenter_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    ctx = _lib.default_context(DEV, "fp32")
    A, K, D = 8, 3, 4                                                        # 3^8 = 6561 > MASG_NOALLOC_MAX_BARYCENTRES = 4096
    pi, mu, sigma, rate = _inputs(A, K, D, 13)
    out = torch.zeros(8, device=DEV)
    a32 = torch.zeros(A * K, dtype=torch.int32, device=DEV)
    h_rate = (ctypes.c_float * A)(*rate)
    rc = ctx.lib.masg_gmm_barycenter_noalloc(ctx.h, ctx.stream(), _lib.ptr(pi), _lib.ptr(mu), _lib.ptr(sigma),
                                             ctypes.cast(h_rate, ctypes.c_void_p), A, K, D, _lib.ptr(out), _lib.ptr(out), _lib.ptr(out),
                                             _lib.ptr(a32))
    assert rc != 0
    msg = ctx.lib.masg_last_error(ctx.h).decode()
    assert "MASG_NOALLOC_MAX_BARYCENTRES" in msg and "4096" in msg
    torch.cuda.synchronize()                                                 # nothing was launched, nothing faulted
