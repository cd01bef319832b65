"""GPU: ``ib200_dot_static`` -- the dot product for callers that must not allocate GPU memory -- against ``ib200_dot``.

Who cannot allocate: a plug-in inside an application whose pool allocator already owns the GPU's memory (a raw
``cudaMalloc`` fails there or goes behind the pool's back), and a solver's time loop on a GPU shared with other
streams (``cudaMalloc`` / ``cudaFree`` stall the whole device).  ``ib200_dot_static`` keeps its block partials and the
staged copy of host vectors in a fixed array inside the library's device code instead; the TEST may allocate (it has
to put vectors on the GPU somehow), the entry point under test does not.

This file sorts last on purpose: ``pytest -x`` stops at the first failure, and a failure in this add-on must not hide
the parity tests.  It builds everything it needs itself, so it also passes alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HOST, DEVICE = 0, 1
UNSUPPORTED = 103  # IB200_ERROR_UNSUPPORTED (include/interp_b200.h)
MAX_HOST_N = 1 << 20  # IB200_DOT_STATIC_MAX_HOST_N


def _vp(x):
    return C.c_void_p(x)


def test_dot_static_matches_dot_two_calls_in_a_row(ctx):
    """Device vectors, two calls back to back with different data and lengths (the second one shorter, so anything
    the first call left in the shared scratch would show): same bits as ``ib200_dot``, and within the suite's 1e-12
    of the NumPy dot product.  The second call must leave the GPU's free memory untouched."""
    import torch

    from interplib_b200._lib import check, lib

    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    rng = np.random.default_rng(20261106)
    free_before = None
    with torch.cuda.device(dev):  # the entry point works on the CURRENT device
        for call, n in enumerate((1_000_003, 4_099)):
            a_h, b_h = rng.standard_normal((2, n))
            a, b = torch.from_numpy(a_h).to(dev), torch.from_numpy(b_h).to(dev)
            out_static = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
            out_ref = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
            torch.cuda.synchronize(dev)  # inputs are in place before either library stream reads them
            if call == 1:
                free_before = torch.cuda.mem_get_info(dev)[0]
            check(lib.ib200_dot_static(n, _vp(a.data_ptr()), _vp(b.data_ptr()), _vp(out_static.data_ptr()), DEVICE,
                                       _vp(stream.cuda_stream)))
            stream.synchronize()
            if call == 1:
                assert torch.cuda.mem_get_info(dev)[0] == free_before, "ib200_dot_static changed the free device memory"
            check(lib.ib200_dot(ctx.handle, n, _vp(a.data_ptr()), _vp(b.data_ptr()), _vp(out_ref.data_ptr()), DEVICE))
            ctx.synchronize()
            got, ref = float(out_static.cpu()[0]), float(out_ref.cpu()[0])
            exact = float(a_h @ b_h)
            assert np.isfinite(got)
            assert got == ref, (call, got, ref)  # same kernels, same reduction tree
            assert abs(got - exact) <= 1e-12 * float(np.abs(a_h) @ np.abs(b_h)), (call, got, exact)


def test_dot_static_host_vectors_and_limit(ctx):
    """Host vectors are staged in the built-in scratch: two calls in a row agree with ``ib200_dot``'s host mode; the
    largest stageable length is accepted, one more is refused with an error that names the limit; n = 0 gives 0."""
    from interplib_b200._lib import check, lib

    rng = np.random.default_rng(7)
    for n in (100_003, 7):
        a, b = rng.random((2, n))
        res = np.full(1, np.nan)
        check(lib.ib200_dot_static(n, _vp(a.ctypes.data), _vp(b.ctypes.data), _vp(res.ctypes.data), HOST, None))
        assert res[0] == ctx.dot(a, b)
        assert abs(res[0] - float(a @ b)) < 1e-12 * float(a @ b)

    a = rng.random(MAX_HOST_N + 1)
    res = np.full(1, np.nan)
    check(lib.ib200_dot_static(MAX_HOST_N, _vp(a.ctypes.data), _vp(a.ctypes.data), _vp(res.ctypes.data), HOST, None))
    assert res[0] == ctx.dot(a[:MAX_HOST_N], a[:MAX_HOST_N])
    res[0] = np.nan
    code = lib.ib200_dot_static(MAX_HOST_N + 1, _vp(a.ctypes.data), _vp(a.ctypes.data), _vp(res.ctypes.data), HOST, None)
    assert code == UNSUPPORTED and np.isnan(res[0])
    assert b"IB200_DOT_STATIC_MAX_HOST_N" in lib.ib200_last_error()

    check(lib.ib200_dot_static(0, None, None, _vp(res.ctypes.data), HOST, None))
    assert res[0] == 0.0
