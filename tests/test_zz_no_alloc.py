"""ddm_dot_noalloc: the dot product for callers that cannot allocate (no counterpart in the reference).
Its scratch is a static device array inside libddm_b200.so instead of the context's cudaMalloc'ed one.
Called through the C ABI and checked against the existing entry point ddm_dot, with the tolerance tests/test_gpu_parity.py uses
for scalar products (1e-13 relative to |x| |y|), on two calls in a row: the array is shared and keeps the previous call's
contents, so the second call (shorter vectors, other data) must not be disturbed by what the first one left behind.

Collected last on purpose: a failure here hides nothing.  Runs alone or after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CHUNK, MAX_PARTIALS = 8192, 16384  # kNoAllocChunk, kNoAllocMaxPartials in csrc/ddm_b200.cu


@pytest.fixture(scope="module")
def ctx():
    from ddm_b200 import api
    from ddm_b200.setup import problem

    # 41^3 = 68 921 DOFs: nine chunks of partial sums, the last one ragged
    P = problem.poisson_q1_cube(40, (2, 2, 2), overlap=1, coarse="none", numeric="device", workers=1)
    c = api.Context.from_products(P, with_coarse=False)
    yield c
    c.close()


def noalloc_dot(ctx, n, x, y, out):
    """-> status of ddm_dot_noalloc(ctx, n, x.d, y.d, out.d) with the vectors' device pointers."""
    ptr = lambda v: C.c_void_p(ctx.lib.ddm_vec_device_ptr(v.h))
    return ctx.lib.ddm_dot_noalloc(ctx.h, C.c_int64(n), ptr(x), ptr(y), ptr(out))


def test_two_calls_in_a_row_agree_with_ddm_dot(ctx):
    from ddm_b200 import api

    sp = api.NonOverlappingScalarProduct(ctx)
    n = ctx.n_own
    assert n > 8 * CHUNK
    rng = np.random.default_rng(42)
    out = ctx.vector()
    for call, m in enumerate((n, n // 3)):  # second call: fewer chunks than the scratch still holds from the first
        a, b = rng.standard_normal(n), rng.standard_normal(n)
        a[m:] = 0.0  # ddm_dot always sums all n_own entries: zero tails make both entry points sum the same m entries
        b[m:] = 0.0
        x, y = ctx.vector(a), ctx.vector(b)
        want = sp.dot(x, y)  # existing entry point (cudaMalloc'ed partials, result fetched to the host)
        assert noalloc_dot(ctx, m, x, y, out) == 0, ctx.lib.ddm_last_error()
        got = out.download()[0]  # the result stays on the device: first entry of `out`
        tol = 1e-13 * np.linalg.norm(a) * np.linalg.norm(b)
        assert abs(got - want) <= tol, (call, got, want)
        assert abs(got - float(a @ b)) <= tol
    # inputs untouched, and a repeat gives the same bits (fixed reduction tree)
    assert noalloc_dot(ctx, m, x, y, out) == 0
    assert out.download()[0] == got
    assert np.array_equal(x.download(), a)


def test_limits_and_empty_input(ctx):
    x, y, out = ctx.vector(np.ones(ctx.n_own)), ctx.vector(np.ones(ctx.n_own)), ctx.vector(np.full(ctx.n_own, 7.0))
    assert noalloc_dot(ctx, 0, x, y, out) == 0
    assert out.download()[0] == 0.0
    assert noalloc_dot(ctx, 5, x, y, out) == 0
    assert out.download()[0] == 5.0 and out.download()[1] == 7.0  # exactly one double is written
    assert noalloc_dot(ctx, CHUNK * MAX_PARTIALS + 1, x, y, out) != 0  # beyond the static scratch: refused, nothing launched
    msg = ctx.lib.ddm_last_error().decode()
    assert "exceeds the maximum" in msg and str(CHUNK * MAX_PARTIALS) in msg
    assert noalloc_dot(ctx, -1, x, y, out) != 0
    assert out.download()[0] == 5.0
