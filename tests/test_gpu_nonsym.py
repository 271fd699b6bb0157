"""GPU parity on the non-symmetric path (config C4 shape): LU subdomain factors through ddm_upload_factor_lu, LU coarse
matrix, restricted additive Schwarz, multiplicative coarse correction, GMRES -- against the CPU oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EPS = np.finfo(float).eps


def rel(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.fixture(scope="module")
def case(cd_dg):
    from ddm_b200 import api
    from oracle import oracle

    P = cd_dg
    ctx = api.Context.from_products(P)
    kap = max(np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs)
    yield P, ctx, oracle.Oracle(P, nthreads=4), max(1e-12, 8 * EPS * kap), max(1e-12, 8 * EPS * np.linalg.cond(P.A0))
    ctx.close()


@pytest.mark.parametrize("typ", ["restricted", "standard"])
def test_schwarz_apply_lu(case, typ):
    from ddm_b200 import api

    P, ctx, O, tol, _ = case
    d = np.random.default_rng(2).standard_normal(P.n)
    x = ctx.vector()
    api.SchwarzPreconditioner(ctx, typ).apply(x, ctx.vector(d))
    assert rel(x.download(), O.schwarz_apply(d, restricted=typ == "restricted")) <= tol


def test_coarse_apply_lu(case):
    from ddm_b200 import api

    P, ctx, O, _, tol0 = case
    d = np.random.default_rng(3).standard_normal(P.n)
    x = ctx.vector()
    api.GalerkinPreconditioner(ctx).apply(x, ctx.vector(d))
    assert rel(x.download(), O.coarse_apply(d)) <= tol0


@pytest.mark.parametrize("mode", ["multiplicative", "additive"])
def test_twolevel_apply_lu(case, mode):
    from ddm_b200 import api

    P, ctx, O, tol, tol0 = case
    d = np.random.default_rng(4).standard_normal(P.n)
    op = api.NonOverlappingOperator(ctx)
    C = api.CombinedPreconditioner(mode)
    C.set_op(op)
    C.add(api.SchwarzPreconditioner(ctx, "restricted"))
    C.add(api.GalerkinPreconditioner(ctx))
    want = O.twolevel_apply(d, restricted=True, mode=mode)
    assert rel(C.apply_host(d), want) <= max(tol, tol0)


def test_gmres_ras_as_shipped(case):
    """convectiondiffusiondg.ini: restricted + multiplicative + GMRES(50), reduction 1e-8."""
    from ddm_b200 import api

    P, ctx, O, _, _ = case
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
    C = api.CombinedPreconditioner("multiplicative")
    C.set_op(op)
    C.add(api.SchwarzPreconditioner(ctx, "restricted"))
    C.add(api.GalerkinPreconditioner(ctx))
    gm = api.RestartedGMResSolver(op, sp, C, reduction=1e-8, restart=50, maxit=300)
    xo, ito, convo, histo = O.gmres(P.b, reduction=1e-8, restart=50, maxit=300, restricted=True, mode="multiplicative")
    assert convo
    x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
    r = gm.solve_device(x, b)
    assert r.converged and abs(r.iterations - ito) <= 1
    assert np.allclose(r.history[:4], histo[:4], rtol=1e-7)
    xh, r2 = gm.solve_host(P.b)
    assert r2.converged and abs(r2.iterations - ito) <= 1
    for got in (x.download(), xh):
        assert np.linalg.norm(got - xo) <= 1e-6 * np.linalg.norm(xo)
        assert np.linalg.norm(P.b - P.A @ got) <= 1e-6 * np.linalg.norm(P.b)
    x1, b1 = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
    r3 = gm.apply(x1, b1)  # literal drop-in loop
    assert r3.converged and abs(r3.iterations - ito) <= 1
