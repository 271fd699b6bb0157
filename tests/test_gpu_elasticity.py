"""GPU parity on config C3's shape: 3x3 block-CSR SpMV (ddm_upload_operator_bcsr3) and the two-level method on P1
elasticity, against the CPU oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EPS = np.finfo(float).eps


def rel(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.fixture(scope="module")
def case(elast):
    from ddm_b200 import api
    from oracle import oracle

    P = elast
    ctx = api.Context.from_products(P)
    kap = max(np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs)
    yield P, ctx, oracle.Oracle(P, nthreads=4), max(1e-12, 8 * EPS * max(kap, np.linalg.cond(P.A0)))
    ctx.close()


def test_bcsr3_spmv(case):
    from ddm_b200 import api

    P, ctx, O, _ = case
    rng = np.random.default_rng(0)
    xh, yh = rng.standard_normal(P.n), rng.standard_normal(P.n)
    op = api.NonOverlappingOperator(ctx)
    x, y = ctx.vector(xh), ctx.vector(yh)
    op.apply(x, y)
    want = O.spmv(xh)
    assert np.max(np.abs(y.download() - want)) <= 1e-14 * np.max(np.abs(want)) * 30  # ~30 terms per row, any order
    y.upload(yh)
    op.applyscaleadd(-0.5, x, y)
    assert rel(y.download(), yh - 0.5 * want) <= 1e-14
    B = P.A.tobsr((3, 3))
    assert ctx.stat("spmv") == 76.0 * B.indices.size + 4.0 * (P.n // 3 + 1) + 48.0 * (P.n // 3)


@pytest.mark.parametrize("typ,mode", [("standard", "additive"), ("restricted", "multiplicative")])
def test_twolevel_apply_elasticity(case, typ, mode):
    from ddm_b200 import api

    P, ctx, O, tol = case
    d = np.random.default_rng(1).standard_normal(P.n)
    op = api.NonOverlappingOperator(ctx)
    C = api.CombinedPreconditioner(mode)
    C.set_op(op)
    C.add(api.SchwarzPreconditioner(ctx, typ))
    C.add(api.GalerkinPreconditioner(ctx))
    assert rel(C.apply_host(d), O.twolevel_apply(d, restricted=typ == "restricted", mode=mode)) <= tol


def test_solvers_elasticity(case):
    from ddm_b200 import api

    P, ctx, O, _ = case
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
    C = api.CombinedPreconditioner("additive")
    C.add(api.SchwarzPreconditioner(ctx, "standard"))
    C.add(api.GalerkinPreconditioner(ctx))
    cg = api.CGSolver(op, sp, C, reduction=1e-8, maxit=300)
    xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=300)
    xh, r = cg.solve_host(P.b)
    assert r.converged and abs(r.iterations - ito) <= 1
    assert np.linalg.norm(P.b - P.A @ xh) <= 1.05e-8 * np.linalg.norm(P.b)
    # as shipped (linearelasticity.ini): restricted + multiplicative + GMRES(100), reduction 1e-6
    Cm = api.CombinedPreconditioner("multiplicative")
    Cm.set_op(op)
    Cm.add(api.SchwarzPreconditioner(ctx, "restricted"))
    Cm.add(api.GalerkinPreconditioner(ctx))
    gm = api.RestartedGMResSolver(op, sp, Cm, reduction=1e-6, restart=100, maxit=1000)
    xg, itg, convg, _ = O.gmres(P.b, reduction=1e-6, restart=100, maxit=1000, restricted=True, mode="multiplicative")
    xh2, r2 = gm.solve_host(P.b)
    assert r2.converged and abs(r2.iterations - itg) <= 1
    assert np.linalg.norm(xh2 - xg) <= 1e-4 * np.linalg.norm(xg)
