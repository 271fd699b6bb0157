"""Edge cases of the Krylov entry points on the device, against the oracle and Dune's conventions
(extern/dune-istl/dune/istl/solver.hh:457-480: converged when def < def0 * reduction or def < 1e-30, SolverAbort on a NaN / infinite
defect; solvers.hh:296-303: a solve that starts converged does no iteration).  Written after the last GPU call of round 1."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(small3d):
    from ddm_b200 import api
    from oracle import oracle

    P = small3d
    ctx = api.Context.from_products(P)
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)

    def prec(mode="additive", typ="standard"):
        Cp = api.CombinedPreconditioner(mode)
        Cp.set_op(op)
        Cp.add(api.SchwarzPreconditioner(ctx, typ))
        Cp.add(api.GalerkinPreconditioner(ctx))
        return Cp

    yield P, ctx, op, sp, prec, oracle.Oracle(P, nthreads=2)
    ctx.close()


def test_zero_right_hand_side_needs_no_iteration(env):
    from ddm_b200 import api

    P, ctx, op, sp, prec, O = env
    for solver in (api.CGSolver(op, sp, prec(), reduction=1e-8, maxit=50),
                   api.RestartedGMResSolver(op, sp, prec("multiplicative", "restricted"), reduction=1e-8, restart=10, maxit=50)):
        x, b = ctx.vector(np.zeros(P.n)), ctx.vector(np.zeros(P.n))
        res = solver.solve_device(x, b)
        assert res.converged and res.iterations == 0
        assert np.all(x.download() == 0.0)
    _, it, conv, _ = O.pcg(np.zeros(P.n), reduction=1e-8, maxit=50)
    assert conv and it == 0


def test_iteration_cap_reports_not_converged(env):
    from ddm_b200 import api

    P, ctx, op, sp, prec, O = env
    xo, ito, convo, histo = O.pcg(P.b, reduction=1e-8, maxit=2)
    assert ito == 2 and not convo
    x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
    res = api.CGSolver(op, sp, prec(), reduction=1e-8, maxit=2).solve_device(x, b)
    assert res.iterations == 2 and not res.converged
    assert np.allclose(res.history[:3], histo[:3], rtol=1e-8)
    assert np.linalg.norm(x.download() - xo) <= 1e-8 * np.linalg.norm(xo)


def test_short_restart_gmres(env):
    from ddm_b200 import api

    P, ctx, op, sp, prec, O = env
    xo, ito, convo, _ = O.gmres(P.b, reduction=1e-8, restart=2, maxit=300, restricted=True, mode="multiplicative")
    assert convo
    x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
    res = api.RestartedGMResSolver(op, sp, prec("multiplicative", "restricted"), reduction=1e-8, restart=2, maxit=300).solve_device(x, b)
    # restarting every second step amplifies rounding differences in the stopping test: a couple of steps either way
    assert res.converged and abs(res.iterations - ito) <= 2
    assert np.linalg.norm(x.download() - xo) <= 1e-6 * np.linalg.norm(xo)


def test_nan_defect_aborts_and_the_context_stays_usable(env):
    from ddm_b200 import api
    from ddm_b200.lib import DDMError

    P, ctx, op, sp, prec, O = env
    bad = P.b.copy()
    bad[P.n // 2] = np.nan
    cg = api.CGSolver(op, sp, prec(), reduction=1e-8, maxit=100)
    with pytest.raises(DDMError, match="NaN"):  # Dune::SolverAbort, solver.hh:458-466
        cg.solve_device(ctx.vector(np.zeros(P.n)), ctx.vector(bad))
    gm = api.RestartedGMResSolver(op, sp, prec("multiplicative", "restricted"), reduction=1e-8, restart=10, maxit=100)
    with pytest.raises(DDMError, match="NaN"):
        gm.solve_device(ctx.vector(np.zeros(P.n)), ctx.vector(bad))
    # the failed solves leave nothing behind: the next one is a normal one
    xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=100)
    x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
    res = cg.solve_device(x, b)
    assert res.converged and abs(res.iterations - ito) <= 1
    assert np.linalg.norm(x.download() - xo) <= 1e-8 * np.linalg.norm(xo)


def test_wrong_vector_length_is_refused(env):
    from ddm_b200.lib import DDMError

    P, ctx, *_ = env
    with pytest.raises(DDMError):
        ctx.vector(np.zeros(P.n + 1))


def test_restricted_schwarz_without_a_partition_of_unity(small3d):
    """schwarz.hh:138-141: the weights are applied only when a partition of unity was given; without one the restricted variant
    is the plain sum over the subdomains, i.e. what the standard variant computes."""
    from ddm_b200 import api
    from oracle import oracle

    P = small3d
    ctx = api.Context.from_products(P, with_pou=False)
    try:
        d = np.random.default_rng(9).standard_normal(P.n)
        x = ctx.vector()
        api.SchwarzPreconditioner(ctx, "restricted").apply(x, ctx.vector(d))
        want = oracle.Oracle(P, nthreads=2).schwarz_apply(d, restricted=False)
        tol = max(1e-12, 8 * np.finfo(float).eps * max(np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs))
        assert np.linalg.norm(x.download() - want) <= tol * np.linalg.norm(want)
    finally:
        ctx.close()
