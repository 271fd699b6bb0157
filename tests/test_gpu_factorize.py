"""Device numeric factorisation (ddm_factorize_subdomain) against the host factorisation of the same analysis, and the
apply path built on it against the oracle.  Reference: cholmod_l_factorize inside Dune::Cholmod::setMatrix
(extern/dune-istl/dune/istl/cholmod.hh:469-472), twolevel_schwarz.hh:99-104 (numeric update with unchanged index sets)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def pending(P):
    """The same products with every subdomain factor replaced by (analysis, A_i): numeric phase left to the device."""
    import copy

    from ddm_b200.setup import factor

    Q = copy.copy(P)
    Q.factors = []
    for s in P.subs:
        Ai = P.A[s][:, s].tocsr()
        Ai.sort_indices()
        co = P.coords[s] if P.coords is not None else None
        Q.factors.append(factor.PendingFactor(factor.analyse(Ai, co, 32), Ai))
    return Q


@pytest.fixture(scope="module")
def brick():
    """Q1 bricks (config C5's family) large enough for wide supernodes (separators of 14 x 14 nodes)."""
    from ddm_b200.setup import problem

    return problem.poisson_q1_cube(24, (2, 2, 2), overlap=2, coarse="pou", workers=1)


@pytest.mark.parametrize("name", ["small3d", "small2d", "brick"])
def test_device_panels_equal_host_panels(request, name):
    from ddm_b200 import api, lib

    assert b"dryrun" not in lib.load().ddm_version()
    P = request.getfixturevalue(name)
    Q = pending(P)
    ctx = api.Context.from_products(Q)
    worst = 0.0
    for i, F in enumerate(Q.factors):
        x = ctx.factorize(i, F, want_panels=True)  # same structure as already there: value update in place
        H = F.host()
        worst = max(worst, rel(x, H.x))
        # backward error of the factor itself: L L^T = P A P^T on a probe vector
        v = np.random.default_rng(i).standard_normal(F.n)
        from ddm_b200.setup import factor

        dev = factor.SupernodalFactor(n=H.n, perm=H.perm, super_=H.super_, pi=H.pi, px=H.px, s=H.s, x=x)
        y = factor.solve_supernodal(dev, F.A @ v)
        assert rel(y, v) <= 1e-9
    # cond-independent: both run the same multifrontal recurrences in FP64, only summation order differs
    assert worst <= 1e-11, worst
    ctx.close()


@pytest.mark.parametrize("name", ["small3d", "brick"])
def test_apply_and_pcg_with_device_factors(request, name):
    from ddm_b200 import api
    from oracle import oracle

    P = request.getfixturevalue(name)
    ctx = api.Context.from_products(pending(P))
    O = oracle.Oracle(P, nthreads=4)
    d = np.random.default_rng(5).standard_normal(P.n)
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
    C = api.CombinedPreconditioner("additive")
    C.set_op(op)
    C.add(api.SchwarzPreconditioner(ctx, "standard"))
    C.add(api.GalerkinPreconditioner(ctx))
    got, want = C.apply_host(d), O.twolevel_apply(d)
    kap = max(np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs) if P.n < 4000 else 1e6
    assert rel(got, want) <= max(1e-12, 8 * np.finfo(float).eps * max(kap, np.linalg.cond(P.A0)))
    cg = api.CGSolver(op, sp, C, reduction=1e-8, maxit=200)
    x, res = cg.solve_host(P.b)
    xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=200)
    assert res.converged and convo and abs(res.iterations - ito) <= 1
    assert rel(x, xo) <= 1e-8


def test_indefinite_matrix_is_refused(small3d):
    from ddm_b200 import api
    from ddm_b200.lib import DDMError

    Q = pending(small3d)
    ctx = api.Context.from_products(Q)
    F = Q.factors[0]
    bad = F.A.copy()
    bad.data = -bad.data
    import dataclasses

    with pytest.raises(DDMError, match="positive definite"):
        ctx.factorize(0, dataclasses.replace(F, A=bad))
    ctx.close()
