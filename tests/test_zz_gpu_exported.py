"""The GPU path on products that came through the dune-ddm exporter: tests/golden/exported_ranks.ddmpack was
written by scripts/make_exported_fixture.py -- four ranks dumped with exportRank from the reference's Dune containers, factors in
CHOLMOD's layouts (supernodal LL^T verbatim; simplicial LL^T / LDL^T regrouped into fundamental supernodes), merged by
ddm_b200/setup/rankpacks.py.  The CPU half pins the fixture with the oracle, the GPU half is the parity test."""
import os

import numpy as np
import pytest

FIXTURE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "exported_ranks.ddmpack")


@pytest.fixture(scope="module")
def exported():
    from ddm_b200.setup import pack

    return pack.load_products(FIXTURE)


def test_fixture_is_a_consistent_problem(exported):
    from ddm_b200.setup import factor
    from oracle import oracle

    P = exported
    assert P.nsub == 4 and P.n == 512 and P.A0.shape[0] == sum(z.shape[1] for z in P.Z)
    rng = np.random.default_rng(0)
    for s, F in zip(P.subs, P.factors):  # every factor, whatever layout it travelled in, inverts R_i A R_i^T
        Ai = P.A[s][:, s]
        x = rng.standard_normal(s.size)
        assert np.linalg.norm(factor.solve_supernodal(F, Ai @ x) - x) <= 1e-9 * np.linalg.norm(x)
    tot = np.zeros(P.n)
    for s, d in zip(P.subs, P.pou):
        np.add.at(tot, s, d)
    assert np.allclose(tot, 1.0, atol=1e-12)  # partition of unity (test_galerkin_coarse_matrix.cc:252-263)
    x, it, conv, hist = oracle.Oracle(P, nthreads=2).pcg(P.b, reduction=1e-8, maxit=100)
    assert conv and it < 40 and np.linalg.norm(P.b - P.A @ x) <= 1e-7 * np.linalg.norm(P.b)


@pytest.mark.gpu
def test_gpu_solves_the_exported_problem(exported):
    from ddm_b200 import api
    from oracle import oracle

    P = exported
    O = oracle.Oracle(P, nthreads=2)
    ctx = api.Context.from_products(P)
    try:
        op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
        d = np.random.default_rng(3).standard_normal(P.n)
        tol = max(1e-12, 8 * np.finfo(float).eps * max(np.linalg.cond(P.A0), max(np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs)))
        for mode, typ in (("additive", "standard"), ("multiplicative", "restricted")):
            Cp = api.CombinedPreconditioner(mode)
            Cp.set_op(op)
            Cp.add(api.SchwarzPreconditioner(ctx, typ))
            Cp.add(api.GalerkinPreconditioner(ctx))
            x = ctx.vector()
            Cp.apply(x, ctx.vector(d))
            want = O.twolevel_apply(d, restricted=typ == "restricted", mode=mode)
            assert np.linalg.norm(x.download() - want) <= tol * np.linalg.norm(want)
        Cp = api.CombinedPreconditioner("additive")
        Cp.add(api.SchwarzPreconditioner(ctx, "standard"))
        Cp.add(api.GalerkinPreconditioner(ctx))
        xv, bv = ctx.vector(), ctx.vector(P.b)
        res = api.CGSolver(op, sp, Cp, reduction=1e-8, maxit=100).solve_device(xv, bv)
        xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=100)
        assert res.converged and convo and abs(res.iterations - ito) <= 1
        assert np.linalg.norm(xv.download() - xo) <= 1e-8 * np.linalg.norm(xo)
    finally:
        ctx.close()
