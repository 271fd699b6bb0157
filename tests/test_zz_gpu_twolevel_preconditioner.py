"""TwoLevelSchwarzPreconditioner (examples/pdelab_schwarz.hh:26-203): one object from a problem description and a parameter tree,
behind the Dune::Preconditioner interface.  CPU half: configuration errors; GPU half: each supported coarse-space type applies like
the oracle and drives CG / GMRES to the oracle's iteration count.  Written after the last GPU call of round 1."""
import numpy as np
import pytest


def _problem(ncell=8, nsub=4):
    from ddm_b200.setup import decomp, fem, mesh

    coords, cells = mesh.structured_simplices((ncell,) * 3)
    cent = coords[cells].mean(axis=1)
    Ke = fem.p1_element_matrices(coords, cells, fem.kappa_islands(cent))
    dirichlet = (coords[:, 0] < 1e-9) | (coords[:, 0] > 1 - 1e-9)
    rhs = fem.assemble(coords.shape[0], cells, Ke) @ np.where(dirichlet, 1 - coords[:, 0], 0.0)
    return dict(coords=coords, cells=cells, Ke=Ke, dirichlet=dirichlet, rhs=rhs, cell_part=decomp.rcb_partition(cent, nsub), nsub=nsub)


def test_unknown_and_unsupported_coarse_spaces_are_refused():
    from ddm_b200 import twolevel
    from ddm_b200.lib import DDMError

    with pytest.raises(DDMError, match="Unknown coarse space type"):  # Dune::NotImplemented, pdelab_schwarz.hh:63
        twolevel.TwoLevelSchwarzPreconditioner(_problem(), {"coarsespace.type": "bogus"})
    with pytest.raises(DDMError, match="not implemented"):
        twolevel.TwoLevelSchwarzPreconditioner(_problem(), {"coarsespace.type": "msgfem_ring"})
    with pytest.raises(DDMError, match="Unknown solver type"):
        twolevel.TwoLevelSchwarzSolver(np.zeros((1, 2)), np.zeros((0, 3), dtype=np.int64), np.zeros(0, dtype=np.int64), 1, {"solver.type": "bicgstab"})


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["geneo", "pou", "none"])
def test_preconditioner_object_matches_the_oracle(kind):
    from ddm_b200 import api, twolevel
    from oracle import oracle

    T = twolevel.TwoLevelSchwarzPreconditioner(_problem(), {"coarsespace.type": kind, "coarse_space.nev": 4, "schwarz.type": "standard"}, workers=1)
    try:
        P, ctx = T.P, T.ctx
        assert T.category == "nonoverlapping" and (P.A0.shape[0] == 0) == (kind == "none")
        O = oracle.Oracle(P, nthreads=2)
        with_coarse = kind != "none"
        d = np.random.default_rng(1).standard_normal(P.n)
        v = ctx.vector()
        T.pre(v, None)
        T.apply(v, ctx.vector(d))
        T.post(v)
        want = O.twolevel_apply(d, with_coarse=with_coarse)
        conds = [np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs] + ([np.linalg.cond(P.A0)] if with_coarse else [])
        tol = max(1e-12, 8 * np.finfo(float).eps * max(conds))
        assert np.linalg.norm(v.download() - want) <= tol * np.linalg.norm(want)
        xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=200, with_coarse=with_coarse)
        cg = api.CGSolver(T.getNonOverlappingOperator(), T.getScalarProduct(), T.combined, reduction=1e-8, maxit=200)
        x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
        res = cg.solve_device(x, b)
        assert res.converged and convo and abs(res.iterations - ito) <= 1
        assert np.linalg.norm(x.download() - xo) <= 1e-8 * np.linalg.norm(xo)
    finally:
        T.close()


@pytest.mark.gpu
def test_ragged_coarse_space_from_an_eigenvalue_threshold():
    """coarse_space.threshold > 0 (eigensolver_params.hh:30, spectra.hh:157-163): subdomains keep different numbers of
    eigenvectors; the device path (ragged k_i in ddm_upload_coarse) applies like the oracle."""
    from ddm_b200 import api, twolevel
    from oracle import oracle

    T = twolevel.TwoLevelSchwarzPreconditioner(_problem(10, 6), {"overlap": 2, "coarse_space.nev": 4, "coarse_space.nev_max": 8, "coarse_space.threshold": 0.95,
                                                                "schwarz.type": "standard"}, workers=1)
    try:
        P, ctx = T.P, T.ctx
        ks = [z.shape[1] for z in P.Z]
        assert len(set(ks)) > 1 and P.A0.shape[0] == sum(ks)
        O = oracle.Oracle(P, nthreads=2)
        d = np.random.default_rng(2).standard_normal(P.n)
        v = ctx.vector()
        T.apply(v, ctx.vector(d))
        want = O.twolevel_apply(d)
        tol = max(1e-12, 8 * np.finfo(float).eps * max([np.linalg.cond(P.A0)] + [np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs]))
        assert np.linalg.norm(v.download() - want) <= tol * np.linalg.norm(want)
        xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=200)
        x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
        res = api.CGSolver(T.getNonOverlappingOperator(), T.getScalarProduct(), T.combined, reduction=1e-8, maxit=200).solve_device(x, b)
        assert res.converged and convo and abs(res.iterations - ito) <= 1
    finally:
        T.close()
