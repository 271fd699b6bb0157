"""TwoLevelSchwarzSolver facade (dune/ddm/twolevel_schwarz.hh:58-146): first apply builds everything, later applies with a matrix
of the same pattern refresh values only (update branch :99-104) -- checked against a fresh build and a direct solve."""
import numpy as np
import pytest
import scipy.sparse.linalg as spla

pytestmark = pytest.mark.gpu


def _dg(scale):
    from ddm_b200.setup import dg

    def kap(cent):
        return scale * dg.kappa_cd(cent)

    return dg.assemble_cd_dg(40, kappa_fn=kap)


def test_facade_first_apply_and_value_update():
    from ddm_b200 import twolevel
    from ddm_b200.setup import decomp

    A, rhs, coords, cell_dofs, cent = _dg(1.0)
    part = decomp.rcb_partition(cent, 6)
    cfg = {"overlap": 1, "mode": "multiplicative", "fine.type": "restricted", "solver.restart": 50, "solver.maxit": 300}
    S = twolevel.TwoLevelSchwarzSolver(coords, cell_dofs, part, 6, config=cfg, symmetric=False, workers=1)
    try:
        z, r = np.zeros(A.shape[0]), rhs.copy()
        res1 = S.apply(A, z, r, 1e-10)
        assert res1.converged
        x_ref = spla.spsolve(A.tocsc(), rhs)
        assert np.linalg.norm(z - x_ref) <= 1e-7 * np.linalg.norm(x_ref)
        launches_setup = S.ctx.launches()
        # "Newton step": same pattern, new values (diffusion scaled by 3)
        A2, rhs2, *_ = _dg(3.0)
        A2 = A2.tocsr()
        assert np.array_equal(A2.indptr, S.P.A.indptr) and np.array_equal(A2.indices, S.P.A.indices)
        z2, r2 = np.zeros(A.shape[0]), rhs2.copy()
        res2 = S.apply(A2, z2, r2, 1e-10)
        assert res2.converged and S.updates == 1
        x2_ref = spla.spsolve(A2.tocsc(), rhs2)
        assert np.linalg.norm(z2 - x2_ref) <= 1e-7 * np.linalg.norm(x2_ref)
        # same iteration count as a solver built from scratch on the new matrix: the in-place update lost nothing
        F = twolevel.TwoLevelSchwarzSolver(coords, cell_dofs, part, 6, config=cfg, symmetric=False, workers=1)
        try:
            z3, r3 = np.zeros(A.shape[0]), rhs2.copy()
            res3 = F.apply(A2, z3, r3, 1e-10)
            assert abs(res3.iterations - res2.iterations) <= 1
            assert np.linalg.norm(z3 - z2) <= 1e-8 * np.linalg.norm(z2)
        finally:
            F.close()
        with pytest.raises(Exception):
            S.apply(A2[:-4, :-4], z2[:-4], r2[:-4], 1e-8)  # pattern change is refused
        assert S.ctx.launches() > launches_setup
    finally:
        S.close()


def test_facade_symmetric_cg():
    from ddm_b200 import twolevel
    from ddm_b200.setup import decomp, fem, mesh

    coords, cells = mesh.structured_simplices((10, 10, 10))
    cent = coords[cells].mean(axis=1)
    Ke = fem.p1_element_matrices(coords, cells, fem.kappa_islands(cent))
    dirichlet = (coords[:, 0] < 1e-6) | (coords[:, 0] > 1 - 1e-6)
    Afull = fem.assemble(coords.shape[0], cells, Ke)
    A = fem.eliminate_dirichlet(Afull, dirichlet)
    b = np.where(dirichlet, 0.0, Afull @ np.where(dirichlet, 1.0 - coords[:, 0], 0.0))
    part = decomp.rcb_partition(cent, 8)
    S = twolevel.TwoLevelSchwarzSolver(coords, cells, part, 8, config={"overlap": 2, "solver.type": "cgsolver", "fine.type": "standard"},
                                       dirichlet=dirichlet, workers=1)
    try:
        z, r = np.zeros(A.shape[0]), b.copy()
        res = S.apply(A, z, r, 1e-8)
        assert res.converged and res.iterations < 60
        assert np.linalg.norm(b - A @ z) <= 1.05e-8 * np.linalg.norm(b)
    finally:
        S.close()
