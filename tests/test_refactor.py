"""Numeric re-factorisation with a kept symbolic analysis ("matrix value update without re-analysis",
twolevel_schwarz.hh:99-104): the numeric phase on new values must give exactly what a factorisation from scratch gives."""
import numpy as np
import pytest
import scipy.sparse as sp


def _same(F, G):
    names = ("perm", "super_", "pi", "px", "s", "x") + (("xu",) if hasattr(F, "xu") else ())
    return all(np.array_equal(getattr(F, a), getattr(G, a)) for a in names)


@pytest.mark.parametrize("name", ["small3d", "cd_dg"])
def test_numeric_phase_equals_full_factorisation(request, name):
    from ddm_b200.setup import factor

    P = request.getfixturevalue(name)
    sym = not hasattr(P.factors[0], "xu")
    s = P.subs[1]
    Ai = P.A[s][:, s].tocsr()
    Ai.sort_indices()
    an = factor.analyse(Ai, None, 32, symmetric=sym)
    numeric, full = (factor.numeric_cholesky, factor.cholesky_supernodal) if sym else (factor.numeric_lu, factor.lu_supernodal)
    assert _same(numeric(an, Ai.data), full(Ai, None, 32))
    # new values on the same pattern (a Newton step changes the coefficient, not the mesh)
    rng = np.random.default_rng(0)
    scale = sp.diags(1.0 + 0.3 * rng.random(Ai.shape[0]))
    Bi = (scale @ Ai @ scale).tocsr()
    Bi.sort_indices()
    assert np.array_equal(Bi.indices, Ai.indices) and np.array_equal(Bi.indptr, Ai.indptr)
    G = numeric(an, Bi.data)
    assert _same(G, full(Bi, None, 32))
    d = rng.standard_normal(Ai.shape[0])
    x = factor.solve_supernodal(G, d) if sym else factor.solve_supernodal_lu(G, d)
    assert np.linalg.norm(Bi @ x - d) <= 1e-10 * np.linalg.norm(d)


@pytest.mark.parametrize("workers", [1, 2])
def test_facade_keeps_the_analysis_across_updates(small3d, workers):
    """TwoLevelSchwarzSolver._factors: first call analyses, later calls only run the numeric phase on values picked out of the
    new operator's data array; both must equal factorising R_i A R_i^T from scratch."""
    from ddm_b200 import twolevel
    from ddm_b200.setup import factor

    P = small3d
    S = twolevel.TwoLevelSchwarzSolver(P.coords, np.zeros((0, 4), dtype=np.int64), np.zeros(0, dtype=np.int64), P.nsub, workers=workers)
    S._subs = P.subs
    F1 = S._factors(P.A)
    assert S._analysis is not None and all(a is not None for a in S._analysis)
    kept = [id(a) for a in S._analysis]
    rng = np.random.default_rng(1)
    scale = sp.diags(1.0 + 0.2 * rng.random(P.n))
    A2 = (scale @ P.A @ scale).tocsr()
    A2.sort_indices()
    F2 = S._factors(A2)
    assert [id(a) for a in S._analysis] == kept  # nothing was re-analysed
    for i, s in enumerate(P.subs):
        for A, F in ((P.A, F1[i]), (A2, F2[i])):
            Ai = A[s][:, s].tocsr()
            Ai.sort_indices()
            assert _same(F, factor.cholesky_supernodal(Ai, P.coords[s], 32))
