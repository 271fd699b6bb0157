"""Non-symmetric path on the CPU side: the supernodal LU restatement (stand-in for UMFPACK), the oracle's LU
subdomain solve and LU coarse solve, on config C4's Q1-DG convection-diffusion problem."""
import numpy as np
import pytest
import scipy.sparse.linalg as spla


def rel(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def test_dg_matrix_properties(cd_dg):
    P = cd_dg
    assert P.n == 4 * 48 * 48
    assert abs(P.A - P.A.T).max() > 1e-3  # convection: non-symmetric
    assert 15.0 < P.A.nnz / P.n < 21.0  # ~20 nnz/row for C4 (structural zeros dropped)
    S = ((P.A + P.A.T) * 0.5).toarray()
    assert np.linalg.eigvalsh(S).min() > 0.0  # coercive (weighted SIPG penalty), so LU without pivoting is safe
    assert P.meta["symmetric"] is False and all(hasattr(F, "xu") for F in P.factors)
    assert abs(P.A0 - P.A0.T).max() > 1e-9 * abs(P.A0).max()  # small next to the 1e5-diffusion block, but not round-off


def test_supernodal_lu_solves(cd_dg):
    from ddm_b200.setup import factor

    P = cd_dg
    rng = np.random.default_rng(0)
    wide = 0
    for i, s in enumerate(P.subs):
        F = P.factors[i]
        wide = max(wide, int(np.diff(F.super_).max()))
        Ai = P.A[s][:, s].tocsc()
        b = rng.standard_normal(s.size)
        x = factor.solve_supernodal_lu(F, b)
        eta = np.linalg.norm(Ai @ x - b) / (spla.norm(Ai) * np.linalg.norm(x) + np.linalg.norm(b))
        assert eta <= 1e-14
    assert wide > 32  # the fixture exercises the wide-supernode (cuBLAS) conversion path on the GPU


def test_oracle_lu_paths(cd_dg):
    from oracle import oracle

    P = cd_dg
    O = oracle.Oracle(P, nthreads=2)
    assert O.nonsym
    d = np.random.default_rng(1).standard_normal(P.n)
    kap = max(np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs)
    tol = 8 * np.finfo(float).eps * kap
    acc = np.zeros(P.n)
    for i, s in enumerate(P.subs):
        xi = spla.spsolve(P.A[s][:, s].tocsc(), d[s])
        assert rel(O.sub_solve(i, d[s]), xi) <= tol
        np.add.at(acc, s, P.pou[i] * xi)
    assert rel(O.schwarz_apply(d, restricted=True), acc) <= tol
    # coarse level with the dense LU of the non-symmetric A0
    off = np.concatenate(([0], np.cumsum([z.shape[1] for z in P.Z])))
    d0 = np.concatenate([Z.T @ d[s] for Z, s in zip(P.Z, P.subs)])
    y = np.linalg.solve(P.A0, d0)
    want = np.zeros(P.n)
    for i, s in enumerate(P.subs):
        np.add.at(want, s, P.Z[i] @ y[off[i] : off[i + 1]])
    assert rel(O.coarse_apply(d), want) <= 8 * np.finfo(float).eps * np.linalg.cond(P.A0)
    x, it, conv, hist = O.gmres(P.b, reduction=1e-8, restart=50, maxit=300, restricted=True, mode="multiplicative")
    assert conv and it < 60
    assert np.linalg.norm(P.b - P.A @ x) <= 1e-6 * np.linalg.norm(P.b)


def test_pack_roundtrip_keeps_lu(cd_dg, tmp_path):
    from ddm_b200.setup import pack

    path = str(tmp_path / "cd.ddmpack")
    pack.save_products(path, cd_dg)
    Q = pack.load_products(path)
    assert all(np.array_equal(a.xu, b.xu) and np.array_equal(a.x, b.x) for a, b in zip(cd_dg.factors, Q.factors))


def test_lu_without_pivoting_refuses_a_matrix_that_needs_it():
    """the restated LU never pivots; a factor that is not backward stable must stop the setup (UMFPACK would pivot)."""
    import scipy.sparse as sp

    from ddm_b200.setup import factor

    n = 40
    A = sp.diags([np.ones(n - 1), np.full(n, 1e-13), 2 * np.ones(n - 1)], [-1, 0, 1]).tocsr()  # tiny pivots, well-conditioned with row swaps
    with pytest.raises(np.linalg.LinAlgError, match="pivot"):
        factor.lu_supernodal(A, None, 8)
    B = sp.diags([-np.ones(n - 1), 4 * np.ones(n), -2 * np.ones(n - 1)], [-1, 0, 1]).tocsr()
    assert factor.check_lu_backward_error(B, factor.numeric_lu(factor.analyse(B, None, 8, symmetric=False), B.data)) < 1e-14
