"""The CPU oracle against (a) the rank-wise restatement of what dune-ddm literally executes and
(b) the mathematical invariants of SURVEY.md section 4 (the reference has no test of the apply path)."""
import numpy as np
import pytest

from oracle import oracle


def rel(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.fixture(scope="module", params=["small3d", "small2d"])
def case(request):
    P = request.getfixturevalue(request.param)
    return P, oracle.Oracle(P, nthreads=2), oracle.RankwiseReference(P, P.n_novlp)


def test_direct_solver_boundary_residual_identity(case):
    # normwise backward error of the direct solve (the identity ||A_i x_i - d_i|| <= 1e-12 ||d_i|| of SURVEY section 4,
    # scaled for the 1e6 coefficient contrast): pins factor + sweeps at the CHOLMOD boundary (cholmod.hh:293)
    P, O, _ = case
    rng = np.random.default_rng(3)
    for i, s in enumerate(P.subs):
        Ai = P.A[s][:, s]
        d = rng.standard_normal(s.size)
        x = O.sub_solve(i, d)
        import scipy.sparse.linalg as spla

        eta = np.linalg.norm(Ai @ x - d) / (spla.norm(Ai) * np.linalg.norm(x) + np.linalg.norm(d))
        assert eta <= 1e-14


def _tol(P):
    """Two different direct solvers (the oracle's supernodal sweeps, LAPACK's dense Cholesky in the rank-wise
    restatement) agree to eps * cond at best; 1e-12 is the bar wherever the conditioning allows it."""
    kap = max(np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs)
    kap = max(kap, np.linalg.cond(P.A0)) if P.A0.size else kap
    return max(1e-12, 8 * np.finfo(float).eps * kap)


def test_unique_dof_formulation_equals_rankwise(case):
    P, O, R = case
    rng = np.random.default_rng(0)
    d = rng.standard_normal(P.n)
    tol = _tol(P)
    for restricted in (False, True):
        a = O.twolevel_apply(d, restricted=restricted)
        b = R.twolevel(d, restricted=restricted)
        assert rel(a, b) <= tol
    dr = R.to_ranks(d)
    assert rel(O.schwarz_apply(d), R.from_ranks(R.schwarz(dr, False))) <= tol
    assert rel(O.coarse_apply(d), R.from_ranks(R.coarse(dr))) <= tol


def test_rank_vectors_are_consistent(case):
    # every rank's copy of a shared DOF holds the same value after an apply
    P, _, R = case
    rng = np.random.default_rng(1)
    xr = R.schwarz(R.to_ranks(rng.standard_normal(P.n)), True)
    g = R.from_ranks(xr)
    for r, (s, m) in enumerate(zip(P.subs, P.n_novlp)):
        assert rel(xr[r], g[s[:m]]) <= 1e-13


def test_pou_sums_to_one_and_vanishes_on_the_boundary(case):
    P, _, _ = case
    tot = np.zeros(P.n)
    for s, w in zip(P.subs, P.pou):
        np.add.at(tot, s, w)
    assert np.max(np.abs(tot - 1.0)) <= 1e-10  # is_pou, examples/poisson.cc:141-156


def test_additive_standard_preconditioner_is_symmetric(case):
    P, O, _ = case
    rng = np.random.default_rng(2)
    u, v = rng.standard_normal(P.n), rng.standard_normal(P.n)
    a = v @ O.twolevel_apply(u)
    b = u @ O.twolevel_apply(v)
    assert abs(a - b) <= 1e-10 * max(abs(a), abs(b))


def test_multiplicative_combine(case):
    P, O, _ = case
    rng = np.random.default_rng(4)
    d = rng.standard_normal(P.n)
    x1 = O.schwarz_apply(d)
    want = x1 + O.coarse_apply(d - P.A @ x1)  # combined_preconditioner.hh:144-158
    assert rel(O.twolevel_apply(d, mode="multiplicative"), want) <= 1e-13


def test_a0_is_galerkin_product(case):
    P, _, _ = case
    import scipy.sparse as sp

    cols = []
    for s, Z in zip(P.subs, P.Z):
        for k in range(Z.shape[1]):
            v = np.zeros(P.n)
            v[s] = Z[:, k]
            cols.append(v)
    Zg = np.stack(cols, 1)
    A0 = Zg.T @ (P.A @ Zg)
    assert np.linalg.norm(A0 - P.A0) <= 1e-12 * np.linalg.norm(A0)
    assert np.linalg.norm(P.A0 - P.A0.T) <= 1e-12 * np.linalg.norm(A0)


def test_pcg_converges_and_coarse_space_helps(case):
    P, O, _ = case
    x2, it2, conv2, h2 = O.pcg(P.b, reduction=1e-8, maxit=300)
    x1, it1, conv1, h1 = O.pcg(P.b, reduction=1e-8, maxit=300, with_coarse=False)
    assert conv1 and conv2
    assert it2 <= it1
    r = P.b - P.A @ x2
    assert np.linalg.norm(r) <= 1.01e-8 * np.linalg.norm(P.b)
    assert rel(x1, x2) <= 1e-3  # both solve A x = b to 1e-8 in the residual; cond(A) ~ 1e7


def test_threaded_oracle_equals_serial(case):
    P, O, _ = case
    rng = np.random.default_rng(5)
    d = rng.standard_normal(P.n)
    O1 = oracle.Oracle(P, nthreads=1)
    assert np.array_equal(O1.twolevel_apply(d), O.twolevel_apply(d))


def test_single_subdomain_is_a_direct_solve():
    """Edge case: one subdomain (no overlap region, so GenEO contributes no coarse vector): the one-level method is the exact
    inverse and CG converges in one step."""
    from ddm_b200.setup import problem
    from oracle import oracle

    P = problem.unitcube_p1(6, 1, overlap=1, nev=3, workers=1)
    assert P.A0.shape == (0, 0) and P.Z[0].shape[1] == 0
    O = oracle.Oracle(P)
    x, it, conv, _ = O.pcg(P.b, reduction=1e-8, maxit=10, with_coarse=False)
    assert conv and it <= 2
    assert np.linalg.norm(P.b - P.A @ x) <= 1e-10 * np.linalg.norm(P.b)


def test_q1_brick_family_converges():
    """Config C5 family at toy size (structured Q1, channels coefficient, brick subdomains): the setup restatement gives a
    consistent problem and the two-level method converges on it."""
    from ddm_b200.setup import problem
    from oracle import oracle

    P = problem.poisson_q1_cube(12, (2, 2, 2), overlap=2, coarse="geneo", nev=6, workers=1)
    assert P.n == 13 ** 3 and P.nsub == 8 and abs(P.A - P.A.T).max() <= 1e-12 * abs(P.A).max()
    tot = np.zeros(P.n)
    for s, d in zip(P.subs, P.pou):
        np.add.at(tot, s, d)
    assert np.allclose(tot, 1.0, atol=1e-12)
    O = oracle.Oracle(P, nthreads=2)
    x, it, conv, hist = O.pcg(P.b, reduction=1e-8, maxit=100)
    assert conv and it < 60 and np.linalg.norm(P.b - P.A @ x) <= 1e-7 * np.linalg.norm(P.b)


@pytest.mark.parametrize("kind", ["standard", "distance", "trivial"])
def test_partition_of_unity_types(kind):
    """The three PartitionOfUnityType values (pou.hh:24-32,79-141): each sums to one over the sharing subdomains; standard and
    distance vanish on the subdomain boundary, trivial is the owner indicator; restricted Schwarz with any of them converges."""
    from ddm_b200.setup import problem
    from oracle import oracle

    P = problem.unitcube_p1(8, 4, overlap=2, nev=4, workers=1, pou_kind=kind)
    tot = np.zeros(P.n)
    for s, w in zip(P.subs, P.pou):
        np.add.at(tot, s, w)
    assert np.max(np.abs(tot - 1.0)) <= 1e-12
    for r, (s, w) in enumerate(zip(P.subs, P.pou)):
        assert np.all(w >= 0.0) and np.all(w <= 1.0 + 1e-15)
        if kind == "trivial":
            assert np.array_equal(w, (P.owner[s] == r).astype(np.float64))
        else:
            assert (w == 0.0).any() and w[: int(P.n_novlp[r])].max() > 0.0  # zero on the subdomain boundary, not inside
        if kind == "standard":  # 1 / (number of subdomains holding the DOF as a non-boundary DOF)
            inv = 1.0 / w[w > 0]
            assert np.allclose(inv, np.round(inv), atol=1e-12)
    x, it, conv, _ = oracle.Oracle(P, nthreads=2).gmres(P.b, reduction=1e-8, restart=30, maxit=200, restricted=True)
    assert conv and np.linalg.norm(P.b - P.A @ x) <= 1e-6 * np.linalg.norm(P.b)
