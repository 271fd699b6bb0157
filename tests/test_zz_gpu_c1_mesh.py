"""Config C1 as BASELINE.json states it -- examples/poisson on the reference's data/unitsquare.msh (7 441 nodes), 16 subdomains,
overlap 2, GenEO nev 16, CG to 1e-8 -- not a structured stand-in.  The mesh-dependent products were produced on the CPU
container from the reference's file (scripts/make_c1_fixture.py) and travel in tests/golden/c1_unitsquare.npz; factors and the
Galerkin matrix are rebuilt from them.  CPU half: the fixture is the mesh's problem and the oracle solves it; GPU half: parity."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
MESH = "/root/reference/data/unitsquare.msh"


def load_c1():
    from ddm_b200.setup import coarse, factor
    from ddm_b200.setup.problem import SetupProducts

    d = np.load(os.path.join(HERE, "golden", "c1_unitsquare.npz"))
    n = d["b"].size
    A = sp.csr_matrix((d["A_data"], d["A_indices"], d["A_indptr"]), shape=(n, n))
    sp_, k = d["sub_ptr"], d["k"]
    subs = [d["sub_dofs"][sp_[i] : sp_[i + 1]].astype(np.int64) for i in range(k.size)]
    pou = [d["pou"][sp_[i] : sp_[i + 1]] for i in range(k.size)]
    Z, o = [], 0
    for i, s in enumerate(subs):
        Z.append(np.asfortranarray(d["Z"][o : o + s.size * k[i]].reshape((s.size, k[i]), order="F")))
        o += s.size * k[i]
    factors = []
    for s in subs:
        Ai = A[s][:, s].tocsr()
        Ai.sort_indices()
        factors.append(factor.cholesky_supernodal(Ai, None, 32))
    A0, _, _ = coarse.galerkin_matrix(A, subs, Z)
    return SetupProducts(A=A, b=d["b"], owner=d["owner"], subs=subs, pou=pou, factors=factors, Z=Z, A0=A0, n_novlp=d["n_novlp"],
                         meta=dict(nsub=k.size, n=n, N0=int(A0.shape[0]), symmetric=True, mesh="unitsquare.msh"))


@pytest.fixture(scope="module")
def c1():
    return load_c1()


def test_fixture_is_config_c1(c1):
    from oracle import oracle

    P = c1
    assert P.n == 7441 and P.nsub == 16 and P.A0.shape[0] == 16 * 16  # 7 441 DOFs, 16 subdomains, nev 16
    x, it, conv, _ = oracle.Oracle(P, nthreads=2).pcg(P.b, reduction=1e-8, maxit=100)  # pdelab_example.ini: CG, 1e-8, maxit 100
    assert conv and it <= 12 and np.linalg.norm(P.b - P.A @ x) <= 1e-7 * np.linalg.norm(P.b)


@pytest.mark.skipif(not os.path.exists(MESH), reason="needs the reference's data/unitsquare.msh")
def test_fixture_comes_from_the_reference_mesh(c1):
    from ddm_b200.setup import problem

    Q = problem.poisson_gmsh(MESH, 2, 16, refine=0, overlap=2, nev=16, workers=1)
    assert np.array_equal(Q.A.indptr, c1.A.indptr) and np.array_equal(Q.A.indices, c1.A.indices) and np.array_equal(Q.A.data, c1.A.data)
    assert np.array_equal(Q.b, c1.b) and all(np.array_equal(a, b) for a, b in zip(Q.subs, c1.subs))
    assert all(np.array_equal(a, b) for a, b in zip(Q.pou, c1.pou))
    for a, b in zip(Q.Z, c1.Z):  # ARPACK is deterministic for a fixed start vector; allow for a rebuilt BLAS all the same
        assert a.shape == b.shape and np.linalg.norm(a - b) <= 1e-8 * np.linalg.norm(b)


@pytest.mark.gpu
def test_gpu_parity_on_config_c1(c1):
    from ddm_b200 import api
    from oracle import oracle

    P = c1
    O = oracle.Oracle(P, nthreads=2)
    ctx = api.Context.from_products(P)
    try:
        op, sp_ = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
        Cp = api.CombinedPreconditioner("additive")
        Cp.set_op(op)
        Cp.add(api.SchwarzPreconditioner(ctx, "standard"))
        Cp.add(api.GalerkinPreconditioner(ctx))
        d = np.random.default_rng(4).standard_normal(P.n)
        x = ctx.vector()
        Cp.apply(x, ctx.vector(d))
        want = O.twolevel_apply(d)
        tol = max(1e-12, 8 * np.finfo(float).eps * max([np.linalg.cond(P.A0)] + [np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs]))
        assert np.linalg.norm(x.download() - want) <= tol * np.linalg.norm(want)
        xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=100)
        xv, bv = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
        res = api.CGSolver(op, sp_, Cp, reduction=1e-8, maxit=100).solve_device(xv, bv)
        assert res.converged and convo and abs(res.iterations - ito) <= 1
        assert np.linalg.norm(xv.download() - xo) <= 1e-8 * np.linalg.norm(xo)
    finally:
        ctx.close()


@pytest.mark.skipif(not os.path.exists("/root/reference/data/unitcube.msh"), reason="needs the reference's data/unitcube.msh")
def test_config_c2_mesh_refines_and_solves():
    """Config C2's input: data/unitcube.msh (1 142 nodes, 4 602 tetrahedra) uniformly refined.  Red refinement keeps the volume and
    stays conforming (every interior face shared by exactly two cells); the two-level method converges on the refined mesh."""
    from ddm_b200.setup import mesh, problem
    from oracle import oracle

    coords, cells = mesh.read_gmsh22("/root/reference/data/unitcube.msh", 3)
    assert coords.shape == (1142, 3) and cells.shape == (4602, 4)
    c1_, t1 = mesh.refine_simplices(coords, cells)
    assert t1.shape[0] == 8 * cells.shape[0]

    def volume(c, t):
        e = c[t[:, 1:]] - c[t[:, :1]]
        return np.abs(np.linalg.det(e)).sum() / 6.0

    assert abs(volume(coords, cells) - 1.0) <= 1e-12 and abs(volume(c1_, t1) - 1.0) <= 1e-12
    faces = np.sort(np.concatenate([t1[:, [0, 1, 2]], t1[:, [0, 1, 3]], t1[:, [0, 2, 3]], t1[:, [1, 2, 3]]]), axis=1)
    _, counts = np.unique(faces, axis=0, return_counts=True)
    assert set(np.unique(counts)) == {1, 2}
    P = problem.poisson_gmsh("/root/reference/data/unitcube.msh", 3, 8, refine=1, overlap=2, nev=8, workers=1)
    x, it, conv, _ = oracle.Oracle(P, nthreads=2).pcg(P.b, reduction=1e-8, maxit=100)
    assert P.n == c1_.shape[0] and conv and np.linalg.norm(P.b - P.A @ x) <= 1e-7 * np.linalg.norm(P.b)
