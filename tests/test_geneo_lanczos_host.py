"""The batched shift-invert Lanczos of setup/geneo_gpu.py with the device calls replaced by host stand-ins
(tests/lanczos_standin.py): its eigenvalues must be those of the per-subdomain ARPACK runs of the host restatement, its vectors
span the same spaces up to clusters cut by nev, and the recurrence must stop by its own convergence test."""
import numpy as np
import pytest


def _eigenproblems(P, shift=1e-3):
    """Per subdomain: C_i = D B_neu D, the Cholesky factor of A_neu + shift C_i, and ARPACK's answer."""
    from ddm_b200.setup import coarse, factor, fem, geneo_gpu, mesh

    coords, cells = mesh.structured_simplices((P.meta_ncell,) * 3)
    cent = coords[cells].mean(axis=1)
    Ke = fem.p1_element_matrices(coords, cells, fem.kappa_islands(cent))
    shared = np.zeros(P.n, dtype=np.int32)
    for s in P.subs:
        shared[s] += 1
    Cs, Ks, ref = [], [], []
    for i, s in enumerate(P.subs):
        An, Bn = coarse.neumann_matrices(P.n, cells, Ke, s, shared, P.dirichlet)
        Cm = geneo_gpu.weighted_overlap_matrix(Bn, P.pou[i])
        Cs.append(Cm)
        Ks.append(factor.cholesky_supernodal((An + shift * Cm).tocsr(), coords[s], 32))
        ref.append(coarse.geneo_basis(An, Bn, P.pou[i], P.dirichlet[s], nev=P.meta["nev"], shift=shift, tol=1e-5))
    return Cs, Ks, ref


@pytest.mark.parametrize("ncell,nsub,nev", [(10, 6, 5), (12, 8, 6)])
def test_host_run_of_the_batched_lanczos(monkeypatch, ncell, nsub, nev):
    from ddm_b200 import api
    from ddm_b200.setup import geneo_gpu, problem

    import lanczos_standin

    P = problem.unitcube_p1(ncell, nsub, overlap=2, nev=nev, workers=1)
    P.meta_ncell = ncell
    Cs, Ks, ref = _eigenproblems(P)
    monkeypatch.setattr(api.Context, "from_products", lanczos_standin.HostContext.from_products)
    Z, lams = geneo_gpu.batched_geneo(Cs, Ks, P.pou, [P.dirichlet[s] for s in P.subs], nev=nev, shift=1e-3, tol=1e-5)
    ctx = lanczos_standin.HostContext.last
    assert ctx.lib.solves <= 160 + 2  # stopped by its convergence test, not by the step limit
    for i in range(nsub):
        Zr, lr = ref[i]
        k = min(lams[i].size, lr.size)
        assert k >= nev - 1
        # eigenvalues to the eigensolver tolerance (the zero eigenvalue of the floating subdomains absolutely)
        assert np.allclose(lams[i][:k], lr[:k], rtol=1e-4, atol=1e-8)
        Qa, Qb = np.linalg.qr(Z[i][:, :k])[0], np.linalg.qr(Zr[:, :k])[0]
        sv = np.linalg.svd(Qa.T @ Qb, compute_uv=False)
        assert (sv > 1.0 - 1e-4).sum() >= k - 1  # one direction may differ where nev cuts a cluster


def test_step_limit_is_reported(monkeypatch):
    from ddm_b200 import api
    from ddm_b200.setup import geneo_gpu, problem

    import lanczos_standin

    P = problem.unitcube_p1(10, 6, overlap=2, nev=5, workers=1)
    P.meta_ncell = 10
    Cs, Ks, _ = _eigenproblems(P)
    monkeypatch.setattr(api.Context, "from_products", lanczos_standin.HostContext.from_products)
    with pytest.warns(RuntimeWarning, match="did not reach tolerance"):
        Z, lams = geneo_gpu.batched_geneo(Cs, Ks, P.pou, [P.dirichlet[s] for s in P.subs], nev=5, shift=1e-3, tol=1e-12, m0=8, mmax=8)
    assert all(z.shape[1] <= 5 for z in Z)


def test_threshold_count_follows_the_reference_retry_loop():
    """spectra.hh:118-163 on plain numbers: nev doubles while the largest computed eigenvalue is below the threshold and
    nev < nev_max; the kept pairs are those below the threshold among the first nconv - 1, at least one."""
    from ddm_b200.setup.coarse import threshold_count

    ev = np.array([0.0, 0.1, 0.2, 0.3, 0.6, 0.7, 0.8, 0.9])
    assert threshold_count(ev, 4, 8, 0.25) == (4, 3)  # ev[3] >= thr: stop at nev, keep 0.0 0.1 0.2
    assert threshold_count(ev, 4, 8, 0.65) == (8, 5)  # ev[3] < thr: double, then five below
    assert threshold_count(ev, 4, 8, 5.0) == (8, 7)  # everything below: nev_max reached, the loop stops one short of nconv
    assert threshold_count(ev + 1.0, 4, 8, 0.5) == (4, 1)  # nothing below: one vector is kept anyway
    assert threshold_count(ev, 2, 8, 0.65) == (8, 5)  # 2 -> 4 -> 8


def test_eigenvalue_threshold_gives_ragged_coarse_spaces(monkeypatch):
    """threshold > 0 (eigensolver_params.hh:30): every subdomain keeps the eigenvectors below it -- different counts per
    subdomain --, host ARPACK path and batched path (host stand-in) select the same numbers, the method converges."""
    from ddm_b200 import api
    from ddm_b200.setup import problem
    from oracle import oracle

    import lanczos_standin

    kw = dict(overlap=2, nev=4, workers=1, eig_threshold=0.95, nev_max=8)
    Ph = problem.unitcube_p1(10, 6, **kw)
    ks = [z.shape[1] for z in Ph.Z]
    assert min(ks) >= 1 and max(ks) <= 7 and len(set(ks)) > 1, ks
    assert Ph.A0.shape[0] == sum(ks)
    monkeypatch.setattr(api.Context, "from_products", lanczos_standin.HostContext.from_products)
    Pg = problem.unitcube_p1(10, 6, eigensolver="gpu", **kw)
    assert [z.shape[1] for z in Pg.Z] == ks
    x, it, conv, _ = oracle.Oracle(Ph, nthreads=2).pcg(Ph.b, reduction=1e-8, maxit=200)
    y, it2, conv2, _ = oracle.Oracle(Pg, nthreads=2).pcg(Pg.b, reduction=1e-8, maxit=200)
    assert conv and conv2 and abs(it - it2) <= 1 and np.linalg.norm(x - y) <= 1e-6 * np.linalg.norm(x)
