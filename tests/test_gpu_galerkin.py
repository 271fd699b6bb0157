"""Galerkin product on the device (ddm_galerkin_assemble): A0 = Z^T A Z built from the solve-phase kernels must
equal the host product of the setup restatement, for symmetric (P1 Poisson, elasticity) and non-symmetric (DG) operators."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["small3d", "cd_dg", "elast"])
def test_device_galerkin_matrix(request, name):
    from ddm_b200 import api
    from oracle import oracle

    P = request.getfixturevalue(name)
    ctx = api.Context.from_products(P, device_galerkin=True)
    try:
        A0 = ctx.A0_device
        assert np.linalg.norm(A0 - P.A0) <= 1e-12 * np.linalg.norm(P.A0)
        # and the coarse level built from it behaves like the uploaded one
        d = np.random.default_rng(0).standard_normal(P.n)
        x = ctx.vector()
        api.GalerkinPreconditioner(ctx).apply(x, ctx.vector(d))
        want = oracle.Oracle(P, nthreads=2).coarse_apply(d)
        tol = max(1e-12, 8 * np.finfo(float).eps * np.linalg.cond(P.A0))
        assert np.linalg.norm(x.download() - want) <= tol * np.linalg.norm(want)
    finally:
        ctx.close()


def test_coarse_apply_needs_the_matrix(small3d):
    import ctypes as C

    from ddm_b200 import api
    from ddm_b200.lib import DDMError

    P = small3d
    ctx = api.Context.from_products(P, device_galerkin=True)
    try:
        # re-upload the basis without a matrix: applies must refuse until ddm_galerkin_assemble has run
        k = np.array([z.shape[1] for z in P.Z], dtype=np.int32)
        Z = np.concatenate([z.ravel(order="F") for z in P.Z])
        api.check(ctx.lib.ddm_upload_coarse(ctx.h, k.ctypes.data_as(C.POINTER(C.c_int32)), Z.ctypes.data_as(C.POINTER(C.c_double)),
                                            C.c_int32(int(P.A0.shape[0])), C.c_int32(0), None))
        api.check(ctx.lib.ddm_finalize(ctx.h))
        with pytest.raises(DDMError):
            api.GalerkinPreconditioner(ctx).apply(ctx.vector(), ctx.vector(P.b))
        api.check(ctx.lib.ddm_galerkin_assemble(ctx.h, None))
        api.GalerkinPreconditioner(ctx).apply(ctx.vector(), ctx.vector(P.b))
    finally:
        ctx.close()


def test_batched_gpu_geneo_matches_host_eigensolver():
    """the lock-step shift-invert Lanczos on the GPU spans the same GenEO coarse space as the per-subdomain ARPACK
    runs of the host restatement -- up to the choice inside an eigenvalue cluster cut by nev (cubic subdomains have degenerate
    pairs: one principal angle per subdomain may differ) -- and PCG converges alike."""
    from ddm_b200 import api
    from ddm_b200.setup import problem

    kw = dict(overlap=2, nev=6, workers=1)
    Ph = problem.unitcube_p1(12, 8, **kw)
    Pg = problem.unitcube_p1(12, 8, eigensolver="gpu", **kw)
    shapes_ok = [z.shape for z in Pg.Z] == [z.shape for z in Ph.Z]
    matched = []
    for Zg, Zh in zip(Pg.Z, Ph.Z):
        # columns are D v normalised: compare the spanned spaces through the principal angles
        Qg, Qh = np.linalg.qr(Zg)[0], np.linalg.qr(Zh)[0]
        sv = np.linalg.svd(Qg.T @ Qh, compute_uv=False)
        matched.append(int((sv > 1.0 - 1e-4).sum()))
    its = []
    for P in (Ph, Pg):
        ctx = api.Context.from_products(P)
        try:
            op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
            C = api.CombinedPreconditioner("additive")
            C.add(api.SchwarzPreconditioner(ctx, "standard"))
            C.add(api.GalerkinPreconditioner(ctx))
            x, r = api.CGSolver(op, sp, C, reduction=1e-8, maxit=200).solve_host(P.b)
            its.append((r.iterations, r.converged))
        finally:
            ctx.close()
    print("geneo gpu vs host: shapes", [z.shape for z in Pg.Z], [z.shape for z in Ph.Z], "matched directions", matched, "pcg", its)
    assert shapes_ok
    assert all(m >= Zh.shape[1] - 1 for m, Zh in zip(matched, Ph.Z))
    assert its[0][1] and its[1][1] and abs(its[0][0] - its[1][0]) <= 2
