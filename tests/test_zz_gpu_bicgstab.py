"""Dune::BiCGSTABSolver, unchanged, on the GPU path through the adaptors (oracle/_ref/dune_krylov arms gpu / gpu-host): the third
solver of the reference's factory (solverfactory.hh:148-167) needs only the operator, scalar product and preconditioner
interfaces plus the vector operations DeviceVector already offers.  Written after the last GPU call of round 1."""
import numpy as np
import pytest

from test_ref_krylov import needs_bin, run

pytestmark = pytest.mark.gpu


@needs_bin
@pytest.mark.parametrize("arm", ["gpu", "gpu-host"])
def test_dune_bicgstab_drives_the_gpu_path(small3d, tmp_path, arm):
    from ddm_b200.setup import pack

    P = small3d
    path = str(tmp_path / "p.ddmpack")
    pack.save_products(path, P)
    rc, xc = run("cpu", path, "bicgstab", str(tmp_path / "xc.bin"), typ="restricted", mode="multiplicative")
    rg, xg = run(arm, path, "bicgstab", str(tmp_path / "xg.bin"), typ="restricted", mode="multiplicative")
    # BiCGSTAB's recurrences amplify rounding differences more than CG's: a couple of (half) steps either way
    assert rc["converged"] and rg["converged"] and abs(rg["iterations"] - rc["iterations"]) <= 2
    assert np.linalg.norm(P.b - P.A @ xg) <= 1e-7 * np.linalg.norm(P.b)
    assert np.linalg.norm(xg - xc) <= 1e-6 * np.linalg.norm(xc)


def test_python_bicgstab_on_the_device_vectors(small3d):
    """api.BiCGSTABSolver (pinned against Dune's loop on the CPU, tests/test_ref_krylov.py) on the real device classes."""
    from ddm_b200 import api

    P = small3d
    ctx = api.Context.from_products(P)
    try:
        op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
        Cp = api.CombinedPreconditioner("multiplicative")
        Cp.set_op(op)
        Cp.add(api.SchwarzPreconditioner(ctx, "restricted"))
        Cp.add(api.GalerkinPreconditioner(ctx))
        x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
        res = api.BiCGSTABSolver(op, sp, Cp, reduction=1e-8, maxit=100).apply(x, b)
        assert res.converged and res.iterations < 40
        assert np.linalg.norm(P.b - P.A @ x.download()) <= 1e-7 * np.linalg.norm(P.b)
        assert np.linalg.norm(b.download() - (P.b - P.A @ x.download())) <= 1e-6 * np.linalg.norm(P.b)  # b holds the defect
    finally:
        ctx.close()
