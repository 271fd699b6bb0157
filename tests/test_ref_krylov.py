"""oracle/_ref/dune_krylov = the reference's own Dune::CGSolver / RestartedGMResSolver (compiled from
/root/reference/extern/dune-istl by oracle/Makefile) driving (a) the CPU oracle and (b) the GPU product through the
Dune adaptors (dune-ddm_b200/dune/ddm_b200/adaptors.hh).  (a) pins the oracle's PCG loop against solvers.hh:279-340;
(b) is the drop-in claim itself: unchanged Dune solvers on the C ABI."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "dune_krylov")

needs_bin = pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/dune_krylov not built (needs /root/reference at build time)")


def run(arm, pack_path, solver, out, reduction=1e-8, maxit=300, restart=30, typ="standard", mode="additive", coarse=1, nt=2):
    r = subprocess.run([BIN, arm, pack_path, solver, repr(reduction), str(maxit), str(restart), typ, mode, str(coarse), str(nt), out],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    return json.loads(r.stdout.strip().splitlines()[-1]), np.fromfile(out)


@pytest.fixture(scope="module")
def packed(small3d, tmp_path_factory):
    from ddm_b200.setup import pack

    path = str(tmp_path_factory.mktemp("pack") / "small3d.ddmpack")
    pack.save_products(path, small3d)
    return small3d, path


@needs_bin
def test_pack_roundtrip(packed):
    from ddm_b200.setup import pack

    P, path = packed
    d = pack.read(path)
    assert d["n"][0] == P.n and np.array_equal(d["A_val"], P.A.data) and np.array_equal(d["F0_x"], P.factors[0].x)


@needs_bin
@pytest.mark.parametrize("coarse", [1, 0])
def test_oracle_pcg_is_dune_cgsolver(packed, tmp_path, coarse):
    from oracle import oracle

    P, path = packed
    O = oracle.Oracle(P, nthreads=2)
    x, it, conv, hist = O.pcg(P.b, reduction=1e-8, maxit=300, with_coarse=bool(coarse))
    r, xd = run("cpu", path, "cg", str(tmp_path / "x.bin"), coarse=coarse)
    assert r["converged"] and conv and r["iterations"] == it
    h = np.array(r["norms"])
    # same recurrences; FMA contraction differs between the two compilations, and CG amplifies it over the iterations
    assert h.size == hist.size and np.max(np.abs(h[:5] - hist[:5]) / hist[:5]) <= 1e-12 and np.max(np.abs(h - hist) / hist) <= 1e-6
    assert np.linalg.norm(xd - x) <= 1e-9 * np.linalg.norm(x)


@needs_bin
def test_dune_gmres_on_the_oracle_converges(packed, tmp_path):
    P, path = packed
    r, xd = run("cpu", path, "gmres", str(tmp_path / "x.bin"), typ="restricted", mode="multiplicative")
    assert r["converged"]
    # GMRES stops on the preconditioned residual (solvers.hh:948,1002); the true one follows within the conditioning
    assert np.linalg.norm(P.b - P.A @ xd) <= 1e-5 * np.linalg.norm(P.b)


@needs_bin
@pytest.mark.gpu
@pytest.mark.parametrize("arm", ["gpu", "gpu-host"])
def test_dune_cgsolver_drives_the_gpu_path(packed, tmp_path, arm):
    P, path = packed
    rc, xc = run("cpu", path, "cg", str(tmp_path / "xc.bin"))
    rg, xg = run(arm, path, "cg", str(tmp_path / "xg.bin"))
    assert rg["converged"] and abs(rg["iterations"] - rc["iterations"]) <= 1
    assert np.linalg.norm(xg - xc) <= 1e-8 * np.linalg.norm(xc)
    m = min(len(rg["norms"]), len(rc["norms"])) - 2
    assert np.allclose(rg["norms"][:m], rc["norms"][:m], rtol=1e-6)


@needs_bin
@pytest.mark.gpu
def test_generic_preconditioner_list_goes_level_by_level(packed, tmp_path):
    """CombinedPreconditioner::add takes any list (combined_preconditioner.hh:90-109): with the coarse level added first
    the drop-in class cannot use its fused pass and applies the levels one by one as the reference does (:132-158).  The
    additive sum commutes, so Dune::CGSolver must behave as with the usual order."""
    P, path = packed
    rc, xc = run("cpu", path, "cg", str(tmp_path / "xc.bin"))
    rg, xg = run("gpu-list", path, "cg", str(tmp_path / "xg.bin"))
    assert rg["converged"] and abs(rg["iterations"] - rc["iterations"]) <= 1
    assert np.linalg.norm(xg - xc) <= 1e-8 * np.linalg.norm(xc)
    # multiplicative with the coarse level first is a different (non-symmetric) method: it must run and converge under GMRES
    rm, xm = run("gpu-list", path, "gmres", str(tmp_path / "xm.bin"), typ="restricted", mode="multiplicative")
    assert rm["converged"] and np.linalg.norm(P.b - P.A @ xm) <= 1e-5 * np.linalg.norm(P.b)


@needs_bin
@pytest.mark.gpu
@pytest.mark.parametrize("arm", ["gpu", "gpu-host"])
def test_dune_gmres_drives_the_gpu_path(packed, tmp_path, arm):
    P, path = packed
    rc, xc = run("cpu", path, "gmres", str(tmp_path / "xc.bin"), typ="restricted", mode="multiplicative")
    rg, xg = run(arm, path, "gmres", str(tmp_path / "xg.bin"), typ="restricted", mode="multiplicative")
    assert rg["converged"] and abs(rg["iterations"] - rc["iterations"]) <= 1
    assert np.linalg.norm(xg - xc) <= 1e-6 * np.linalg.norm(xc)


@needs_bin
def test_oracle_gmres_is_dune_gmres_on_the_nonsymmetric_problem(cd_dg, tmp_path):
    """Config C4 (LU subdomain factors, non-symmetric A0): the oracle's GMRES loop against Dune::RestartedGMResSolver
    (solvers.hh:880-1080) on the same CPU preconditioner -- same iteration count, same defect history, same solution."""
    from ddm_b200.setup import pack
    from oracle import oracle

    path = str(tmp_path / "cd.ddmpack")
    pack.save_products(path, cd_dg)
    O = oracle.Oracle(cd_dg, nthreads=2)
    assert O.nonsym
    x, it, conv, hist = O.gmres(cd_dg.b, reduction=1e-8, restart=30, maxit=300, restricted=True)
    r, xd = run("cpu", path, "gmres", str(tmp_path / "x.bin"), typ="restricted")
    assert r["converged"] and conv and r["iterations"] == it
    # the recorded norm() calls of GMRES are the Arnoldi norms h_{j+1,j}, not defects: compare the first one (the initial
    # preconditioned defect) and the final reduction the solver reports from its Givens recurrence
    assert abs(r["norms"][0] - hist[0]) <= 1e-9 * hist[0]  # two dense LUs of A0 (here: LAPACK, there: a plain loop), cond(A0) eps apart
    assert abs(r["reduction"] - hist[it] / hist[0]) <= 1e-6 * r["reduction"]
    assert np.linalg.norm(xd - x) <= 1e-8 * np.linalg.norm(x)
    # stopping is on the preconditioned defect; with the 1e7 coefficient jump the true defect follows three digits behind
    assert np.linalg.norm(cd_dg.b - cd_dg.A @ xd) <= 1e-3 * np.linalg.norm(cd_dg.b)


@needs_bin
@pytest.mark.parametrize("which", ["small3d", "cd_dg"])
def test_dune_bicgstab_on_the_oracle_converges(request, tmp_path, which):
    """The third solver the reference's factory hands out (solverfactory.hh:148-167, examples/*.ini `type = bicgstabsolver`) needs
    nothing beyond the same three interfaces: Dune::BiCGSTABSolver on the CPU preconditioner, symmetric and non-symmetric problem."""
    from ddm_b200.setup import pack

    P = request.getfixturevalue(which)
    path = str(tmp_path / "p.ddmpack")
    pack.save_products(path, P)
    r, xd = run("cpu", path, "bicgstab", str(tmp_path / "x.bin"), typ="restricted", mode="multiplicative")
    assert r["converged"] and r["iterations"] < 60
    assert np.linalg.norm(P.b - P.A @ xd) <= (1e-7 if which == "small3d" else 1e-3) * np.linalg.norm(P.b)


class _HostVec:
    """numpy stand-in for api.DeviceVector (the operations the Krylov loops use), to run api.BiCGSTABSolver without a GPU."""

    def __init__(self, ctx, values=None):
        self.ctx, self.a = ctx, np.zeros(ctx.n) if values is None else np.array(values, dtype=np.float64)

    def fill(self, v):
        self.a[:] = v
        return self

    def assign(self, o):
        self.a[:] = o.a
        return self

    def axpy(self, s, o):
        self.a += s * o.a
        return self

    def __imul__(self, s):
        self.a *= s
        return self

    def __iadd__(self, o):
        self.a += o.a
        return self


class _HostCtx:
    def __init__(self, n):
        self.n = n

    def vector(self, values=None):
        return _HostVec(self, values)


@needs_bin
def test_python_bicgstab_loop_is_dune_bicgstab(packed, tmp_path):
    """api.BiCGSTABSolver restates solvers.hh:419-590; on host stand-ins for operator, scalar product and preconditioner (the
    oracle) it must walk through the same defect norms as Dune::BiCGSTABSolver on the same CPU preconditioner."""
    from ddm_b200 import api
    from oracle import oracle

    P, path = packed
    O = oracle.Oracle(P, nthreads=2)

    class Op:
        def apply(self, x, y):
            y.a[:] = P.A @ x.a

        def applyscaleadd(self, alpha, x, y):
            y.a += alpha * (P.A @ x.a)

    class Sp:
        def dot(self, x, y):
            return float(x.a @ y.a)

        def norm(self, x):
            return float(np.sqrt(x.a @ x.a))

    class Prec:
        def pre(self, x, b):
            pass

        def post(self, x):
            pass

        def apply(self, v, d):
            v.a[:] = O.twolevel_apply(d.a, restricted=True, mode="multiplicative")

    ctx = _HostCtx(P.n)
    x, b = ctx.vector(), ctx.vector(P.b)
    res = api.BiCGSTABSolver(Op(), Sp(), Prec(), reduction=1e-8, maxit=300).apply(x, b)
    r, xd = run("cpu", path, "bicgstab", str(tmp_path / "x.bin"), typ="restricted", mode="multiplicative")
    assert res.converged and r["converged"] and res.iterations == r["iterations"]
    h = np.array(r["norms"])
    assert h.size == res.history.size
    assert np.max(np.abs(h[:4] - res.history[:4]) / h[:4]) <= 1e-10 and np.max(np.abs(h - res.history) / h) <= 1e-4
    assert np.linalg.norm(x.a - xd) <= 1e-8 * np.linalg.norm(xd)
    # iteration cap: the count is the last half step the stopping test saw
    x, b = ctx.vector(), ctx.vector(P.b)
    res = api.BiCGSTABSolver(Op(), Sp(), Prec(), reduction=1e-8, maxit=2).apply(x, b)
    r, _ = run("cpu", path, "bicgstab", str(tmp_path / "x2.bin"), maxit=2, typ="restricted", mode="multiplicative")
    assert not res.converged and not r["converged"] and res.iterations == r["iterations"] and res.history.size == len(r["norms"])


@needs_bin
@pytest.mark.parametrize("solver", ["cg", "gmres"])
def test_python_dropin_loops_are_dune_solvers(packed, tmp_path, solver):
    """api.CGSolver.apply / api.RestartedGMResSolver.apply (the literal drop-in loops of the Python mirror) on host stand-ins
    against Dune's solvers on the same CPU preconditioner: same iteration count, same solution."""
    from ddm_b200 import api
    from oracle import oracle

    P, path = packed
    O = oracle.Oracle(P, nthreads=2)
    typ, mode = ("standard", "additive") if solver == "cg" else ("restricted", "multiplicative")

    class Op:
        def apply(self, x, y):
            y.a[:] = P.A @ x.a

        def applyscaleadd(self, alpha, x, y):
            y.a += alpha * (P.A @ x.a)

    class Sp:
        def dot(self, x, y):
            return float(x.a @ y.a)

        def norm(self, x):
            return float(np.sqrt(x.a @ x.a))

    class Prec:
        def pre(self, x, b):
            pass

        def post(self, x):
            pass

        def apply(self, v, d):
            v.a[:] = O.twolevel_apply(d.a, restricted=typ == "restricted", mode=mode)

    ctx = _HostCtx(P.n)
    x, b = ctx.vector(), ctx.vector(P.b)
    S = api.CGSolver(Op(), Sp(), Prec(), reduction=1e-8, maxit=300) if solver == "cg" else \
        api.RestartedGMResSolver(Op(), Sp(), Prec(), reduction=1e-8, restart=30, maxit=300)
    res = S.apply(x, b)
    r, xd = run("cpu", path, solver, str(tmp_path / "x.bin"), typ=typ, mode=mode)
    assert res.converged and r["converged"] and res.iterations == r["iterations"]
    assert abs(res.reduction - r["reduction"]) <= 1e-4 * r["reduction"]
    assert np.linalg.norm(x.a - xd) <= 1e-8 * np.linalg.norm(xd)
