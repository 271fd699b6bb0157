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
@pytest.mark.parametrize("arm", ["gpu", "gpu-host"])
def test_dune_gmres_drives_the_gpu_path(packed, tmp_path, arm):
    P, path = packed
    rc, xc = run("cpu", path, "gmres", str(tmp_path / "xc.bin"), typ="restricted", mode="multiplicative")
    rg, xg = run(arm, path, "gmres", str(tmp_path / "xg.bin"), typ="restricted", mode="multiplicative")
    assert rg["converged"] and abs(rg["iterations"] - rc["iterations"]) <= 1
    assert np.linalg.norm(xg - xc) <= 1e-6 * np.linalg.norm(xc)
