"""Parity as numbers, at configuration size, per subdomain ("do this" 5).

For C2 (3-D P1 Poisson, 64 subdomains, 0.46 M DOFs), C3 as shipped (elasticity, 28 431 DOFs), C4 (Q1-DG, 512^2 cells, 16
subdomains, LU) and one brick case of the C5 family with 45^3-node subdomains, each through the C ABI:
  * every subdomain solve x_i = A_i^{-1} d_i (all at once on a block-diagonal context, ddm_blockdiag_solve) against the
    oracle's solve with the same factor (schwarz.hh:130-134 -> cholmod.hh:263-306), and its backward error
    ||A_i x_i - d_i|| / (||A_i|| ||x_i|| + ||d_i||) (infinity norms);
  * one Schwarz, one coarse and one two-level application against the oracle, measured on every subdomain's index set
    (schwarz.hh:115-149, galerkin_preconditioner.hh:151-194, combined_preconditioner.hh:127-163).
The numbers go to parity_report.json (gpurun_out/ when it exists) and are printed.  Bounds: north_star's 1e-12 where it
holds; where it does not, ten times the measured value with the term that causes it named in BOUNDS."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# (sub-solve vs oracle, backward error, schwarz, coarse, two-level); None = not applicable
BOUNDS = {
    # Subdomain solves, their backward errors and the Schwarz level hold north_star's 1e-12 everywhere (measured <= 3e-15).
    # Coarse and two-level applications are limited by the Galerkin solve: two FP64 solves of A0 y = d0 by different
    # factorisations (cuSOLVER potrf + explicit L^{-1} on the device, LAPACK potrf + substitution in the oracle) differ by up
    # to cond(A0) eps in the forward error whatever their quality -- the report's coarse_gpu_vs_refined and
    # coarse_oracle_vs_refined (both against a solve refined in extended precision) sit at the same distance.  Bounds below
    # are ten times the values measured on a B200 (profiles/r02_parity_report.json), cond(A0) beside them:
    "C2": dict(solve=1e-12, berr=1e-14, schwarz=1e-12, coarse=3e-8, twolevel=2e-8),  # cond(A0) 4.4e8: measured 2.7e-9 / 1.1e-9
    # C3 and C4 combine multiplicatively: the coarse level is fed d - A x1, whose last bits change from run to run with the order of
    # the sweeps' atomics, and cond(A0) amplifies that -- two-level values of 7.6e-11 and 9.0e-10 were both measured for C4
    "C3": dict(solve=1e-12, berr=1e-14, schwarz=1e-12, coarse=7e-11, twolevel=1e-9),  # cond(A0) 9.2e6: measured 6.6e-12 / 6.4e-12
    "C4": dict(solve=1e-12, berr=1e-14, schwarz=1e-12, coarse=1.5e-10, twolevel=1e-8),  # cond(A0) 7.7e8: measured 1.5e-11 / 7.6e-11 ... 9.0e-10
    "C5brick": dict(solve=1e-12, berr=1e-14, schwarz=1e-12, coarse=1e-12, twolevel=1e-12),  # cond(A0) 6.9e6: measured 2.2e-14 / 3.5e-15
}


def build(name):
    from ddm_b200.setup import problem

    if name == "C2":
        return problem.unitcube_p1(76, 64, overlap=2, nev=16), dict(restricted=False, mode="additive")
    if name == "C3":
        return problem.elasticity_box(8, 16, overlap=1, nev=12), dict(restricted=True, mode="multiplicative")
    if name == "C4":
        return problem.convdiff_dg(512, 16), dict(restricted=True, mode="multiplicative")
    if name == "C5brick":
        return problem.poisson_q1_cube(80, (2, 2, 2), overlap=2, coarse="geneo", nev=16, numeric="device", eigensolver="gpu"), dict(restricted=False, mode="additive")
    raise KeyError(name)


def blockdiag(P):
    """Block-diagonal products: operator blockdiag(A_i), subdomain i = its own contiguous range (as setup/geneo_gpu.py builds them)."""
    import scipy.sparse as sp

    from ddm_b200.setup.problem import SetupProducts

    sizes = np.array([s.size for s in P.subs], dtype=np.int64)
    off = np.concatenate(([0], np.cumsum(sizes)))
    Ais = [P.A[s][:, s].tocsr() for s in P.subs]
    big = sp.block_diag(Ais, format="csr")
    big.sort_indices()
    Q = SetupProducts(A=big, b=np.zeros(int(off[-1])), owner=np.repeat(np.arange(P.nsub, dtype=np.int32), sizes),
                      subs=[np.arange(off[i], off[i + 1], dtype=np.int64) for i in range(P.nsub)], pou=[np.ones(int(k)) for k in sizes],
                      factors=P.factors, Z=[np.zeros((int(k), 0), order="F") for k in sizes], A0=np.zeros((0, 0)), n_novlp=sizes.copy(), meta=dict(kind="blockdiag"))
    return Q, Ais, off


def coarse_reference(P, d):
    """x = sum_i R_i^T Z_i y_i with A0 y = (Z_i^T R_i d)_i solved to working accuracy: LU solve plus iterative refinement with
    the residual in long double (galerkin_preconditioner.hh:151-194 in exact arithmetic, for all practical purposes)."""
    import scipy.linalg as sla

    off = np.concatenate(([0], np.cumsum([z.shape[1] for z in P.Z])))
    d0 = np.concatenate([z.T @ d[s] for z, s in zip(P.Z, P.subs)])
    lu = sla.lu_factor(P.A0)
    y = sla.lu_solve(lu, d0)
    A0l, d0l = P.A0.astype(np.longdouble), d0.astype(np.longdouble)
    for _ in range(4):
        r = (d0l - A0l @ y.astype(np.longdouble)).astype(np.float64)
        y = y + sla.lu_solve(lu, r)
    x = np.zeros(P.n)
    for i, (z, s) in enumerate(zip(P.Z, P.subs)):
        np.add.at(x, s, z @ y[off[i] : off[i + 1]])
    return x


def inf(v):
    return float(np.abs(v).max()) if v.size else 0.0


@pytest.mark.parametrize("name", ["C2", "C3", "C4", "C5brick"])
def test_parity_numbers_at_config_size(name):
    import ctypes as C

    from ddm_b200 import api, lib
    from ddm_b200.lib import check
    from ddm_b200.setup import factor
    from oracle import oracle

    assert b"dryrun" not in lib.load().ddm_version()
    P, kw = build(name)
    rng = np.random.default_rng(11)
    rep = dict(config=name, dofs=int(P.n), subdomains=int(P.nsub), coarse_dim=int(P.A0.shape[0]), n_i_max=int(max(s.size for s in P.subs)))
    # ---- subdomain solves on a block-diagonal context
    Q, Ais, off = blockdiag(P)
    ctxb = api.Context.from_products(Q, with_coarse=False)
    host_factors = []
    for i, F in enumerate(P.factors):
        if hasattr(F, "an"):  # device factorisation: the oracle gets the panels the device produced
            x = ctxb.factorize(i, F, want_panels=True)
            an = F.an
            F = factor.SupernodalFactor(n=an.n, perm=an.perm.astype(np.int32), super_=an.first.astype(np.int32), pi=an.pi, px=an.px, s=an.row_lists(), x=x)
        host_factors.append(F)
    P.factors = host_factors
    O = oracle.Oracle(P, nthreads=8)
    d = rng.standard_normal(int(off[-1]))
    dv, xv = ctxb.vector(d), ctxb.vector()
    check(ctxb.lib.ddm_blockdiag_solve(ctxb.h, dv.h, xv.h))
    xg = xv.download()
    e_solve, e_berr = [], []
    for i, Ai in enumerate(Ais):
        di, xi = d[off[i] : off[i + 1]], xg[off[i] : off[i + 1]]
        xo = O.sub_solve(i, di)
        e_solve.append(np.linalg.norm(xi - xo) / np.linalg.norm(xo))
        e_berr.append(inf(Ai @ xi - di) / (abs(Ai).sum(axis=1).max() * inf(xi) + inf(di)))
    del dv, xv
    ctxb.close()
    rep.update(solve_vs_oracle_max=float(max(e_solve)), solve_vs_oracle_median=float(np.median(e_solve)),
               backward_error_max=float(max(e_berr)), worst_subdomain=int(np.argmax(e_solve)))
    # ---- applications on the real context, measured per subdomain index set
    ctx = api.Context.from_products(P)
    dg = rng.standard_normal(P.n)
    S = api.SchwarzPreconditioner(ctx, "restricted" if kw["restricted"] else "standard")
    G = api.GalerkinPreconditioner(ctx)
    dvec, xvec = ctx.vector(dg), ctx.vector()

    def per_sub(got, want):
        return float(max(np.linalg.norm(got[s] - want[s]) / max(np.linalg.norm(want[s]), 1e-300) for s in P.subs))

    S.apply(xvec, dvec)
    rep["schwarz_max"] = per_sub(xvec.download(), O.schwarz_apply(dg, restricted=kw["restricted"]))
    dvec.upload(dg)
    G.apply(xvec, dvec)
    xc_gpu, xc_orc = xvec.download(), O.coarse_apply(dg)
    rep["coarse_max"] = per_sub(xc_gpu, xc_orc)
    # Two FP64 solves of A0 y = d0 by different factorisations differ by up to cond(A0) eps in the forward error, whatever
    # their quality: measure both against the coarse correction computed with a refined solve (residuals in extended precision)
    xc_ref = coarse_reference(P, dg)
    rep["coarse_gpu_vs_refined"] = per_sub(xc_gpu, xc_ref)
    rep["coarse_oracle_vs_refined"] = per_sub(xc_orc, xc_ref)
    op = api.NonOverlappingOperator(ctx)
    Cp = api.CombinedPreconditioner(kw["mode"])
    Cp.set_op(op)
    Cp.add(S)
    Cp.add(G)
    rep["twolevel_max"] = per_sub(Cp.apply_host(dg), O.twolevel_apply(dg, restricted=kw["restricted"], mode=kw["mode"]))
    rep["cond_A0"] = float(np.linalg.cond(P.A0))
    del dvec, xvec
    ctx.close()
    out = os.path.join(ROOT, "gpurun_out") if os.path.isdir(os.path.join(ROOT, "gpurun_out")) else ROOT
    path = os.path.join(out, "parity_report.json")
    allrep = json.load(open(path)) if os.path.exists(path) else {}
    allrep[name] = rep
    json.dump(allrep, open(path, "w"), indent=1)
    print("PARITY", json.dumps(rep))
    B = BOUNDS[name]
    assert rep["backward_error_max"] <= B["berr"], rep
    assert rep["solve_vs_oracle_max"] <= B["solve"], rep
    assert rep["schwarz_max"] <= B["schwarz"], rep
    assert rep["coarse_max"] <= B["coarse"], rep
    # the device's coarse solve is as accurate as the oracle's (distance to the refined solve within a factor of ten)
    assert rep["coarse_gpu_vs_refined"] <= 10 * max(rep["coarse_oracle_vs_refined"], 1e-13), rep
    assert rep["twolevel_max"] <= B["twolevel"], rep
