"""Config C4 with 64 subdomains (Q1-DG convection-diffusion 512^2, LU subdomain factors, restricted Schwarz + multiplicative
coarse level, GMRES(50)): why two runs that both stop at a reduction of 1e-8 need not agree to 1e-8 in the solution.
GMRES stops on the PRECONDITIONED residual (solvers.hh:948,1002); the solution error left at that point is
cond(M^-1 A) times larger.  The script measures, for the CPU oracle and (with --gpu) for the device path, the distance between
the iterate at 1e-8 and the iterate at 1e-12, and the distance between the two paths at equal reduction.

  python scripts/c4_solution_check.py [--gpu] [--ncell 512] [--nsub 64]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))
sys.path.insert(0, ROOT)
import numpy as np

from ddm_b200.setup import problem
from oracle import oracle

ap = argparse.ArgumentParser()
ap.add_argument("--gpu", action="store_true")
ap.add_argument("--ncell", type=int, default=512)
ap.add_argument("--nsub", type=int, default=64)
a = ap.parse_args()
P = problem.convdiff_dg(a.ncell, a.nsub)
O = oracle.Oracle(P, nthreads=oracle.lib().oracle_max_threads())
rel = lambda u, v: float(np.linalg.norm(u - v) / np.linalg.norm(v))
kw = dict(restart=50, maxit=400, restricted=True, mode="multiplicative")
out = dict(config=f"C4 {a.ncell}^2 cells, {a.nsub} subdomains", dofs=int(P.n))
xo = {}
for red in (1e-8, 1e-12):
    x, it, conv, _ = O.gmres(P.b, reduction=red, **kw)
    xo[red] = x
    out[f"oracle_iterations_{red:g}"] = int(it)
    out[f"oracle_true_residual_{red:g}"] = rel(P.A @ x, P.b)
out["oracle_1e-8_vs_oracle_1e-12"] = rel(xo[1e-8], xo[1e-12])
if a.gpu:
    from ddm_b200 import api

    ctx = api.Context.from_products(P)
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
    C = api.CombinedPreconditioner("multiplicative")
    C.set_op(op)
    C.add(api.SchwarzPreconditioner(ctx, "restricted"))
    C.add(api.GalerkinPreconditioner(ctx))
    xg = {}
    for red in (1e-8, 1e-12):
        S = api.RestartedGMResSolver(op, sp, C, reduction=red, restart=50, maxit=400)
        x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
        res = S.solve_device(x, b)
        xg[red] = x.download()
        out[f"gpu_iterations_{red:g}"] = int(res.iterations)
    out["gpu_1e-8_vs_gpu_1e-12"] = rel(xg[1e-8], xg[1e-12])
    out["gpu_vs_oracle_at_1e-8"] = rel(xg[1e-8], xo[1e-8])
    out["gpu_vs_oracle_at_1e-12"] = rel(xg[1e-12], xo[1e-12])
    ctx.close()
print(json.dumps(out), flush=True)
