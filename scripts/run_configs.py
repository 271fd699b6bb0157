"""Runs BASELINE.json's configurations C1, C3, C4 (C2 and the C5 family are bench.py workloads) on one GPU next to the CPU
oracle and prints one JSON line per configuration: iterations (GPU / oracle), time to the requested reduction on the GPU
(device-resident loop, CUDA events, best of 3 after a warm-up solve), the oracle's wall time on all host threads, and the
preconditioner-apply rate.  Results go into BASELINE.md section 4.

  python scripts/run_configs.py [c1 c3 c3r c4 c4b ...]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))
sys.path.insert(0, ROOT)
import numpy as np

from ddm_b200 import api
from ddm_b200.setup import problem
from oracle import oracle

CONFIGS = {
    # name: (builder, solver, kwargs)
    "c1": (lambda: problem.unitsquare_p1(85, 16, overlap=2, nev=16), "cg", dict(typ="standard", mode="additive", reduction=1e-8, maxit=100)),
    "c3": (lambda: problem.elasticity_box(8, 16, overlap=1, nev=12), "gmres", dict(typ="restricted", mode="multiplicative", reduction=1e-6, maxit=1000, restart=100)),
    "c3cg": (lambda: problem.elasticity_box(8, 16, overlap=1, nev=12), "cg", dict(typ="standard", mode="additive", reduction=1e-8, maxit=500)),
    "c3r": (lambda: problem.elasticity_box(16, 32, overlap=1, nev=12), "cg", dict(typ="standard", mode="additive", reduction=1e-8, maxit=500)),
    "c4": (lambda: problem.convdiff_dg(512, 16), "gmres", dict(typ="restricted", mode="multiplicative", reduction=1e-8, maxit=300, restart=50)),
    "c4b": (lambda: problem.convdiff_dg(512, 64), "gmres", dict(typ="restricted", mode="multiplicative", reduction=1e-8, maxit=300, restart=50)),
}


def run(name):
    build, solver, kw = CONFIGS[name]
    t0 = time.perf_counter()
    P = build()
    t_setup = time.perf_counter() - t0
    nthreads = oracle.lib().oracle_max_threads()
    O = oracle.Oracle(P, nthreads=nthreads)
    restricted = kw["typ"] == "restricted"
    t0 = time.perf_counter()
    if solver == "cg":
        xo, ito, convo, _ = O.pcg(P.b, reduction=kw["reduction"], maxit=kw["maxit"], restricted=restricted, mode=kw["mode"])
    else:
        xo, ito, convo, _ = O.gmres(P.b, reduction=kw["reduction"], restart=kw["restart"], maxit=kw["maxit"], restricted=restricted, mode=kw["mode"])
    t_cpu = time.perf_counter() - t0
    ctx = api.Context.from_products(P)
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
    C = api.CombinedPreconditioner(kw["mode"])
    C.set_op(op)
    C.add(api.SchwarzPreconditioner(ctx, kw["typ"]))
    C.add(api.GalerkinPreconditioner(ctx))
    S = (api.CGSolver(op, sp, C, reduction=kw["reduction"], maxit=kw["maxit"]) if solver == "cg"
         else api.RestartedGMResSolver(op, sp, C, reduction=kw["reduction"], restart=kw["restart"], maxit=kw["maxit"]))
    best, res = None, None
    for rep in range(4):
        x, b = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
        res = S.solve_device(x, b)
        if rep:
            best = res.elapsed if best is None else min(best, res.elapsed)
    xg = x.download()
    ctx.timers(True)
    d, y = ctx.vector(P.b), ctx.vector()
    for _ in range(10):
        C.apply(y, d)
    ms, calls = ctx.timer("CombinedPreconditioner/apply")
    ctx.timers(False)
    bytes_apply = sum(ctx.stat(k) for k in ("trisolve_fwd", "trisolve_bwd", "restrict", "prolong", "coarse")) + (ctx.stat("spmv") if kw["mode"] == "multiplicative" else 0.0)
    line = dict(config=name, problem=P.meta.get("problem"), dofs=int(P.n), subdomains=int(P.nsub), coarse_dim=int(P.A0.shape[0]), solver=solver, **{k: v for k, v in kw.items()},
                iterations_gpu=int(res.iterations), iterations_oracle=int(ito), converged=bool(res.converged and convo),
                gpu_seconds=best, cpu_oracle_seconds=t_cpu, cpu_threads=nthreads,
                rel_diff_solution=float(np.linalg.norm(xg - xo) / np.linalg.norm(xo)),
                true_residual=float(np.linalg.norm(P.b - P.A @ xg) / np.linalg.norm(P.b)),
                apply_ms=ms / max(calls, 1), apply_GB=bytes_apply / 1e9, apply_GBps=bytes_apply / (ms / max(calls, 1) * 1e-3) / 1e9,
                factor_nnz=ctx.stat("factor_nnz"), host_setup_seconds=t_setup)
    print(json.dumps(line), flush=True)
    ctx.close()


if __name__ == "__main__":
    for nm in (sys.argv[1:] or list(CONFIGS)):
        run(nm)
