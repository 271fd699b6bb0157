"""First-light check on a GPU box: every ABI entry against the oracle on a small problem."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200")); sys.path.insert(0, ROOT)
import numpy as np
from ddm_b200.setup import problem
from ddm_b200 import api
from oracle import oracle

def rel(a, b): return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)

N = int(os.environ.get("N", 16)); NS = int(os.environ.get("NS", 8))
t = time.time(); P = problem.unitcube_p1(N, NS, overlap=2, nev=8); print("setup %.1fs" % (time.time() - t), P.meta, flush=True)
O = oracle.Oracle(P, nthreads=8)
t = time.time(); ctx = api.Context.from_products(P); print("upload %.1fs" % (time.time() - t), flush=True)
rng = np.random.default_rng(0)
d = rng.standard_normal(P.n)
op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
x, y = ctx.vector(d), ctx.vector()
op.apply(x, y); print("spmv", rel(y.download(), O.spmv(d)), flush=True)
print("dot", abs(sp.dot(x, x) - O.dot(d, d)) / O.dot(d, d), flush=True)
for typ in ("standard", "restricted"):
    S = api.SchwarzPreconditioner(ctx, typ)
    S.apply(y, x); ctx.sync()
    print("schwarz", typ, rel(y.download(), O.schwarz_apply(d, restricted=typ == "restricted")), flush=True)
G = api.GalerkinPreconditioner(ctx); G.apply(y, x); print("coarse", rel(y.download(), O.coarse_apply(d)), flush=True)
for mode in ("additive", "multiplicative"):
    Cp = api.CombinedPreconditioner(mode); Cp.set_op(op); Cp.add(api.SchwarzPreconditioner(ctx, "standard")); Cp.add(G)
    Cp.apply(y, x); print("twolevel", mode, rel(y.download(), O.twolevel_apply(d, mode=mode)), flush=True)
    print("twolevel host", mode, rel(Cp.apply_host(d), O.twolevel_apply(d, mode=mode)), flush=True)
Cp = api.CombinedPreconditioner("additive"); Cp.add(api.SchwarzPreconditioner(ctx, "standard")); Cp.add(G)
cg = api.CGSolver(op, sp, Cp, reduction=1e-8, maxit=200)
xo, ito, convo, histo = O.pcg(P.b, reduction=1e-8, maxit=200)
xv, bv = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
r1 = cg.apply(xv, bv); print("dropin CG its", r1.iterations, "oracle", ito, "sol diff", rel(xv.download(), xo), flush=True)
xv, bv = ctx.vector(np.zeros(P.n)), ctx.vector(P.b)
r2 = cg.solve_device(xv, bv); print("device CG its", r2.iterations, "elapsed", r2.elapsed, "sol diff", rel(xv.download(), xo), flush=True)
xh, r3 = cg.solve_host(P.b); print("host CG its", r3.iterations, rel(xh, xo), flush=True)
print("launches", ctx.launches())
ctx.timers(True)
for _ in range(5): Cp.apply(y, x)
for ev in ("Schwarz/get defect", "Schwarz/local solve", "Schwarz/add solution", "GalerkinPrec/coarse solve", "CombinedPreconditioner/apply"):
    print(ev, ctx.timer(ev))
for k in ("trisolve_fwd", "trisolve_stored", "spmv", "coarse", "n_tasks_fwd", "n_chunks"): print(k, ctx.stat(k))
