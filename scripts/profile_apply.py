"""Minimal driver for ncu: loads a .ddmpack (made by `bench.py --setup-only --pack ...`), uploads it and runs a
few two-level applies and one PCG solve.  No torch, no subprocess, no fork: nothing for the profiler to trip over."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))
import numpy as np

from ddm_b200 import api
from ddm_b200.setup import pack

P = pack.load_products(sys.argv[1])
napply = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ctx = api.Context.from_products(P)
op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
C = api.CombinedPreconditioner("additive")
C.add(api.SchwarzPreconditioner(ctx, "standard"))
C.add(api.GalerkinPreconditioner(ctx))
d, x = ctx.vector(P.b), ctx.vector()
ctx.sync()
import ctypes

ctypes.CDLL("libcuda.so.1").cuProfilerStart()  # ncu --profile-from-start off: skip the upload's conversion kernels
ctx.timers(True)
for _ in range(napply):
    C.apply(x, d)
    op.apply(d, x)
ctx.sync()
for ev in ("Schwarz/local solve fwd", "Schwarz/local solve bwd", "CombinedPreconditioner/apply"):
    ms, n = ctx.timer(ev)
    print(ev, ms / max(n, 1), "ms")
ctx.timers(False)
if len(sys.argv) > 3:
    cg = api.CGSolver(op, sp, C, reduction=1e-8, maxit=500)
    xv, bv = ctx.vector(), ctx.vector(P.b)
    r = cg.solve_device(xv, bv)
    print("pcg", r.iterations, r.converged, r.elapsed)
print("launches", ctx.launches())
