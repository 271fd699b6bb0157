"""Multi-GPU parity: run under torchrun (one rank per GPU); every rank uploads its shard, the assembled results are
compared with the single-process CPU oracle on rank 0.  Exit code 0 = all checks passed.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/multi_gpu_check.py
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

from ddm_b200 import api, lib
from ddm_b200.setup import problem


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    # torch.distributed only moves host data here (unique id, assembled results): gloo, as in bench.py.  The data path's NCCL
    # communicator lives inside libddm_b200.so; a second (torch) NCCL communicator in the same process is avoided.
    dist.init_process_group("gloo")
    ncell, nsub = int(os.environ.get("NCELL", 16)), int(os.environ.get("NSUB", 4 * world))
    P = problem.unitcube_p1(ncell, nsub, overlap=2, nev=6, workers=1)
    buf = (C.c_uint8 * 128)()
    if rank == 0:
        api.check(lib.load().ddm_comm_unique_id(buf))
    t = torch.tensor(list(buf), dtype=torch.uint8)
    dist.broadcast(t, 0)
    ctx = api.Context.from_products(P, device=local, rank=rank, world=world, unique_id=t.tolist())
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
    rng = np.random.default_rng(11)
    d = rng.standard_normal(P.n)

    def gather(vec_own):
        out = torch.zeros(P.n, dtype=torch.float64)
        out[torch.from_numpy(ctx.shard.owned)] = torch.from_numpy(vec_own)
        dist.all_reduce(out)
        return out.numpy()

    results = {}
    x, y = ctx.vector(ctx.owned(d)), ctx.vector()
    op.apply(x, y)
    results["spmv"] = gather(y.download())
    results["dot"] = sp.dot(x, x)
    for mode in ("additive", "multiplicative"):
        for typ in ("standard", "restricted"):
            Cp = api.CombinedPreconditioner(mode)
            Cp.set_op(op)
            Cp.add(api.SchwarzPreconditioner(ctx, typ))
            Cp.add(api.GalerkinPreconditioner(ctx))
            Cp.apply(y, x)
            results[f"{mode}/{typ}"] = gather(y.download())
    Cp = api.CombinedPreconditioner("additive")
    Cp.add(api.SchwarzPreconditioner(ctx, "standard"))
    Cp.add(api.GalerkinPreconditioner(ctx))
    cg = api.CGSolver(op, sp, Cp, reduction=1e-8, maxit=300)
    xv, bv = ctx.vector(), ctx.vector(ctx.owned(P.b))
    r = cg.solve_device(xv, bv)
    results["pcg_x"] = gather(xv.download())
    ok = True
    if rank == 0:
        from oracle import oracle

        O = oracle.Oracle(P, nthreads=4)
        rel = lambda a, b: np.linalg.norm(a - b) / np.linalg.norm(b)
        tol = max(1e-12, 8 * np.finfo(float).eps * max(np.linalg.cond(P.A0), max(np.linalg.cond(P.A[s][:, s].toarray()) for s in P.subs)))
        checks = [("spmv", rel(results["spmv"], O.spmv(d)), 1e-14), ("dot", abs(results["dot"] - d @ d) / (d @ d), 1e-13)]
        for mode in ("additive", "multiplicative"):
            for typ in ("standard", "restricted"):
                checks.append((f"{mode}/{typ}", rel(results[f"{mode}/{typ}"], O.twolevel_apply(d, restricted=typ == "restricted", mode=mode)), tol))
        xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=300)
        checks.append(("pcg iterations", abs(r.iterations - ito), 1))
        checks.append(("pcg solution", rel(results["pcg_x"], xo), 1e-8))
        for name, err, t_ in checks:
            good = err <= t_
            ok &= bool(good)
            print(f"[{world} GPUs] {name}: {err:.3e} (tol {t_:.1e}) {'ok' if good else 'FAIL'}", flush=True)
    flag = torch.tensor([1 if ok else 0])
    dist.broadcast(flag, 0)
    ctx.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
