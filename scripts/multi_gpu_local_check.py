"""Multi-GPU parity of the rank-local path (bench.py's default): under torchrun every rank builds only its own region of the
Q1 brick problem (setup/local.py), factors on its device, joins the Galerkin product; rank 0 also builds the whole problem in
one process and compares the assembled results with the CPU oracle.  Exit code 0 = all checks passed.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/multi_gpu_local_check.py
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
from ddm_b200.setup import local, problem


def main():
    rank, world, dev = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo")  # host plumbing only; the data path's NCCL communicator lives inside libddm_b200.so
    g = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[world]
    ncell, p = int(os.environ.get("NCELL", 12)), int(os.environ.get("P", 2))
    shape, pg = tuple(ncell * k for k in g), tuple(p * k for k in g)
    ok = True
    for coarse in ("pou", "geneo"):
        L = local.poisson_q1_cube_local(shape, pg, g, rank, overlap=2, coarse=coarse, nev=6, workers=2, numeric="device", eigensolver="gpu", eig_device=dev)
        ks = [None] * world
        dist.all_gather_object(ks, [int(z.shape[1]) for z in L.Z])
        L.k_all = np.concatenate([np.asarray(k, dtype=np.int64) for k in ks])
        buf = (C.c_uint8 * 128)()
        if rank == 0:
            api.check(lib.load().ddm_comm_unique_id(buf))
        t = torch.tensor(list(buf), dtype=torch.uint8)
        dist.broadcast(t, 0)
        ctx = api.Context.from_local(L, device=dev, unique_id=t.tolist() if world > 1 else None, keep_A0=True)
        op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
        n = L.n_global
        d = np.random.default_rng(11).standard_normal(n)
        owned = L.shard.owned

        def gather(vec_own):
            out = torch.zeros(n, dtype=torch.float64)
            out[torch.from_numpy(owned)] = torch.from_numpy(vec_own)
            dist.all_reduce(out)
            return out.numpy()

        res = {}
        x, y = ctx.vector(d[owned]), ctx.vector()
        op.apply(x, y)
        res["spmv"] = gather(y.download())
        res["dot"] = sp.dot(x, x)
        Cp = api.CombinedPreconditioner("additive")
        Cp.set_op(op)
        Cp.add(api.SchwarzPreconditioner(ctx, "standard"))
        Cp.add(api.GalerkinPreconditioner(ctx))
        Cp.apply(y, x)
        res["apply"] = gather(y.download())
        cg = api.CGSolver(op, sp, Cp, reduction=1e-8, maxit=300)
        xv, bv = ctx.vector(), ctx.vector(L.b_own)
        r = cg.solve_device(xv, bv)
        res["pcg_x"] = gather(xv.download())
        bg = gather(L.b_own)
        if rank == 0:
            from oracle import oracle

            P = problem.poisson_q1_cube(shape if world > 1 else shape[0], pg, group=g, overlap=2, coarse=coarse, nev=6, workers=2)
            O = oracle.Oracle(P, nthreads=4)
            rel = lambda a, b: np.linalg.norm(a - b) / np.linalg.norm(b)
            checks = [("rhs", rel(bg, P.b), 1e-13), ("spmv", rel(res["spmv"], O.spmv(d)), 1e-13), ("dot", abs(res["dot"] - d @ d) / (d @ d), 1e-13)]
            if coarse == "pou":  # same basis on both sides: the application itself is comparable
                tol = max(1e-12, 8 * np.finfo(float).eps * np.linalg.cond(P.A0))
                checks.append(("two-level apply", rel(res["apply"], O.twolevel_apply(d)), tol))
                checks.append(("Galerkin matrix", np.abs(ctx.A0_device - P.A0).max() / np.abs(P.A0).max(), 1e-12))
            xo, ito, convo, _ = O.pcg(P.b, reduction=1e-8, maxit=300)
            checks.append(("pcg iterations", abs(r.iterations - ito), 1))
            checks.append(("pcg solution", rel(res["pcg_x"], xo), 1e-8 if coarse == "pou" else 5e-8))
            checks.append(("true residual", np.linalg.norm(P.b - P.A @ res["pcg_x"]) / np.linalg.norm(P.b), 1.05e-8))
            for name, err, t_ in checks:
                good = err <= t_
                ok &= bool(good)
                print(f"[{world} GPUs, local setup, {coarse}] {name}: {err:.3e} (tol {t_:.1e}) {'ok' if good else 'FAIL'}", flush=True)
        del x, y, xv, bv
        ctx.close()
        dist.barrier()
    flag = torch.tensor([1 if ok else 0])
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
