#!/usr/bin/env python
"""Benchmark of the solve phase: two-level additive Schwarz PCG to 1e-8 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--ncell C] [--nsub S] [--coarse geneo|pou]

A step = one complete solve phase of the workload: PCG (Dune::CGSolver recurrences) preconditioned with the two-level
additive Schwarz method, from x0 = 0 to a defect reduction of 1e-8.  The workload is BASELINE.json configs[1]
("3D Poisson, heterogeneous coefficient, 64 subdomains, 2-level GenEO PCG") on a Kuhn-triangulated unit cube
(the reference's data/unitcube.msh does not exist on the GPU box); with N GPUs the mesh of the same cube is refined to
64 subdomains and ncell^3 cells per GPU (weak scaling).  Setup (mesh, assembly, factorisation, GenEO) runs on the host and is not timed, as in the reference's
own "Total/Linear solve" event (examples/poisson.cc:171-174).
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "schwarz_pcg_solved_dofs_per_second_to_1e-8"
UNIT = "MDOF/s"
REDUCTION = 1e-8
MAXIT = 500


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ncell", type=int, default=int(os.environ.get("DDM_BENCH_NCELL", 112)), help="cells per unit length")
    ap.add_argument("--nsub", type=int, default=64, help="subdomains per GPU")
    ap.add_argument("--workload", default="p1cube", choices=["p1cube", "q1cube"],
                    help="p1cube: BASELINE.json configs[1] (default, the metric's configuration); q1cube: configs[4]'s family (structured Q1, "
                         "27-point, channels coefficient, brick subdomains; --ncell cells per axis and GPU, nsub must be a cube)")
    ap.add_argument("--coarse", default="geneo", choices=["geneo", "pou"])
    ap.add_argument("--nev", type=int, default=16)
    ap.add_argument("--eigensolver", default=os.environ.get("DDM_BENCH_EIGENSOLVER", "host"), choices=["host", "gpu"],
                    help="GenEO eigenproblems (setup, untimed): per-subdomain ARPACK on the host, or the batched device Lanczos")
    ap.add_argument("--overlap", type=int, default=2)
    ap.add_argument("--eig-threshold", type=float, default=-0.5,
                    help="GenEO: keep the eigenvectors below this eigenvalue instead of a fixed count (eigensolver_params.hh:30; <= 0: off, the default)")
    ap.add_argument("--nev-max", type=int, default=None, help="GenEO with --eig-threshold: upper limit of the doubling of nev (default 2 nev)")
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pack", default=None, help=".ddmpack cache of the setup products: loaded if present, else built and saved")
    ap.add_argument("--strong", action="store_true",
                    help="strong scaling: the N = 1 problem (ncell^3 cells, nsub subdomains in total) split over the GPUs instead of a problem that grows with N")
    ap.add_argument("--setup-only", action="store_true", help="build (and save) the setup products, then exit")
    return ap.parse_args()


def gpu_grid(n):
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}.get(n) or (n, 1, 1)


def job_shape(a, world):
    """(cells per axis, subdomains in total): weak scaling refines the unit cube with N and hosts nsub subdomains per GPU,
    strong scaling keeps the N = 1 problem and deals its nsub subdomains out."""
    if a.strong:
        if a.nsub % world:
            raise SystemExit(f"--strong: {a.nsub} subdomains do not split evenly over {world} GPUs")
        return (a.ncell,) * 3, a.nsub
    return tuple(a.ncell * k for k in gpu_grid(world)), a.nsub * world


def brick_grids(a, world):
    """q1cube: (brick grid of the whole job, GPU grid that groups the bricks).  nsub = p^3 bricks per GPU (weak) or in total (strong)."""
    p = round(a.nsub ** (1 / 3))
    g = gpu_grid(world)
    if p ** 3 != a.nsub:
        raise SystemExit("--workload q1cube needs a cubic --nsub")
    if a.strong:
        if any(p % k for k in g):
            raise SystemExit(f"--strong: a {p}x{p}x{p} brick grid does not split over the {g} GPU grid")
        return (p, p, p), g
    return tuple(p * k for k in g), g


def workload_name(a, world):
    if a.workload == "q1cube":
        shape, _ = job_shape(a, world)
        pg, _ = brick_grids(a, world)
        return (f"3D Q1 Poisson (27-point) on the unit cube, channels coefficient 1e6, structured {shape[0]}x{shape[1]}x{shape[2]} cells, "
                f"{pg[0]}x{pg[1]}x{pg[2]} brick subdomains, overlap {a.overlap}, {a.coarse} coarse space nev {a.nev}, additive, CG 1e-8")
    shape, nsub = job_shape(a, world)
    thr = f" threshold {a.eig_threshold:g} (nev <= {a.nev_max or 2 * a.nev})" if a.eig_threshold > 0 and a.coarse == "geneo" else ""
    return (f"3D P1 Poisson on the unit cube, islands coefficient (poisson_coefficient.lua), Kuhn mesh {shape[0]}x{shape[1]}x{shape[2]} cells, "
            f"{nsub} subdomains (RCB), overlap {a.overlap}, {a.coarse} coarse space nev {a.nev}{thr}, additive, CG 1e-8")


def build_problem(a, world, rank=None, gather=None):
    """rank/gather given: this process builds only its own subdomains' factors and bases (one process per GPU)."""
    from ddm_b200.setup import fem, mesh, pack, problem

    if a.pack and os.path.exists(a.pack) and gather is None:
        return pack.load_products(a.pack)
    kw = {}
    if a.eig_threshold > 0 and a.coarse == "geneo":
        kw.update(eig_threshold=a.eig_threshold, nev_max=a.nev_max)
    if a.eigensolver == "gpu" and a.coarse == "geneo":
        kw.update(eigensolver="gpu", eig_device=int(os.environ.get("LOCAL_RANK", 0)))
    if gather is not None:
        from ddm_b200.partition import sub_ranges

        s0, s1 = sub_ranges(job_shape(a, world)[1], world)[rank]
        kw.update(only=range(s0, s1), gather=gather, workers=max(1, (os.cpu_count() or 1) // world), assemble_part=(rank, world))
    P = _build_problem(a, world, fem, mesh, problem, **kw)
    if a.pack and gather is None:
        pack.save_products(a.pack, P)
    return P


def _build_problem(a, world, fem, mesh, problem, **kw):
    if a.workload == "q1cube":
        # config C5's family: with N GPUs the unit cube is refined by the GPU grid and every GPU hosts a p x p x p block of bricks
        # (8x8x8 subdomains -> 2x2x2 GPU bricks of 4x4x4); CPU oracle: 28 -> 29 iterations from 1 to 8 GPUs at 32 cells per GPU axis
        shape, _ = job_shape(a, world)
        pg, g = brick_grids(a, world)
        return problem.poisson_q1_cube(shape if world > 1 else a.ncell, pg, group=g, overlap=a.overlap, coarse=a.coarse, nev=a.nev, **kw)
    # Weak scaling = the same problem (unit cube, same coefficient) on a finer mesh with more subdomains, 64 per GPU, as the
    # reference scales (grid.refine + one rank per subdomain).  A first version tiled the coefficient pattern over a growing
    # box instead: with 8x8x8 subdomains the fixed nev = 16 GenEO space no longer captured all high-contrast inclusions and
    # PCG (GPU and CPU oracle alike) stagnated near 1e-6.
    shape, nsub = job_shape(a, world)
    coords, cells = mesh.structured_simplices(shape, lengths=(1.0, 1.0, 1.0))
    return problem.poisson_simplex(coords, cells, nsub, kappa_fn=fem.kappa_islands, xmax=1.0, overlap=a.overlap,
                                   coarse=a.coarse, nev=a.nev, meta=dict(mesh=f"kuhn-{shape}"), **kw)


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.rows.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.12)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    return rank, world, local


def cpu_baseline_sample(P, its_full, budget_s, nthreads):
    """The oracle (CPU restatement, all host threads) on a bounded sample: the first m PCG iterations of the same
    solve, extrapolated to the full iteration count.  oracle/ is used here only as the timed baseline."""
    from oracle import oracle

    O = oracle.Oracle(P, nthreads=nthreads)
    t0 = time.perf_counter()
    O.pcg(P.b, reduction=REDUCTION, maxit=1)
    t1 = time.perf_counter() - t0  # 2 applies + 2 SpMV
    m = int(max(1, min(its_full, (budget_s - t1) / max(t1 / 2, 1e-9))))
    t0 = time.perf_counter()
    _, it, conv, _ = O.pcg(P.b, reduction=REDUCTION, maxit=m)
    t = time.perf_counter() - t0
    t_full = t * (its_full + 1) / (it + 1)  # it iterations cost it+1 applies
    return {"value": P.n / t_full / 1e6, "unit": UNIT, "cores": nthreads, "kind": "port",
            "sample": f"first {it} of {its_full} PCG iterations of the same solve ({t:.1f} s measured, extrapolated by apply count)",
            "seconds_full_solve_est": t_full}


def run_reference(a):
    """Reference arm: the reference's CPU algorithm for the path (oracle port; the reference binary itself needs
    MPI + SuiteSparse + Eigen, absent from this image) on all host threads, same config, metric and unit."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from oracle import oracle

    P = build_problem(a, max(world, a.gpus))
    nthreads = oracle.lib().oracle_max_threads()
    O = oracle.Oracle(P, nthreads=nthreads)
    t0 = time.perf_counter()
    _, its_full, conv, hist = O.pcg(P.b, reduction=REDUCTION, maxit=MAXIT)  # full solve once: iteration count + warm-up
    t_full_once = time.perf_counter() - t0
    m = int(max(1, min(its_full, 30.0 / max(t_full_once / (its_full + 1), 1e-9))))
    for _ in range(max(0, a.warmup - 1)):
        O.pcg(P.b, reduction=REDUCTION, maxit=min(m, 2))
    ts = []
    for _ in range(a.steps):
        t0 = time.perf_counter()
        _, it, _, _ = O.pcg(P.b, reduction=REDUCTION, maxit=m)
        ts.append((time.perf_counter() - t0) * (its_full + 1) / (it + 1))
    t = float(np.mean(ts))
    val = P.n / t / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong" if a.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a, max(world, a.gpus)), "dofs": int(P.n), "iterations": int(its_full), "converged": bool(conv)},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": nthreads, "kind": "port",
                             "sample": f"each step = first {m} of {its_full} PCG iterations, extrapolated by apply count; full solve measured once: {t_full_once:.1f} s"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_b200(a):
    rank, world, local = dist_env()
    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the solve phase has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    import torch.distributed as dist

    uid = None
    gather = None
    if world > 1:
        # torch.distributed is host-side plumbing only (setup exchange, barrier, max over ranks): gloo.  The data path's
        # collectives (halo exchange, coarse all-reduce, Krylov scalars) are NCCL calls inside libddm_b200.so.
        dist.init_process_group("gloo")

        def gather(obj):
            out = [None] * world
            dist.all_gather_object(out, obj)
            return out

    from ddm_b200 import api, lib

    t0 = time.perf_counter()
    P = build_problem(a, world, rank, gather)
    if rank == 0:
        print(f"[bench] host setup {time.perf_counter() - t0:.1f} s: n={P.n} nsub={P.nsub} N0={P.A0.shape[0]} "
              f"nnz(L) on rank 0 = {sum(F.nnz_lower() for F in P.factors if F is not None) / 1e6:.1f} M", file=sys.stderr, flush=True)
    t0 = time.perf_counter()
    if world > 1:
        import ctypes as C

        buf = (C.c_uint8 * 128)()
        if rank == 0:
            api.check(lib.load().ddm_comm_unique_id(buf))
        t = torch.tensor(list(buf), dtype=torch.uint8)
        dist.broadcast(t, 0)
        uid = t.tolist()
    ctx = api.Context.from_products(P, device=local, rank=rank, world=world, unique_id=uid)
    if rank == 0:
        print(f"[bench] upload + device schedules {time.perf_counter() - t0:.1f} s", file=sys.stderr, flush=True)
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
    Cp = api.CombinedPreconditioner("additive")
    Cp.set_op(op)
    Cp.add(api.SchwarzPreconditioner(ctx, "standard"))
    Cp.add(api.GalerkinPreconditioner(ctx))
    cg = api.CGSolver(op, sp, Cp, reduction=REDUCTION, maxit=MAXIT)
    b_own = ctx.owned(P.b)
    x, b = ctx.vector(), ctx.vector()

    def barrier():
        if world > 1:
            dist.barrier()
        ctx.sync()
        torch.cuda.synchronize()

    def step():
        x.fill(0.0)
        b.upload(b_own)  # b is overwritten by the defect (solvers.hh:284): restore it, device resident from here on
        return cg.solve_device(x, b)

    for _ in range(a.warmup):
        res = step()
    # ---- timed region: K solves, inputs resident in HBM, CUDA events on the library's stream
    ctx.timers(True)
    l0 = ctx.launches()
    barrier()
    with ClockSampler(local) as clk:
        t_wall = time.perf_counter()
        dev_ms = 0.0
        for _ in range(a.steps):
            x.fill(0.0)
            b.upload(b_own)
            ctx.stopwatch_start()
            res = cg.solve_device(x, b)
            dev_ms += ctx.stopwatch_stop()
        barrier()
        t_wall = time.perf_counter() - t_wall
    launches = ctx.launches() - l0
    timers = {ev: ctx.timer(ev) for ev in ("Schwarz/local solve fwd", "Schwarz/local solve bwd", "Schwarz/local solve", "Schwarz/get defect",
                                          "Schwarz/add solution", "GalerkinPrec/coarse solve", "CombinedPreconditioner/apply", "Operator/apply")}
    ctx.timers(False)
    # ---- e2e: the same solve through the host-buffer entry (H2D of b and x0, D2H of x and the defect, every step)
    bw, xw = np.empty_like(b_own), np.empty_like(b_own)  # the caller's vectors, reused across solves as a driver would

    def solve_e2e():
        np.copyto(bw, b_own)  # b is overwritten by the defect: restore the right-hand side, start from x0 = 0
        xw.fill(0.0)
        return cg.solve_host(bw, xw, overwrite=True)

    for _ in range(min(2, a.warmup)):
        solve_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        xh, rh = solve_e2e()
    barrier()
    t_e2e = time.perf_counter() - t0
    # ---- reduce over ranks
    vals = torch.tensor([dev_ms, t_wall, t_e2e], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
    dev_ms, t_wall, t_e2e = vals.tolist()
    stats = {k: ctx.stat(k) for k in ("trisolve_fwd", "trisolve_bwd", "trisolve_stored", "spmv", "restrict", "prolong", "coarse", "factor_nnz")}
    per_rank = [None] * world  # load balance of the sharded path: factor entries and sweep times per rank
    fwd_ms, fwd_calls = timers["Schwarz/local solve fwd"]
    bwd_ms, bwd_calls = timers["Schwarz/local solve bwd"]
    mine = {"factor_nnz": stats["factor_nnz"], "fwd_ms": fwd_ms / max(fwd_calls, 1), "bwd_ms": bwd_ms / max(bwd_calls, 1), "dofs_owned": int(ctx.n_own)}
    if world > 1:
        dist.all_gather_object(per_rank, mine)
    else:
        per_rank = [mine]
    if rank != 0:
        if world > 1:
            dist.barrier()
        return
    n = P.n
    ms_step = dev_ms / a.steps
    value = n / (ms_step * 1e-3) / 1e6
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
    # dominant kernel: the forward triangular sweep (the backward sweep is reported beside it)
    kern = {}
    for name, ev in (("trisolve_fwd", "Schwarz/local solve fwd"), ("trisolve_bwd", "Schwarz/local solve bwd")):
        ms, calls = timers[ev]
        kern[name] = {"ms": ms / max(calls, 1), "GBps": stats[name] / (ms / max(calls, 1) * 1e-3) / 1e9 if ms > 0 else None, "launches": calls}
    dom = max(kern, key=lambda k: kern[k]["ms"])
    apply_ms, apply_calls = timers["CombinedPreconditioner/apply"]
    apply_bytes = stats["trisolve_fwd"] + stats["trisolve_bwd"] + stats["restrict"] + stats["prolong"] + stats["coarse"]
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        traffic = tr.get(f"{'' if a.workload == 'p1cube' else a.workload + '_'}ncell{a.ncell}_nsub{a.nsub}_{a.coarse}_nev{a.nev}_overlap{a.overlap}", {}).get(f"{dom}_kernel")
    except OSError:
        pass
    if a.strong and world > 1:
        traffic = None  # the capture is of one GPU holding all nsub subdomains
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong" if a.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a, world), "dofs": int(n), "subdomains": int(P.nsub), "coarse_dim": int(P.A0.shape[0]),
                   "factor_nnz": stats["factor_nnz"], "iterations": int(res.iterations), "converged": bool(res.converged),
                   "l2_policy": "inputs larger than L2 (factors %.1f GB per sweep vs 126 MB L2)" % (stats["trisolve_stored"] / 1e9)},
        "time_to_1e-8_s": ms_step * 1e-3, "wall_ms_per_step": t_wall / a.steps * 1e3,
        "precond_apply": {"ms": apply_ms / max(apply_calls, 1), "algorithmic_GB": apply_bytes / 1e9,
                          "GBps": apply_bytes / (apply_ms / max(apply_calls, 1) * 1e-3) / 1e9 if apply_ms > 0 else None,
                          "frac_of_peak": apply_bytes / (apply_ms / max(apply_calls, 1) * 1e-3) / 1e9 / peak if apply_ms > 0 else None},
        "e2e": {"value": n / (t_e2e / a.steps) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(2 * 8 * ctx.n_own), "d2h_bytes_per_step": int(2 * 8 * ctx.n_own)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": f"{dom}_kernel", "achieved": kern[dom]["GBps"], "peak": peak, "unit": "GB/s",
                     "frac": kern[dom]["GBps"] / peak if kern[dom]["GBps"] else None, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": stats[dom], "ms_per_launch": kern[dom]["ms"], "kernels": kern},
        "timers_ms_per_call": {k: (v[0] / max(v[1], 1)) for k, v in timers.items()},
        "clocks": clk.summary(),
        "per_rank": per_rank,
    }
    if not a.no_cpu_baseline:
        from oracle import oracle

        line["cpu_baseline"] = cpu_baseline_sample(P, int(res.iterations), a.cpu_seconds, oracle.lib().oracle_max_threads()) if world == 1 else None
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()


def main():
    a = parse()
    if a.setup_only:
        build_problem(a, max(a.gpus, dist_env()[1]))
        return
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
