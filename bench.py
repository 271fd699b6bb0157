#!/usr/bin/env python
"""Benchmark of the solve phase: two-level additive Schwarz PCG to 1e-8 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload q1cube|p1cube] [--ncell C] [--nsub S]

A step = one complete solve phase of the workload: PCG (Dune::CGSolver recurrences) preconditioned with the two-level additive
Schwarz method (GenEO coarse space), from x0 = 0 to a defect reduction of 1e-8.  Default workload = the configuration
BASELINE.json's metric is quoted on, configs[4]: structured 3-D Q1 Poisson (27-point), channels coefficient 1e6, 4 x 4 x 4 brick
subdomains per GPU, overlap 2, GenEO nev 16; with N GPUs the unit cube is refined by the GPU grid (weak scaling: 8 GPUs =
8 x 8 x 8 = 512 subdomains).  Setup is not timed, as in the reference's own "Total/Linear solve" event (examples/poisson.cc:171-174):
every process builds only its own GPU's region of the mesh (setup/local.py), numeric factorisations, GenEO eigensolves and the
Galerkin product run on the device.  `--workload p1cube` = configs[1] (3-D P1, islands coefficient, 64 RCB subdomains).
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
# q1cube: config C5's own brick size -- 58 cells owned per brick and axis, overlapping subdomains of up to 63^3 = 250 k nodes,
# 12.6 M DOFs and 88 GB of factor per GPU, 464^3 cells = 100.5 M DOFs and 512 subdomains on eight GPUs
DEFAULT_NCELL = {"q1cube": 232, "p1cube": 112}
# PCG iteration counts of the q1cube family measured by the GPU arm (profiles/r02c_*, r02d_*): 28 on one GPU and 28 on eight;
# used by the CPU arm only where no CPU solve of the whole job can be run (bounded sample at N > 1)
ASSUMED_ITERATIONS = 28


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("DDM_BENCH_WORKLOAD", "q1cube"), choices=["p1cube", "q1cube"],
                    help="q1cube: BASELINE.json configs[4], the configuration the metric is quoted on (default); p1cube: configs[1]")
    ap.add_argument("--ncell", type=int, default=int(os.environ.get("DDM_BENCH_NCELL", 0)), help="cells per axis and GPU (0: the workload's default)")
    ap.add_argument("--nsub", type=int, default=64, help="subdomains per GPU")
    ap.add_argument("--coarse", default="geneo", choices=["geneo", "pou"])
    ap.add_argument("--nev", type=int, default=16)
    ap.add_argument("--eigensolver", default=os.environ.get("DDM_BENCH_EIGENSOLVER", "gpu"), choices=["host", "gpu"],
                    help="GenEO eigenproblems (setup, untimed): batched device Lanczos (default) or per-subdomain ARPACK on the host")
    ap.add_argument("--numeric", default="device", choices=["host", "device"], help="numeric factorisation of the subdomain matrices (setup, untimed)")
    ap.add_argument("--setup", default="local", choices=["local", "global"],
                    help="q1cube weak scaling: local = every process builds its own region only (default); global = one global mesh per process")
    ap.add_argument("--overlap", type=int, default=2)
    ap.add_argument("--eig-threshold", type=float, default=-0.5)
    ap.add_argument("--nev-max", type=int, default=None)
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--deterministic", action="store_true", help="reproducible sweeps (ddm_set_deterministic): no floating-point atomics")
    ap.add_argument("--pack", default=None, help=".ddmpack cache of the setup products (global setup only)")
    ap.add_argument("--strong", action="store_true", help="strong scaling: the N = 1 problem split over the GPUs")
    ap.add_argument("--setup-only", action="store_true")
    a = ap.parse_args()
    if a.ncell <= 0:
        a.ncell = DEFAULT_NCELL[a.workload]
    return a


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


def config_of(a, world, n, nsub, N0):
    """The same dictionary in both arms (the driver compares them)."""
    return {"workload": workload_name(a, world), "dofs": int(n), "subdomains": int(nsub), "coarse_dim": int(N0),
            "l2_policy": "inputs larger than L2: every sweep streams all subdomain factors (GBs) once, 126 MB L2"}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def use_local_setup(a):
    return a.workload == "q1cube" and not a.strong and a.setup == "local" and not a.pack and a.eig_threshold <= 0


def build_problem(a, world, rank=None, gather=None):
    """Global setup path.  rank/gather given: this process builds only its own subdomains' factors and bases (one process per GPU)."""
    from ddm_b200.setup import fem, mesh, pack, problem

    if a.pack and os.path.exists(a.pack) and gather is None:
        return pack.load_products(a.pack)
    kw = dict(numeric=a.numeric)
    if a.eig_threshold > 0 and a.coarse == "geneo":
        kw.update(eig_threshold=a.eig_threshold, nev_max=a.nev_max)
    if a.eigensolver == "gpu" and a.coarse == "geneo":
        kw.update(eigensolver="gpu", eig_device=int(os.environ.get("LOCAL_RANK", 0)))
    if gather is not None:
        from ddm_b200.partition import sub_ranges

        s0, s1 = sub_ranges(job_shape(a, world)[1], world)[rank]
        kw.update(only=range(s0, s1), gather=gather, workers=max(1, host_cores() // world), assemble_part=(rank, world))
    else:
        kw.update(workers=host_cores())
    P = _build_problem(a, world, fem, mesh, problem, **kw)
    if a.pack and gather is None:
        pack.save_products(a.pack, P)
    return P


def _build_problem(a, world, fem, mesh, problem, **kw):
    if a.workload == "q1cube":
        shape, _ = job_shape(a, world)
        pg, g = brick_grids(a, world)
        return problem.poisson_q1_cube(shape if world > 1 else a.ncell, pg, group=g, overlap=a.overlap, coarse=a.coarse, nev=a.nev, **kw)
    shape, nsub = job_shape(a, world)
    coords, cells = mesh.structured_simplices(shape, lengths=(1.0, 1.0, 1.0))
    return problem.poisson_simplex(coords, cells, nsub, kappa_fn=fem.kappa_islands, xmax=1.0, overlap=a.overlap,
                                   coarse=a.coarse, nev=a.nev, meta=dict(mesh=f"kuhn-{shape}"), **kw)


def build_local(a, world, rank, device, alone=False):
    """alone: this process is the only one working (reference arm: the other ranks have exited), so it takes all host cores."""
    from ddm_b200.setup import local

    shape, _ = job_shape(a, world)
    pg, g = brick_grids(a, world)
    return local.poisson_q1_cube_local(shape, pg, g, rank, overlap=a.overlap, coarse=a.coarse, nev=a.nev,
                                       workers=host_cores() if alone else max(1, host_cores() // world),
                                       numeric=a.numeric, eigensolver=a.eigensolver, eig_device=device)


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


def mem_available_bytes():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable"):
                return int(ln.split()[1]) * 1024
    except OSError:
        pass
    return 16 << 30


# ----------------------------------------------------------------------------- CPU arm (oracle: test infrastructure, timed here only)
def host_products(ctx, L, subset=None):
    """SetupProducts for the CPU oracle from one rank's local products (N = 1: the whole problem).  Subdomain factors are the
    LL^T panels the device factorisation produced (ddm_factorize_subdomain's x_out: what ddm_upload_factor would have been
    given), so both arms work with the same setup products; subset: keep only these subdomains (bounded sample)."""
    import scipy.sparse as sp

    from ddm_b200.setup import coarse as coarse_mod
    from ddm_b200.setup import factor
    from ddm_b200.setup.problem import SetupProducts

    S = L.shard
    nloc = S.n_own + S.n_ghost
    A = sp.csr_matrix((S.val, S.col, S.rowptr), shape=(S.n_own, nloc))
    if S.n_ghost:  # N > 1: the sample lives on this rank's shard; ghost columns are dropped (rows stay the same length: same bytes)
        A = sp.csr_matrix((S.val, np.minimum(S.col, S.n_own - 1), S.rowptr), shape=(S.n_own, S.n_own))
    ids = list(range(L.nsub)) if subset is None else list(subset)
    subs, pou, Z, factors = [], [], [], []
    for j in ids:
        s = S.sub_dofs[S.sub_ptr[j] : S.sub_ptr[j + 1]].astype(np.int64)
        if S.n_ghost:
            s = np.minimum(s, S.n_own - 1)
        subs.append(s)
        pou.append(L.pou[j])
        Z.append(L.Z[j])
        F = L.factors[j]
        if hasattr(F, "an"):
            x = ctx.factorize(j, F, want_panels=True)
            an = F.an
            F = factor.SupernodalFactor(n=an.n, perm=an.perm.astype(np.int32), super_=an.first.astype(np.int32), pi=an.pi, px=an.px, s=an.row_lists(), x=x)
        factors.append(F)
    if subset is None and not S.n_ghost and getattr(ctx, "A0_device", None) is not None:
        A0 = ctx.A0_device
    else:  # sample: Galerkin matrix of the kept subdomains only
        A0 = coarse_mod.galerkin_matrix(A, subs, Z)[0] if not S.n_ghost else np.eye(sum(z.shape[1] for z in Z))
    b = np.zeros(S.n_own)
    b[:] = L.b_own
    return SetupProducts(A=A, b=b, owner=np.zeros(S.n_own, dtype=np.int32), subs=subs, pou=pou, factors=factors, Z=Z, A0=A0, meta=dict(L.meta))


def cpu_arm(a, world, ctx, L=None, P=None, its_hint=None, budget_s=30.0, nthreads=None, steps=1):
    """Times the CPU oracle (all host threads) on the workload.  Whole problem when its factors fit in host memory (N = 1):
    PCG itself, first m iterations, extrapolated by apply count.  Otherwise a bounded sample: the subdomain solves of a subset
    of this rank's subdomains plus full-length vector passes, extrapolated by factor entries and by the number of GPUs' shares.
    steps: how many samples are timed (their mean is returned); the budget is shared between them.
    Returns (seconds per solve, iterations or None, converged or None, description, oracle)."""
    from oracle import oracle

    nthreads = nthreads or host_cores()
    if P is None:
        nnz = np.array([F.nnz_lower() for F in L.factors], dtype=np.float64)
        panel_bytes = 8.0 * sum(float(F.an.px[-1]) if hasattr(F, "an") else float(F.x.size) for F in L.factors)
        full = world == 1 and panel_bytes < 0.6 * mem_available_bytes()
        if full:
            P = host_products(ctx, L)
    else:
        full = True
    if full:
        O = oracle.Oracle(P, nthreads=nthreads)
        t0 = time.perf_counter()
        _, it1, _, _ = O.pcg(P.b, reduction=REDUCTION, maxit=1)
        t1 = time.perf_counter() - t0  # 2 applies + 2 SpMV
        if its_hint is None:  # the reference arm: one full solve for the iteration count
            t0 = time.perf_counter()
            _, its, conv, _ = O.pcg(P.b, reduction=REDUCTION, maxit=MAXIT)
            t_once = time.perf_counter() - t0
        else:
            its, conv, t_once = int(its_hint), True, None
        steps = max(1, int(steps))
        m = int(max(1, min(its, budget_s / steps / max(t1 / 2, 1e-9) - 1)))  # m iterations cost m + 1 applies
        ts, t_sum = [], 0.0
        for _ in range(steps):
            t0 = time.perf_counter()
            _, it, _, _ = O.pcg(P.b, reduction=REDUCTION, maxit=m)
            t = time.perf_counter() - t0
            t_sum += t
            ts.append(t * (its + 1) / (it + 1))  # it iterations cost it + 1 applies
        t_full = float(np.mean(ts))
        each = "" if steps == 1 else f"each of {steps} steps = "
        desc = f"{each}first {it} of {its} PCG iterations of the same solve ({t_sum:.1f} s measured in all, extrapolated by apply count)"
        if t_once is not None:
            desc += f"; full solve measured once: {t_once:.1f} s"
        return t_full, its, bool(conv), desc, O
    # bounded sample on this rank's shard
    budget_bytes = min(0.35 * mem_available_bytes(), 24e9)
    order = np.argsort(-nnz)
    pick, tot = [], 0.0
    step = max(1, len(order) // 8)
    for j in list(order[::step]) + list(order):
        pb = 8.0 * float(L.factors[j].an.px[-1]) if hasattr(L.factors[j], "an") else 8.0 * L.factors[j].x.size
        if j in pick or (pick and tot + pb > budget_bytes):
            continue
        pick.append(int(j))
        tot += pb
        if len(pick) >= max(8, nthreads):
            break
    Ps = host_products(ctx, L, subset=sorted(pick))
    O = oracle.Oracle(Ps, nthreads=nthreads)
    d = np.random.default_rng(0).standard_normal(Ps.n)
    O.twolevel_apply(d)
    reps, t_app = 0, 0.0
    while (t_app < budget_s / 2 and reps < 20) or reps < min(int(steps), 20):  # at least `steps` timed applications (capped at 20)
        t0 = time.perf_counter()
        O.twolevel_apply(d)
        t_app += time.perf_counter() - t0
        reps += 1
    t_app /= reps
    t0 = time.perf_counter()
    for _ in range(3):
        y = O.spmv(d)
        O.dot(d, y)
        O.dot(d, d)
        d2 = d + 0.5 * y
        d2 += 0.25 * d
    t_vec = (time.perf_counter() - t0) / 3
    share = nnz.sum() / nnz[pick].sum()
    t_iter = (t_app * share + t_vec) * world  # the other GPUs' shares are congruent blocks of the same cube
    its = int(its_hint) if its_hint is not None else None
    desc = (f"{reps} timed applications; subdomain solves of {len(pick)} of {L.nsub * world} subdomains ({100 / share / world:.1f} % of the factor entries) + SpMV and vector passes of one GPU's "
            f"share, extrapolated by factor entries and x{world} shares; {t_iter:.2f} s per PCG iteration")
    return t_iter, its, None, desc, O


def run_reference(a):
    """Reference arm: the reference's CPU algorithm for the path (oracle port; the reference binary itself needs MPI +
    SuiteSparse + Eigen, absent from this image) on all host threads, same config, metric and unit.  The setup products are
    made exactly as in the GPU arm (device-accelerated setup, untimed: the host-only setup does not reach this size in the
    driver's window); the timed solve is pure CPU."""
    rank, world, local = dist_env()
    world = max(world, a.gpus) if world == 1 else world
    if rank != 0:
        return
    from ddm_b200 import api

    nthreads = host_cores()
    if use_local_setup(a):
        L = build_local(a, world, 0, local, alone=True)
        L.k_all = np.array([z.shape[1] for z in L.Z] * world, dtype=np.int64)  # the other blocks carry the same counts (nev each)
        n, nsub, N0 = L.n_global, L.nsub_global, int(L.k_all.sum())
        if world == 1:
            ctx = api.Context.from_local(L, device=local, keep_A0=True)
        else:  # sample on rank 0's shard: no collectives, the factorisations run through a one-GPU context of the same shard
            import copy

            L1 = copy.copy(L)
            L1.shard = copy.copy(L.shard)
            L1.shard.world, L1.shard.rank = 1, 0
            ctx = api.Context.from_local(L1, device=local, with_coarse=False)
        t_solve, its, conv, desc, _ = cpu_arm(a, world, ctx, L=L, budget_s=60.0, nthreads=nthreads, steps=a.steps)
        ctx.close()
        if its is None:  # bounded sample: no CPU solve of the whole job exists; iteration count of this family (flat from 1 to 8 GPUs)
            its = int(os.environ.get("DDM_BENCH_ASSUME_ITS", ASSUMED_ITERATIONS))
            desc += (f"; iteration count {its} taken from the GPU arm's measurements of this family (28 at 1 and at 8 GPUs; the CPU solve at N = 1 "
                     "gives the same count): the whole job's factors do not fit one host, so no CPU solve of it was run")
            t_solve = t_solve * (its + 1)
    else:
        P = build_problem(a, world)
        n, nsub, N0 = P.n, P.nsub, P.A0.shape[0]
        if any(hasattr(F, "an") for F in P.factors):
            P.factors = [F.host() if hasattr(F, "an") else F for F in P.factors]
        t_solve, its, conv, desc, _ = cpu_arm(a, world, None, P=P, budget_s=60.0, nthreads=nthreads, steps=a.steps)
    val = n / t_solve / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": t_solve * 1e3, "higher_is_better": True, "scaling": "strong" if a.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(a, world, n, nsub, N0), "iterations": int(its), "converged": conv,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": nthreads, "kind": "port", "sample": desc,
                             "setup": "same products as the GPU arm (device-accelerated setup, untimed); timed solve on the host cores only"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- GPU arm
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
    L = P = None
    if use_local_setup(a):
        L = build_local(a, world, rank, local)
        mine = [int(z.shape[1]) for z in L.Z]
        L.k_all = np.concatenate([np.asarray(k, dtype=np.int64) for k in (gather(mine) if world > 1 else [mine])])
        n, nsub, N0 = L.n_global, L.nsub_global, int(L.k_all.sum())
        nnz_r0 = sum(F.nnz_lower() for F in L.factors)
    else:
        P = build_problem(a, world, rank, gather)
        n, nsub, N0 = P.n, P.nsub, P.A0.shape[0]
        nnz_r0 = sum(F.nnz_lower() for F in P.factors if F is not None)
    t_setup = time.perf_counter() - t0
    if rank == 0:
        print(f"[bench] host setup {t_setup:.1f} s: n={n} nsub={nsub} N0={N0} nnz(L) on rank 0 = {nnz_r0 / 1e6:.1f} M", file=sys.stderr, flush=True)
    t0 = time.perf_counter()
    if world > 1:
        import ctypes as C

        buf = (C.c_uint8 * 128)()
        if rank == 0:
            api.check(lib.load().ddm_comm_unique_id(buf))
        t = torch.tensor(list(buf), dtype=torch.uint8)
        dist.broadcast(t, 0)
        uid = t.tolist()
    if L is not None:
        ctx = api.Context.from_local(L, device=local, unique_id=uid, keep_A0=(world == 1 and not a.no_cpu_baseline), deterministic=a.deterministic)
        b_own = L.b_own
    else:
        ctx = api.Context.from_products(P, device=local, rank=rank, world=world, unique_id=uid, deterministic=a.deterministic)
        b_own = ctx.owned(P.b)
    t_upload = time.perf_counter() - t0
    if rank == 0:
        print(f"[bench] upload, device factorisation, schedules, Galerkin product {t_upload:.1f} s", file=sys.stderr, flush=True)
    op, sp = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
    Cp = api.CombinedPreconditioner("additive")
    Cp.set_op(op)
    Cp.add(api.SchwarzPreconditioner(ctx, "standard"))
    Cp.add(api.GalerkinPreconditioner(ctx))
    cg = api.CGSolver(op, sp, Cp, reduction=REDUCTION, maxit=MAXIT)
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
    # ---- timed region: K solves, inputs resident in HBM, CUDA events on the library's stream.  The per-kernel event timers
    # stay on: their events come from a pool (two cudaEventRecord per scope, nothing is created inside the solve).
    ctx.timers(True)
    l0 = ctx.launches()
    barrier()
    profiled = bool(os.environ.get("DDM_PROFILE_REGION"))  # ncu --profile-from-start off: only the timed solves are captured
    if profiled:
        api.check(ctx.lib.ddm_profiler_range(ctx.h, 1))
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
    if profiled:
        api.check(ctx.lib.ddm_profiler_range(ctx.h, 0))
    launches = ctx.launches() - l0
    timers = {ev: ctx.timer(ev) for ev in ("Schwarz/local solve fwd", "Schwarz/local solve bwd", "Schwarz/local solve", "Schwarz/get defect",
                                          "Schwarz/add solution", "GalerkinPrec/coarse solve", "CombinedPreconditioner/apply", "Operator/apply")}
    ctx.timers(False)
    # ---- parity block (untimed): true residual of the last timed solve, ||b - A x|| / ||b|| over all ranks
    r = ctx.vector(b_own)
    bnorm = sp.norm(r)
    op.applyscaleadd(-1.0, x, r)
    true_res = sp.norm(r) / bnorm if bnorm > 0 else 0.0
    # ---- e2e: the same solve through the host-buffer entry (H2D of b and x0, D2H of x and the defect, every step); the
    # caller's two vectors are page-locked once (ddm_host_register), as a driver does for its BlockVector storage
    bw, xw = np.empty_like(b_own), np.empty_like(b_own)
    ctx.register_host(bw)
    ctx.register_host(xw)

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
    ctx.unregister_host(bw)
    ctx.unregister_host(xw)
    # ---- reduce over ranks
    its_all = torch.tensor([float(res.iterations), float(res.converged)], dtype=torch.float64)
    vals = torch.tensor([dev_ms, t_wall, t_e2e], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        lo = its_all.clone()
        dist.all_reduce(its_all, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        same_its = bool((lo == its_all).all())
    else:
        same_its = True
    dev_ms, t_wall, t_e2e = vals.tolist()
    stats = {k: ctx.stat(k) for k in ("trisolve_fwd", "trisolve_bwd", "trisolve_stored", "trisolve_stored_bwd", "spmv", "restrict", "prolong", "coarse", "factor_nnz")}
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
    ms_step = dev_ms / a.steps
    ok = bool(res.converged) and same_its and true_res <= 1.05 * REDUCTION
    value = n / (ms_step * 1e-3) / 1e6 if ok else None  # a solve that did not reach 1e-8 has no time-to-1e-8
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
        kern[name] = {"ms": ms / max(calls, 1), "GBps": stats[name] / (ms / max(calls, 1) * 1e-3) / 1e9 if ms > 0 else None, "launches": calls,
                      "algorithmic_bytes": stats[name]}
    dom = max(kern, key=lambda k: kern[k]["ms"])
    apply_ms, apply_calls = timers["CombinedPreconditioner/apply"]
    apply_bytes = stats["trisolve_fwd"] + stats["trisolve_bwd"] + stats["restrict"] + stats["prolong"] + stats["coarse"]
    traffic = traffic_note = None
    try:  # ncu dram bytes of the dominant kernel: only a capture of a factor with exactly this many entries counts
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        for ent in tr.get("captures", []):
            if int(ent.get("factor_nnz", -1)) == int(stats["factor_nnz"]) and world == 1:
                traffic, traffic_note = ent.get(f"{dom}_kernel"), ent.get("source")
    except (OSError, ValueError):
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong" if a.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(a, world, n, nsub, N0),
        "iterations": int(res.iterations), "converged": bool(res.converged), "factor_nnz": stats["factor_nnz"], "deterministic_sweeps": bool(a.deterministic),
        "parity": {"true_residual": true_res, "bound": 1.05 * REDUCTION, "iterations_equal_on_all_ranks": same_its, "oracle_iterations": None},
        "time_to_1e-8_s": ms_step * 1e-3, "wall_ms_per_step": t_wall / a.steps * 1e3,
        "setup_s": {"host": t_setup, "upload_and_device_setup": t_upload},
        "precond_apply": {"ms": apply_ms / max(apply_calls, 1), "algorithmic_GB": apply_bytes / 1e9,
                          "GBps": apply_bytes / (apply_ms / max(apply_calls, 1) * 1e-3) / 1e9 if apply_ms > 0 else None,
                          "frac_of_peak": apply_bytes / (apply_ms / max(apply_calls, 1) * 1e-3) / 1e9 / peak if apply_ms > 0 else None},
        "e2e": {"value": n / (t_e2e / a.steps) / 1e6 if ok else None, "unit": UNIT, "h2d_bytes_per_step": int(2 * 8 * ctx.n_own), "d2h_bytes_per_step": int(2 * 8 * ctx.n_own)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": f"{dom}_kernel", "achieved": kern[dom]["GBps"], "peak": peak, "unit": "GB/s",
                     "frac": kern[dom]["GBps"] / peak if kern[dom]["GBps"] else None, "traffic": traffic, "traffic_source": traffic_note,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": stats[dom], "stored_bytes_per_launch": stats["trisolve_stored" if dom == "trisolve_fwd" else "trisolve_stored_bwd"],
                     "ms_per_launch": kern[dom]["ms"], "kernels": kern},
        "timers_ms_per_call": {k: (v[0] / max(v[1], 1)) for k, v in timers.items()},
        "clocks": clk.summary(),
        "per_rank": per_rank,
    }
    if not a.no_cpu_baseline and world == 1:
        t_solve, its_cpu, _, desc, O = cpu_arm(a, world, ctx, L=L, P=P if L is None else None, its_hint=int(res.iterations), budget_s=a.cpu_seconds)
        if its_cpu is None:
            t_solve *= int(res.iterations) + 1
        line["cpu_baseline"] = {"value": n / t_solve / 1e6, "unit": UNIT, "cores": host_cores(), "kind": "port", "sample": desc, "seconds_full_solve_est": t_solve}
    else:
        line["cpu_baseline"] = None
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
