#!/usr/bin/env python
"""CPU check of bench.py's weak-scaling family: builds the N-GPU problem of `bench.py --gpus N` in one process (no GPU, no
torchrun) and solves it with the oracle, so that the iteration count the GPUs should reproduce is known beforehand.

  python scripts/oracle_family_check.py --gpus 2 [--ncell 112] [--workload p1cube|q1cube] [bench.py options]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))

import bench  # noqa: E402


def main():
    a = bench.parse()
    t0 = time.perf_counter()
    P = bench.build_problem(a, a.gpus)
    t_setup = time.perf_counter() - t0
    from oracle import oracle

    O = oracle.Oracle(P, nthreads=os.cpu_count() or 1)
    t0 = time.perf_counter()
    x, it, conv, hist = O.pcg(P.b, reduction=bench.REDUCTION, maxit=bench.MAXIT)
    t_solve = time.perf_counter() - t0
    print(json.dumps({"gpus": a.gpus, "workload": bench.workload_name(a, a.gpus), "dofs": int(P.n), "subdomains": int(P.nsub),
                      "coarse_dim": int(P.A0.shape[0]), "iterations": int(it), "converged": bool(conv), "reduction": float(hist[-1] / hist[0]),
                      "setup_s": round(t_setup, 1), "oracle_solve_s": round(t_solve, 1), "cores": os.cpu_count()}))


if __name__ == "__main__":
    main()
