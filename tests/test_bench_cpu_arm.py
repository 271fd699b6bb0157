"""bench.py's CPU arm (oracle on the setup products of the rank-local path): the whole-problem mode, the bounded-sample mode it
falls back to when the factors exceed host memory, and the N > 1 mode (rank 0's shard, scaled).  CPU only: host factors, no device."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class NoDevice:
    A0_device = None


@pytest.fixture(scope="module")
def setting():
    import bench
    from ddm_b200.setup import local

    argv, sys.argv = sys.argv, ["bench.py", "--ncell", "16", "--nsub", "8", "--eigensolver", "host", "--numeric", "host"]
    try:
        a = bench.parse()
    finally:
        sys.argv = argv
    L = local.poisson_q1_cube_local((16, 16, 16), (2, 2, 2), (1, 1, 1), 0, overlap=2, coarse="geneo", nev=4, workers=1, numeric="host", eigensolver="host")
    L.k_all = np.array([z.shape[1] for z in L.Z])
    return bench, a, L


def test_whole_problem_mode_runs_the_real_solve(setting):
    bench, a, L = setting
    t, its, conv, desc, O = bench.cpu_arm(a, 1, NoDevice(), L=L, budget_s=1.0, nthreads=2)
    assert conv and its > 5 and t > 0 and "PCG iterations of the same solve" in desc
    # the products handed to the oracle are the global problem: its solve satisfies the operator the shard carries
    P = bench.host_products(NoDevice(), L)
    x, it2, conv2, _ = O.pcg(P.b, reduction=1e-8, maxit=200)
    assert conv2 and it2 == its and np.linalg.norm(P.b - P.A @ x) <= 1.05e-8 * np.linalg.norm(P.b)


def test_bounded_sample_mode(setting, monkeypatch):
    bench, a, L = setting
    monkeypatch.setattr(bench, "mem_available_bytes", lambda: 0)  # nothing fits: sample
    t_iter, its, conv, desc, _ = bench.cpu_arm(a, 1, NoDevice(), L=L, its_hint=12, budget_s=1.0, nthreads=2)
    assert its == 12 and conv is None and t_iter > 0 and "subdomain solves of" in desc and "of 8 subdomains" in desc


def test_sample_on_rank_zero_of_two(setting):
    bench, a, _ = setting
    from ddm_b200.setup import local

    L2 = local.poisson_q1_cube_local((32, 16, 16), (4, 2, 2), (2, 1, 1), 0, overlap=2, coarse="geneo", nev=4, workers=1, numeric="host", eigensolver="host")
    t_iter, its, conv, desc, _ = bench.cpu_arm(a, 2, NoDevice(), L=L2, budget_s=1.0, nthreads=2)
    assert its is None and t_iter > 0 and "of 16 subdomains" in desc and "x2 shares" in desc


def test_both_arms_emit_the_same_config(setting):
    bench, a, L = setting
    c1 = bench.config_of(a, 1, L.n_global, L.nsub_global, int(L.k_all.sum()))
    c2 = bench.config_of(a, 1, L.n_global, L.nsub_global, int(L.k_all.sum()))
    assert c1 == c2 and set(c1) == {"workload", "dofs", "subdomains", "coarse_dim", "l2_policy"} and "brick subdomains" in c1["workload"]


def test_reference_arm_times_as_many_samples_as_steps(setting, monkeypatch):
    """--impl reference --steps K: K bounded samples are timed (whole-problem mode: K short PCG runs; sample mode: at least K
    applications), so the line's `steps` says what was done."""
    bench, a, L = setting
    t, its, conv, desc, _ = bench.cpu_arm(a, 1, NoDevice(), L=L, budget_s=1.0, nthreads=2, steps=3)
    assert conv and t > 0 and "each of 3 steps" in desc
    monkeypatch.setattr(bench, "mem_available_bytes", lambda: 0)
    t_iter, _, _, desc, _ = bench.cpu_arm(a, 1, NoDevice(), L=L, its_hint=12, budget_s=0.0, nthreads=2, steps=4)
    assert t_iter > 0 and desc.startswith("4 timed applications")
