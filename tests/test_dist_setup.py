"""Setup sharded over processes (bench.py with N > 1): every rank factors and builds the coarse basis of its own subdomains,
bases and Galerkin columns are exchanged over gloo; the assembled products must equal the single-process setup."""
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))

WORKER = r"""
import os, sys, pickle
sys.path.insert(0, os.path.join({root!r}, "dune-ddm_b200"))
import numpy as np, torch.distributed as dist
from ddm_b200.setup import problem
from ddm_b200 import partition
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=int(sys.argv[1]), world_size=2)
rank, world = dist.get_rank(), dist.get_world_size()
def gather(obj):
    out = [None] * world
    dist.all_gather_object(out, obj)
    return out
a, b = partition.sub_ranges(6, world)[rank]
P = problem.unitcube_p1(10, 6, overlap=2, nev=4, workers=1, only=range(a, b), gather=gather, assemble_part=(rank, world))
have = [i for i, F in enumerate(P.factors) if F is not None]
pickle.dump(dict(A0=P.A0, Z=P.Z, have=have, x={{i: P.factors[i].x for i in have}}), open(sys.argv[2], "wb"))
dist.destroy_process_group()
"""


@pytest.mark.timeout(300)
def test_sharded_setup_equals_single_process(tmp_path):
    import pickle
    import socket

    from ddm_b200.setup import problem

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=os.path.dirname(HERE), port=port))
    outs = [str(tmp_path / f"r{r}.pkl") for r in range(2)]
    procs = [subprocess.Popen([sys.executable, str(script), str(r), outs[r]]) for r in range(2)]
    for p in procs:
        assert p.wait(timeout=280) == 0
    P = problem.unitcube_p1(10, 6, overlap=2, nev=4, workers=1)
    R = [pickle.load(open(o, "rb")) for o in outs]
    assert R[0]["have"] == [0, 1, 2] and R[1]["have"] == [3, 4, 5]
    # the matrix is summed in a different order (rank slices of the cells), so values agree to rounding, not bit for bit;
    # eigenvectors of a cluster may rotate, the Galerkin matrix is compared through its spectrum
    assert np.array_equal(R[0]["A0"], R[1]["A0"])
    for r in R:
        assert np.allclose(np.linalg.eigvalsh(r["A0"]), np.linalg.eigvalsh(P.A0), rtol=1e-6)
        assert [z.shape for z in r["Z"]] == [z.shape for z in P.Z]
        for i in r["have"]:
            assert r["x"][i].shape == P.factors[i].x.shape and np.allclose(r["x"][i], P.factors[i].x, rtol=1e-9, atol=1e-9 * np.abs(P.factors[i].x).max())


BENCH_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "dune-ddm_b200"))
import torch.distributed as dist, numpy as np
import bench
from ddm_b200 import partition
sys.argv = ["bench.py", "--ncell", "8", "--nsub", "8", "--gpus", os.environ["WORLD_SIZE"], "--eigensolver", "host", "--numeric", "host",
            "--setup", "global"] + {extra!r}
a = bench.parse()
rank, world, local = bench.dist_env()
dist.init_process_group("gloo")
def gather(obj):
    out = [None] * world
    dist.all_gather_object(out, obj)
    return out
P = bench.build_problem(a, world, rank, gather)
S = partition.shard(P, rank, world)
have = [i for i, F in enumerate(P.factors) if F is not None]
assert have == list(range(*S.sub_range)), (have, S.sub_range)
assert P.nsub == (8 if a.strong else 8 * world) and (P.n == 9 ** 3 or not a.strong)
if a.workload == "q1cube":  # every rank's bricks form one block of the brick grid
    cent = np.array([P.coords[s[: int(m)]].mean(axis=0) for s, m in zip(P.subs, P.n_novlp)])
    lo, hi = S.sub_range
    ext = cent[lo:hi].max(axis=0) - cent[lo:hi].min(axis=0)
    assert (ext <= 0.5 + 1e-9).all(), ext
assert P.A0.shape[0] == sum(z.shape[1] for z in P.Z) and (a.workload == "q1cube" or bench.workload_name(a, world).count(f"{{P.nsub}} subdomains") == 1)
assert abs(P.A0 - P.A0.T).max() <= 1e-12 * abs(P.A0).max()
sums = [None] * world
dist.all_gather_object(sums, (float(abs(P.A0).sum()), float(abs(P.A).sum()), S.n_own))
assert len(set(s[:2] for s in sums)) == 1 and sum(s[2] for s in sums) == P.n  # identical products on every rank
dist.barrier()
"""


@pytest.mark.timeout(300)
@pytest.mark.parametrize("extra", [["--workload", "p1cube"], ["--workload", "p1cube", "--strong"], ["--workload", "q1cube"], ["--workload", "q1cube", "--strong"]],
                         ids=["weak", "strong", "q1-weak", "q1-strong"])
def test_bench_sharded_setup_under_torchrun(tmp_path, extra):
    """bench.py's N > 1 setup wiring (weak: refined unit cube, 64 -> here 8 subdomains per rank; strong: the N = 1 problem dealt
    out), gloo exchange, 4 ranks on the CPU."""
    import socket

    script = tmp_path / "bench_worker.py"
    script.write_text(BENCH_WORKER.format(root=os.path.dirname(HERE), extra=extra))
    for attempt in range(2):  # the port is picked before torchrun binds it: a second try covers the rare race for it
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=4", "--master-addr", "127.0.0.1",
                            "--master-port", str(port), str(script)], capture_output=True, text=True, timeout=140)
        if r.returncode == 0 or "AssertionError" in r.stderr:
            break
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]


LOCAL_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "dune-ddm_b200"))
import torch.distributed as dist, numpy as np
import bench
sys.argv = ["bench.py", "--ncell", "8", "--nsub", "8", "--gpus", os.environ["WORLD_SIZE"], "--eigensolver", "host", "--numeric", "host", "--nev", "4"]
a = bench.parse()
rank, world, local = bench.dist_env()
assert bench.use_local_setup(a)  # the bench's default path for the Q1 brick family
dist.init_process_group("gloo")
L = bench.build_local(a, world, rank, 0)
mine = [int(z.shape[1]) for z in L.Z]
ks = [None] * world
dist.all_gather_object(ks, mine)
L.k_all = np.concatenate([np.asarray(k, dtype=np.int64) for k in ks])
S = L.shard
assert S.rank == rank and S.world == world and len(L.factors) == 8 and L.nsub_global == 8 * world
assert L.k_all.size == L.nsub_global and int(L.k_all[S.sub_range[0]:S.sub_range[1]].sum()) == sum(mine)
# every DOF has exactly one owner, and what rank r sends to q is what q expects from r, id by id
info = [None] * world
sends = {{int(p): S.owned[S.send_idx[S.send_ptr[i]:S.send_ptr[i + 1]]] for i, p in enumerate(S.peers)}}
recvs = {{int(p): S.ghosts[S.recv_ptr[i]:S.recv_ptr[i + 1]] for i, p in enumerate(S.peers)}}
dist.all_gather_object(info, dict(owned=S.owned, sends=sends, recvs=recvs, n=L.n_global))
allowned = np.concatenate([i["owned"] for i in info])
assert allowned.size == L.n_global and np.unique(allowned).size == L.n_global
for q, other in enumerate(info):
    if q == rank:
        continue
    mine_to_q = sends.get(q, np.zeros(0, dtype=np.int64))
    q_from_me = other["recvs"].get(rank, np.zeros(0, dtype=np.int64))
    assert np.array_equal(mine_to_q, q_from_me), (rank, q)
name = bench.workload_name(a, world)
assert "brick subdomains" in name
dist.barrier()
"""


@pytest.mark.timeout(300)
def test_bench_rank_local_setup_under_torchrun(tmp_path):
    """bench.py's default N > 1 wiring on the CPU: four ranks (gloo) each build their own region (setup/local.py), exchange only
    the coarse-vector counts, and their halo plans fit together: every DOF has one owner, and the ids rank r sends to rank q
    are exactly, and in the same order, the ghost ids q expects from r."""
    import socket

    script = tmp_path / "local_worker.py"
    script.write_text(LOCAL_WORKER.format(root=os.path.dirname(HERE)))
    for attempt in range(2):
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=4", "--master-addr", "127.0.0.1",
                            "--master-port", str(port), str(script)], capture_output=True, text=True, timeout=200)
        if r.returncode == 0 or "AssertionError" in r.stderr:
            break
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
