"""The N>1 path on CPU: two processes, torch.distributed/gloo standing in for NCCL, each replaying its rank's
sequence (tests/shard_emulator.py) on its shard; the assembled result must equal the single-process oracle."""
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))

WORKER = r"""
import os, sys, pickle
sys.path.insert(0, os.path.join({root!r}, "dune-ddm_b200")); sys.path.insert(0, {root!r}); sys.path.insert(0, {here!r})
import numpy as np, torch, torch.distributed as dist
from ddm_b200.setup import problem
from ddm_b200 import partition
from oracle import oracle
from shard_emulator import RankEmu

dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=int(sys.argv[1]), world_size=2)
rank, world = dist.get_rank(), dist.get_world_size()
P = problem.unitcube_p1(10, 6, overlap=2, nev=4, workers=1)
S = partition.shard(P, rank, world)
E = RankEmu(P, S, oracle.Oracle(P))

def exchange(send, direction):
    out, reqs = {{}}, []
    for k, p in enumerate(S.peers):
        p = int(p)
        n_in = int(S.recv_ptr[k + 1] - S.recv_ptr[k]) if direction == "fwd" else int(S.send_ptr[k + 1] - S.send_ptr[k])
        buf = torch.zeros(n_in, dtype=torch.float64)
        out[p] = buf
        reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(send[p])), p))
        reqs.append(dist.irecv(buf, p))
    for r in reqs: r.wait()
    return {{p: b.numpy() for p, b in out.items()}}

def allreduce(v):
    t = torch.from_numpy(v.copy()); dist.all_reduce(t); return t.numpy()

rng = np.random.default_rng(7)
d = rng.standard_normal(P.n)
res = {{}}
for name, kw in (("additive", dict()), ("restricted", dict(restricted=True)), ("onelevel", dict(with_coarse=False))):
    v = E.local(d[S.owned])
    res[name] = E.twolevel(v, exchange, allreduce, **kw)
# distributed dot: owned parts only, then allreduce (NonOverlappingScalarProduct)
res["dot"] = allreduce(np.array([d[S.owned] @ d[S.owned]]))[0]
res["owned"] = S.owned
pickle.dump(res, open(sys.argv[2], "wb"))
dist.destroy_process_group()
"""


@pytest.mark.timeout(300)
def test_two_rank_gloo_matches_oracle(tmp_path):
    import pickle
    import socket

    from ddm_b200.setup import problem
    from oracle import oracle

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    root = os.path.dirname(HERE)
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=root, here=HERE, port=port))
    outs = [str(tmp_path / f"r{r}.pkl") for r in range(2)]
    procs = [subprocess.Popen([sys.executable, str(script), str(r), outs[r]]) for r in range(2)]
    for p in procs:
        assert p.wait(timeout=280) == 0
    P = problem.unitcube_p1(10, 6, overlap=2, nev=4, workers=1)
    O = oracle.Oracle(P)
    d = np.random.default_rng(7).standard_normal(P.n)
    R = [pickle.load(open(o, "rb")) for o in outs]
    want = {"additive": O.twolevel_apply(d), "restricted": O.twolevel_apply(d, restricted=True),
            "onelevel": O.twolevel_apply(d, with_coarse=False)}
    for name, w in want.items():
        got = np.zeros(P.n)
        for r in R:
            got[r["owned"]] = r[name]
        err = np.linalg.norm(got - w) / np.linalg.norm(w)
        assert err <= 1e-10, (name, err)  # the emulator solves A0 with LAPACK LU, the oracle with Cholesky
    assert abs(R[0]["dot"] - d @ d) <= 1e-12 * (d @ d) and R[0]["dot"] == R[1]["dot"]
