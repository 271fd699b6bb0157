"""Host-side logic without a GPU: the C-ABI library's exported surface, error behaviour of the host mirror,
and the sharding / halo plan of ddm_b200/partition.py (what replaces OwnerOverlapCopyCommunication)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from ddm_b200 import api, lib, partition

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    h = open(os.path.join(ROOT, "include", "ddm_b200.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(ddm_[a-z0-9_]+)\s*\(", h)))


def test_library_exports_every_declared_symbol():
    L = lib.load()
    names = _declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(L, n), f"libddm_b200.so does not export {n}"
    assert sorted(lib.SYMBOLS) == names
    assert b"sm_100a" in L.ddm_version()


def test_no_cpu_fallback_without_a_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(lib.DDMError, match="CUDA device"):
        api.Context(0)


def test_reference_error_conventions():
    with pytest.raises(lib.DDMError, match="Unknown Schwarz type"):  # schwarz.hh:83
        api.SchwarzPreconditioner(None, "hybrid")
    with pytest.raises(lib.DDMError, match="Unknown apply mode"):  # combined_preconditioner.hh:66
        api.CombinedPreconditioner("both")
    c = api.CombinedPreconditioner("additive")
    with pytest.raises(lib.DDMError, match="No preconditioners"):
        c.apply(None, None)

    class Wrong:
        category = "overlapping"

    with pytest.raises(lib.DDMError, match="category"):  # combined_preconditioner.hh:92-96
        c.add(Wrong())


@pytest.mark.parametrize("world", [1, 2, 3, 4])
def test_shards_partition_the_problem(small3d, world):
    P = small3d
    shards = [partition.shard(P, r, world) for r in range(world)]
    owned = np.concatenate([S.owned for S in shards])
    assert np.array_equal(np.sort(owned), np.arange(P.n))  # every DOF owned exactly once
    assert sum(S.sub_range[1] - S.sub_range[0] for S in shards) == P.nsub
    koff = 0
    for S in shards:
        a, b = S.sub_range
        assert S.k_offset == koff
        koff += sum(P.Z[i].shape[1] for i in range(a, b))
        # owner = lowest subdomain that sees the DOF, so a rank owns what its subdomains own
        assert ((P.owner[S.owned] >= a) & (P.owner[S.owned] < b)).all()
        g = np.concatenate([S.owned, S.ghosts])
        assert np.unique(g).size == g.size
        # subdomain lists and operator columns resolve to the right global ids
        for j, i in enumerate(range(a, b)):
            assert np.array_equal(g[S.sub_dofs[S.sub_ptr[j] : S.sub_ptr[j + 1]]], P.subs[i])
        rows = P.A[S.owned]
        assert np.array_equal(g[S.col], rows.indices)
        assert np.array_equal(S.val, rows.data)
    assert koff == P.A0.shape[0]


@pytest.mark.parametrize("world", [2, 4])
def test_halo_plan_is_consistent(small3d, world):
    """send lists of rank r towards p name, in p's ghost order, exactly the DOFs p keeps as ghosts of r."""
    P = small3d
    shards = [partition.shard(P, r, world) for r in range(world)]
    for S in shards:
        for k, p in enumerate(S.peers):
            T = shards[int(p)]
            kk = int(np.flatnonzero(T.peers == S.rank)[0])
            sent = S.owned[S.send_idx[S.send_ptr[k] : S.send_ptr[k + 1]]]
            kept = T.ghosts[T.recv_ptr[kk] : T.recv_ptr[kk + 1]]
            assert np.array_equal(sent, kept)


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_spmv_and_halo_sum_in_lockstep(small3d, world):
    from oracle import oracle
    from shard_emulator import RankEmu

    P = small3d
    O = oracle.Oracle(P)
    emus = [RankEmu(P, partition.shard(P, r, world), O) for r in range(world)]
    rng = np.random.default_rng(0)
    x = rng.standard_normal(P.n)
    vs = [e.local(x[e.S.owned]) for e in emus]
    # forward halo, all ranks in lock-step
    send = [{int(p): v[e.S.send_idx[e.S.send_ptr[k] : e.S.send_ptr[k + 1]]] for k, p in enumerate(e.S.peers)} for e, v in zip(emus, vs)]
    for e, v in zip(emus, vs):
        e.halo_forward(v, lambda _s, _d, me=e.S.rank: {p: send[p][me] for p in send[me]})
        g = np.concatenate([e.S.owned, e.S.ghosts])
        assert np.array_equal(v, x[g])
    y = np.zeros(P.n)
    for e, v in zip(emus, vs):
        y[e.S.owned] = e.spmv(v, lambda _s, _d: {int(p): v[e.S.n_own + e.S.recv_ptr[k] : e.S.n_own + e.S.recv_ptr[k + 1]] for k, p in enumerate(e.S.peers)})
    assert np.allclose(y, P.A @ x, rtol=0, atol=1e-13 * np.abs(P.A @ x).max())
    # reverse halo add: sum of all local contributions equals the global scatter-add
    contrib = [rng.standard_normal(e.nloc) for e in emus]
    want = np.zeros(P.n)
    for e, c in zip(emus, contrib):
        np.add.at(want, np.concatenate([e.S.owned, e.S.ghosts]), c)
    send = [{int(p): c[e.S.n_own + e.S.recv_ptr[k] : e.S.n_own + e.S.recv_ptr[k + 1]] for k, p in enumerate(e.S.peers)} for e, c in zip(emus, contrib)]
    got = np.zeros(P.n)
    for e, c in zip(emus, contrib):
        c = c.copy()
        e.halo_reverse_add(c, lambda _s, _d, me=e.S.rank: {p: send[p][me] for p in send[me]})
        got[e.S.owned] = c[: e.S.n_own]
    assert np.allclose(got, want, rtol=0, atol=1e-13)


def test_eight_ranks_all_to_all_halo_and_apply():
    """The 2x2x2 rank layout of an 8-GPU box: every rank has all seven others as peers.  Halo plan symmetry and the rank-local
    two-level sequence (emulated with threads exchanging in lock step) against the single-process oracle."""
    import threading

    from ddm_b200.setup import problem
    from oracle import oracle
    from shard_emulator import RankEmu

    world = 8
    P = problem.unitcube_p1(12, 32, overlap=2, nev=4, workers=1)
    O = oracle.Oracle(P)
    shards = [partition.shard(P, r, world) for r in range(world)]
    assert max(S.peers.size for S in shards) == world - 1
    for S in shards:
        for k, p in enumerate(S.peers):
            T = shards[int(p)]
            kk = np.flatnonzero(T.peers == S.rank)
            assert kk.size == 1  # peers are mutual: every send has its receive
            sent = S.owned[S.send_idx[S.send_ptr[k] : S.send_ptr[k + 1]]]
            assert np.array_equal(sent, T.ghosts[T.recv_ptr[kk[0]] : T.recv_ptr[kk[0] + 1]])
    emus = [RankEmu(P, S, O) for S in shards]
    bar, box = threading.Barrier(world), {}

    def comm(rank):
        def exchange(send, direction):
            box[rank] = send
            bar.wait()
            out = {p: box[p][rank] for p in send}
            bar.wait()
            return out

        def allreduce(v):
            box[("a", rank)] = v.copy()
            bar.wait()
            tot = sum(box[("a", r)] for r in range(world))
            bar.wait()
            return tot

        return exchange, allreduce

    d = np.random.default_rng(3).standard_normal(P.n)
    res = [None] * world

    def run(r):
        ex, ar = comm(r)
        res[r] = emus[r].twolevel(emus[r].local(d[shards[r].owned]), ex, ar)

    threads = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    got = np.zeros(P.n)
    for S, xr in zip(shards, res):
        got[S.owned] = xr
    want = O.twolevel_apply(d)
    assert np.linalg.norm(got - want) <= 1e-10 * np.linalg.norm(want)
