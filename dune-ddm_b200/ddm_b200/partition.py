"""Sharding of the setup products over GPUs (host logic, integer work only).

The reference runs one MPI rank per subdomain; here one process drives one GPU that
hosts a contiguous block of subdomains.  A rank owns the DOFs whose owning subdomain it hosts
(owner = lowest subdomain that sees the DOF, /root/reference/dune/ddm/pdelab_helper.hh:44-82), keeps
ghost copies of every other DOF its subdomains or matrix rows touch, and exchanges them with the
halo plan built here (the role of the reference's OwnerOverlapCopyCommunication interfaces,
/root/reference/extern/dune-istl/dune/istl/owneroverlapcopy.hh:311-387).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class RankShard:
    rank: int
    world: int
    sub_range: tuple  # hosted subdomains [s0, s1)
    owned: np.ndarray  # int64 global ids owned by this rank, ascending (local index = position)
    ghosts: np.ndarray  # int64 global ids of ghost copies, grouped by owning rank
    peers: np.ndarray  # int32 peer ranks (ascending)
    send_ptr: np.ndarray  # int64[npeers+1]
    send_idx: np.ndarray  # int32 owned local indices sent to each peer, in that peer's ghost order
    recv_ptr: np.ndarray  # int64[npeers+1] ghost block boundaries
    # operator rows of the owned DOFs, columns in local numbering (owned | ghost)
    rowptr: np.ndarray
    col: np.ndarray
    val: np.ndarray
    sub_ptr: np.ndarray
    sub_dofs: np.ndarray  # local numbering
    k_offset: int

    @property
    def n_own(self):
        return int(self.owned.size)

    @property
    def n_ghost(self):
        return int(self.ghosts.size)


def sub_ranges(nsub, world):
    """Contiguous blocks of subdomains per rank, sizes differing by at most one."""
    cuts = [(nsub * r) // world for r in range(world + 1)]
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def dof_rank(P, world):
    """Owning rank of every DOF."""
    ranges = sub_ranges(P.nsub, world)
    sub2rank = np.empty(P.nsub, dtype=np.int32)
    for r, (a, b) in enumerate(ranges):
        sub2rank[a:b] = r
    return sub2rank[P.owner], ranges


def _needed(P, drank, ranges, r):
    """Global ids rank r touches: its subdomains' index lists and the columns of its owned rows."""
    a, b = ranges[r]
    owned = np.flatnonzero(drank == r).astype(np.int64)
    parts = [P.subs[i] for i in range(a, b)]
    A = P.A
    rows = A[owned]
    parts.append(rows.indices.astype(np.int64))
    touched = np.unique(np.concatenate(parts)) if parts else np.zeros(0, dtype=np.int64)
    ghosts = touched[drank[touched] != r]
    order = np.lexsort((ghosts, drank[ghosts]))
    return owned, ghosts[order], rows


def shard(P, rank, world):
    """Rank-local view of the setup products."""
    drank, ranges = dof_rank(P, world)
    owned, ghosts, rows = _needed(P, drank, ranges, rank)
    n = P.n
    g2l = -np.ones(n, dtype=np.int64)
    g2l[owned] = np.arange(owned.size)
    g2l[ghosts] = owned.size + np.arange(ghosts.size)
    gr = drank[ghosts]
    peers_recv = np.unique(gr)
    # what every other rank wants from me
    send_lists = {}
    for q in range(world):
        if q == rank:
            continue
        _, gq, _ = _needed(P, drank, ranges, q)
        mine = gq[drank[gq] == rank]
        if mine.size:
            send_lists[q] = mine
    peers = np.array(sorted(set(peers_recv.tolist()) | set(send_lists.keys())), dtype=np.int32)
    send_ptr, recv_ptr, send_idx = [0], [0], []
    for p in peers:
        lst = send_lists.get(int(p), np.zeros(0, dtype=np.int64))
        send_idx.append(g2l[lst])
        send_ptr.append(send_ptr[-1] + lst.size)
        recv_ptr.append(recv_ptr[-1] + int((gr == p).sum()))
    a, b = ranges[rank]
    sub_ptr = np.concatenate(([0], np.cumsum([P.subs[i].size for i in range(a, b)]))).astype(np.int64)
    sub_dofs = g2l[np.concatenate([P.subs[i] for i in range(a, b)])] if b > a else np.zeros(0, dtype=np.int64)
    assert (sub_dofs >= 0).all()
    col = g2l[rows.indices]
    assert (col >= 0).all()
    ks = np.array([z.shape[1] for z in P.Z], dtype=np.int64)
    return RankShard(
        rank=rank, world=world, sub_range=(a, b), owned=owned, ghosts=ghosts, peers=peers,
        send_ptr=np.asarray(send_ptr, dtype=np.int64),
        send_idx=(np.concatenate(send_idx) if send_idx else np.zeros(0)).astype(np.int32),
        recv_ptr=np.asarray(recv_ptr, dtype=np.int64),
        rowptr=rows.indptr.astype(np.int64), col=col.astype(np.int32), val=rows.data.astype(np.float64),
        sub_ptr=sub_ptr, sub_dofs=sub_dofs.astype(np.int32), k_offset=int(ks[:a].sum()),
    )
