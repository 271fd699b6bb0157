"""Domain decomposition on the host: element partition, ownership, overlap growth,
boundary identification and partition of unity.

Restates, for "P virtual ranks in one process", what the reference does with MPI:
  * non-overlapping index set, owner = lowest rank that sees a DOF
    (/root/reference/dune/ddm/pdelab_helper.hh:16-94),
  * overlap growth by `overlap` layers of the matrix graph
    (/root/reference/dune/ddm/overlap_extension.hh:54-285): the first entries of an
    overlapping index list are the rank's non-overlapping DOFs, new layers are appended,
  * boundary identification (/root/reference/dune/ddm/datahandles.hh:123-192),
  * partition of unity (/root/reference/dune/ddm/pou.hh:58-141).
All integer outputs are deterministic functions of (A's graph, element partition).
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp


def rcb_partition(points, nparts):
    """Recursive coordinate bisection of `points` into `nparts` parts (any nparts >= 1).

    Stands in for ParMETIS (/root/reference/examples/poisson.cc:129), which is not
    available; ties are broken by index so the result is deterministic.
    """
    part = np.zeros(points.shape[0], dtype=np.int32)
    cols = [np.ascontiguousarray(points[:, d]) for d in range(points.shape[1])]

    def rec(idx, p0, np_):  # idx ascending
        if np_ == 1:
            part[idx] = p0
            return
        nl = np_ // 2
        vs = [c[idx] for c in cols]
        ax = int(np.argmax([x.max() - x.min() for x in vs]))
        cut = (idx.size * nl) // np_
        # the `cut` smallest under the order (coordinate, index), by selection instead of a full sort: everything below the
        # cut value goes left, entries equal to it go left in index order until the count is reached
        v = vs[ax]
        if cut == 0:
            left = np.zeros(idx.size, dtype=bool)
        else:
            kth = np.partition(v, cut - 1)[cut - 1]
            left = v < kth
            eq = np.flatnonzero(v == kth)
            left[eq[: cut - int(left.sum())]] = True
        rec(idx[left], p0, nl)
        rec(idx[~left], p0 + nl, np_ - nl)

    rec(np.arange(points.shape[0]), 0, nparts)
    return part


def box_partition(shape, pgrid, group=None):
    """Partition the cells of a structured grid into a pgrid[0] x pgrid[1] x ... brick layout.

    group = (gx, gy, ...): bricks are numbered so that each of the gx * gy * ... equal blocks of the brick grid owns a
    contiguous range of part ids (block-major, then brick-lexicographic inside the block) -- one block per GPU, the
    partitioning of SURVEY 8(e): 8 x 8 x 8 subdomains -> 2 x 2 x 2 GPU bricks of 4 x 4 x 4."""
    idx = np.meshgrid(*[np.arange(s) for s in shape], indexing="ij")
    bricks = [(idx[ax] * p) // s for ax, (s, p) in enumerate(zip(shape, pgrid))]
    if group is None:
        group = (1,) * len(shape)
    per = [p // g for p, g in zip(pgrid, group)]
    if any(p % g for p, g in zip(pgrid, group)):
        raise ValueError(f"brick grid {tuple(pgrid)} does not split into {tuple(group)} equal blocks")
    blk = np.zeros(shape, dtype=np.int32)
    loc = np.zeros(shape, dtype=np.int32)
    for ax in range(len(shape)):
        blk = blk * group[ax] + bricks[ax] // per[ax]
        loc = loc * per[ax] + bricks[ax] % per[ax]
    return (blk * int(np.prod(per)) + loc).ravel().astype(np.int32)


def nonoverlapping_sets(n, cells, cell_part, nparts):
    """Per rank: sorted DOFs of its elements (owner and border copies); owner = lowest rank."""
    k = cells.shape[1]
    inc = sp.coo_matrix(
        (np.ones(cells.size, dtype=np.int8), (np.repeat(cell_part, k), cells.ravel())), shape=(nparts, n)
    ).tocsr()
    inc.sum_duplicates()
    novlp = [np.sort(inc.indices[inc.indptr[r] : inc.indptr[r + 1]]).astype(np.int64) for r in range(nparts)]
    owner = np.full(n, nparts, dtype=np.int32)
    for r in range(nparts - 1, -1, -1):
        owner[novlp[r]] = r
    return novlp, owner


def grow_overlap(A, novlp, overlap):
    """Overlapping index lists: novlp[r] first, then `overlap` graph layers in discovery order.

    Within a layer new indices are appended in ascending global id.
    """
    G = (A != 0).astype(np.int8).tocsr()
    n = A.shape[0]
    out = []
    for base in novlp:
        mask = np.zeros(n, dtype=bool)
        mask[base] = True
        parts = [base]
        front = base
        for _ in range(overlap):
            ind = sp.csr_matrix(
                (np.ones(front.size, dtype=np.int8), (np.zeros(front.size, dtype=np.int64), front)), shape=(1, n)
            )
            nb = (ind @ G).indices
            new = np.sort(nb[~mask[nb]]).astype(np.int64)
            mask[new] = True
            parts.append(new)
            front = new
        out.append(np.concatenate(parts))
    return out


def boundary_masks(A, subs):
    """Boundary of each overlapping subdomain.

    datahandles.hh:159-175 marks local index i when a sharing rank reports a matrix
    neighbour of i that this rank does not know.  With overlap >= 1 the owner of any
    neighbour j of i also holds i, so the condition is "i has a graph neighbour outside
    the subdomain".
    """
    n = A.shape[0]
    G = (A != 0).astype(np.int32).tocsr()
    masks = []
    for s in subs:
        inside = np.zeros(n, dtype=np.int32)
        inside[s] = 1
        deg_in = G[s] @ inside
        deg = np.diff(G.indptr)[s]
        masks.append(deg_in < deg)
    return masks


def graph_distance(Ai, seeds, maxdist):
    """BFS distance in the graph of the local matrix, capped at maxdist+1."""
    n = Ai.shape[0]
    dist = np.full(n, np.iinfo(np.int32).max - 1, dtype=np.int64)
    dist[seeds] = 0
    G = (Ai != 0).astype(np.int8).tocsr()
    front = np.flatnonzero(seeds)
    d = 0
    while front.size and d <= maxdist:
        d += 1
        ind = sp.csr_matrix(
            (np.ones(front.size, dtype=np.int8), (np.zeros(front.size, dtype=np.int64), front)), shape=(1, n)
        )
        nb = (ind @ G).indices
        new = nb[dist[nb] > d]
        dist[new] = d
        front = new
    return dist


def pou_raw(A, subs, bmasks, owner, kind="distance", shrink=0, overlap=1, ids=None):
    """Un-normalised weights of the subdomains `ids` (default: all), pou.hh:79-127."""
    ids = range(len(subs)) if ids is None else ids
    raw = {}
    for r in ids:
        s, bm = subs[r], bmasks[r]
        if kind == "trivial":
            w = (owner[s] == r).astype(np.float64)
        elif kind == "standard":
            w = np.where(bm, 0.0, 1.0)
        elif kind == "distance":
            Ai = A[s][:, s]
            dist = graph_distance(Ai, bm, 4 * overlap)
            w = np.ones(s.size)
            near = dist <= 4 * overlap
            w[near] = np.maximum(dist[near] - shrink, 0).astype(np.float64)
        else:
            raise ValueError(f"unknown partition of unity type {kind!r}")
        raw[r] = w
    return raw


def pou_normalise(n, subs, raw, bmasks, kind="distance"):
    """Divide by the sum over all subdomains that see a DOF (the halo sum of pou.hh:128-140)."""
    if kind == "trivial":
        return list(raw)
    total = np.zeros(n)
    for s, w in zip(subs, raw):
        np.add.at(total, s, w)
    pou = []
    for s, w, bm in zip(subs, raw, bmasks):
        with np.errstate(divide="ignore", invalid="ignore"):
            d = np.where(bm, 0.0, w / total[s])
        pou.append(d)
    return pou


def partition_of_unity(A, subs, bmasks, owner, kind="distance", shrink=0, overlap=1):
    """Partition-of-unity weights D_i per overlapping subdomain (pou.hh:79-140).

    The reference computes the distance with 4*overlap+1 Gauss-Seidel sweeps; every value
    <= 4*overlap it can produce equals the true graph distance, and larger values all map
    to weight 1, so a breadth-first search gives identical weights.
    """
    raw = pou_raw(A, subs, bmasks, owner, kind, shrink, overlap)
    return pou_normalise(A.shape[0], subs, [raw[r] for r in range(len(subs))], bmasks, kind)
