"""Rank-local setup of the structured Q1 brick family (config C5): every process builds only the part of the mesh its own
GPU needs, so no process ever holds the global problem (100 M DOFs at 8 GPUs).

The reference works the same way: each MPI rank assembles on its own grid partition and learns about its neighbours through
the overlap extension (/root/reference/dune/ddm/overlap_extension.hh:54-285, pdelab_helper.hh:16-94).  Here a rank's region is
the block of bricks its GPU hosts, extended by `overlap + 1` cell layers: inside it the generic host routines (decomp.py,
coarse.py, factor.py -- the same code the one-process setup runs on the whole mesh) give exactly the global result for the
rank's own subdomains, because everything they look at (matrix rows, neighbours' overlap, distance to a subdomain boundary up
to 4 * overlap) lies within the region; neighbouring bricks appear truncated and only lend their index sets and raw
partition-of-unity weights.  tests/test_local_setup.py compares every product with the one-process setup of the same problem.

Unique global DOF id = lexicographic lattice index (x slowest), as in mesh.structured_q1.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

import numpy as np
import scipy.sparse as sp

from . import coarse as coarse_mod
from . import decomp, factor, fem, mesh
from ..partition import RankShard


@dataclass
class LocalProducts:
    """What one rank uploads (api.Context.from_local): its shard of the operator and the products of its own subdomains."""

    shard: RankShard
    n_global: int
    nsub_global: int
    b_own: np.ndarray  # right-hand side at the owned DOFs
    pou: list  # per hosted subdomain
    factors: list
    Z: list
    k_all: np.ndarray | None = None  # coarse vectors per subdomain of the whole job (filled after the exchange of the counts)
    meta: dict = field(default_factory=dict)

    @property
    def nsub(self):
        return len(self.pou)


def _first_cell(b, ncell, pg):
    return -((-b * ncell) // pg)  # cell c belongs to brick (c * pg) // ncell


def brick_ids(ncell, pgrid, group, lo, hi):
    """Global part id (decomp.box_partition numbering) of every cell of the box [lo, hi) of cells, C order."""
    per = [p // g for p, g in zip(pgrid, group)]
    idx = np.meshgrid(*[np.arange(a, b) for a, b in zip(lo, hi)], indexing="ij")
    blk = np.zeros(idx[0].shape, dtype=np.int64)
    loc = np.zeros(idx[0].shape, dtype=np.int64)
    for ax in range(3):
        br = (idx[ax] * pgrid[ax]) // ncell[ax]
        blk = blk * group[ax] + br // per[ax]
        loc = loc * per[ax] + br % per[ax]
    return (blk * int(np.prod(per)) + loc).ravel()


def brick_coloring(pgrid, group):
    """Colouring of all bricks of the job for the batched Galerkin product (ddm_set_coarse_coloring): colour = brick coordinates
    modulo 3, so two bricks of one colour are at least three bricks apart along some axis and no brick neighbours both (bricks
    are coupled through A with their 26 neighbours only: an overlapping subdomain plus one matrix layer stays inside the
    neighbouring bricks as long as overlap + 1 < brick width).  Brick ids follow decomp.box_partition's block-major numbering.
    Returns (color int32[nsub], 27, owner_of int32[nsub, 27])."""
    per = [p // g for p, g in zip(pgrid, group)]
    b = np.stack(np.meshgrid(*[np.arange(p) for p in pgrid], indexing="ij"), -1).reshape(-1, 3)
    blk = np.zeros(b.shape[0], dtype=np.int64)
    loc = np.zeros(b.shape[0], dtype=np.int64)
    for ax in range(3):
        blk = blk * group[ax] + b[:, ax] // per[ax]
        loc = loc * per[ax] + b[:, ax] % per[ax]
    ids = blk * int(np.prod(per)) + loc
    nsub = b.shape[0]
    idgrid = np.empty(tuple(pgrid), dtype=np.int64)
    idgrid[b[:, 0], b[:, 1], b[:, 2]] = ids
    col_of = lambda q: (q[..., 0] % 3) * 9 + (q[..., 1] % 3) * 3 + (q[..., 2] % 3)
    color = np.empty(nsub, dtype=np.int32)
    color[ids] = col_of(b)
    owner = np.full((nsub, 27), -1, dtype=np.int32)
    for d in np.stack(np.meshgrid(*[np.arange(-1, 2)] * 3, indexing="ij"), -1).reshape(-1, 3):
        q = b + d
        ok = ((q >= 0) & (q < np.asarray(pgrid))).all(axis=1)
        owner[ids[ok], col_of(q[ok])] = idgrid[q[ok, 0], q[ok, 1], q[ok, 2]]
    return color, 27, owner


def block_cells(ncell, pgrid, group, rank):
    """Cell box [lo, hi) of the block of bricks rank `rank` hosts."""
    per = [p // g for p, g in zip(pgrid, group)]
    bc = np.unravel_index(rank, group)
    lo = [_first_cell(int(bc[ax]) * per[ax], ncell[ax], pgrid[ax]) for ax in range(3)]
    hi = [_first_cell((int(bc[ax]) + 1) * per[ax], ncell[ax], pgrid[ax]) for ax in range(3)]
    return lo, hi


def poisson_q1_cube_local(ncell, pgrid, group, rank, overlap=2, coarse="geneo", nev=16, shift=1e-3, eig_tol=1e-5, leaf_size=32,
                          pou_kind="distance", shrink=0, kappa_fn=fem.kappa_channels, workers=None, numeric="device",
                          eigensolver="gpu", eig_device=0):
    """Setup products of rank `rank` of prod(group) for the Q1 Poisson problem on ncell[0] x ncell[1] x ncell[2] cells with
    pgrid bricks (problem.poisson_q1_cube on the whole mesh, then partition.shard, gives the same arrays)."""
    import time

    verbose, t0 = bool(os.environ.get("DDM_SETUP_VERBOSE")), [time.time()]

    def tick(what):
        if verbose:
            print(f"[setup r{rank}] {what}: {time.time() - t0[0]:.2f}s", flush=True)
        t0[0] = time.time()

    ncell = tuple(int(c) for c in ((ncell,) * 3 if np.ndim(ncell) == 0 else ncell))
    world = int(np.prod(group))
    halo = overlap + 1
    blo, bhi = block_cells(ncell, pgrid, group, rank)
    lo = [max(0, a - halo) for a in blo]
    hi = [min(n, b + halo) for n, b in zip(ncell, bhi)]
    shape = tuple(b - a for a, b in zip(lo, hi))
    nn = tuple(s + 1 for s in shape)  # local node lattice
    NN = tuple(c + 1 for c in ncell)  # global node lattice
    # ---- local mesh: same coordinates, centroids and coefficient values as the global mesh
    _, cells = mesh.structured_q1(shape)
    axes = [np.linspace(0.0, 1.0, NN[ax])[lo[ax] : hi[ax] + 1] for ax in range(3)]
    g = np.meshgrid(*axes, indexing="ij")
    coords = np.stack([q.ravel() for q in g], 1)
    del g
    gi = np.meshgrid(*[np.arange(lo[ax], hi[ax] + 1, dtype=np.int64) for ax in range(3)], indexing="ij")
    gid = ((gi[0] * NN[1] + gi[1]) * NN[2] + gi[2]).ravel()  # global id of every local node (ascending in local order)
    lat = [q.ravel() for q in gi]
    del gi
    cent = coords[cells[:, 0]]
    for k in range(1, 8):
        cent = cent + coords[cells[:, k]]
    cent = cent / 8
    kap = kappa_fn(cent)
    del cent
    h = 1.0 / ncell[0] if len(set(ncell)) == 1 else tuple(1.0 / c for c in ncell)

    def ke_fn(idx):
        return fem.q1_element_matrices(h, kap[idx])

    nloc = coords.shape[0]
    dirichlet = (coords[:, 0] < 1e-9) | (coords[:, 0] > 1.0 - 1e-9)
    workers = workers or max(1, (os.cpu_count() or 1) // world)
    A_full = fem.assemble_chunked(nloc, cells, ke_fn, chunk=1_000_000, workers=workers)
    x0 = np.where(dirichlet, 1.0 - coords[:, 0], 0.0)
    rhs = A_full @ x0
    A = fem.eliminate_dirichlet(A_full, dirichlet)
    b = np.where(dirichlet, 0.0, rhs)
    tick("mesh + assembly")
    # ---- decomposition inside the region: bricks present here, compact ids in ascending global id (owner = lowest id)
    part_g = brick_ids(ncell, pgrid, group, lo, hi)
    present = np.unique(part_g)
    part = np.searchsorted(present, part_g).astype(np.int32)
    nsub_loc = present.size
    nsub_global = int(np.prod(pgrid))
    per_rank = nsub_global // world
    mine = np.flatnonzero((present >= rank * per_rank) & (present < (rank + 1) * per_rank))
    assert mine.size == per_rank
    novlp, owner = decomp.nonoverlapping_sets(nloc, cells, part, nsub_loc)
    subs = decomp.grow_overlap(A_full, novlp, overlap)
    tick("overlap")
    bmasks = decomp.boundary_masks(A_full, subs)
    raw = decomp.pou_raw(A_full, subs, bmasks, owner, pou_kind, shrink, overlap)
    pou_all = decomp.pou_normalise(nloc, subs, [raw[i] for i in range(nsub_loc)], bmasks, pou_kind)
    shared = np.zeros(nloc, dtype=np.int32)
    for s in subs:
        shared[s] += 1
    tick("pou")
    # ---- shard: owned DOFs, ghosts grouped by owning rank, halo plan (partition.shard on the global products gives the same)
    owner_rank = (present[owner] // per_rank).astype(np.int64)  # bricks are numbered block-major: one block per rank
    own_loc = np.flatnonzero(owner_rank == rank)  # ascending local index = ascending global id
    rows = A[own_loc]
    touched = np.zeros(nloc, dtype=bool)
    for i in mine:
        touched[subs[i]] = True
    touched[rows.indices] = True
    gh_loc = np.flatnonzero(touched & (owner_rank != rank))
    gh_loc = gh_loc[np.lexsort((gid[gh_loc], owner_rank[gh_loc]))]
    l2s = -np.ones(nloc, dtype=np.int64)  # region-local -> shard-local (owned | ghost)
    l2s[own_loc] = np.arange(own_loc.size)
    l2s[gh_loc] = own_loc.size + np.arange(gh_loc.size)
    gr = owner_rank[gh_loc]
    # what every other rank touches = the node box of its block grown by `overlap` (union of its subdomains; the columns of
    # its owned rows lie inside): my owned nodes in that box are what I send, in ascending global id
    send_lists = {}
    for q in range(world):
        if q == rank:
            continue
        qlo, qhi = block_cells(ncell, pgrid, group, q)
        inside = np.ones(own_loc.size, dtype=bool)
        for ax in range(3):
            c = lat[ax][own_loc]
            inside &= (c >= qlo[ax] - overlap) & (c <= qhi[ax] + overlap)
        if inside.any():
            send_lists[q] = own_loc[inside]
    peers = np.array(sorted(set(np.unique(gr).tolist()) | set(send_lists.keys())), dtype=np.int32)
    send_ptr, recv_ptr, send_idx = [0], [0], []
    for p in peers:
        lst = send_lists.get(int(p), np.zeros(0, dtype=np.int64))
        send_idx.append(l2s[lst])
        send_ptr.append(send_ptr[-1] + lst.size)
        recv_ptr.append(recv_ptr[-1] + int((gr == p).sum()))
    sub_ptr = np.concatenate(([0], np.cumsum([subs[i].size for i in mine]))).astype(np.int64)
    sub_dofs = l2s[np.concatenate([subs[i] for i in mine])]
    assert (sub_dofs >= 0).all()
    col = l2s[rows.indices]
    assert (col >= 0).all()
    S = RankShard(rank=rank, world=world, sub_range=(rank * per_rank, (rank + 1) * per_rank), owned=gid[own_loc], ghosts=gid[gh_loc],
                  peers=peers, send_ptr=np.asarray(send_ptr, dtype=np.int64),
                  send_idx=(np.concatenate(send_idx) if send_idx else np.zeros(0)).astype(np.int32),
                  recv_ptr=np.asarray(recv_ptr, dtype=np.int64), rowptr=rows.indptr.astype(np.int64), col=col.astype(np.int32),
                  val=rows.data.astype(np.float64), sub_ptr=sub_ptr, sub_dofs=sub_dofs.astype(np.int32), k_offset=-1)
    tick("shard")
    # ---- own subdomains: matrices and coarse-basis inputs in worker processes, then one symbolic analysis per pattern
    # (congruent bricks share it: their index lists are ordered alike, so their matrices have identical CSR patterns)
    in_mine = np.zeros(nloc, dtype=bool)
    for i in mine:
        in_mine[subs[i]] = True
    cand_all = np.flatnonzero(in_mine[cells[:, 0]])  # cells that can lie inside one of my subdomains
    _L.clear()
    _L.update(A=A, subs=subs, coords=coords, cells=cells, ke_fn=ke_fn, dirichlet=dirichlet, shared=shared, pou=pou_all, nloc=nloc,
              coarse=coarse, eigensolver=eigensolver, nev=nev, shift=shift, eig_tol=eig_tol, cand_all=cand_all)
    todo = [int(i) for i in mine]
    try:
        if workers > 1 and len(todo) > 1:
            import multiprocessing as mp

            with mp.get_context("fork").Pool(min(workers, len(todo))) as pool:
                res = dict(pool.imap_unordered(_local_job, todo))
        else:
            res = dict(_local_job(i) for i in todo)
        tick("subdomain matrices")
        mats = {}  # pattern key -> (matrix, coordinates) of one representative
        for i in todo:
            for M in (res[i]["Ai"], res[i].get("K")):
                if M is not None:
                    mats.setdefault(_pattern_key(M), (M, coords[subs[i]]))
        keys = list(mats)
        _L.update(mats=mats, leaf_size=leaf_size)
        if workers > 1 and len(keys) > 1:
            import multiprocessing as mp

            with mp.get_context("fork").Pool(min(workers, len(keys))) as pool:
                ans = dict(zip(keys, pool.map(_analyse_job, keys)))
        else:
            ans = {k: _analyse_job(k) for k in keys}
    finally:
        _L.clear()
    tick(f"symbolic analysis ({len(keys)} patterns for {len(todo)} subdomains)")

    def fac(M):
        an = ans[_pattern_key(M)]
        return factor.PendingFactor(an, M) if numeric == "device" else factor.numeric_cholesky(an, M.data)

    factors, Zs, geneo_in = [], [], []
    for i in todo:
        r = res[i]
        factors.append(fac(r["Ai"]))
        Zs.append(r.get("Z"))
        geneo_in.append((r["Cm"], fac(r["K"])) if r.get("K") is not None else None)
    if coarse == "geneo" and eigensolver == "gpu":
        from . import geneo_gpu

        batch = [j for j, gin in enumerate(geneo_in) if gin is not None]
        if batch:
            Zb, _ = geneo_gpu.batched_geneo([geneo_in[j][0] for j in batch], [geneo_in[j][1] for j in batch], [pou_all[todo[j]] for j in batch],
                                            [dirichlet[subs[todo[j]]] for j in batch], nev=nev, shift=shift, tol=eig_tol, device=eig_device)
            for j, Zi in zip(batch, Zb):
                Zs[j] = Zi
        tick("GenEO eigensolve (device)")
    meta = dict(problem="poisson-q1", dim=3, ncell=ncell, pgrid=tuple(pgrid), group=tuple(group), overlap=overlap, coarse=coarse, nev=nev,
                n=int(np.prod(NN)), nsub=nsub_global, symmetric=True)
    return LocalProducts(shard=S, n_global=int(np.prod(NN)), nsub_global=nsub_global, b_own=np.ascontiguousarray(b[own_loc]),
                         pou=[pou_all[i] for i in mine], factors=factors, Z=Zs, meta=meta)


_L: dict = {}


def _pattern_key(M):
    return (M.shape[0], M.nnz, hash(M.indptr.tobytes()), hash(M.indices.tobytes()))


def _analyse_job(key):
    M, co = _L["mats"][key]
    return factor.analyse(M, co, _L["leaf_size"], symmetric=True)


def _local_job(i):
    """A_i = R_i A R_i^T and the inputs of subdomain i's coarse basis (worker process; everything is read from _L)."""
    from threadpoolctl import threadpool_limits

    with threadpool_limits(1):
        g = _L
        s = g["subs"][i]
        Ai = g["A"][s][:, s].tocsr()
        Ai.sort_indices()
        out = dict(Ai=Ai)
        dl = g["dirichlet"][s]
        if g["coarse"] == "geneo":
            in_s = np.zeros(g["nloc"], dtype=bool)
            in_s[s] = True
            cand = g["cand_all"][in_s[g["cells"][g["cand_all"], 0]]]
            An, Bn = coarse_mod.neumann_matrices(g["nloc"], g["cells"], _KeView(g["ke_fn"]), s, g["shared"], g["dirichlet"], cand=cand)
            if g["eigensolver"] == "gpu":
                from . import geneo_gpu

                Cm = geneo_gpu.weighted_overlap_matrix(Bn, g["pou"][i])
                if Cm.nnz == 0 or abs(Cm).max() == 0.0:  # no overlap region: no coarse vector from this subdomain
                    out["Z"] = np.zeros((s.size, 0), order="F")
                else:
                    K = (An + g["shift"] * Cm).tocsr()
                    K.sort_indices()
                    out.update(Cm=Cm, K=K)
            else:
                out["Z"] = coarse_mod.geneo_basis(An, Bn, g["pou"][i], dl, nev=g["nev"], shift=g["shift"], tol=g["eig_tol"])[0]
        elif g["coarse"] == "pou":
            out["Z"] = coarse_mod.pou_template_basis(g["coords"][s], g["pou"][i], dl)
        elif g["coarse"] == "none":
            out["Z"] = np.zeros((s.size, 0), order="F")
        else:
            raise ValueError(f"unknown coarse space {g['coarse']!r}")
        return i, out


class _KeView:
    """Element matrices on demand, indexed like an array by coarse.neumann_matrices (Ke[sel])."""

    def __init__(self, ke_fn):
        self.ke_fn = ke_fn

    def __getitem__(self, sel):
        idx = np.flatnonzero(sel) if getattr(sel, "dtype", None) == bool else np.asarray(sel)
        return self.ke_fn(idx)
