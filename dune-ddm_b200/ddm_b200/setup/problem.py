"""Builds the setup products of a two-level Schwarz solve on the host and stores them.

This is the wiring order of the reference drivers
(/root/reference/examples/poisson.cc:184-306 and /root/reference/examples/pdelab_schwarz.hh:36-180):
grid -> assembly -> non-overlapping index set -> overlap growth -> overlapping Dirichlet
matrices -> partition of unity -> subdomain factorisation -> coarse basis -> Galerkin matrix.
Everything here runs once; the products are uploaded through the C ABI (include/ddm_b200.h)
and the solve phase never comes back to this module.
"""
from __future__ import annotations

import multiprocessing as mp
import os
from dataclasses import dataclass, field

import numpy as np
import scipy.sparse as sp

from . import coarse as coarse_mod
from . import decomp, factor, fem, mesh


@dataclass
class SetupProducts:
    """Everything the solve phase needs (the ".ddmpack" content)."""

    A: sp.csr_matrix  # global assembled, Dirichlet-eliminated operator on unique DOFs
    b: np.ndarray  # right-hand side (assembled residual at the Dirichlet lift)
    owner: np.ndarray  # int32[n] owning subdomain (lowest rank that sees the DOF)
    subs: list  # per subdomain: int64 global ids of the overlapping index list (R_i)
    pou: list  # per subdomain: float64 weights D_i
    factors: list  # per subdomain: factor.SupernodalFactor of A_i = R_i A R_i^T
    Z: list  # per subdomain: float64[n_i, k_i] (Fortran order) coarse basis
    A0: np.ndarray  # dense Galerkin matrix Z^T A Z
    n_novlp: np.ndarray | None = None  # per subdomain: the first n_novlp entries of subs[i] are its non-overlapping DOFs
    coords: np.ndarray | None = None
    dirichlet: np.ndarray | None = None
    meta: dict = field(default_factory=dict)

    @property
    def n(self):
        return self.A.shape[0]

    @property
    def nsub(self):
        return len(self.subs)

    def save(self, path):
        d = dict(
            A_indptr=self.A.indptr.astype(np.int64), A_indices=self.A.indices.astype(np.int32), A_data=self.A.data,
            b=self.b, owner=self.owner, A0=self.A0, nsub=np.int64(self.nsub), n_novlp=np.asarray(self.n_novlp, dtype=np.int64),
            coords=self.coords if self.coords is not None else np.zeros((0, 0)),
            dirichlet=self.dirichlet if self.dirichlet is not None else np.zeros(0, dtype=bool),
            meta=np.array(repr(self.meta)),
        )
        for i in range(self.nsub):
            F = self.factors[i]
            d.update({f"sub{i}": self.subs[i], f"pou{i}": self.pou[i], f"Z{i}": self.Z[i],
                      f"F{i}_perm": F.perm, f"F{i}_super": F.super_, f"F{i}_pi": F.pi, f"F{i}_px": F.px,
                      f"F{i}_s": F.s, f"F{i}_x": F.x})
        np.savez(path, **d)

    @staticmethod
    def load(path):
        d = np.load(path, allow_pickle=False)
        n = d["b"].size
        A = sp.csr_matrix((d["A_data"], d["A_indices"], d["A_indptr"]), shape=(n, n))
        ns = int(d["nsub"])
        subs, pou, Z, factors = [], [], [], []
        for i in range(ns):
            subs.append(d[f"sub{i}"]); pou.append(d[f"pou{i}"]); Z.append(np.asfortranarray(d[f"Z{i}"]))
            factors.append(factor.SupernodalFactor(n=subs[-1].size, perm=d[f"F{i}_perm"], super_=d[f"F{i}_super"],
                                                   pi=d[f"F{i}_pi"], px=d[f"F{i}_px"], s=d[f"F{i}_s"], x=d[f"F{i}_x"]))
        import ast
        return SetupProducts(A=A, b=d["b"], owner=d["owner"], subs=subs, pou=pou, factors=factors, Z=Z, A0=d["A0"],
                             n_novlp=d["n_novlp"],
                             coords=d["coords"] if d["coords"].size else None,
                             dirichlet=d["dirichlet"] if d["dirichlet"].size else None,
                             meta=ast.literal_eval(str(d["meta"])))


# ----------------------------------------------------------------------------- per-subdomain work (process pool)
_G: dict = {}


def _sub_job(i):
    from threadpoolctl import threadpool_limits

    with threadpool_limits(1):
        g = _G
        s = g["subs"][i]
        A = g["A"]
        Ai = A[s][:, s].tocsr()
        Ai.sort_indices()
        co = g["coords"][s] if g["coords"] is not None else None
        if g["numeric"] == "device" and g["symmetric"]:  # symbolic analysis here, numeric phase by ddm_factorize_subdomain at upload
            F = factor.PendingFactor(factor.analyse(Ai, co, g["leaf_size"], symmetric=True), Ai)
        else:
            F = factor.cholesky_supernodal(Ai, co, g["leaf_size"]) if g["symmetric"] else factor.lu_supernodal(Ai, co, g["leaf_size"])
        kind = g["coarse"]
        dl = g["dirichlet"][s]
        if kind == "geneo":
            An, Bn = coarse_mod.neumann_matrices(A.shape[0], g["cells"], g["Ke"], s, g["shared"], g["dirichlet"])
            if g["eigensolver"] == "gpu":
                # batched device eigensolve (setup/geneo_gpu.py): this worker only factors K_i = A_neu + shift C_i
                from . import geneo_gpu

                Cm = geneo_gpu.weighted_overlap_matrix(Bn, g["pou"][i])
                if Cm.nnz == 0 or abs(Cm).max() == 0.0:  # no overlap region: no coarse vector from this subdomain
                    return i, F, np.zeros((s.size, 0), order="F")
                K = (An + g["shift"] * Cm).tocsr()
                K.sort_indices()
                FK = (factor.PendingFactor(factor.analyse(K, co, g["leaf_size"], symmetric=True), K) if g["numeric"] == "device"
                      else factor.cholesky_supernodal(K, co, g["leaf_size"]))
                return i, F, (Cm, FK)
            Zi, ev = coarse_mod.geneo_basis(An, Bn, g["pou"][i], dl, nev=g["nev"], shift=g["shift"], tol=g["eig_tol"],
                                            threshold=g["eig_threshold"], nev_max=g["nev_max"])
        elif kind == "pou":
            Zi = coarse_mod.pou_template_basis(g["coords"][s], g["pou"][i], dl)
        elif kind == "none":
            Zi = np.zeros((s.size, 0), order="F")
        else:
            raise ValueError(f"unknown coarse space {kind!r}")
        return i, F, Zi


def build(coords, cells, Ke, dirichlet, rhs, cell_part, nsub, overlap=2, pou_kind="distance", shrink=0,
          coarse="geneo", nev=16, shift=1e-3, eig_tol=1e-5, leaf_size=32, workers=None, meta=None, only=None, gather=None,
          A_full=None, symmetric=True, ke_fn=None, eigensolver="host", eig_device=0, eig_threshold=-0.5, nev_max=None, numeric="host"):
    """Generic pipeline from an element-wise description to SetupProducts.

    only / gather: with several processes (one per GPU) every process factors and builds the coarse basis of its own
    subdomains `only` (the reference does the same: one rank, one subdomain; with ke_fn(cell ids) -> element matrices
    given instead of Ke, element matrices are kept only for the cells that touch those subdomains) and `gather(obj) -> [obj of rank 0, ...]`
    exchanges the bases and the column blocks of the Galerkin matrix (the reference's neighbour exchange and
    gatherMatrixFromRows, galerkin_preconditioner.hh:246-328).  Factors of other ranks' subdomains stay None."""
    import time

    verbose = bool(os.environ.get("DDM_SETUP_VERBOSE"))
    t0 = [time.time()]

    def tick(what):
        if verbose:
            print(f"[setup] {what}: {time.time() - t0[0]:.2f}s", flush=True)
        t0[0] = time.time()

    n = coords.shape[0]
    if A_full is None:  # continuous FE: assemble from the element matrices; DG passes its matrix (cells = DOFs per cell)
        A_full = fem.assemble(n, cells, Ke)
    A = fem.eliminate_dirichlet(A_full, dirichlet)
    b = np.where(dirichlet, 0.0, rhs)
    tick("assemble")
    novlp, owner = decomp.nonoverlapping_sets(n, cells, cell_part, nsub)
    todo = list(range(nsub)) if only is None else list(only)

    def everyone(local):  # {subdomain: value} of this process -> list over all subdomains
        if gather is None:
            return [local[i] for i in range(nsub)]
        full = [None] * nsub
        for part in gather(local):
            for i, v in part.items():
                full[i] = v
        return full

    # overlap growth, boundary masks and raw partition-of-unity weights are per-subdomain work: each process does its own
    subs = everyone(dict(zip(todo, decomp.grow_overlap(A_full, [novlp[i] for i in todo], overlap))))
    tick("overlap")
    bmasks = everyone(dict(zip(todo, decomp.boundary_masks(A_full, [subs[i] for i in todo]))))
    raw = everyone(decomp.pou_raw(A_full, subs, bmasks, owner, pou_kind, shrink, overlap, ids=todo))
    pou = decomp.pou_normalise(n, subs, raw, bmasks, pou_kind)
    tick("pou")
    shared = np.zeros(n, dtype=np.int32)
    for s in subs:
        shared[s] += 1
    if Ke is None and ke_fn is not None and coarse == "geneo":
        # element matrices for the Neumann problems of the subdomains this process handles
        mine = np.zeros(n, dtype=bool)
        for i in todo:
            mine[subs[i]] = True
        keep = np.flatnonzero(mine[cells].all(axis=1))
        cells, Ke = cells[keep], ke_fn(keep)
    _G.clear()
    _G.update(A=A, subs=subs, coords=coords, cells=cells, Ke=Ke, dirichlet=dirichlet, shared=shared, pou=pou,
              coarse=coarse, nev=nev, shift=shift, eig_tol=eig_tol, leaf_size=leaf_size, symmetric=symmetric,
              eigensolver=eigensolver if coarse == "geneo" else "host", eig_threshold=eig_threshold, nev_max=nev_max, numeric=numeric)
    workers = workers or min(max(len(todo), 1), os.cpu_count() or 1)
    factors = [None] * nsub
    Z = [None] * nsub
    if workers > 1 and len(todo) > 1:
        with mp.get_context("fork").Pool(workers) as pool:
            for i, F, Zi in pool.imap_unordered(_sub_job, todo):
                factors[i], Z[i] = F, Zi
    else:
        for i in todo:
            _, factors[i], Z[i] = _sub_job(i)
    _G.clear()
    if coarse == "geneo" and eigensolver == "gpu" and todo:
        tick("factor A_i and K_i")
        from . import geneo_gpu

        batch = [i for i in todo if isinstance(Z[i], tuple)]
        if batch:
            # eigenvalue threshold (eigensolver_params.hh:30,51): the lock-step solve computes the largest count the reference's
            # doubling could reach, the per-subdomain selection then follows its retry loop (spectra.hh:118-163)
            cap, top = nev, (2 * nev if nev_max is None else nev_max)
            while eig_threshold > 0 and cap < top:
                cap *= 2
            Zs, evs = geneo_gpu.batched_geneo([Z[i][0] for i in batch], [Z[i][1] for i in batch], [pou[i] for i in batch],
                                              [dirichlet[subs[i]] for i in batch], nev=cap, shift=shift, tol=eig_tol, device=eig_device)
            for i, Zi, ev in zip(batch, Zs, evs):
                if eig_threshold > 0 and ev.size:
                    Zi = np.asfortranarray(Zi[:, : coarse_mod.threshold_count(ev, nev, top, eig_threshold)[1]])
                Z[i] = Zi
    tick("factor + coarse basis")
    if gather is None:
        A0, off, _ = coarse_mod.galerkin_matrix(A, subs, Z)
    else:
        for part in gather({i: Z[i] for i in todo}):
            for i, Zi in part.items():
                Z[i] = Zi
        # column blocks travel sparse: only the neighbour blocks are non-zero (a few percent of N0 x k at hundreds of subdomains)
        mine_cols = sp.csr_matrix(coarse_mod.galerkin_columns(A, subs, Z, todo))
        A0 = np.hstack([blk.toarray() for blk in gather(mine_cols)])
    tick("galerkin")
    m = dict(meta or {})
    m.update(nsub=nsub, overlap=overlap, pou=pou_kind, coarse=coarse, nev=nev, n=int(n), N0=int(A0.shape[0]), symmetric=bool(symmetric))
    return SetupProducts(A=A, b=b, owner=owner, subs=subs, pou=pou, factors=factors, Z=Z, A0=A0, coords=coords,
                         n_novlp=np.array([v.size for v in novlp], dtype=np.int64),
                         dirichlet=dirichlet, meta=m)


# ----------------------------------------------------------------------------- named configurations
def _assemble_global(n, cells, ke_fn, kw, chunk=4_000_000):
    """The assembled matrix before Dirichlet elimination.  kw["assemble_part"] = (rank, world) together with kw["gather"]: every
    process assembles its slice of the cells, the partial matrices are exchanged and summed in rank order; otherwise one process
    assembles chunk by chunk (element matrices never all at once), the chunks in parallel workers."""
    part_of = kw.pop("assemble_part", None)
    if part_of is not None and kw.get("gather") is not None:
        r, w = part_of
        lo, hi = cells.shape[0] * r // w, cells.shape[0] * (r + 1) // w
        A_loc = fem.assemble_chunked(n, cells[lo:hi], lambda idx: ke_fn(idx + lo), chunk=chunk)
        A_full = None
        for ip, ix, da in kw["gather"]((A_loc.indptr, A_loc.indices, A_loc.data)):
            Ap = sp.csr_matrix((da, ix, ip), shape=(n, n))
            A_full = Ap if A_full is None else A_full + Ap
        A_full.sum_duplicates()
        A_full.sort_indices()
        return A_full
    return fem.assemble_chunked(n, cells, ke_fn, chunk=chunk, workers=kw.get("workers") or (os.cpu_count() or 1))


def poisson_simplex(coords, cells, nsub, kappa_fn=fem.kappa_islands, xmax=1.0, **kw):
    """P1 Poisson with Dirichlet data g = 1 - x/xmax on x in {0, xmax} (poisson_coefficient.lua:1-3,38-44)."""
    import time

    verbose, t0 = bool(os.environ.get("DDM_SETUP_VERBOSE")), [time.time()]

    def tick(what):
        if verbose:
            print(f"[setup] {what}: {time.time() - t0[0]:.2f}s", flush=True)
        t0[0] = time.time()

    d = coords.shape[1]
    cent = coords[cells[:, 0]]  # centroids vertex by vertex: same sums as coords[cells].mean(axis=1) without the (m, d+1, d) temporary
    for k in range(1, cells.shape[1]):
        cent = cent + coords[cells[:, k]]
    cent = cent / cells.shape[1]
    kap = kappa_fn(cent)

    def ke_fn(idx):
        return fem.p1_element_matrices(coords, cells[idx], kap[idx])

    dirichlet = (coords[:, 0] < 1e-6) | (coords[:, 0] > xmax - 1e-6)
    n = coords.shape[0]
    tick("coefficient")
    A_full = _assemble_global(n, cells, ke_fn, kw)
    # residual at the Dirichlet lift x0 (g on the boundary, 0 inside): d = A_full x0 - f, f = 0
    x0 = np.where(dirichlet, 1.0 - coords[:, 0] / xmax, 0.0)
    rhs = A_full @ x0
    tick("assemble A_full")
    part = decomp.rcb_partition(cent, nsub)
    tick("partition")
    meta = dict(kw.pop("meta", {}), problem="poisson-p1", dim=d)
    return build(coords, cells, None, dirichlet, rhs, part, nsub, meta=meta, A_full=A_full, ke_fn=ke_fn, **kw)


def poisson_q1_cube(ncell, pgrid, kappa_fn=fem.kappa_channels, group=None, **kw):
    """Structured Q1 Poisson on the unit cube, brick decomposition (config C5 family).  ncell: cells per axis, or one count per
    axis (the multi-GPU family refines the cube by the GPU grid, so 2 and 4 GPUs work on bricks instead of cubes); group: GPU
    grid for the brick numbering (decomp.box_partition).  Element matrices are produced chunk by chunk (64 doubles per cell:
    6 GB at the 231^3 cells of one C5 GPU share if held at once)."""
    shape = (int(ncell),) * 3 if np.ndim(ncell) == 0 else tuple(int(c) for c in ncell)
    coords, cells = mesh.structured_q1(shape)
    cent = coords[cells[:, 0]]
    for k in range(1, cells.shape[1]):
        cent = cent + coords[cells[:, k]]
    cent = cent / cells.shape[1]
    kap = kappa_fn(cent)
    del cent
    h = 1.0 / shape[0] if len(set(shape)) == 1 else tuple(1.0 / c for c in shape)

    def ke_fn(idx):
        return fem.q1_element_matrices(h, kap[idx])

    n = coords.shape[0]
    dirichlet = (coords[:, 0] < 1e-9) | (coords[:, 0] > 1.0 - 1e-9)
    A_full = _assemble_global(n, cells, ke_fn, kw, chunk=1_000_000)
    x0 = np.where(dirichlet, 1.0 - coords[:, 0], 0.0)
    rhs = A_full @ x0
    part = decomp.box_partition(shape, pgrid, group)
    nsub = int(np.prod(pgrid))
    meta = dict(kw.pop("meta", {}), problem="poisson-q1", dim=3, ncell=shape, pgrid=tuple(pgrid))
    return build(coords, cells, None, dirichlet, rhs, part, nsub, meta=meta, A_full=A_full, ke_fn=ke_fn, **kw)


def unitcube_p1(ncell, nsub, **kw):
    """Config C2 stand-in that needs no mesh file: Kuhn-triangulated unit cube, islands coefficient."""
    coords, cells = mesh.structured_simplices((ncell,) * 3)
    return poisson_simplex(coords, cells, nsub, meta=dict(mesh=f"kuhn-{ncell}^3"), **kw)


def poisson_gmsh(path, dim, nsub, refine=0, **kw):
    """Configs C1 / C2 on the reference's own meshes: Gmsh 2.2 file (data/unitsquare.msh, data/unitcube.msh), `refine` uniform
    refinements (grid->globalRefine, examples/poisson.cc:87-139), islands coefficient, Dirichlet data on x in {0, 1}."""
    coords, cells = mesh.read_gmsh22(path, dim)
    for _ in range(int(refine)):
        coords, cells = mesh.refine_simplices(coords, cells)
    return poisson_simplex(coords, cells, nsub, meta=dict(mesh=f"{os.path.basename(path)} refined {int(refine)}x"), **kw)


def unitsquare_p1(ncell, nsub, **kw):
    coords, cells = mesh.structured_simplices((ncell,) * 2)
    return poisson_simplex(coords, cells, nsub, meta=dict(mesh=f"kuhn-{ncell}^2"), **kw)


def convdiff_dg(ncell, nsub, overlap=1, **kw):
    """Config C4: Q1-DG convection-diffusion on ncell^2 cells (examples/convectiondiffusiondg.{cc,hh,ini}): non-symmetric,
    LU subdomain factors, coarse space = partition of unity times the template vectors {1, x, y, xy}
    (dune/ddm/twolevel_schwarz.hh:47-51,107); meant for restricted Schwarz + multiplicative coarse + GMRES."""
    from . import dg

    A, rhs, coords, cell_dofs, cent = dg.assemble_cd_dg(ncell)
    part = decomp.rcb_partition(cent, nsub)
    dirichlet = np.zeros(A.shape[0], dtype=bool)  # boundary conditions are weak in DG
    meta = dict(kw.pop("meta", {}), problem="convdiff-q1dg", dim=2, ncell=int(ncell))
    kw.setdefault("coarse", "pou")
    return build(coords, cell_dofs, None, dirichlet, rhs, part, nsub, overlap=overlap, meta=meta, A_full=A, symmetric=False, **kw)


def elasticity_box(gridsize=8, nsub=16, overlap=1, nev=12, **kw):
    """Config C3: P1 linear elasticity on the 10 x 1 x 1.5 simplex box with steel bars in rubber
    (examples/linearelasticity.cc:39-41, coefficient.lua), clamped at x = 0, gravity load, GenEO nev 12
    (linearelasticity.ini:27-30).  DOFs are node-major: the operator is a 3x3 block-CSR matrix."""
    from . import elasticity as el

    g = int(gridsize)
    shape = (10 * g, g, int(1.5 * g))
    coords, cells = mesh.structured_simplices(shape, lengths=(10.0, 1.0, 1.5))
    cent = coords[cells].mean(axis=1)
    lam, mu = el.lame_bars(cent)
    Ke, vol = el.p1_elasticity_element_matrices(coords, cells, lam, mu)
    dcells = el.dof_cells(cells)
    dcoords = np.repeat(coords, 3, axis=0)
    dirichlet = np.repeat(coords[:, 0] < 1e-9, 3)
    rhs = el.body_force(coords.shape[0], cells, vol)
    part = decomp.rcb_partition(cent, nsub)
    meta = dict(kw.pop("meta", {}), problem="elasticity-p1", dim=3, block=3, mesh=f"kuhn-{shape}")
    return build(dcoords, dcells, Ke, dirichlet, rhs, part, nsub, overlap=overlap, nev=nev, meta=meta, **kw)
