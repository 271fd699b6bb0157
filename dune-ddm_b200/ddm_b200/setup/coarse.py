"""Coarse space on the host (setup, runs once): GenEO basis and Galerkin matrix.

GenEO follows /root/reference/dune/ddm/coarsespaces/coarse_spaces.hh:319-331:
    A_neu v = lambda (D B_neu D) v,   smallest eigenvalues by shift-invert,
then finalize_eigenvectors (:53-61: v <- D v, normalised to unit 2-norm) and zeroing at
Dirichlet DOFs (/root/reference/examples/poisson.cc:235-238).  The reference drives
Spectra's SymGEigsShiftSolver (/root/reference/dune/ddm/eigensolvers/spectra.hh:112-215,
nev 16, ncv 2 nev, shift 1e-3, tolerance 1e-5, eigensolver_params.hh:43-51); Spectra needs
Eigen, which is absent, so ARPACK's generalised shift-invert Lanczos (scipy eigsh, mode 3)
plays that role.  Eigenvector signs/rotations are solver dependent; span(Z) is what
matters.

The Galerkin matrix is A0 = Z^T A Z in the reference's block ordering (offsets = exclusive
scan of the per-rank basis sizes, /root/reference/dune/ddm/galerkin_preconditioner.hh:246-256).
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


def neumann_matrices(n, cells, Ke, sub, shared_count, dirichlet, cand=None):
    """A_neu (elements inside the subdomain) and B_neu (elements inside its overlap region),
    both in the subdomain's local numbering, Dirichlet rows/cols eliminated.  cand: indices of the cells worth looking at
    (a superset of the cells inside the subdomain; default all) -- Ke is indexed with cell indices or a mask over all cells."""
    loc = -np.ones(n, dtype=np.int64)
    loc[sub] = np.arange(sub.size)
    if cand is not None:
        cells = cells[cand]
    lc = loc[cells]
    inside = (lc >= 0).all(axis=1)
    ovl_node = np.zeros(n, dtype=bool)
    ovl_node[sub] = shared_count[sub] >= 2
    in_ovl = inside & ovl_node[cells].all(axis=1)
    k = cells.shape[1]

    def asm(sel):
        c = lc[sel]
        rows = np.repeat(c, k, axis=1).ravel()
        cols = np.tile(c, (1, k)).ravel()
        M = sp.coo_matrix((Ke[sel if cand is None else cand[sel]].ravel(), (rows, cols)), shape=(sub.size, sub.size)).tocsr()
        M.sum_duplicates()
        return M

    A_neu, B_neu = asm(inside), asm(in_ovl)
    dl = dirichlet[sub]
    free = sp.diags((~dl).astype(np.float64))
    A_neu = (free @ A_neu @ free + sp.diags(dl.astype(np.float64))).tocsr()
    B_neu = (free @ B_neu @ free).tocsr()
    return A_neu, B_neu


def threshold_count(ev, nev, nev_max, threshold):
    """How many of the ascending eigenvalues `ev` the reference's retry loop ends up keeping (spectra.hh:118-163): it solves
    for nev pairs, doubles nev while the largest one is still below the threshold and nev < nev_max, then keeps the pairs
    below the threshold among the first nconv - 1 (at least one).  Returns (pairs computed in the last round, pairs kept)."""
    n = int(nev)
    while True:
        nconv = min(n, len(ev))
        if nconv == len(ev) or ev[nconv - 1] >= threshold or n >= nev_max:
            break
        n *= 2
    cnt = 0
    while cnt < nconv - 1 and ev[cnt] < threshold:
        cnt += 1
    return nconv, max(cnt, 1)


def geneo_basis(A_neu, B_neu, pou, dirichlet_local, nev=16, shift=1e-3, tol=1e-5, seed=1, threshold=-0.5, nev_max=None):
    """Columns: pou-scaled, normalised eigenvectors for the nev smallest eigenvalues -- or, with a positive threshold, for the
    eigenvalues below it (nev doubled until the spectrum is covered or nev_max = 2 nev is reached, eigensolver_params.hh:20-30)."""
    n = A_neu.shape[0]
    D = sp.diags(pou)
    C = (D @ B_neu @ D).tocsc()
    rng = np.random.default_rng(seed)
    v0 = rng.standard_normal(n)
    # Shift to the left of the spectrum (all eigenvalues are >= 0): A + shift*C is SPD-like
    # and the eigenvalues nearest the shift are the smallest ones the reference asks for.
    if C.nnz == 0 or abs(C).max() == 0.0:
        # no overlap region (a single subdomain): the eigenproblem has no finite eigenvalue, the subdomain adds no coarse vector
        return np.zeros((n, 0), order="F"), np.zeros(0)
    K = (A_neu + shift * C).tocsc()
    lu = spla.splu(K, permc_spec="MMD_AT_PLUS_A", diag_pivot_thresh=0.0, options=dict(SymmetricMode=True))
    OPinv = spla.LinearOperator((n, n), matvec=lu.solve, dtype=np.float64)
    Ac = A_neu.tocsc()
    nev_max = 2 * nev if nev_max is None else nev_max
    want = nev
    while True:
        k = min(want, n - 2)
        ncv = min(max(2 * k, k + 2), n - 1)
        try:
            w, V = spla.eigsh(Ac, k=k, M=C, sigma=-shift, which="LM", OPinv=OPinv, tol=tol, ncv=ncv, v0=v0, maxiter=1000)
        except spla.ArpackNoConvergence as e:  # keep what converged, as the reference's retry loop would
            w, V = e.eigenvalues, e.eigenvectors
        order = np.argsort(w)
        w, V = w[order], V[:, order]
        if threshold <= 0 or w.size == 0 or w[-1] >= threshold or want >= nev_max or k < want:
            break
        want *= 2  # spectra.hh:194-197: not enough eigenvalues below the threshold yet
    if threshold > 0 and w.size:
        keep_n = threshold_count(w, w.size, w.size, threshold)[1]
        w, V = w[:keep_n], V[:, :keep_n]
    V = V * pou[:, None]
    V[dirichlet_local, :] = 0.0
    nrm = np.linalg.norm(V, axis=0)
    keep = nrm > 1e-14
    V = V[:, keep] / nrm[keep]
    return np.asfortranarray(V), w[keep]


def pou_template_basis(coords_local, pou, dirichlet_local, degree=1):
    """POU coarse space: pou * {1, x, y, (z), xy...} as in
    /root/reference/dune/ddm/twolevel_schwarz.hh:47-51,107 (1, x, y, xy)."""
    d = coords_local.shape[1]
    cols = [np.ones(coords_local.shape[0])]
    if degree >= 1:
        for k in range(d):
            cols.append(coords_local[:, k])
        cols.append(np.prod(coords_local[:, :2], axis=1))
    V = np.stack(cols, 1) * pou[:, None]
    V[dirichlet_local, :] = 0.0
    nrm = np.linalg.norm(V, axis=0)
    V = V[:, nrm > 1e-14] / nrm[nrm > 1e-14]
    return np.asfortranarray(V)


def galerkin_matrix(A, subs, bases):
    """Dense A0 = Z^T A Z, Z = [R_0^T Z_0 | R_1^T Z_1 | ...] and the column offsets."""
    n = A.shape[0]
    ks = np.array([b.shape[1] for b in bases], dtype=np.int64)
    off = np.concatenate(([0], np.cumsum(ks)))
    rows = np.concatenate([np.repeat(s[:, None], b.shape[1], 1).ravel() for s, b in zip(subs, bases)])
    cols = np.concatenate(
        [np.tile(np.arange(o, o + b.shape[1]), (s.size, 1)).ravel() for s, b, o in zip(subs, bases, off[:-1])]
    )
    vals = np.concatenate([np.ascontiguousarray(b).ravel() for b in bases])
    Z = sp.coo_matrix((vals, (rows, cols)), shape=(n, int(off[-1]))).tocsr()
    A0 = (Z.T @ (A @ Z)).toarray()
    return A0, off, Z


def galerkin_columns(A, subs, bases, mine):
    """The columns of A0 = Z^T A Z that belong to the subdomains `mine` (a contiguous block), computed block by block from
    the neighbours only -- block (i, j) = Z_i^T (A R_j^T Z_j) restricted to the rows both touch -- the way the reference's
    build_solver exchanges A Z_j with the neighbouring ranks (galerkin_preconditioner.hh:258-328).  Never forms the global
    n x N0 basis matrix."""
    n = A.shape[0]
    ks = np.array([b.shape[1] for b in bases], dtype=np.int64)
    off = np.concatenate(([0], np.cumsum(ks)))
    N0 = int(off[-1])
    mine = list(mine)
    if not mine:
        return np.zeros((N0, 0))
    c0, c1 = int(off[min(mine)]), int(off[max(mine) + 1])
    out = np.zeros((N0, c1 - c0))
    A = A.tocsr()
    At = A.T.tocsr()
    memb = sp.csr_matrix((np.ones(sum(s.size for s in subs), dtype=np.int8), (np.concatenate(subs), np.repeat(np.arange(len(subs)), [s.size for s in subs]))),
                         shape=(n, len(subs)))
    pos = np.full(n, -1, dtype=np.int64)
    for j in mine:
        if ks[j] == 0:
            continue
        sj = subs[j]
        R = np.unique(At[sj].indices)  # rows of A with an entry in the columns s_j
        Y = A[R][:, sj] @ bases[j]  # (|R|, k_j): the non-zero rows of A R_j^T Z_j
        pos[R] = np.arange(R.size)
        for i in np.unique(memb[R].indices):
            if ks[i] == 0:
                continue
            idx = pos[subs[i]]
            m = idx >= 0
            if m.any():
                out[off[i] : off[i + 1], off[j] - c0 : off[j + 1] - c0] = bases[i][m].T @ Y[idx[m]]
        pos[R] = -1
    return out


def galerkin_matrix_operands(subs, bases, n):
    ks = np.array([b.shape[1] for b in bases], dtype=np.int64)
    off = np.concatenate(([0], np.cumsum(ks)))
    rows = np.concatenate([np.repeat(s[:, None], b.shape[1], 1).ravel() for s, b in zip(subs, bases)])
    cols = np.concatenate(
        [np.tile(np.arange(o, o + b.shape[1]), (s.size, 1)).ravel() for s, b, o in zip(subs, bases, off[:-1])]
    )
    vals = np.concatenate([np.ascontiguousarray(b).ravel() for b in bases])
    return ks, off, sp.coo_matrix((vals, (rows, cols)), shape=(n, int(off[-1]))).tocsc()


def coupling_coloring(A, subs):
    """Colouring for the batched Galerkin product on the device (ddm_set_coarse_coloring): subdomains m and i are coupled when
    A has an entry between their index sets (then block (m, i) of A0 = Z^T A Z can be non-zero -- the neighbour structure the
    reference's build_solver exchanges, galerkin_preconditioner.hh:258-328); two subdomains may share a colour only if no
    subdomain is coupled to both.  Greedy colouring of that "distance two" graph in subdomain order.
    Returns (color int32[nsub], ncolors, owner_of int32[nsub, ncolors]: the subdomain of colour c coupled to m, or -1)."""
    ns = len(subs)
    n = A.shape[0]
    memb = sp.csr_matrix((np.ones(sum(s.size for s in subs), dtype=np.float64), (np.concatenate(subs), np.repeat(np.arange(ns), [s.size for s in subs]))),
                         shape=(n, ns))
    pat = sp.csr_matrix((np.ones(A.nnz), A.indices, A.indptr), shape=A.shape)
    C = ((memb.T @ (pat @ memb)) > 0).tocsr()  # coupled, self included
    C = ((C + C.T) > 0).tocsr()
    G2 = ((C.T @ C) > 0).tocsr()
    color = np.full(ns, -1, dtype=np.int32)
    for i in range(ns):
        used = set(int(c) for c in color[G2.indices[G2.indptr[i] : G2.indptr[i + 1]]] if c >= 0)
        c = 0
        while c in used:
            c += 1
        color[i] = c
    ncolors = int(color.max()) + 1 if ns else 0
    owner = np.full((ns, ncolors), -1, dtype=np.int32)
    for m in range(ns):
        nb = C.indices[C.indptr[m] : C.indptr[m + 1]]
        assert np.unique(color[nb]).size == nb.size  # by construction: at most one coupled subdomain per colour
        owner[m, color[nb]] = nb
    return color, ncolors, owner
