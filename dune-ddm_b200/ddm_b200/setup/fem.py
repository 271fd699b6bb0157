"""Finite-element assembly for the setup restatement (host side, runs once).

Stands in for the PDELab assembly the reference does in
/root/reference/examples/poisson.hh:279-319 (P1/Q1 ConvectionDiffusionFEM with a
scalar permeability) and for the Dirichlet elimination of
/root/reference/examples/pdelab_helper.hh:34-46 (Dirichlet rows become identity,
Dirichlet columns are zeroed).
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp


# ----------------------------------------------------------------------------- coefficients
def kappa_islands(x):
    """Permeability of /root/reference/examples/poisson_coefficient.lua:9-36.

    `x` is (m, d); for d = 3 the 2-D pattern is evaluated on (x, y) and the
    checkerboard also alternates in z.
    """
    X, Y = np.ascontiguousarray(x[:, 0]), np.ascontiguousarray(x[:, 1])
    k = np.ones(x.shape[0])
    m = (X > 0.3) & (X < 0.9) & (Y > 0.6 - (X - 0.3) / 6) & (Y < 0.8 - (X - 0.3) / 6)
    k[m] = 1e5 * (X[m] + Y[m]) * 10.0
    m = (X > 0.1) & (X < 0.5) & (Y > 0.1 + X) & (Y < 0.25 + X)
    k[m] = 1e5 * (1.0 + 7.0 * Y[m])
    m = (X > 0.5) & (X < 0.9) & (Y > 0.15 - (X - 0.5) * 0.25) & (Y < 0.35 - (X - 0.5) * 0.25)
    k[m] = 1e5 * 2.5
    ix = np.floor(15.0 * X).astype(np.int64)
    iy = np.floor(15.0 * Y).astype(np.int64)
    m = (ix % 2 == 0) & (iy % 2 == 0)
    if x.shape[1] == 3:
        iz = np.floor(15.0 * x[:, 2]).astype(np.int64)
        m &= iz % 2 == 0
        k[m] = 1e5 * (1.0 + ix[m] + iy[m] + iz[m])
    else:
        k[m] = 1e5 * (1.0 + ix[m] + iy[m])
    return k


def kappa_channels(x, contrast=1e6):
    """High-permeability beams through the cube (config C5), after the channel
    pattern of /root/reference/examples/poisson_problems.hh:62-95 extruded to 3-D."""
    k = np.ones(x.shape[0])
    fy = np.floor(8.0 * x[:, 1]).astype(np.int64)
    fz = np.floor(8.0 * x[:, 2]).astype(np.int64) if x.shape[1] == 3 else np.zeros_like(fy)
    m = (fy % 2 == 1) & (fz % 2 == 1)
    k[m] = contrast
    return k


# ----------------------------------------------------------------------------- element matrices
def p1_element_matrices(coords, cells, kappa):
    """Stiffness matrices kappa * int grad phi_i . grad phi_j for simplices; (m, d+1, d+1).

    Gradients from the closed-form inverse of the edge matrix (cofactors), component arrays over all elements at once: no
    per-element LAPACK call, no small batched products."""
    d = coords.shape[1]
    m = cells.shape[0]
    if d not in (2, 3):
        raise ValueError("p1_element_matrices: d must be 2 or 3")
    X = [np.ascontiguousarray(coords[:, k]) for k in range(d)]
    V = [[X[k][cells[:, j]] for k in range(d)] for j in range(d + 1)]  # V[j][k]: component k of vertex j
    e = [[V[j][k] - V[0][k] for k in range(d)] for j in range(1, d + 1)]  # edge vectors
    if d == 2:
        (a, b) = e
        det = a[0] * b[1] - a[1] * b[0]
        g = [None, [b[1] / det, -b[0] / det], [-a[1] / det, a[0] / det]]
    else:
        (a, b, c) = e
        cross = lambda u, w: [u[1] * w[2] - u[2] * w[1], u[2] * w[0] - u[0] * w[2], u[0] * w[1] - u[1] * w[0]]
        bc, ca, ab = cross(b, c), cross(c, a), cross(a, b)
        det = a[0] * bc[0] + a[1] * bc[1] + a[2] * bc[2]
        g = [None, [x / det for x in bc], [x / det for x in ca], [x / det for x in ab]]
    g[0] = [-sum(g[j][k] for j in range(1, d + 1)) for k in range(d)]  # gradients sum to zero
    scale = kappa * (np.abs(det) / {2: 2.0, 3: 6.0}[d])
    Ke = np.empty((m, d + 1, d + 1))
    for i in range(d + 1):
        for j in range(i, d + 1):
            v = g[i][0] * g[j][0]
            for k in range(1, d):
                v = v + g[i][k] * g[j][k]
            v *= scale
            Ke[:, i, j] = v
            if j != i:
                Ke[:, j, i] = v
    return Ke


def _q1_reference_matrix():
    """8x8 Laplace stiffness of the unit cube (2-point Gauss), node order as structured_q1."""
    gp = np.array([0.5 - 0.5 / np.sqrt(3.0), 0.5 + 0.5 / np.sqrt(3.0)])
    K = np.zeros((8, 8))
    corners = [(a, b, c) for a in (0, 1) for b in (0, 1) for c in (0, 1)]
    for qx in gp:
        for qy in gp:
            for qz in gp:
                g = np.zeros((8, 3))
                for n, (a, b, c) in enumerate(corners):
                    fx, dfx = (qx, 1.0) if a else (1 - qx, -1.0)
                    fy, dfy = (qy, 1.0) if b else (1 - qy, -1.0)
                    fz, dfz = (qz, 1.0) if c else (1 - qz, -1.0)
                    g[n] = (dfx * fy * fz, fx * dfy * fz, fx * fy * dfz)
                K += g @ g.T / 8.0
    return K


def _q1_directional_matrices():
    """8x8 matrices int d_a phi_i d_a phi_j on the unit cube for a = x, y, z (2-point Gauss), node order as structured_q1."""
    gp = np.array([0.5 - 0.5 / np.sqrt(3.0), 0.5 + 0.5 / np.sqrt(3.0)])
    K = np.zeros((3, 8, 8))
    corners = [(a, b, c) for a in (0, 1) for b in (0, 1) for c in (0, 1)]
    for qx in gp:
        for qy in gp:
            for qz in gp:
                g = np.zeros((8, 3))
                for n, (a, b, c) in enumerate(corners):
                    fx, dfx = (qx, 1.0) if a else (1 - qx, -1.0)
                    fy, dfy = (qy, 1.0) if b else (1 - qy, -1.0)
                    fz, dfz = (qz, 1.0) if c else (1 - qz, -1.0)
                    g[n] = (dfx * fy * fz, fx * dfy * fz, fx * fy * dfz)
                for ax in range(3):
                    K[ax] += np.outer(g[:, ax], g[:, ax]) / 8.0
    return K


def q1_element_matrices(h, kappa):
    """Q1 stiffness on bricks; (m, 8, 8).  h scalar: cubes of edge h, kappa * h * K_ref.  h = (hx, hy, hz): bricks,
    kappa * (hy hz / hx Kx + hx hz / hy Ky + hx hy / hz Kz) (the refined unit cube of the multi-GPU family is split unevenly
    over the axes for 2 and 4 GPUs)."""
    if np.ndim(h) == 0:
        return (kappa * h)[:, None, None] * _q1_reference_matrix()[None]
    hx, hy, hz = (float(v) for v in h)
    Kd = _q1_directional_matrices()
    Kb = (hy * hz / hx) * Kd[0] + (hx * hz / hy) * Kd[1] + (hx * hy / hz) * Kd[2]
    return kappa[:, None, None] * Kb[None]


def assemble(n, cells, Ke, subset=None):
    """Sum element matrices into an (n, n) CSR matrix (sorted columns, duplicates summed)."""
    if subset is not None:
        cells, Ke = cells[subset], Ke[subset]
    k = cells.shape[1]
    rows = np.repeat(cells, k, axis=1).ravel()
    cols = np.tile(cells, (1, k)).ravel()
    A = sp.coo_matrix((Ke.ravel(), (rows, cols)), shape=(n, n)).tocsr()
    A.sum_duplicates()
    A.sort_indices()
    return A


_ASM: dict = {}


def _assemble_chunk(lo):
    n, cells, ke_fn, chunk = _ASM["n"], _ASM["cells"], _ASM["ke_fn"], _ASM["chunk"]
    idx = np.arange(lo, min(lo + chunk, cells.shape[0]))
    return assemble(n, cells[idx], ke_fn(idx))


def assemble_chunked(n, cells, ke_fn, chunk=4_000_000, workers=1):
    """assemble() over chunks of cells, element matrices produced on the fly by ke_fn(index array): bounds the transient
    memory of large meshes (several ranks build the same global matrix side by side).  workers > 1: the chunks are assembled
    by forked processes and summed in chunk order, so the result does not depend on the worker count."""
    starts = list(range(0, cells.shape[0], chunk))
    _ASM.update(n=n, cells=cells, ke_fn=ke_fn, chunk=chunk)
    try:
        if workers > 1 and len(starts) > 1:
            import multiprocessing as mp

            with mp.get_context("fork").Pool(min(workers, len(starts))) as pool:
                parts = pool.imap(_assemble_chunk, starts)
                A = None
                for part in parts:
                    A = part if A is None else A + part
        else:
            A = None
            for lo in starts:
                part = _assemble_chunk(lo)
                A = part if A is None else A + part
    finally:
        _ASM.clear()
    if A is None:
        A = sp.csr_matrix((n, n))
    A = A.tocsr()
    A.sum_duplicates()
    A.sort_indices()
    return A


def eliminate_dirichlet(A, dirichlet):
    """Identity rows / zero columns at constrained DOFs; pattern is kept.

    Follows /root/reference/examples/pdelab_helper.hh:34-46.
    """
    A = A.tocsr().copy()
    free = (~dirichlet).astype(np.float64)
    Dfree = sp.diags(free)
    A = (Dfree @ A @ Dfree + sp.diags(dirichlet.astype(np.float64))).tocsr()
    A.sort_indices()
    return A
