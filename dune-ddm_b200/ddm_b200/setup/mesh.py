"""Mesh input for the host-side setup restatement.

The reference reads Gmsh 2.2 ASCII meshes through dune-grid's GmshReader
(/root/reference/examples/poisson.cc:87-139) and refines them uniformly
(`grid->globalRefine`).  Here: a small reader for the same files, red refinement
of simplices, and structured generators for the synthetic configurations.
"""
from __future__ import annotations

import numpy as np


def read_gmsh22(path: str, dim: int):
    """Read nodes and the `dim`-dimensional simplices of a Gmsh 2.2 ASCII file.

    Returns (coords[n, dim] float64, cells[m, dim+1] int64), zero-based.
    """
    with open(path) as fh:
        lines = fh.read().split("\n")
    i = lines.index("$Nodes")
    n = int(lines[i + 1])
    nodes = np.array([ln.split()[1:4] for ln in lines[i + 2 : i + 2 + n]], dtype=np.float64)
    j = lines.index("$Elements")
    m = int(lines[j + 1])
    want = {2: 2, 3: 4}[dim]  # gmsh type: 2 = triangle, 4 = tetrahedron
    cells = []
    for ln in lines[j + 2 : j + 2 + m]:
        t = ln.split()
        if int(t[1]) == want:
            ntags = int(t[2])
            cells.append([int(v) - 1 for v in t[3 + ntags : 3 + ntags + dim + 1]])
    cells = np.asarray(cells, dtype=np.int64)
    used = np.unique(cells)
    remap = -np.ones(n, dtype=np.int64)
    remap[used] = np.arange(used.size)
    return nodes[used, :dim].copy(), remap[cells]


def _edge_midpoints(coords, cells, local_edges):
    """Unique edge midpoints; returns new coords and per-cell midpoint ids."""
    n = coords.shape[0]
    e = np.stack([cells[:, list(le)] for le in local_edges], axis=1)  # (m, ne, 2)
    e.sort(axis=2)
    key = e[..., 0] * n + e[..., 1]
    uniq, inv = np.unique(key.ravel(), return_inverse=True)
    a, b = uniq // n, uniq % n
    mid = 0.5 * (coords[a] + coords[b])
    return np.vstack([coords, mid]), n + inv.reshape(key.shape)


def refine_simplices(coords, cells):
    """One level of uniform (red) refinement: triangle -> 4, tetrahedron -> 8."""
    d = cells.shape[1] - 1
    if d == 2:
        le = [(0, 1), (1, 2), (0, 2)]
        newc, m = _edge_midpoints(coords, cells, le)
        v0, v1, v2 = cells.T
        m01, m12, m02 = m.T
        out = np.concatenate(
            [
                np.stack([v0, m01, m02], 1),
                np.stack([m01, v1, m12], 1),
                np.stack([m02, m12, v2], 1),
                np.stack([m01, m12, m02], 1),
            ]
        )
        return newc, out
    le = [(0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3)]
    newc, m = _edge_midpoints(coords, cells, le)
    v0, v1, v2, v3 = cells.T
    m01, m02, m03, m12, m13, m23 = m.T
    out = np.concatenate(
        [
            np.stack([v0, m01, m02, m03], 1),
            np.stack([m01, v1, m12, m13], 1),
            np.stack([m02, m12, v2, m23], 1),
            np.stack([m03, m13, m23, v3], 1),
            # inner octahedron split along the m02-m13 diagonal
            np.stack([m01, m02, m03, m13], 1),
            np.stack([m01, m02, m12, m13], 1),
            np.stack([m02, m03, m13, m23], 1),
            np.stack([m02, m12, m13, m23], 1),
        ]
    )
    return newc, out


def structured_simplices(shape, lengths=None):
    """Kuhn triangulation of a structured box grid with `shape` cells per axis."""
    d = len(shape)
    lengths = lengths or (1.0,) * d
    axes = [np.linspace(0.0, lengths[k], shape[k] + 1) for k in range(d)]
    grid = np.meshgrid(*axes, indexing="ij")
    coords = np.stack([g.ravel() for g in grid], 1)
    npts = tuple(s + 1 for s in shape)
    strides = np.array([int(np.prod(npts[k + 1 :])) for k in range(d)], dtype=np.int64)
    cidx = np.meshgrid(*[np.arange(s) for s in shape], indexing="ij")
    base = sum(c.ravel().astype(np.int64) * st for c, st in zip(cidx, strides))
    import itertools

    cells = []
    for perm in itertools.permutations(range(d)):
        verts = [base]
        cur = base
        for ax in perm:
            cur = cur + strides[ax]
            verts.append(cur)
        cells.append(np.stack(verts, 1))
    cells = np.concatenate(cells)
    # fix orientation so volumes are positive (not needed for stiffness, kept tidy)
    return coords, cells


def structured_q1(shape):
    """Structured Q1 grid on the unit cube: nodes and hexahedral cell->node table.

    Node (i,j,k) has id (i*ny + j)*nz + k with nx,ny,nz = shape+1 (x slowest).
    """
    nx, ny, nz = (s + 1 for s in shape)
    ax = [np.linspace(0.0, 1.0, n) for n in (nx, ny, nz)]
    g = np.meshgrid(*ax, indexing="ij")
    coords = np.stack([q.ravel() for q in g], 1)
    ci, cj, ck = np.meshgrid(*[np.arange(s) for s in shape], indexing="ij")
    base = ((ci * ny + cj) * nz + ck).ravel().astype(np.int64)
    offs = [(a * ny + b) * nz + c for a in (0, 1) for b in (0, 1) for c in (0, 1)]
    cells = base[:, None] + np.asarray(offs, dtype=np.int64)[None, :]
    return coords, cells
