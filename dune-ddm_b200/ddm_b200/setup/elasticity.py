"""P1 linear elasticity on simplices (host side, runs once): config C3.

Stands in for PDELab's LinearElasticity local operator as driven by /root/reference/examples/linearelasticity.{cc,hh}:
structured simplex box [0,10] x [0,1] x [0,1.5] (linearelasticity.cc:39-41), Lame parameters from
/root/reference/examples/coefficient.lua (steel bars E = 2e11, nu = 0.3 inside rubber E = 2e7, nu = 0.45), body force
(0, 0, -20000) (linearelasticity.hh:66-72), clamped at x = 0 (coefficient.lua:40-42).

DOF numbering is node-major (DOF 3 i + c = component c of node i), i.e. the memory layout of
Dune::BlockVector<FieldVector<double,3>> and of a BCRSMatrix with 3x3 blocks: the assembled scalar CSR matrix is exactly
the 3x3 block-CSR operator the GPU SpMV streams (ddm_upload_operator_bcsr3).
"""
from __future__ import annotations

import numpy as np


def lame_bars(cent):
    """lambda, mu per cell centre (coefficient.lua:20-71)."""
    x, y, z = cent[:, 0], cent[:, 1], cent[:, 2]
    steel = np.zeros(cent.shape[0], dtype=bool)
    for by in (0.25, 0.75):
        for bz in (0.3, 0.6, 0.9, 1.2):
            steel |= np.sqrt((y - by) ** 2 + (z - bz) ** 2) <= 0.04
    steel &= (x >= 0.0) & (x <= 3.0)
    E = np.where(steel, 2e11, 2e7)
    nu = np.where(steel, 0.3, 0.45)
    return E * nu / (1.0 + nu) / (1.0 - 2.0 * nu), E / 2.0 / (1.0 + nu)


def p1_elasticity_element_matrices(coords, cells, lam, mu):
    """(m, 12, 12) element stiffness of int lambda div u div v + 2 mu eps(u):eps(v), node-major local DOFs."""
    v = coords[cells]
    J = v[:, 1:, :] - v[:, :1, :]
    vol = np.abs(np.linalg.det(J)) / 6.0
    Jinv = np.linalg.inv(J)
    G = np.empty((cells.shape[0], 4, 3))
    G[:, 1:, :] = np.transpose(Jinv, (0, 2, 1))
    G[:, 0, :] = -G[:, 1:, :].sum(axis=1)
    # K[(a,i),(b,j)] = vol (lam G[a,i] G[b,j] + mu G[a,j] G[b,i] + mu delta_ij G[a].G[b])
    K = lam[:, None, None, None, None] * np.einsum("mai,mbj->maibj", G, G)
    K += mu[:, None, None, None, None] * np.einsum("maj,mbi->maibj", G, G)
    gg = np.einsum("mad,mbd->mab", G, G)
    K += (mu[:, None, None] * gg)[:, :, None, :, None] * np.eye(3)[None, None, :, None, :]
    K *= vol[:, None, None, None, None]
    return K.reshape(cells.shape[0], 12, 12), vol


def dof_cells(cells):
    """(m, 4) node table -> (m, 12) DOF table, node-major."""
    return (3 * cells[:, :, None] + np.arange(3)[None, None, :]).reshape(cells.shape[0], -1)


def body_force(n_nodes, cells, vol, f=(0.0, 0.0, -20000.0)):
    """P1 load vector of a constant body force: vol/4 per vertex."""
    b = np.zeros((n_nodes, 3))
    w = np.repeat(vol / 4.0, 4)
    for c in range(3):
        if f[c] != 0.0:
            np.add.at(b[:, c], cells.ravel(), w * f[c])
    return b.ravel()
