"""Q1-DG discretisation of the convection-diffusion problem of config C4 (host side, runs once).

Stands in for PDELab's ConvectionDiffusionDG local operator as the reference uses it in
/root/reference/examples/convectiondiffusiondg.cc:60-75 on a structured YaspGrid with the problem class
/root/reference/examples/convectiondiffusiondg.hh:22-83: kappa = 0.01 (1e5 in [0.3,0.4]^2), b = (1,1), Gaussian source at
(0.2,0.2), Dirichlet g = 0 on the inflow sides x = 0 / y = 0, outflow on x = 1 / y = 1.  Weighted SIPG with upwinding:

  volume    int (kappa grad u - b u) . grad v
  interior  int (b.n) u_upwind [v]  -  int {kappa grad u . n}_w [v]  -  int {kappa grad v . n}_w [u]  +  gamma int [u][v]
  Dirichlet the same with the outside state g, outflow  int (b.n) u v
  gamma = alpha * kappa_harm * p (p + d - 1) / h,  weights w_s = kappa_n / (kappa_s + kappa_n).

DOFs are cell-major, four nodal Q1 functions per cell (corner order (0,0), (1,0), (0,1), (1,1)).  The matrix is
non-symmetric with a symmetric pattern; subdomain matrices need LU (the reference picks UMFPack, convectiondiffusiondg.ini:14).
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp

_GP = np.array([0.5 - 0.5 / np.sqrt(3.0), 0.5 + 0.5 / np.sqrt(3.0)])
_GW = np.array([0.5, 0.5])


def _phi(x, y):
    return np.array([(1 - x) * (1 - y), x * (1 - y), (1 - x) * y, x * y])


def _dphi(x, y):  # (4, 2) reference gradients
    return np.array([[-(1 - y), -(1 - x)], [(1 - y), -x], [-y, (1 - x)], [y, x]])


def _reference_matrices():
    """Volume: K = int grad phi_j . grad phi_i, Cx/Cy = int phi_j d_x/y phi_i, M = int phi_j phi_i (unit cell)."""
    K, Cx, Cy = np.zeros((4, 4)), np.zeros((4, 4)), np.zeros((4, 4))
    for xq, wx in zip(_GP, _GW):
        for yq, wy in zip(_GP, _GW):
            p, g = _phi(xq, yq), _dphi(xq, yq)
            K += wx * wy * g @ g.T
            Cx += wx * wy * np.outer(g[:, 0], p)
            Cy += wx * wy * np.outer(g[:, 1], p)
    return K, Cx, Cy


def _face_matrices(axis):
    """Reference 8x8 matrices of one interior face with normal e_axis from cell s (first four DOFs) to cell n.

    J  = int [phi_j][phi_i],   Us = int phi^s_j [phi_i]  (upwind value taken from s),
    Cs = int (d_n phi^s_j) [phi_i],  Cn = int (d_n phi^n_j) [phi_i]   (test index first).  [v] = v_s - v_n.
    """
    J, Us, Cs, Cn = (np.zeros((8, 8)) for _ in range(4))
    for q, w in zip(_GP, _GW):
        ps = _phi(1.0, q) if axis == 0 else _phi(q, 1.0)
        pn = _phi(0.0, q) if axis == 0 else _phi(q, 0.0)
        gs = (_dphi(1.0, q) if axis == 0 else _dphi(q, 1.0))[:, axis]
        gn = (_dphi(0.0, q) if axis == 0 else _dphi(q, 0.0))[:, axis]
        jump = np.concatenate((ps, -pn))
        J += w * np.outer(jump, jump)
        Us += w * np.outer(jump, np.concatenate((ps, np.zeros(4))))
        Cs += w * np.outer(jump, np.concatenate((gs, np.zeros(4))))
        Cn += w * np.outer(jump, np.concatenate((np.zeros(4), gn)))
    return J, Us, Cs, Cn


def _boundary_matrices(axis, side):
    """4x4 matrices of a boundary face (side 0: low, outward normal -e_axis; side 1: high, +e_axis):
    M = int phi_j phi_i, C = int (d_n phi_j) phi_i with the outward normal derivative."""
    M, C = np.zeros((4, 4)), np.zeros((4, 4))
    sgn = 1.0 if side else -1.0
    for q, w in zip(_GP, _GW):
        xy = (float(side), q) if axis == 0 else (q, float(side))
        p, g = _phi(*xy), sgn * _dphi(*xy)[:, axis]
        M += w * np.outer(p, p)
        C += w * np.outer(p, g)
    return M, C


def kappa_cd(cent):
    k = np.full(cent.shape[0], 0.01)
    m = (cent[:, 0] > 0.3) & (cent[:, 0] < 0.4) & (cent[:, 1] > 0.3) & (cent[:, 1] < 0.4)
    k[m] = 1e5
    return k


def assemble_cd_dg(ncell, b=(1.0, 1.0), alpha=2.0, kappa_fn=kappa_cd):
    """Returns (A csr, rhs, dof_coords (n,2), cell_dofs (m,4), cell_centres (m,2)) for the ncell x ncell grid."""
    N, h = int(ncell), 1.0 / ncell
    m = N * N
    cid = np.arange(m).reshape(N, N)  # cid[ix, iy]
    ix, iy = np.meshgrid(np.arange(N), np.arange(N), indexing="ij")
    cent = np.stack(((ix.ravel() + 0.5) * h, (iy.ravel() + 0.5) * h), 1)
    kap = kappa_fn(cent)
    bx, by = b
    assert bx >= 0 and by >= 0, "upwinding below assumes a flow towards +x, +y"
    K, Cx, Cy = _reference_matrices()
    dofs = 4 * np.arange(m)[:, None] + np.arange(4)[None, :]
    rows, cols, vals = [], [], []

    def add(rd, cd, blocks):
        rows.append(np.repeat(rd[:, :, None], cd.shape[1], 2).ravel())
        cols.append(np.repeat(cd[:, None, :], rd.shape[1], 1).ravel())
        vals.append(blocks.ravel())

    # volume: kappa K - h (bx Cx + by Cy)
    add(dofs, dofs, kap[:, None, None] * K[None] - h * (bx * Cx + by * Cy)[None])
    pfac = alpha * 1 * (1 + 2 - 1)  # alpha p (p + d - 1)
    for axis, bn in ((0, bx), (1, by)):
        J, Us, Cs, Cn = _face_matrices(axis)
        s = (cid[:-1, :] if axis == 0 else cid[:, :-1]).ravel()
        n_ = (cid[1:, :] if axis == 0 else cid[:, 1:]).ravel()
        ks, kn = kap[s], kap[n_]
        ws, wn = kn / (ks + kn), ks / (ks + kn)
        kh = 2.0 * ks * kn / (ks + kn)
        fd = np.concatenate((dofs[s], dofs[n_]), 1)
        blk = (h * bn) * Us[None] - (ws * ks)[:, None, None] * (Cs + Cs.T)[None] - (wn * kn)[:, None, None] * (Cn + Cn.T)[None] \
            + (pfac * kh)[:, None, None] * J[None]
        add(fd, fd, blk)
        # boundary faces: low side = Dirichlet inflow (g = 0), high side = outflow
        Mlo, Clo = _boundary_matrices(axis, 0)
        Mhi, _ = _boundary_matrices(axis, 1)
        lo = (cid[0, :] if axis == 0 else cid[:, 0]).ravel()
        hi = (cid[-1, :] if axis == 0 else cid[:, -1]).ravel()
        # Dirichlet, b.n = -bn < 0: the upwind value is g = 0; diffusion: -kappa (C + C^T) + gamma M, gamma h = pfac kappa
        add(dofs[lo], dofs[lo], kap[lo][:, None, None] * (pfac * Mlo - Clo - Clo.T)[None])
        add(dofs[hi], dofs[hi], np.repeat((h * bn * Mhi)[None], hi.size, 0))
    n = 4 * m
    A = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(n, n)).tocsr()
    A.sum_duplicates()
    A.sort_indices()
    # load vector: Gaussian source (convectiondiffusiondg.hh:56-64), 2x2 Gauss points per cell
    rhs = np.zeros((m, 4))
    for xq, wx in zip(_GP, _GW):
        for yq, wy in zip(_GP, _GW):
            gx, gy = (ix.ravel() + xq) * h, (iy.ravel() + yq) * h
            f = 100.0 * np.exp(-((gx - 0.2) ** 2 + (gy - 0.2) ** 2) / (0.05 * 0.05))
            rhs += (wx * wy * h * h) * f[:, None] * _phi(xq, yq)[None, :]
    corner = np.array([[0, 0], [1, 0], [0, 1], [1, 1]], dtype=np.float64)
    coords = ((np.stack((ix.ravel(), iy.ravel()), 1)[:, None, :] + corner[None]) * h).reshape(n, 2)
    return A, rhs.ravel(), coords, dofs, cent
