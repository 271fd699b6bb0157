"""GenEO eigenproblems of all hosted subdomains solved together on the GPU.

The reference solves, per rank, A_neu v = lambda (D B_neu D) v for the nev smallest eigenvalues with Spectra's shift-invert
Lanczos (/root/reference/dune/ddm/coarsespaces/coarse_spaces.hh:319-331, eigensolvers/spectra.hh:112-215: shift 1e-3,
tolerance 1e-5, seed 1).  Here the same shift-invert Lanczos recurrence runs for all subdomains in lock step on one
block-diagonal device context: the operator is blockdiag(C_i), C_i = D_i B_neu,i D_i, the "subdomain factors" are the Cholesky
factors of K_i = A_neu,i + shift C_i, so one pass of the triangular-sweep kernels applies every K_i^{-1} at once
(ddm_blockdiag_solve) and the per-subdomain inner products / updates are segmented kernels (ddm_seg_dot, ddm_seg_axpby).
Lanczos vectors are kept C-orthogonal by full re-orthogonalisation; Ritz pairs come from the per-subdomain tridiagonal
matrices on the host.  Eigenvector signs and rotations inside clusters are solver dependent; span(Z_i) is what matters.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import scipy.sparse as sp


def weighted_overlap_matrix(B_neu, pou):
    D = sp.diags(pou)
    Cm = (D @ B_neu @ D).tocsr()
    Cm.sort_indices()
    return Cm


def batched_geneo(Cs, K_factors, pous, dirichlets, nev=16, shift=1e-3, tol=1e-5, seed=1, device=0, m0=None, mmax=160):
    """Cs[i]: C_i (csr), K_factors[i]: supernodal Cholesky factor of A_neu,i + shift C_i.  Returns (Z list, eigenvalue list)."""
    from .. import api
    from ..lib import check
    from .problem import SetupProducts

    ns = len(Cs)
    sizes = np.array([c.shape[0] for c in Cs], dtype=np.int64)
    off = np.concatenate(([0], np.cumsum(sizes)))
    P = int(off[-1])
    Cbig = sp.block_diag(Cs, format="csr")
    Cbig.sort_indices()
    prod = SetupProducts(A=Cbig, b=np.zeros(P), owner=np.repeat(np.arange(ns, dtype=np.int32), sizes),
                         subs=[np.arange(off[i], off[i + 1], dtype=np.int64) for i in range(ns)], pou=[np.ones(int(k)) for k in sizes],
                         factors=K_factors, Z=[np.zeros((int(k), 0), order="F") for k in sizes], A0=np.zeros((0, 0)),
                         n_novlp=sizes.copy(), meta=dict(kind="geneo-eigen"))
    ctx = api.Context.from_products(prod, device=device, with_coarse=False)
    lib, h = ctx.lib, ctx.h
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))

    def seg_dot(x, y):
        out = np.empty(ns)
        check(lib.ddm_seg_dot(h, x.h, y.h, dp(out)))
        return out

    def seg_axpby(alpha, x, beta, y):
        a = np.ascontiguousarray(alpha, dtype=np.float64)
        b = None if beta is None else np.ascontiguousarray(beta, dtype=np.float64)
        check(lib.ddm_seg_axpby(h, dp(a), x.h, None if b is None else dp(b), y.h))

    try:
        # start vectors: seed 1 per subdomain (spectra.hh: random initial residual with fixed seed)
        v0 = np.concatenate([np.random.default_rng(seed + 0 * i).standard_normal(int(k)) for i, k in enumerate(sizes)])
        # one application of K^{-1} C purifies the start vector: Lanczos vectors must lie in K^{-1} range(C), a component
        # in null(C) has zero C-norm, is never orthogonalised away and would pollute the Ritz vectors
        r0, t0 = ctx.vector(v0), ctx.vector()
        check(lib.ddm_op_apply(h, r0.h, t0.h))
        V, U = [ctx.vector()], [ctx.vector()]
        check(lib.ddm_blockdiag_solve(h, t0.h, V[0].h))
        check(lib.ddm_op_apply(h, V[0].h, U[0].h))
        b0 = np.sqrt(np.maximum(seg_dot(V[0], U[0]), 1e-300))
        zero = np.zeros(ns)
        seg_axpby(zero, t0, 1.0 / b0, V[0])
        seg_axpby(zero, t0, 1.0 / b0, U[0])
        alphas, betas = [], []  # betas[j] couples v_j and v_{j+1}
        alive = np.ones(ns, dtype=bool)
        m = int(m0 or min(max(3 * nev, nev + 16), mmax))
        w = ctx.vector()
        j = 0
        while True:
            while j < m:
                check(lib.ddm_blockdiag_solve(h, U[j].h, w.h))  # w = K^{-1} C v_j for every subdomain
                a = seg_dot(w, U[j])
                alphas.append(a)
                seg_axpby(-a, V[j], None, w)
                if j > 0:
                    seg_axpby(-betas[j - 1], V[j - 1], None, w)
                for k in range(j + 1):  # full re-orthogonalisation in the C inner product
                    hk = seg_dot(w, U[k])
                    seg_axpby(-hk, V[k], None, w)
                un = ctx.vector()
                check(lib.ddm_op_apply(h, w.h, un.h))
                b2 = seg_dot(w, un)
                b = np.sqrt(np.maximum(b2, 0.0))
                dead = b <= 1e-12 * np.maximum(np.abs(a), 1e-300)  # invariant subspace reached on that subdomain
                alive &= ~dead
                inv = np.where(dead, 0.0, 1.0 / np.where(dead, 1.0, b))
                betas.append(np.where(dead, 0.0, b))
                vn = ctx.vector()
                seg_axpby(inv, w, zero, vn)
                seg_axpby(zero, w, inv, un)
                V.append(vn)
                U.append(un)
                j += 1
            # Ritz pairs per subdomain
            A_ = np.array(alphas)  # (m, ns)
            B_ = np.array(betas)
            thetas, ys, done, late = [], [], True, []
            for i in range(ns):
                T = np.diag(A_[:, i]) + np.diag(B_[: m - 1, i], 1) + np.diag(B_[: m - 1, i], -1)
                th, y = np.linalg.eigh(T)
                order = np.argsort(-th)[:nev]
                order = order[th[order] > 1e-14 * max(th.max(), 1e-300)]  # zero Ritz values: padding of exhausted subdomains
                th, y = th[order], y[:, order]
                res = np.abs(B_[m - 1, i] * y[m - 1, :])
                if (res > tol * np.abs(th)).any():
                    done = False
                    late.append(i)
                thetas.append(th)
                ys.append(y)
            if done or m >= mmax:
                if not done:  # as the host path does with ArpackNoConvergence: keep the Ritz pairs reached, but say so
                    import warnings

                    warnings.warn(f"batched GenEO eigensolve: {len(late)} of {ns} subdomains did not reach tolerance {tol:g} in {m} Lanczos steps "
                                  f"(subdomains {late[:8]}{' ...' if len(late) > 8 else ''})", RuntimeWarning, stacklevel=2)
                break
            m = min(m + 16, mmax)
        Vh = np.stack([V[k].download() for k in range(m)], axis=1)  # (P, m)
    finally:
        for vec in list(locals().get("V", [])) + list(locals().get("U", [])) + [locals().get(k) for k in ("w", "r0", "t0")]:
            if vec is not None and vec.h:  # device vectors must go before their context
                lib.ddm_vec_destroy(vec.h)
                vec.h = None
        ctx.close()
    Z, lams = [], []
    for i in range(ns):
        Vi = Vh[off[i] : off[i + 1]] @ ys[i]
        lam = 1.0 / thetas[i] - shift
        order = np.argsort(lam)
        Vi, lam = Vi[:, order], lam[order]
        Vi = Vi * pous[i][:, None]  # finalize_eigenvectors (coarse_spaces.hh:53-61): v <- D v, unit 2-norm
        Vi[dirichlets[i], :] = 0.0
        nrm = np.linalg.norm(Vi, axis=0)
        keep = nrm > 1e-14
        Z.append(np.asfortranarray(Vi[:, keep] / nrm[keep]))
        lams.append(lam[keep])
    return Z, lams
