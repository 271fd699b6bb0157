"""Batched block solve on the device against one-at-a-time host solves: the property the reference checks for its own
multi-right-hand-side solver (/root/reference/dune/ddm/test/test_blocksolver.cc:32-93: block solve equals the single solves to
1e-12), here for ddm_blockdiag_solve -- all subdomain systems A_i x_i = d_i in one pass of the sweep kernels -- and for the
segmented inner products / updates the batched GenEO eigensolve is built from (ddm_seg_dot, ddm_seg_axpby)."""
import ctypes as C

import numpy as np
import pytest
import scipy.sparse as sp

pytestmark = pytest.mark.gpu


def _blockdiag_context(P):
    from ddm_b200 import api
    from ddm_b200.setup.problem import SetupProducts

    mats = []
    for s in P.subs:
        Ai = P.A[s][:, s].tocsr()
        Ai.sort_indices()
        mats.append(Ai)
    sizes = np.array([m.shape[0] for m in mats], dtype=np.int64)
    off = np.concatenate(([0], np.cumsum(sizes)))
    big = sp.block_diag(mats, format="csr")
    big.sort_indices()
    n, ns = int(off[-1]), len(mats)
    prod = SetupProducts(A=big, b=np.zeros(n), owner=np.repeat(np.arange(ns, dtype=np.int32), sizes),
                         subs=[np.arange(off[i], off[i + 1], dtype=np.int64) for i in range(ns)], pou=[np.ones(int(k)) for k in sizes],
                         factors=P.factors, Z=[np.zeros((int(k), 0), order="F") for k in sizes], A0=np.zeros((0, 0)),
                         n_novlp=sizes.copy(), meta=dict(kind="blockdiag"))
    return api.Context.from_products(prod, device=0, with_coarse=False), mats, off


@pytest.mark.parametrize("name", ["small3d", "cd_dg"])
def test_batched_block_solve_equals_single_solves(request, name):
    from ddm_b200.lib import check
    from ddm_b200.setup import factor

    P = request.getfixturevalue(name)
    ctx, mats, off = _blockdiag_context(P)
    vecs = []
    try:
        lib, h = ctx.lib, ctx.h
        rng = np.random.default_rng(5)
        for trial in range(3):
            d = rng.standard_normal(int(off[-1]))
            dv, xv = ctx.vector(d), ctx.vector()
            vecs += [dv, xv]
            check(lib.ddm_blockdiag_solve(h, dv.h, xv.h))
            x = xv.download()
            for i, Ai in enumerate(mats):
                xi, di = x[off[i] : off[i + 1]], d[off[i] : off[i + 1]]
                F = P.factors[i]
                single = factor.solve_supernodal_lu(F, di) if hasattr(F, "xu") else factor.solve_supernodal(F, di)
                # through the residual: the device sweeps multiply by explicit inverses of the diagonal blocks, so their backward
                # error carries the conditioning of those blocks (not of A_i); a wrong solve would be O(1) here
                scale = abs(Ai).sum(axis=1).max() * np.linalg.norm(xi) + np.linalg.norm(di)
                assert np.linalg.norm(Ai @ xi - di) <= 1e-10 * scale
                # ... and directly, with the room conditioning gives two FP64 evaluations of one factorisation
                tol = max(1e-12, 64 * np.finfo(float).eps * np.linalg.cond(Ai.toarray()))
                assert np.linalg.norm(xi - single) <= tol * np.linalg.norm(single)
    finally:
        for v in vecs:
            if v.h:
                ctx.lib.ddm_vec_destroy(v.h)
                v.h = None
        ctx.close()


def test_segmented_dot_and_update(small3d):
    from ddm_b200.lib import check

    P = small3d
    ctx, mats, off = _blockdiag_context(P)
    vecs = []
    try:
        lib, h = ctx.lib, ctx.h
        ns = len(mats)
        dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
        rng = np.random.default_rng(6)
        x, y = rng.standard_normal(int(off[-1])), rng.standard_normal(int(off[-1]))
        xv, yv = ctx.vector(x), ctx.vector(y)
        vecs += [xv, yv]
        out = np.empty(ns)
        check(lib.ddm_seg_dot(h, xv.h, yv.h, dp(out)))
        want = np.array([x[off[i] : off[i + 1]] @ y[off[i] : off[i + 1]] for i in range(ns)])
        bound = np.array([np.abs(x[off[i] : off[i + 1]]) @ np.abs(y[off[i] : off[i + 1]]) for i in range(ns)])
        assert np.all(np.abs(out - want) <= 1e-13 * bound)
        alpha, beta = rng.standard_normal(ns), rng.standard_normal(ns)
        seg = np.repeat(np.arange(ns), np.diff(off))
        check(lib.ddm_seg_axpby(h, dp(alpha), xv.h, dp(beta), yv.h))  # y_i <- alpha_i x_i + beta_i y_i
        assert np.allclose(yv.download(), alpha[seg] * x + beta[seg] * y, rtol=1e-14, atol=1e-14)
        y2 = yv.download()
        check(lib.ddm_seg_axpby(h, dp(alpha), xv.h, None, yv.h))  # beta omitted: y_i <- y_i + alpha_i x_i
        assert np.allclose(yv.download(), y2 + alpha[seg] * x, rtol=1e-14, atol=1e-14)
    finally:
        for v in vecs:
            if v.h:
                ctx.lib.ddm_vec_destroy(v.h)
                v.h = None
        ctx.close()
