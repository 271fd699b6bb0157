"""Dry-run stand-in for the ctypes library object -- TEST INFRASTRUCTURE, never loaded by the product or by a normal test run.

GPU tests written after a round's last GPU call cannot be executed in the CPU container.  To catch mistakes in their *Python
logic* (argument order, attribute names, error texts, history lengths, tolerances between two CPU evaluations) before they reach a
B200, this pytest plugin replaces `ddm_b200.lib.load()` by an object whose entry points are answered by the CPU oracle:

    PYTHONPATH=tests python -m pytest -p dryrun_standin tests/test_zz_gpu_edges.py

It is opt-in through `-p` only, implements just the solve-phase entries those tests use (no uploads: `Context.from_products`
is replaced too), and says nothing about the CUDA path: a test that passes here has only been proof-read.  Nothing under
dune-ddm_b200/ imports this file."""

import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))
sys.path.insert(0, ROOT)


def val(a):
    return a.value if hasattr(a, "value") else a


def arr(p, n):
    return np.ctypeslib.as_array(p, shape=(n,))


class FakeCtx:
    def __init__(self):
        self.P = None
        self.O = None
        self.with_pou = True


class FakeLib:
    def __init__(self):
        self.ctxs, self.vecs, self.next, self.err = {}, {}, 1000, b""

    def fail(self, msg):
        self.err = msg.encode()
        return 1

    def ddm_last_error(self):
        return self.err

    def ddm_version(self):
        return b"dryrun"

    def ddm_ctx_create(self, dev, hp):
        self.next += 8
        hp._obj.value = self.next
        self.ctxs[self.next] = FakeCtx()
        return 0

    def ddm_ctx_destroy(self, h):
        self.ctxs.pop(val(h), None)
        return 0

    def ctx(self, h):
        return self.ctxs[val(h)]

    # vectors
    def ddm_vec_create(self, ch, vp):
        self.next += 8
        vp._obj.value = self.next
        c = self.ctx(ch)
        self.vecs[self.next] = np.zeros(c.P.n if c.P is not None else 0)
        if c.P is None:
            self.vecs[self.next] = None
            self._pending = getattr(self, "_pending", []) + [(self.next, c)]
        return 0

    def v(self, h):
        k = val(h)
        if self.vecs[k] is None:
            for kk, c in self._pending:
                if kk == k:
                    self.vecs[k] = np.zeros(c.P.n)
        return self.vecs[k]

    def ddm_vec_destroy(self, h):
        self.vecs.pop(val(h), None)
        return 0

    def ddm_vec_upload(self, h, p):
        v = self.v(h)
        v[:] = arr(p, v.size)
        return 0

    def ddm_vec_download(self, h, p):
        v = self.v(h)
        arr(p, v.size)[:] = v
        return 0

    def ddm_vec_fill(self, h, a):
        self.v(h)[:] = val(a)
        return 0

    def ddm_vec_copy(self, h, o):
        self.v(h)[:] = self.v(o)
        return 0

    def ddm_vec_axpy(self, h, a, x):
        self.v(h)[:] += val(a) * self.v(x)
        return 0

    def ddm_vec_scale(self, h, a):
        self.v(h)[:] *= val(a)
        return 0

    def ddm_vec_add(self, h, x):
        self.v(h)[:] += self.v(x)
        return 0

    # operator, products
    def ddm_op_apply(self, ch, x, y):
        self.v(y)[:] = self.ctx(ch).P.A @ self.v(x)
        return 0

    def ddm_op_applyscaleadd(self, ch, a, x, y):
        self.v(y)[:] += val(a) * (self.ctx(ch).P.A @ self.v(x))
        return 0

    def ddm_dot(self, ch, x, y, out):
        out._obj.value = float(self.v(x) @ self.v(y))
        return 0

    def ddm_norm(self, ch, x, out):
        out._obj.value = float(np.sqrt(self.v(x) @ self.v(x)))
        return 0

    # preconditioner
    def _restricted(self, c, typ):
        return bool(val(typ)) and c.with_pou

    def ddm_schwarz_apply(self, ch, typ, d, x):
        c = self.ctx(ch)
        self.v(x)[:] = c.O.schwarz_apply(self.v(d).copy(), restricted=self._restricted(c, typ))
        return 0

    def ddm_coarse_apply(self, ch, d, x):
        c = self.ctx(ch)
        self.v(x)[:] = c.O.coarse_apply(self.v(d).copy())
        return 0

    def ddm_twolevel_apply(self, ch, typ, mode, wc, d, x):
        c = self.ctx(ch)
        self.v(x)[:] = c.O.twolevel_apply(self.v(d).copy(), restricted=self._restricted(c, typ), mode="multiplicative" if val(mode) else "additive",
                                          with_coarse=bool(val(wc)))
        return 0

    def _result(self, resp, histp, it, conv, hist, maxit):
        r = resp._obj
        r.iterations, r.converged = int(it), int(conv)
        r.reduction = float(hist[-1] / hist[0]) if hist[0] > 0 else 0.0
        r.conv_rate = r.reduction ** (1.0 / it) if it else 0.0
        r.elapsed, r.def0 = 0.0, float(hist[0])
        h = arr(histp, val(maxit) + 1)
        h[: len(hist)] = hist

    def ddm_pcg_solve(self, ch, b, x, red, maxit, typ, mode, wc, resp, histp):
        c = self.ctx(ch)
        bv, xv = self.v(b), self.v(x)
        r0 = bv - c.P.A @ xv
        if not np.isfinite(np.linalg.norm(r0)):
            return self.fail("CG: defect is infinite or NaN")
        xs, it, conv, hist = c.O.pcg(bv.copy(), x0=xv.copy(), reduction=val(red), maxit=val(maxit), restricted=self._restricted(c, typ),
                                     mode="multiplicative" if val(mode) else "additive", with_coarse=bool(val(wc)))
        xv[:] = xs
        bv[:] = bv - c.P.A @ xs
        self._result(resp, histp, it, conv, np.asarray(hist), maxit)
        return 0

    def ddm_gmres_solve(self, ch, b, x, red, restart, maxit, typ, mode, wc, resp, histp):
        c = self.ctx(ch)
        bv, xv = self.v(b), self.v(x)
        if not np.isfinite(np.linalg.norm(bv - c.P.A @ xv)):
            return self.fail("GMRes: defect is infinite or NaN")
        xs, it, conv, hist = c.O.gmres(bv.copy(), x0=xv.copy(), reduction=val(red), restart=val(restart), maxit=val(maxit),
                                       restricted=self._restricted(c, typ), mode="multiplicative" if val(mode) else "additive", with_coarse=bool(val(wc)))
        xv[:] = xs
        self._result(resp, histp, it, conv, np.asarray(hist), maxit)
        return 0

    # batched block solve, segmented operations
    def ddm_blockdiag_solve(self, ch, d, x):
        c = self.ctx(ch)
        dv, xv = self.v(d), self.v(x)
        xv[:] = 0.0
        for i, s in enumerate(c.P.subs):
            xv[s] = c.O.sub_solve(i, dv[s].copy())
        return 0

    def ddm_seg_dot(self, ch, x, y, out):
        c = self.ctx(ch)
        o = arr(out, c.P.nsub)
        for i, s in enumerate(c.P.subs):
            o[i] = self.v(x)[s] @ self.v(y)[s]
        return 0

    def ddm_seg_axpby(self, ch, alpha, x, beta, y):
        c = self.ctx(ch)
        a = arr(alpha, c.P.nsub).copy()
        b = arr(beta, c.P.nsub).copy() if beta else np.ones(c.P.nsub)
        for i, s in enumerate(c.P.subs):
            self.v(y)[s] = a[i] * self.v(x)[s] + b[i] * self.v(y)[s]
        return 0

    def ddm_sync(self, ch):
        return 0

    def ddm_kernel_launches(self, ch):
        return 0

    def ddm_timers_enable(self, ch, on):
        return 0

    def ddm_timers_reset(self, ch):
        return 0


FAKE = FakeLib()


def install():
    from ddm_b200 import api, lib as L
    from ddm_b200.partition import shard
    from oracle import oracle

    L.load = lambda: FAKE
    api._l.load = L.load

    def from_products(cls, P, device=0, rank=0, world=1, unique_id=None, with_coarse=True, device_galerkin=False, with_pou=True):
        self = cls(device)
        S = shard(P, rank, world)
        self.shard, self.world, self.rank = S, world, rank
        self.n_own = S.n_own
        self.N0 = int(P.A0.shape[0]) if with_coarse else 0
        c = FAKE.ctx(self.h)
        c.P, c.with_pou = P, with_pou
        c.O = oracle.Oracle(P, nthreads=2)
        if device_galerkin:
            self.A0_device = np.array(P.A0)
        return self

    api.Context.from_products = classmethod(from_products)


def pytest_configure(config):
    install()
