"""Host stand-in for the device context used by setup/geneo_gpu.py -- only the calls the batched Lanczos makes (vector
create / download / destroy, ddm_op_apply, ddm_blockdiag_solve, ddm_seg_dot, ddm_seg_axpby) -- so that the recurrence, the
re-orthogonalisation, the restart logic and the Ritz extraction can be exercised without a GPU.  Test infrastructure only: the
product never sees this module."""
import numpy as np


class _Vec:
    def __init__(self, ctx, values=None):
        self.ctx = ctx
        self.a = np.zeros(ctx.n_own) if values is None else np.array(values, dtype=np.float64)
        self.h = self  # the "handle" the fake library receives

    def download(self):
        return self.a.copy()


class _Lib:
    def __init__(self, ctx):
        self.c = ctx
        self.solves = 0

    def ddm_op_apply(self, h, x, y):
        y.a[:] = self.c.A @ x.a
        return 0

    def ddm_blockdiag_solve(self, h, d, x):
        from ddm_b200.setup import factor

        c = self.c
        for i, F in enumerate(c.factors):
            x.a[c.off[i] : c.off[i + 1]] = factor.solve_supernodal(F, d.a[c.off[i] : c.off[i + 1]])
        self.solves += 1
        return 0

    def ddm_seg_dot(self, h, x, y, out):
        c = self.c
        o = np.ctypeslib.as_array(out, shape=(c.ns,))
        o[:] = np.add.reduceat(x.a * y.a, c.off[:-1])
        return 0

    def ddm_seg_axpby(self, h, alpha, x, beta, y):
        c = self.c
        a = np.repeat(np.ctypeslib.as_array(alpha, shape=(c.ns,)), np.diff(c.off))
        b = 1.0 if beta is None else np.repeat(np.ctypeslib.as_array(beta, shape=(c.ns,)), np.diff(c.off))
        y.a[:] = a * x.a + b * y.a
        return 0

    def ddm_vec_destroy(self, h):
        return 0


class HostContext:
    last = None

    def __init__(self, prod):
        self.A, self.factors = prod.A.tocsr(), prod.factors
        sizes = np.array([s.size for s in prod.subs], dtype=np.int64)
        self.off = np.concatenate(([0], np.cumsum(sizes)))
        self.ns, self.n_own = len(prod.subs), int(self.off[-1])
        self.lib, self.h = _Lib(self), self
        HostContext.last = self

    @classmethod
    def from_products(cls, prod, **kw):
        return cls(prod)

    def vector(self, values=None):
        return _Vec(self, values)

    def close(self):
        pass
