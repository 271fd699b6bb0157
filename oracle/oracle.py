"""CPU ORACLE -- TEST INFRASTRUCTURE, NOT THE PRODUCT (see the header of oracle/oracle.c).

ctypes front end of oracle/liboracle.so plus a numpy restatement of the reference's *rank-wise*
formulation (one consistent local vector per virtual rank, halo copies and halo sums), which is
what dune-ddm literally executes.  tests/test_oracle.py checks that both agree, i.e. that the
unique-DOF formulation used by the C oracle and by the GPU path is the reference's algorithm.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np
import scipy.linalg as sla

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    """Compile liboracle.so (and oracle/_ref when the reference tree is present)."""
    subprocess.run(["make", "-s", "-C", _HERE, "liboracle.so"], check=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.oracle_dot.restype = C.c_double
        _LIB.oracle_pcg.restype = C.c_int
        _LIB.oracle_gmres.restype = C.c_int
        _LIB.oracle_max_threads.restype = C.c_int
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class _Setup(C.Structure):
    _fields_ = [
        ("nsub", C.c_int32), ("n", C.c_int64), ("sub_ptr", C.POINTER(C.c_int64)), ("sub_dofs", C.POINTER(C.c_int32)),
        ("pou", C.POINTER(C.c_double)), ("nsuper", C.POINTER(C.c_int32)), ("super", C.POINTER(C.POINTER(C.c_int32))),
        ("pi", C.POINTER(C.POINTER(C.c_int64))), ("px", C.POINTER(C.POINTER(C.c_int64))),
        ("s", C.POINTER(C.POINTER(C.c_int32))), ("x", C.POINTER(C.POINTER(C.c_double))),
        ("perm", C.POINTER(C.POINTER(C.c_int32))), ("z_ptr", C.POINTER(C.c_int64)), ("k_off", C.POINTER(C.c_int32)),
        ("Z", C.POINTER(C.c_double)), ("A0chol", C.POINTER(C.c_double)), ("rowptr", C.POINTER(C.c_int64)),
        ("col", C.POINTER(C.c_int32)), ("val", C.POINTER(C.c_double)),
        ("xu", C.POINTER(C.POINTER(C.c_double))), ("A0lu", C.POINTER(C.c_double)), ("A0piv", C.POINTER(C.c_int32)),
    ]


class Oracle:
    """The reference algorithm on unique-DOF vectors, all subdomains in one process."""

    def __init__(self, P, nthreads=1):
        self.P = P
        self.nthreads = int(nthreads)
        ns = P.nsub
        k = self._keep = {}
        k["sub_ptr"] = np.concatenate(([0], np.cumsum([s.size for s in P.subs]))).astype(np.int64)
        k["sub_dofs"] = np.concatenate(P.subs).astype(np.int32)
        k["pou"] = np.concatenate(P.pou).astype(np.float64)
        k["nsuper"] = np.array([F.nsuper for F in P.factors], dtype=np.int32)
        for name, t in (("super_", C.c_int32), ("pi", C.c_int64), ("px", C.c_int64), ("s", C.c_int32), ("x", C.c_double), ("perm", C.c_int32)):
            arrs = [np.ascontiguousarray(getattr(F, name)) for F in P.factors]
            k["arr_" + name] = arrs
            k["ptr_" + name] = (C.POINTER(t) * ns)(*[_p(a, t) for a in arrs])
        ks = np.array([z.shape[1] for z in P.Z], dtype=np.int64)
        k["k_off"] = np.concatenate(([0], np.cumsum(ks))).astype(np.int32)
        k["z_ptr"] = np.concatenate(([0], np.cumsum([z.size for z in P.Z]))).astype(np.int64)
        k["Z"] = np.concatenate([z.ravel(order="F") for z in P.Z]) if ns else np.zeros(0)
        if k["Z"].size == 0:
            k["Z"] = np.zeros(1)
        N0 = int(k["k_off"][-1])
        self.nonsym = any(hasattr(F, "xu") for F in P.factors)
        if self.nonsym:  # LU factors and a non-symmetric coarse matrix (restricted Schwarz + GMRES configurations)
            xus = [np.ascontiguousarray(F.xu) for F in P.factors]
            k["arr_xu"] = xus
            k["ptr_xu"] = (C.POINTER(C.c_double) * ns)(*[_p(a, C.c_double) for a in xus])
            k["A0chol"] = np.zeros((1, 1))
            if N0:
                lu, piv = sla.lu_factor(P.A0)
                k["A0lu"], k["A0piv"] = np.ascontiguousarray(lu), np.ascontiguousarray(piv, dtype=np.int32)
        else:
            k["A0chol"] = np.ascontiguousarray(np.linalg.cholesky(P.A0)) if N0 else np.zeros((1, 1))
        k["rowptr"] = P.A.indptr.astype(np.int64)
        k["col"] = P.A.indices.astype(np.int32)
        k["val"] = P.A.data.astype(np.float64)
        S = _Setup()
        S.nsub, S.n = ns, P.n
        S.sub_ptr, S.sub_dofs, S.pou = _p(k["sub_ptr"], C.c_int64), _p(k["sub_dofs"], C.c_int32), _p(k["pou"], C.c_double)
        S.nsuper = _p(k["nsuper"], C.c_int32)
        S.super, S.pi, S.px, S.s, S.x, S.perm = (k["ptr_super_"], k["ptr_pi"], k["ptr_px"], k["ptr_s"], k["ptr_x"], k["ptr_perm"])
        S.z_ptr, S.k_off, S.Z, S.A0chol = _p(k["z_ptr"], C.c_int64), _p(k["k_off"], C.c_int32), _p(k["Z"], C.c_double), _p(k["A0chol"], C.c_double)
        S.rowptr, S.col, S.val = _p(k["rowptr"], C.c_int64), _p(k["col"], C.c_int32), _p(k["val"], C.c_double)
        if self.nonsym:
            S.xu = k["ptr_xu"]
            if N0:
                S.A0lu, S.A0piv = _p(k["A0lu"], C.c_double), _p(k["A0piv"], C.c_int32)
        self.S = S
        self.N0 = N0

    # --- operator and scalar product (NonOverlappingOperator / NonOverlappingScalarProduct)
    def spmv(self, x):
        y = np.empty(self.P.n)
        k = self._keep
        lib().oracle_spmv(C.c_int64(self.P.n), _p(k["rowptr"], C.c_int64), _p(k["col"], C.c_int32), _p(k["val"], C.c_double),
                          _p(np.ascontiguousarray(x), C.c_double), _p(y, C.c_double), C.c_int(self.nthreads))
        return y

    def dot(self, x, y):
        return lib().oracle_dot(C.c_int64(x.size), _p(np.ascontiguousarray(x), C.c_double), _p(np.ascontiguousarray(y), C.c_double))

    # --- preconditioners
    def sub_solve(self, i, b):
        F = self.P.factors[i]
        out, wk = np.empty(F.n), np.empty(F.n)
        b = np.ascontiguousarray(b, dtype=np.float64)
        if hasattr(F, "xu"):
            lib().oracle_super_solve_lu(C.c_int32(F.n), C.c_int32(F.nsuper), _p(F.super_, C.c_int32), _p(F.pi, C.c_int64), _p(F.px, C.c_int64),
                                        _p(F.s, C.c_int32), _p(F.x, C.c_double), _p(F.xu, C.c_double), _p(F.perm, C.c_int32),
                                        _p(b, C.c_double), _p(out, C.c_double), _p(wk, C.c_double))
            return out
        lib().oracle_super_solve(C.c_int32(F.n), C.c_int32(F.nsuper), _p(F.super_, C.c_int32), _p(F.pi, C.c_int64), _p(F.px, C.c_int64),
                                 _p(F.s, C.c_int32), _p(F.x, C.c_double), _p(F.perm, C.c_int32), _p(b, C.c_double), _p(out, C.c_double), _p(wk, C.c_double))
        return out

    def schwarz_apply(self, d, restricted=False):
        x = np.empty(self.P.n)
        lib().oracle_schwarz_apply(C.byref(self.S), C.c_int(int(restricted)), _p(np.ascontiguousarray(d), C.c_double), _p(x, C.c_double), C.c_int(self.nthreads))
        return x

    def coarse_apply(self, d):
        x = np.empty(self.P.n)
        lib().oracle_coarse_apply(C.byref(self.S), _p(np.ascontiguousarray(d), C.c_double), _p(x, C.c_double), C.c_int(self.nthreads))
        return x

    def twolevel_apply(self, d, restricted=False, mode="additive", with_coarse=True):
        x = np.empty(self.P.n)
        lib().oracle_twolevel_apply(C.byref(self.S), C.c_int(int(restricted)), C.c_int(0 if mode == "additive" else 1), C.c_int(int(with_coarse)),
                                    _p(np.ascontiguousarray(d), C.c_double), _p(x, C.c_double), C.c_int(self.nthreads))
        return x

    def pcg(self, b, x0=None, reduction=1e-8, maxit=1000, restricted=False, mode="additive", with_coarse=True):
        x = np.zeros(self.P.n) if x0 is None else np.array(x0, dtype=np.float64)
        r = np.array(b, dtype=np.float64)
        hist = np.zeros(maxit + 1)
        conv = C.c_int(0)
        it = lib().oracle_pcg(C.byref(self.S), C.c_int(int(restricted)), C.c_int(0 if mode == "additive" else 1), C.c_int(int(with_coarse)),
                              _p(x, C.c_double), _p(r, C.c_double), C.c_double(reduction), C.c_int(maxit), _p(hist, C.c_double), C.byref(conv), C.c_int(self.nthreads))
        return x, it, bool(conv.value), hist[: it + 1]


    def gmres(self, b, x0=None, reduction=1e-8, restart=30, maxit=1000, restricted=False, mode="additive", with_coarse=True):
        x = np.zeros(self.P.n) if x0 is None else np.array(x0, dtype=np.float64)
        r = np.array(b, dtype=np.float64)
        hist = np.zeros(maxit + 1)
        conv = C.c_int(0)
        it = lib().oracle_gmres(C.byref(self.S), C.c_int(int(restricted)), C.c_int(0 if mode == "additive" else 1), C.c_int(int(with_coarse)),
                                _p(x, C.c_double), _p(r, C.c_double), C.c_double(reduction), C.c_int(restart), C.c_int(maxit),
                                _p(hist, C.c_double), C.byref(conv), C.c_int(self.nthreads))
        return x, it, bool(conv.value), hist[: it + 1]


# ----------------------------------------------------------------------------- rank-wise restatement (numpy)
class RankwiseReference:
    """dune-ddm as written: every virtual rank r keeps consistent local vectors on its
    non-overlapping index list (the first n_r entries of subs[r]) and work vectors on subs[r].

    Halo primitives follow OwnerOverlapCopyCommunication
    (/root/reference/extern/dune-istl/dune/istl/owneroverlapcopy.hh:311-387); on the overlapping
    index set every index is owner or copy (/root/reference/dune/ddm/overlap_extension.hh:261).
    """

    def __init__(self, P, n_novlp):
        self.P, self.nn = P, n_novlp
        self.solve = [None] * P.nsub
        for i, s in enumerate(P.subs):
            Ai = P.A[s][:, s].toarray()
            self.solve[i] = sla.cho_factor(Ai)
        self.A0f = sla.cho_factor(P.A0) if P.A0.size else None

    def to_ranks(self, v):
        return [v[s[:m]].copy() for s, m in zip(self.P.subs, self.nn)]

    def from_ranks(self, vr):
        v = np.zeros(self.P.n)
        for r, (s, m) in enumerate(zip(self.P.subs, self.nn)):
            own = self.P.owner[s[:m]] == r
            v[s[:m][own]] = vr[r][own]
        return v

    def copy_owner_to_all(self, ovl):
        g = np.zeros(self.P.n)
        for r, s in enumerate(self.P.subs):
            own = self.P.owner[s] == r
            g[s[own]] = ovl[r][own]
        return [g[s].copy() for s in self.P.subs]

    def add_all(self, ovl):
        g = np.zeros(self.P.n)
        for r, s in enumerate(self.P.subs):
            np.add.at(g, s, ovl[r])
        return [g[s].copy() for s in self.P.subs]

    def schwarz(self, dr, restricted):
        d_ovlp = []
        for r, s in enumerate(self.P.subs):
            v = np.zeros(s.size)
            v[: self.nn[r]] = dr[r]
            d_ovlp.append(v)
        d_ovlp = self.copy_owner_to_all(d_ovlp)
        x_ovlp = [sla.cho_solve(self.solve[r], d_ovlp[r]) for r in range(self.P.nsub)]
        if restricted:
            x_ovlp = [x * w for x, w in zip(x_ovlp, self.P.pou)]
        x_ovlp = self.add_all(x_ovlp)
        return [x[:m].copy() for x, m in zip(x_ovlp, self.nn)]

    def coarse(self, dr):
        d_ovlp = []
        for r, s in enumerate(self.P.subs):
            v = np.zeros(s.size)
            v[: self.nn[r]] = dr[r]
            d_ovlp.append(v)
        d_ovlp = self.copy_owner_to_all(d_ovlp)
        d0 = np.concatenate([Z.T @ d for Z, d in zip(self.P.Z, d_ovlp)])  # gatherv
        x0 = sla.cho_solve(self.A0f, d0)
        off = np.concatenate(([0], np.cumsum([Z.shape[1] for Z in self.P.Z])))
        x_ovlp = [Z @ x0[off[r] : off[r + 1]] for r, Z in enumerate(self.P.Z)]  # scatterv + prolongate
        x_ovlp = self.add_all(x_ovlp)
        return [x[:m].copy() for x, m in zip(x_ovlp, self.nn)]

    def twolevel(self, d, restricted=False):
        dr = self.to_ranks(d)
        x1 = self.schwarz(dr, restricted)
        x2 = self.coarse(dr)
        return self.from_ranks([a + b for a, b in zip(x1, x2)])
