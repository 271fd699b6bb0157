"""Host-side mirror of the reference's solve-phase interfaces over the C ABI.

Class names, argument meaning and error behaviour follow the reference:
  NonOverlappingOperator / NonOverlappingScalarProduct  (dune/ddm/nonoverlapping_operator.hh:12-96)
  SchwarzPreconditioner                                  (dune/ddm/schwarz.hh:55-220)
  GalerkinPreconditioner                                 (dune/ddm/galerkin_preconditioner.hh:48-363)
  CombinedPreconditioner                                 (dune/ddm/combined_preconditioner.hh:41-180)
  CGSolver (drop-in loop)                                (extern/dune-istl/dune/istl/solvers.hh:279-340)
All arithmetic runs in libddm_b200.so on the GPU; this module only moves pointers.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import lib as _l
from .lib import DDMError, Result, check
from .partition import shard

STANDARD, RESTRICTED = 0, 1
ADDITIVE, MULTIPLICATIVE = 0, 1


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class DeviceVector:
    """The vector type X the Krylov loops work on (BlockVector<FieldVector<double,1>> on the device)."""

    def __init__(self, ctx, values=None):
        self.ctx = ctx
        self.h = C.c_void_p()
        check(ctx.lib.ddm_vec_create(ctx.h, C.byref(self.h)))
        if values is not None:
            self.upload(values)

    def upload(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        if v.size != self.ctx.n_own:
            raise DDMError(f"vector has {v.size} entries, context owns {self.ctx.n_own}")
        check(self.ctx.lib.ddm_vec_upload(self.h, _p(v, C.c_double)))
        return self

    def download(self):
        out = np.empty(self.ctx.n_own)
        check(self.ctx.lib.ddm_vec_download(self.h, _p(out, C.c_double)))
        return out

    def fill(self, a):
        check(self.ctx.lib.ddm_vec_fill(self.h, C.c_double(a)))
        return self

    def assign(self, other):
        check(self.ctx.lib.ddm_vec_copy(self.h, other.h))
        return self

    def axpy(self, a, x):
        check(self.ctx.lib.ddm_vec_axpy(self.h, C.c_double(a), x.h))
        return self

    def __imul__(self, a):
        check(self.ctx.lib.ddm_vec_scale(self.h, C.c_double(a)))
        return self

    def __iadd__(self, x):
        check(self.ctx.lib.ddm_vec_add(self.h, x.h))
        return self

    def __del__(self):
        try:
            if self.h and self.ctx.h:
                self.ctx.lib.ddm_vec_destroy(self.h)
        except Exception:
            pass


class Context:
    """One GPU's share of a two-level Schwarz solve (one ddm_ctx)."""

    def __init__(self, device=0):
        self.lib = _l.load()
        self.h = C.c_void_p()
        check(self.lib.ddm_ctx_create(C.c_int(device), C.byref(self.h)))
        self.n_own = 0
        self.shard = None
        self.world = 1

    @classmethod
    def from_products(cls, P, device=0, rank=0, world=1, unique_id=None, with_coarse=True, device_galerkin=False, with_pou=True,
                      deterministic=False):
        """Upload the setup products (whole problem, or this rank's shard).  device_galerkin: do not upload P.A0 but
        assemble Z^T A Z on the GPU (ddm_galerkin_assemble); the result is kept in self.A0_device.  with_pou=False uploads no
        partition of unity: restricted Schwarz then applies no weights (schwarz.hh:140, pou == nullptr)."""
        S = shard(P, rank, world)
        a, b = S.sub_range
        N0 = int(P.A0.shape[0])
        coloring = None
        if device_galerkin and with_coarse and N0 > 0:  # batched Galerkin product: colours from the coupling graph of the subdomains
            from .setup import coarse as coarse_mod

            color, ncol, owner = coarse_mod.coupling_coloring(P.A, P.subs)
            coloring = (color, ncol, owner, np.array([z.shape[1] for z in P.Z], dtype=np.int32))
        return cls._from_shard(S, [P.pou[i] for i in range(a, b)], [P.factors[i] for i in range(a, b)], [P.Z[i] for i in range(a, b)],
                               N0, P.A0, device, unique_id, with_coarse, device_galerkin, with_pou, block=P.meta.get("block"),
                               deterministic=deterministic, coloring=coloring)

    @classmethod
    def from_local(cls, L, device=0, unique_id=None, with_coarse=True, with_pou=True, keep_A0=False, deterministic=False):
        """Upload one rank's own products (setup.local.LocalProducts: nothing global exists on any rank).  L.k_all must hold
        the coarse-vector counts of all subdomains of the job; the Galerkin matrix is assembled on the devices."""
        S = L.shard
        k_all = np.asarray(L.k_all, dtype=np.int64)
        S.k_offset = int(k_all[: S.sub_range[0]].sum())
        coloring = None
        if with_coarse and L.meta.get("pgrid") is not None and min(L.meta["ncell"][ax] // L.meta["pgrid"][ax] for ax in range(3)) > 2 * L.meta["overlap"]:
            # (bricks wider than twice the overlap: overlapping subdomains two bricks apart share no interior node, so the
            # Galerkin blocks couple the 26 neighbours only)
            from .setup import local

            color, ncol, owner = local.brick_coloring(L.meta["pgrid"], L.meta["group"])
            coloring = (color, ncol, owner, k_all.astype(np.int32))
        return cls._from_shard(S, L.pou, L.factors, L.Z, int(k_all.sum()), None, device, unique_id, with_coarse, True, with_pou,
                               keep_A0=keep_A0, deterministic=deterministic, coloring=coloring)

    @classmethod
    def _from_shard(cls, S, pou_list, factors, Z_list, N0, A0, device, unique_id, with_coarse, device_galerkin, with_pou, block=None,
                    keep_A0=True, deterministic=False, coloring=None):
        self = cls(device)
        if deterministic:  # reproducible sweeps (no floating-point atomics): schedules are built accordingly by ddm_finalize
            check(self.lib.ddm_set_deterministic(self.h, C.c_int(1)))
        rank, world = S.rank, S.world
        self.shard, self.world, self.rank = S, world, rank
        lib = self.lib
        if world > 1:
            uid = (C.c_uint8 * 128)(*unique_id)
            check(lib.ddm_comm_init(self.h, C.c_int(rank), C.c_int(world), uid))
        if block == 3:  # node-major 3x3 blocks: BCRSMatrix<FieldMatrix<double,3,3>>
            import scipy.sparse as sp

            if S.n_own % 3 or S.n_ghost % 3 or (S.owned[::3] % 3).any() or (S.ghosts[::3] % 3).any():
                raise DDMError("3x3 block operator: owned and ghost DOFs must come in whole nodes")
            B = sp.csr_matrix((S.val, S.col, S.rowptr), shape=(S.n_own, S.n_own + S.n_ghost)).tobsr((3, 3))
            B.sort_indices()
            check(lib.ddm_upload_operator_bcsr3(self.h, C.c_int64(S.n_own // 3), C.c_int64(S.n_ghost // 3),
                                                _p(B.indptr.astype(np.int64), C.c_int64), _p(B.indices.astype(np.int32), C.c_int32),
                                                _p(np.ascontiguousarray(B.data, dtype=np.float64), C.c_double)))
        else:
            check(lib.ddm_upload_operator(self.h, C.c_int64(S.n_own), C.c_int64(S.n_ghost), _p(S.rowptr, C.c_int64),
                                          _p(S.col, C.c_int32), _p(S.val, C.c_double)))
        self.n_own = S.n_own
        if world > 1:
            check(lib.ddm_upload_halo(self.h, C.c_int32(S.peers.size), _p(S.peers, C.c_int32), _p(S.send_ptr, C.c_int64),
                                      _p(S.send_idx, C.c_int32), _p(S.recv_ptr, C.c_int64)))
        ns = len(factors)
        pou = np.concatenate(pou_list).astype(np.float64) if ns else np.zeros(0)
        check(lib.ddm_upload_decomposition(self.h, C.c_int32(ns), _p(S.sub_ptr, C.c_int64), _p(S.sub_dofs, C.c_int32),
                                           _p(pou, C.c_double) if with_pou else None))
        self._upload_factor_list(factors)
        self.N0 = N0 if with_coarse else 0
        if with_coarse and N0 > 0:
            k = np.array([z.shape[1] for z in Z_list], dtype=np.int32)
            Z = np.concatenate([z.ravel(order="F") for z in Z_list]) if ns else np.zeros(1)
            if Z.size == 0:
                Z = np.zeros(1)
            A0c = None if device_galerkin else np.ascontiguousarray(A0, dtype=np.float64)
            check(lib.ddm_upload_coarse(self.h, _p(k, C.c_int32), _p(np.ascontiguousarray(Z), C.c_double), C.c_int32(N0),
                                        C.c_int32(S.k_offset), None if device_galerkin else _p(A0c, C.c_double)))
        check(lib.ddm_finalize(self.h))
        if with_coarse and N0 > 0 and device_galerkin and coloring is not None and not os.environ.get("DDM_NO_COLORING"):
            color, ncol, owner, k_glob = coloring
            check(lib.ddm_set_coarse_coloring(self.h, C.c_int32(color.size), C.c_int32(ncol), _p(np.ascontiguousarray(color, dtype=np.int32), C.c_int32),
                                              _p(np.ascontiguousarray(k_glob, dtype=np.int32), C.c_int32),
                                              _p(np.ascontiguousarray(owner, dtype=np.int32), C.c_int32)))
        if with_coarse and N0 > 0 and device_galerkin:
            self.A0_device = np.empty((N0, N0)) if keep_A0 else None
            check(lib.ddm_galerkin_assemble(self.h, _p(self.A0_device, C.c_double) if keep_A0 else None))
        return self

    def upload_factors(self, P, a, b):
        """ddm_upload_factor / ddm_upload_factor_lu for the hosted subdomains [a, b).  Re-uploading a factor with the same
        symbolic structure only converts the new values in place (device schedules stay valid)."""
        self._upload_factor_list([P.factors[i] for i in range(a, b)])

    def _upload_factor_list(self, factors):
        lib = self.lib
        a = 0
        for i, F in enumerate(factors):
            if hasattr(F, "an"):  # factor.PendingFactor: numeric phase on the device, with the host's symbolic analysis
                self.factorize(i, F)
                continue
            if hasattr(F, "xu"):  # non-symmetric subdomain matrix: LU factor (the reference's UMFPack branch, schwarz.hh:85-92)
                check(lib.ddm_upload_factor_lu(self.h, C.c_int32(i - a), C.c_int32(F.n), C.c_int32(F.nsuper),
                                               _p(np.ascontiguousarray(F.super_, dtype=np.int32), C.c_int32),
                                               _p(np.ascontiguousarray(F.pi, dtype=np.int64), C.c_int64),
                                               _p(np.ascontiguousarray(F.px, dtype=np.int64), C.c_int64),
                                               _p(np.ascontiguousarray(F.s, dtype=np.int32), C.c_int32),
                                               _p(np.ascontiguousarray(F.x, dtype=np.float64), C.c_double),
                                               _p(np.ascontiguousarray(F.xu, dtype=np.float64), C.c_double),
                                               _p(np.ascontiguousarray(F.perm, dtype=np.int32), C.c_int32)))
                continue
            check(lib.ddm_upload_factor(self.h, C.c_int32(i - a), C.c_int32(F.n), C.c_int32(F.nsuper),
                                        _p(np.ascontiguousarray(F.super_, dtype=np.int32), C.c_int32),
                                        _p(np.ascontiguousarray(F.pi, dtype=np.int64), C.c_int64),
                                        _p(np.ascontiguousarray(F.px, dtype=np.int64), C.c_int64),
                                        _p(np.ascontiguousarray(F.s, dtype=np.int32), C.c_int32),
                                        _p(np.ascontiguousarray(F.x, dtype=np.float64), C.c_double),
                                        _p(np.ascontiguousarray(F.perm, dtype=np.int32), C.c_int32)))

    def factorize(self, sub, F, want_panels=False):
        """ddm_factorize_subdomain for local subdomain `sub`: F = factor.PendingFactor (analysis + A_i).  want_panels: also
        return the LL^T panels in cholmod layout (the array ddm_upload_factor would have taken)."""
        an, A = F.an, F.A
        x = np.empty(int(an.px[-1])) if want_panels else None
        check(self.lib.ddm_factorize_subdomain(
            self.h, C.c_int32(sub), C.c_int32(an.n), C.c_int32(an.first.size - 1),
            _p(np.ascontiguousarray(an.first, dtype=np.int32), C.c_int32), _p(np.ascontiguousarray(an.pi, dtype=np.int64), C.c_int64),
            _p(np.ascontiguousarray(an.px, dtype=np.int64), C.c_int64), _p(an.row_lists(), C.c_int32),
            _p(np.ascontiguousarray(an.perm, dtype=np.int32), C.c_int32), _p(np.ascontiguousarray(A.indptr, dtype=np.int64), C.c_int64),
            _p(np.ascontiguousarray(A.indices, dtype=np.int32), C.c_int32), _p(np.ascontiguousarray(A.data, dtype=np.float64), C.c_double),
            _p(x, C.c_double) if want_panels else None))
        return x

    # -- helpers
    def vector(self, values=None):
        return DeviceVector(self, values)

    def register_host(self, arr):
        """Page-lock a caller's float64 array in place (ddm_host_register): the *_host entries then copy from / to it directly."""
        check(self.lib.ddm_host_register(self.h, _p(arr, C.c_double), C.c_int64(arr.size)))

    def unregister_host(self, arr):
        check(self.lib.ddm_host_unregister(self.h, _p(arr, C.c_double)))

    def owned(self, global_vec):
        """Restrict a unique-DOF vector to this rank's owned entries."""
        return np.ascontiguousarray(np.asarray(global_vec)[self.shard.owned])

    def stat(self, key):
        v = C.c_double()
        check(self.lib.ddm_get_stat(self.h, key.encode(), C.byref(v)))
        return v.value

    def timers(self, on=True):
        check(self.lib.ddm_timers_enable(self.h, C.c_int(int(on))))
        check(self.lib.ddm_timers_reset(self.h))

    def timer(self, event):
        ms, n = C.c_double(), C.c_int64()
        check(self.lib.ddm_get_timer(self.h, event.encode(), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def stopwatch_start(self):
        check(self.lib.ddm_stopwatch_start(self.h))

    def stopwatch_stop(self):
        ms = C.c_double()
        check(self.lib.ddm_stopwatch_stop(self.h, C.byref(ms)))
        return ms.value

    def launches(self):
        return int(self.lib.ddm_kernel_launches(self.h))

    def sync(self):
        check(self.lib.ddm_sync(self.h))

    def close(self):
        if self.h:
            self.lib.ddm_ctx_destroy(self.h)
            self.h = C.c_void_p()


class NonOverlappingOperator:
    """y = A x and y += alpha A x (nonoverlapping_operator.hh:34-50)."""

    category = "nonoverlapping"

    def __init__(self, ctx):
        self.ctx = ctx

    def apply(self, x, y):
        check(self.ctx.lib.ddm_op_apply(self.ctx.h, x.h, y.h))

    def applyscaleadd(self, alpha, x, y):
        check(self.ctx.lib.ddm_op_applyscaleadd(self.ctx.h, C.c_double(alpha), x.h, y.h))

    def apply_host(self, x):
        y = np.empty(self.ctx.n_own)
        check(self.ctx.lib.ddm_op_apply_host(self.ctx.h, _p(np.ascontiguousarray(x, dtype=np.float64), C.c_double), _p(y, C.c_double)))
        return y


class NonOverlappingScalarProduct:
    """Owner-masked dot and norm (nonoverlapping_operator.hh:76-83)."""

    category = "nonoverlapping"

    def __init__(self, ctx):
        self.ctx = ctx

    def dot(self, x, y):
        out = C.c_double()
        check(self.ctx.lib.ddm_dot(self.ctx.h, x.h, y.h, C.byref(out)))
        return out.value

    def norm(self, x):
        out = C.c_double()
        check(self.ctx.lib.ddm_norm(self.ctx.h, x.h, C.byref(out)))
        return out.value


class SchwarzPreconditioner:
    """One-level overlapping Schwarz (schwarz.hh:115-149); type 'standard' or 'restricted' (:80-83)."""

    category = "nonoverlapping"

    def __init__(self, ctx, type="restricted"):
        if type not in ("standard", "restricted"):
            raise DDMError(f"Unknown Schwarz type '{type}'")  # schwarz.hh:83
        self.ctx, self.type = ctx, STANDARD if type == "standard" else RESTRICTED

    def pre(self, x, b):
        pass

    def post(self, x):
        pass

    def apply(self, x, d):
        check(self.ctx.lib.ddm_schwarz_apply(self.ctx.h, C.c_int(self.type), d.h, x.h))


class GalerkinPreconditioner:
    """Coarse correction Z A0^{-1} Z^T (galerkin_preconditioner.hh:151-194)."""

    category = "nonoverlapping"

    def __init__(self, ctx):
        self.ctx = ctx

    def pre(self, x, b):
        pass

    def post(self, x):
        pass

    def apply(self, x, d):
        check(self.ctx.lib.ddm_coarse_apply(self.ctx.h, d.h, x.h))


class CombinedPreconditioner:
    """Additive or multiplicative combination (combined_preconditioner.hh:127-163).

    add() takes a SchwarzPreconditioner first and optionally a GalerkinPreconditioner, the order the
    reference drivers use (examples/poisson.cc:305-306); both levels then run as one fused GPU pass.
    """

    category = "nonoverlapping"

    def __init__(self, mode="additive"):
        if mode not in ("additive", "multiplicative"):
            raise DDMError(f"Unknown apply mode '{mode}'")  # combined_preconditioner.hh:66
        self.mode = ADDITIVE if mode == "additive" else MULTIPLICATIVE
        self.precs = []
        self.op = None

    def set_op(self, op):
        self.op = op

    def add(self, prec):
        if getattr(prec, "category", None) != self.category:
            raise DDMError("Preconditioner category mismatch")  # combined_preconditioner.hh:92-96
        if not self.precs and not isinstance(prec, SchwarzPreconditioner):
            raise DDMError("the first preconditioner must be the Schwarz level")
        if self.precs and not isinstance(prec, GalerkinPreconditioner):
            raise DDMError("the second preconditioner must be the Galerkin level")
        self.precs.append(prec)

    def pre(self, x, b):
        pass

    def post(self, x):
        pass

    def apply(self, x, d):
        if not self.precs:
            raise DDMError("No preconditioners added")
        if self.mode == MULTIPLICATIVE and len(self.precs) > 1 and self.op is None:
            raise DDMError("Multiplicative mode needs the operator, call set_op")  # combined_preconditioner.hh:147
        s = self.precs[0]
        check(s.ctx.lib.ddm_twolevel_apply(s.ctx.h, C.c_int(s.type), C.c_int(self.mode), C.c_int(int(len(self.precs) > 1)), d.h, x.h))

    def apply_host(self, d):
        s = self.precs[0]
        x = np.empty(s.ctx.n_own)
        check(s.ctx.lib.ddm_twolevel_apply_host(s.ctx.h, C.c_int(s.type), C.c_int(self.mode), C.c_int(int(len(self.precs) > 1)),
                                                _p(np.ascontiguousarray(d, dtype=np.float64), C.c_double), _p(x, C.c_double)))
        return x


class InverseOperatorResult:
    def __init__(self, iterations=0, converged=False, reduction=0.0, conv_rate=0.0, elapsed=0.0, history=None):
        self.iterations, self.converged, self.reduction = iterations, converged, reduction
        self.conv_rate, self.elapsed, self.history = conv_rate, elapsed, history


class CGSolver:
    """Dune::CGSolver (solvers.hh:279-340) over the three interfaces above.

    apply() is the literal drop-in loop: every scalar comes back to the host exactly where
    Dune::CGSolver reads it.  solve_device() runs the same recurrences with the scalars kept on the
    GPU (ddm_pcg_solve).
    """

    def __init__(self, op, sp, prec, reduction=1e-8, maxit=100, verbose=0):
        self.op, self.sp, self.prec = op, sp, prec
        self.reduction, self.maxit, self.verbose = reduction, maxit, verbose

    def apply(self, x, b):
        hist = []
        self.prec.pre(x, b)
        self.op.applyscaleadd(-1.0, x, b)
        d = self.sp.norm(b)
        def0 = d
        hist.append(d)
        if not np.isfinite(d):
            raise DDMError("CGSolver: defect is infinite or NaN")  # SolverAbort, solver.hh:458-466

        def conv(v):
            return v < def0 * self.reduction or v < 1e-30

        i = 0
        if not conv(d):
            ctx = x.ctx
            p, q = ctx.vector(), ctx.vector()
            p.fill(0.0)
            self.prec.apply(p, b)
            rholast = self.sp.dot(p, b)
            for i in range(1, self.maxit + 1):
                self.op.apply(p, q)
                alpha = self.sp.dot(p, q)
                lam = 0.0 if d == 0.0 else rholast / alpha
                x.axpy(lam, p)
                b.axpy(-lam, q)
                d = self.sp.norm(b)
                hist.append(d)
                if not np.isfinite(d):
                    raise DDMError("CGSolver: defect is infinite or NaN")
                if conv(d):
                    break
                q.fill(0.0)
                self.prec.apply(q, b)
                rho = self.sp.dot(q, b)
                beta = 0.0 if d == 0.0 else rho / rholast
                p *= beta
                p += q
                rholast = rho
        self.prec.post(x)
        red = d / def0 if def0 > 0 else 0.0
        return InverseOperatorResult(i, conv(d), red, red ** (1.0 / i) if i else 0.0, 0.0, np.array(hist))

    def solve_device(self, x, b):
        s = self.prec.precs[0]
        res = Result()
        hist = np.zeros(self.maxit + 1)
        check(s.ctx.lib.ddm_pcg_solve(s.ctx.h, b.h, x.h, C.c_double(self.reduction), C.c_int32(self.maxit), C.c_int(s.type),
                                      C.c_int(self.prec.mode), C.c_int(int(len(self.prec.precs) > 1)), C.byref(res), _p(hist, C.c_double)))
        return InverseOperatorResult(res.iterations, bool(res.converged), res.reduction, res.conv_rate, res.elapsed, hist[: res.iterations + 1])

    def solve_host(self, b_host, x_host=None, overwrite=False):
        """End-to-end call with host buffers (H2D and D2H inside).  overwrite=True works in the caller's float64 arrays, as
        Dune::CGSolver::apply does (b becomes the defect, x the solution); otherwise copies are made first."""
        s = self.prec.precs[0]
        if overwrite and x_host is not None:
            b, x = b_host, x_host
            if not (b.dtype == np.float64 and x.dtype == np.float64 and b.flags.c_contiguous and x.flags.c_contiguous):
                raise DDMError("solve_host(overwrite=True) needs contiguous float64 arrays")
        else:
            b = np.array(b_host, dtype=np.float64)
            x = np.zeros_like(b) if x_host is None else np.array(x_host, dtype=np.float64)
        res = Result()
        hist = np.zeros(self.maxit + 1)
        check(s.ctx.lib.ddm_pcg_solve_host(s.ctx.h, _p(b, C.c_double), _p(x, C.c_double), C.c_double(self.reduction), C.c_int32(self.maxit),
                                           C.c_int(s.type), C.c_int(self.prec.mode), C.c_int(int(len(self.prec.precs) > 1)), C.byref(res), _p(hist, C.c_double)))
        return x, InverseOperatorResult(res.iterations, bool(res.converged), res.reduction, res.conv_rate, res.elapsed, hist[: res.iterations + 1])


class BiCGSTABSolver:
    """Dune::BiCGSTABSolver (solvers.hh:419-590) over the same three interfaces: the literal drop-in loop, every dot and norm
    read on the host where the reference reads it.  Iterations count in halves as in the reference (the stopping test runs
    after each of the two corrections of a step); the reported count is truncated the way InverseOperatorResult stores it."""

    EPSILON = 1e-80

    def __init__(self, op, sp, prec, reduction=1e-8, maxit=100, verbose=0):
        self.op, self.sp, self.prec = op, sp, prec
        self.reduction, self.maxit, self.verbose = reduction, maxit, verbose

    def apply(self, x, b):
        ctx = x.ctx
        r = b
        self.prec.pre(x, r)
        self.op.applyscaleadd(-1.0, x, r)  # b becomes the defect
        rt = ctx.vector().assign(r)
        norm = self.sp.norm(r)
        hist = [norm]
        def0 = norm

        def step(v):
            if not np.isfinite(v):
                raise DDMError("BiCGSTABSolver: defect is infinite or NaN")  # SolverAbort, solver.hh:458-466
            return v < def0 * self.reduction or v < 1e-30

        it = 0.0
        if not step(norm):
            p, v, t, y = (ctx.vector().fill(0.0) for _ in range(4))
            rho = alpha = omega = 1.0
            it = 0.5
            while it < self.maxit:
                rho_new = self.sp.dot(rt, r)
                if abs(rho) <= self.EPSILON:
                    raise DDMError(f"breakdown in BiCGSTAB - rho {rho} <= EPSILON {self.EPSILON} after {it} iterations")
                if abs(omega) <= self.EPSILON:
                    raise DDMError(f"breakdown in BiCGSTAB - omega {omega} <= EPSILON {self.EPSILON} after {it} iterations")
                if it < 1:
                    p.assign(r)
                else:
                    beta = 0.0 if norm == 0.0 else (rho_new / rho) * (alpha / omega)
                    p.axpy(-omega, v)
                    p *= beta
                    p += r
                y.fill(0.0)
                self.prec.apply(y, p)
                self.op.apply(y, v)
                h = self.sp.dot(rt, v)
                if abs(h) < self.EPSILON:
                    raise DDMError(f"abs(h) < EPSILON in BiCGSTAB - abs(h) {abs(h)} < EPSILON {self.EPSILON} after {it} iterations")
                alpha = 0.0 if norm == 0.0 else rho_new / h
                x.axpy(alpha, y)
                r.axpy(-alpha, v)
                norm = self.sp.norm(r)
                hist.append(norm)
                if step(norm):
                    break
                it += 0.5
                y.fill(0.0)
                self.prec.apply(y, r)
                self.op.apply(y, t)
                h = self.sp.dot(t, t)
                omega = 0.0 if norm == 0.0 else self.sp.dot(t, r) / h
                x.axpy(omega, y)
                r.axpy(-omega, t)
                rho = rho_new
                norm = self.sp.norm(r)
                hist.append(norm)
                if step(norm):
                    break
                it += 0.5
            else:
                it -= 0.5  # the loop ran out: the last completed half step is what Iteration recorded
        self.prec.post(x)
        red = norm / def0 if def0 > 0 else 0.0
        return InverseOperatorResult(int(it), step(norm), red, red ** (1.0 / it) if it else 0.0, 0.0, np.array(hist))


class RestartedGMResSolver:
    """Dune::RestartedGMResSolver (solvers.hh:923-1035) over the same three interfaces.

    apply() is the literal drop-in loop (every dot and norm returns to the host where the reference reads it);
    solve_device() / solve_host() run ddm_gmres_solve: basis and vector work on the GPU, one column of H per step
    comes back for the Givens update.
    """

    def __init__(self, op, sp, prec, reduction=1e-8, restart=30, maxit=100, verbose=0):
        self.op, self.sp, self.prec = op, sp, prec
        self.reduction, self.restart, self.maxit, self.verbose = reduction, restart, maxit, verbose

    @staticmethod
    def _gen_rot(dx, dy):
        eps, ndx, ndy = 1e-15, abs(dx), abs(dy)
        if ndy < eps:
            return 1.0, 0.0
        if ndx < eps:
            return 0.0, 1.0
        temp = min(ndx, ndy) / max(ndx, ndy)
        if ndy > ndx:
            return 1.0 / np.sqrt(1.0 + temp * temp) * temp, 1.0 / np.sqrt(1.0 + temp * temp) * dx * dy / ndx / ndy
        return 1.0 / np.sqrt(1.0 + temp * temp), 1.0 / np.sqrt(1.0 + temp * temp) * (dy / dx)

    def apply(self, x, b):
        ctx, m = x.ctx, self.restart
        hist = []
        s, sn, cs = np.zeros(m + 1), np.zeros(m), np.zeros(m)
        H = np.zeros((m + 1, m))
        b2, w = ctx.vector().assign(b), ctx.vector()
        v = [ctx.vector() for _ in range(m + 1)]
        self.prec.pre(x, b)
        self.op.applyscaleadd(-1.0, x, b)
        v[0].fill(0.0)
        self.prec.apply(v[0], b)
        norm = self.sp.norm(v[0])
        def0 = norm
        hist.append(norm)
        if not np.isfinite(norm):
            raise DDMError("GMRes: defect is infinite or NaN")

        def conv(val):
            return val < def0 * self.reduction or val < 1e-30

        done = conv(norm)
        j = 1
        while j <= self.maxit and not done:
            v[0] *= 0.0 if norm == 0.0 else 1.0 / norm
            s[:] = 0.0
            s[0] = norm
            i = 0
            while i < m and j <= self.maxit and not done:
                w.fill(0.0)
                v[i + 1].fill(0.0)
                self.op.apply(v[i], v[i + 1])
                self.prec.apply(w, v[i + 1])
                for k in range(i + 1):
                    H[k, i] = self.sp.dot(v[k], w)
                    w.axpy(-H[k, i], v[k])
                H[i + 1, i] = self.sp.norm(w)
                if abs(H[i + 1, i]) < 1e-80:
                    raise DDMError(f"breakdown in GMRes - |w| == 0.0 after {j} iterations")  # SolverAbort, solvers.hh:983
                v[i + 1].assign(w)
                v[i + 1] *= 0.0 if norm == 0.0 else 1.0 / H[i + 1, i]
                for k in range(i):
                    H[k, i], H[k + 1, i] = cs[k] * H[k, i] + sn[k] * H[k + 1, i], -sn[k] * H[k, i] + cs[k] * H[k + 1, i]
                cs[i], sn[i] = self._gen_rot(H[i, i], H[i + 1, i])
                H[i, i], H[i + 1, i] = cs[i] * H[i, i] + sn[i] * H[i + 1, i], -sn[i] * H[i, i] + cs[i] * H[i + 1, i]
                s[i], s[i + 1] = cs[i] * s[i] + sn[i] * s[i + 1], -sn[i] * s[i] + cs[i] * s[i + 1]
                norm = abs(s[i + 1])
                hist.append(norm)
                if not np.isfinite(norm):
                    raise DDMError("GMRes: defect is infinite or NaN")
                done = conv(norm)
                i += 1
                j += 1
            w.fill(0.0)
            y = s.copy()
            for a in range(i - 1, -1, -1):
                rhs = s[a] - H[a, a + 1 : i] @ y[a + 1 : i]
                y[a] = 0.0 if rhs == 0.0 else rhs / H[a, a]
                w.axpy(y[a], v[a])
            x += w
            if not done and j < self.maxit:
                b.assign(b2)
                self.op.applyscaleadd(-1.0, x, b)
                v[0].fill(0.0)
                self.prec.apply(v[0], b)
                norm = self.sp.norm(v[0])
        self.prec.post(x)
        red = norm / def0 if def0 > 0 else 0.0
        return InverseOperatorResult(j - 1, done, red, red ** (1.0 / (j - 1)) if j > 1 else 0.0, 0.0, np.array(hist))

    def _args(self):
        s = self.prec.precs[0]
        return s, (C.c_double(self.reduction), C.c_int32(self.restart), C.c_int32(self.maxit), C.c_int(s.type), C.c_int(self.prec.mode),
                   C.c_int(int(len(self.prec.precs) > 1)))

    def solve_device(self, x, b):
        s, args = self._args()
        res = Result()
        hist = np.zeros(self.maxit + 1)
        check(s.ctx.lib.ddm_gmres_solve(s.ctx.h, b.h, x.h, *args, C.byref(res), _p(hist, C.c_double)))
        return InverseOperatorResult(res.iterations, bool(res.converged), res.reduction, res.conv_rate, res.elapsed, hist[: res.iterations + 1])

    def solve_host(self, b_host, x_host=None):
        s, args = self._args()
        b = np.array(b_host, dtype=np.float64)
        x = np.zeros_like(b) if x_host is None else np.array(x_host, dtype=np.float64)
        res = Result()
        hist = np.zeros(self.maxit + 1)
        check(s.ctx.lib.ddm_gmres_solve_host(s.ctx.h, _p(b, C.c_double), _p(x, C.c_double), *args, C.byref(res), _p(hist, C.c_double)))
        return x, InverseOperatorResult(res.iterations, bool(res.converged), res.reduction, res.conv_rate, res.elapsed, hist[: res.iterations + 1])
