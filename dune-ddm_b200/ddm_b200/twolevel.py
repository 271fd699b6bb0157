"""One-stop two-level Schwarz solver over the C ABI: the counterpart of the reference's public façade
TwoLevelSchwarzSolver (/root/reference/dune/ddm/twolevel_schwarz.hh:27-146).

The first apply() builds everything from the assembled matrix: non-overlapping sets (owner = lowest rank,
pdelab_helper.hh:44-82), overlap growth, partition of unity, the POU coarse space from the template vectors {1, x, y, xy}
(twolevel_schwarz.hh:47-51,107), subdomain factorisation, upload; the Galerkin matrix is assembled on the GPU
(ddm_galerkin_assemble).  Every later apply() with a matrix of the same pattern only refreshes values
(twolevel_schwarz.hh:99-104): operator values, numeric re-factorisation into the existing device tiles, coarse matrix -- index
sets, orderings and symbolic factorisations, bases, schedules and device allocations are kept.  Solver and preconditioner options follow the reference's
ParameterTree keys (overlap, mode, fine.type, solver.type/restart/maxit; defaults :117-123: GMRES(30), maxit 1000).
One GPU per solver object.
"""
from __future__ import annotations

import ctypes as C
import multiprocessing as mp
import os

import numpy as np
import scipy.sparse as sp

from . import api
from .lib import DDMError, check
from .setup import coarse as coarse_mod
from .setup import decomp, factor
from .setup.problem import SetupProducts

_G: dict = {}


def _factor_job(i):
    """First call: extract A_i = R_i A R_i^T, analyse its pattern, factor.  Later calls (pattern unchanged): numeric phase only,
    on the values picked straight out of the new operator's data array."""
    from threadpoolctl import threadpool_limits

    with threadpool_limits(1):
        A, sym = _G["A"], _G["symmetric"]
        numeric = factor.numeric_cholesky if sym else factor.numeric_lu
        if _G["analysis"] is not None:
            an, src = _G["analysis"][i], _G["src"][i]
            return i, numeric(an, A.data[src]), None, None
        s = _G["subs"][i]
        tag = sp.csr_matrix((np.arange(1, A.nnz + 1, dtype=np.float64), A.indices, A.indptr), shape=A.shape)
        Ti = tag[s][:, s].tocsr()
        Ti.sort_indices()
        src = np.rint(Ti.data).astype(np.int64) - 1  # A_i.data = A.data[src]
        Ai = sp.csr_matrix((A.data[src], Ti.indices, Ti.indptr), shape=Ti.shape)
        an = factor.analyse(Ai, _G["coords"][s], 32, symmetric=sym)
        F = numeric(an, Ai.data)
        if not sym:
            factor.check_lu_backward_error(Ai, F)  # the restated LU does not pivot: refuse a factor that is not backward stable
        return i, F, an, src


class TwoLevelSchwarzSolver:
    def __init__(self, coords, dof_cells, cell_part, nsub, config=None, symmetric=True, dirichlet=None, device=0, workers=None):
        """coords: (n, d) DOF coordinates; dof_cells: (m, k) DOFs of every element; cell_part: subdomain of every element."""
        cfg = dict(overlap=1, mode="additive", **{"fine.type": "restricted", "solver.type": "restartedgmressolver",
                                                   "solver.restart": 30, "solver.maxit": 1000, "pou.type": "distance", "pou.shrink": 0})
        cfg.update(config or {})
        if cfg["solver.type"] not in ("restartedgmressolver", "cgsolver", "bicgstabsolver"):
            raise DDMError(f"Unknown solver type '{cfg['solver.type']}'")
        self.cfg, self.coords, self.cells, self.part, self.nsub = cfg, np.asarray(coords), np.asarray(dof_cells), np.asarray(cell_part), int(nsub)
        self.symmetric, self.device = bool(symmetric), device
        self.dirichlet = np.zeros(self.coords.shape[0], dtype=bool) if dirichlet is None else np.asarray(dirichlet, dtype=bool)
        self.workers = workers or min(self.nsub, os.cpu_count() or 1)
        self.ctx = None
        self.P = None
        self.res = None
        self.updates = 0
        self._analysis = None  # per subdomain: factor.Analysis, kept across value updates (twolevel_schwarz.hh:99-104)
        self._src = None  # per subdomain: positions of A_i's values in the operator's data array

    # -- setup pieces
    def _factors(self, A):
        first = self._analysis is None
        _G.clear()
        _G.update(A=A, subs=self.P.subs if self.P else self._subs, coords=self.coords, symmetric=self.symmetric,
                  analysis=self._analysis, src=self._src)
        out = [None] * self.nsub
        ans, srcs = [None] * self.nsub, [None] * self.nsub

        def keep(i, F, an, src):
            out[i], ans[i], srcs[i] = F, an, src

        if self.workers > 1:
            with mp.get_context("fork").Pool(self.workers) as pool:
                for res in pool.imap_unordered(_factor_job, range(self.nsub)):
                    keep(*res)
        else:
            for i in range(self.nsub):
                keep(*_factor_job(i))
        _G.clear()
        if first:
            self._analysis, self._src = ans, srcs
        return out

    def _first(self, A, b):
        n = A.shape[0]
        cfg = self.cfg
        novlp, owner = decomp.nonoverlapping_sets(n, self.cells, self.part, self.nsub)
        self._subs = subs = decomp.grow_overlap(A, novlp, cfg["overlap"])
        bm = decomp.boundary_masks(A, subs)
        pou = decomp.partition_of_unity(A, subs, bm, owner, cfg["pou.type"], cfg["pou.shrink"], cfg["overlap"])
        Z = [coarse_mod.pou_template_basis(self.coords[s], pou[i], self.dirichlet[s]) for i, s in enumerate(subs)]
        factors = self._factors(A)
        N0 = sum(z.shape[1] for z in Z)
        self.P = SetupProducts(A=A, b=b, owner=owner, subs=subs, pou=pou, factors=factors, Z=Z, A0=np.zeros((N0, N0)),
                               n_novlp=np.array([v.size for v in novlp], dtype=np.int64), coords=self.coords, dirichlet=self.dirichlet,
                               meta=dict(nsub=self.nsub, symmetric=self.symmetric))
        self.ctx = api.Context.from_products(self.P, device=self.device, device_galerkin=True)
        self.P.A0 = self.ctx.A0_device

    def _update(self, A):
        P, ctx = self.P, self.ctx
        if A.shape != P.A.shape or not (np.array_equal(A.indptr, P.A.indptr) and np.array_equal(A.indices, P.A.indices)):
            raise DDMError("TwoLevelSchwarzSolver: the matrix pattern changed; build a new solver")
        P.A = A
        check(ctx.lib.ddm_update_operator_values(ctx.h, api._p(np.ascontiguousarray(A.data, dtype=np.float64), C.c_double)))
        P.factors = self._factors(A)
        ctx.upload_factors(P, 0, self.nsub)
        check(ctx.lib.ddm_update_coarse_matrix(ctx.h, None))  # A0 re-assembled on the device from the new operator values
        self.updates += 1

    # -- the reference's entry point: apply(A, z, r, reduction)
    def apply(self, A, z, r, reduction):
        """Solve A z = r to the given defect reduction; z holds the initial guess and receives the solution, r is left as the
        defect (Dune::InverseOperator convention).  Returns the InverseOperatorResult (also kept in self.res)."""
        A = sp.csr_matrix(A)
        A.sort_indices()
        if self.ctx is None:
            self._first(A, np.array(r, dtype=np.float64))
        else:
            self._update(A)
        ctx, cfg = self.ctx, self.cfg
        op, spr = api.NonOverlappingOperator(ctx), api.NonOverlappingScalarProduct(ctx)
        prec = api.CombinedPreconditioner(cfg["mode"])
        prec.set_op(op)
        prec.add(api.SchwarzPreconditioner(ctx, cfg["fine.type"]))
        prec.add(api.GalerkinPreconditioner(ctx))
        if cfg["solver.type"] == "cgsolver":
            solver = api.CGSolver(op, spr, prec, reduction=reduction, maxit=cfg["solver.maxit"])
        elif cfg["solver.type"] == "bicgstabsolver":
            solver = api.BiCGSTABSolver(op, spr, prec, reduction=reduction, maxit=cfg["solver.maxit"])
        else:
            solver = api.RestartedGMResSolver(op, spr, prec, reduction=reduction, restart=cfg["solver.restart"], maxit=cfg["solver.maxit"])
        x, b = ctx.vector(np.asarray(z, dtype=np.float64)), ctx.vector(np.asarray(r, dtype=np.float64))
        self.res = solver.solve_device(x, b) if hasattr(solver, "solve_device") else solver.apply(x, b)
        z[:] = x.download()
        r[:] = b.download()
        return self.res

    def close(self):
        if self.ctx is not None:
            self.ctx.close()
            self.ctx = None


class TwoLevelSchwarzPreconditioner:
    """The reference's one-object preconditioner (examples/pdelab_schwarz.hh:26-203): built from a problem description and a
    parameter tree, it hides index sets, partition of unity, coarse space, both levels and their combination behind the
    Dune::Preconditioner interface (pre / apply / post / category) and hands out the matching operator and scalar product.

    problem: dict with the keys of setup.problem.build -- coords, cells, Ke (element matrices, or ke_fn), dirichlet, rhs,
    cell_part, nsub; optional A_full, symmetric.  ptree keys (defaults as in the reference): overlap (1, pdelab_schwarz.hh:44),
    coarsespace.type (geneo; supported here: geneo, pou, none -- the other types of :56-63 raise like Dune::NotImplemented),
    coarse_space.nev / shift / tolerance / threshold / nev_max (eigensolver_params.hh:20-51), pou.type / pou.shrink, schwarz.type (restricted,
    schwarz.hh:80), combined_preconditioner.mode (additive, combined_preconditioner.hh:57).
    """

    category = "nonoverlapping"
    _known = ("geneo", "pou", "none")
    _other = ("constraint_geneo", "msgfem", "geneo_ring", "msgfem_ring", "harmonic_extension", "algebraic_geneo", "algebraic_msgfem",
              "svd", "msgfem_euclid")

    def __init__(self, problem, ptree=None, device=0, workers=None):
        from .setup import problem as setup_problem

        cfg = {"overlap": 1, "coarsespace.type": "geneo", "coarse_space.nev": 16, "coarse_space.shift": 1e-3, "coarse_space.tolerance": 1e-5,
               "coarse_space.threshold": -0.5, "coarse_space.nev_max": None,
               "pou.type": "distance", "pou.shrink": 0, "schwarz.type": "restricted", "combined_preconditioner.mode": "additive"}
        cfg.update(ptree or {})
        kind = cfg["coarsespace.type"]
        if kind not in self._known:
            if kind in self._other:
                raise DDMError(f"coarse space type '{kind}' is not implemented on this path")
            raise DDMError("Unknown coarse space type: " + str(kind))  # pdelab_schwarz.hh:63
        self.cfg = cfg
        p = dict(problem)
        self.P = setup_problem.build(p["coords"], p["cells"], p.get("Ke"), p["dirichlet"], p["rhs"], p["cell_part"], p["nsub"],
                                     overlap=cfg["overlap"], pou_kind=cfg["pou.type"], shrink=cfg["pou.shrink"], coarse=kind,
                                     nev=cfg["coarse_space.nev"], shift=cfg["coarse_space.shift"], eig_tol=cfg["coarse_space.tolerance"],
                                     eig_threshold=cfg["coarse_space.threshold"], nev_max=cfg["coarse_space.nev_max"],
                                     workers=workers, A_full=p.get("A_full"), symmetric=p.get("symmetric", True), ke_fn=p.get("ke_fn"))
        has_coarse = self.P.A0.shape[0] > 0
        self.ctx = api.Context.from_products(self.P, device=device, with_coarse=has_coarse)
        self._op = api.NonOverlappingOperator(self.ctx)
        self._prec = api.CombinedPreconditioner(cfg["combined_preconditioner.mode"])
        self._prec.set_op(self._op)
        self._prec.add(api.SchwarzPreconditioner(self.ctx, cfg["schwarz.type"]))
        if has_coarse:
            self._prec.add(api.GalerkinPreconditioner(self.ctx))

    # Dune::Preconditioner
    def pre(self, x, b):
        self._prec.pre(x, b)

    def apply(self, v, d):
        self._prec.apply(v, d)

    def post(self, x):
        self._prec.post(x)

    # what the reference's getters hand to the solver factory (pdelab_schwarz.hh:190-197)
    def getNonOverlappingOperator(self):
        return self._op

    def getScalarProduct(self):
        return api.NonOverlappingScalarProduct(self.ctx)

    def get_basis(self):
        return self.P.Z

    @property
    def combined(self):
        """The CombinedPreconditioner inside (what the device-resident Krylov loops of api.CGSolver / RestartedGMResSolver take)."""
        return self._prec

    def close(self):
        if self.ctx is not None:
            self.ctx.close()
            self.ctx = None
