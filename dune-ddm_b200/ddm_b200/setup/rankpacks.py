"""Merge the per-rank dumps of a real dune-ddm run (dune/ddm_b200/exporter.hh) into SetupProducts.

Every MPI rank of the reference holds its share in local numbering: an *additive* local matrix on its
non-overlapping index set (owner and border copies), the overlapping index set (non-overlapping indices first, then the added
layers: overlap_extension.hh:100-104,260-262), D_i, the CHOLMOD factor of A_i, its coarse vectors.  The GPU library works on
one entry per unique DOF, so the merge is a renumbering:

  unique DOFs   = the distinct global indices of all non-overlapping sets, in ascending order of the reference's ids
  owner         = the rank whose index set marks the DOF `owner` (for conforming spaces the lowest rank that sees it,
                  pdelab_helper.hh:44-82; ghost-cell copies of a DG space never own)
  A             = sum over ranks of P_r^T A_local,r P_r          (mv + border sum, nonoverlapping_operator.hh:34-39)
  R_i, D_i, Z_i = as exported, indices renumbered
  factor of A_i = CHOLMOD's own L and permutation where exported (supernodal arrays as ddm_upload_factor takes them), else a
                  factorisation of the exported A_ovlp (non-symmetric problems: UMFPack's row pivoting has no counterpart in
                  the symmetric-structure LU layout), else of R_i A R_i^T
  A0            = Z^T A Z (the reference keeps its coarse matrix private, galerkin_preconditioner.hh:197,331)
"""
from __future__ import annotations

import glob
import os

import numpy as np
import scipy.sparse as sp

from . import coarse as coarse_mod
from . import factor, pack
from .problem import SetupProducts

OWNER, OVERLAP, COPY = 1, 2, 3  # Dune::OwnerOverlapCopyAttributeSet::AttributeSet (owneroverlapcopy.hh:59-61)


def read_rank_packs(directory):
    paths = sorted(glob.glob(os.path.join(directory, "rank*.ddmpack")), key=lambda p: int(os.path.basename(p)[4:-8]))
    if not paths:
        raise FileNotFoundError(f"no rank*.ddmpack files in {directory}")
    packs = [pack.read(p) for p in paths]
    nr = int(packs[0]["nranks"][0])
    if len(packs) != nr or [int(d["rank"][0]) for d in packs] != list(range(nr)):
        raise ValueError(f"{directory}: expected the dumps of ranks 0..{nr - 1}, found {[int(d['rank'][0]) for d in packs]}")
    return packs


def merge(packs, symmetric=None, leaf_size=32, check=True):
    """packs: list of dicts (pack.read) in rank order, or a directory.  Returns SetupProducts on unique DOFs."""
    if isinstance(packs, (str, os.PathLike)):
        packs = read_rank_packs(packs)
    nr = len(packs)
    block = int(packs[0]["A_block"][0])
    gids = np.unique(np.concatenate([d["novlp_gid"] for d in packs]))
    n = gids.size

    def renumber(g, what):
        idx = np.searchsorted(gids, g)
        if (idx >= n).any() or (gids[np.minimum(idx, n - 1)] != g).any():
            raise ValueError(f"{what}: global index not present in any non-overlapping index set")
        return idx.astype(np.int64)

    owner = np.full(n, -1, dtype=np.int32)
    rows, cols, vals = [], [], []
    b = np.zeros(n)
    have_b = all("rhs" in d for d in packs)
    novlp = []
    for r, d in enumerate(packs):
        loc = renumber(d["novlp_gid"], f"rank {r}")
        if np.unique(loc).size != loc.size:
            raise ValueError(f"rank {r}: repeated global index in the non-overlapping set")
        novlp.append(loc)
        own = d["novlp_attr"] == OWNER
        if (owner[loc[own]] >= 0).any():
            raise ValueError(f"rank {r}: a DOF has two owners")
        owner[loc[own]] = r
        rp, ci, va = d["A_rowptr"], d["A_col"], d["A_val"]
        if rp.size != loc.size + 1:
            raise ValueError(f"rank {r}: local matrix has {rp.size - 1} rows, index set {loc.size}")
        rows.append(np.repeat(loc, np.diff(rp)))
        cols.append(loc[ci])
        vals.append(va)
        if have_b:
            b[loc] = d["rhs"]  # consistent vectors: every sharer holds the same value (twolevel_schwarz.hh:134-136)
    if (owner < 0).any():
        raise ValueError(f"{int((owner < 0).sum())} DOFs have no owner")
    A = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(n, n)).tocsr()
    A.sum_duplicates()
    A.sort_indices()
    if symmetric is None:
        D = (A - A.T).tocsr()
        symmetric = D.nnz == 0 or abs(D).max() <= 1e-12 * abs(A).max()

    subs, pou, factors, Z = [], [], [], []
    for r, d in enumerate(packs):
        s = renumber(d["ovlp_gid"], f"rank {r} (overlapping set)")
        ni = s.size
        if check and not np.array_equal(s[: novlp[r].size], novlp[r]):
            raise ValueError(f"rank {r}: the overlapping index set does not start with the non-overlapping one")
        subs.append(s)
        pou.append(np.array(d["pou"], dtype=np.float64) if "pou" in d else np.ones(ni))
        Zr = d["Z"]
        Z.append(np.asfortranarray(Zr.reshape(Zr.shape[0], ni).T) if Zr.size else np.zeros((ni, 0), order="F"))
        Ai = None
        if "Aovlp_rowptr" in d:
            Ai = sp.csr_matrix((d["Aovlp_val"], d["Aovlp_col"], d["Aovlp_rowptr"]), shape=(ni, ni))
            Ai.sort_indices()
        if "F_x" in d:
            F = factor.SupernodalFactor(n=ni, perm=d["F_perm"].astype(np.int32), super_=d["F_super"].astype(np.int32), pi=d["F_pi"].astype(np.int64),
                                        px=d["F_px"].astype(np.int64), s=d["F_s"].astype(np.int32), x=np.array(d["F_x"], dtype=np.float64))
            if check:  # CHOLMOD's factor must solve the subdomain system the unique-DOF operator defines
                ref = Ai if Ai is not None else A[s][:, s]
                x = np.random.default_rng(r).standard_normal(ni)
                y = factor.solve_supernodal(F, ref @ x)
                if np.linalg.norm(y - x) > 1e-6 * np.linalg.norm(x):
                    raise ValueError(f"rank {r}: the exported factor does not invert the subdomain matrix")
        else:
            if Ai is None:
                Ai = A[s][:, s].tocsr()
                Ai.sort_indices()
            F = factor.cholesky_supernodal(Ai, None, leaf_size) if symmetric else factor.lu_supernodal(Ai, None, leaf_size)
        factors.append(F)
    A0, _, _ = coarse_mod.galerkin_matrix(A, subs, Z)
    meta = dict(nsub=nr, n=int(n), N0=int(A0.shape[0]), symmetric=bool(symmetric), source="dune-ddm rank dumps", block=block)
    return SetupProducts(A=A, b=b, owner=owner, subs=subs, pou=pou, factors=factors, Z=Z, A0=A0,
                         n_novlp=np.array([v.size for v in novlp], dtype=np.int64), meta=meta)


def main(argv=None):
    """python -m ddm_b200.setup.rankpacks <directory with rank*.ddmpack> <out.ddmpack>: merged products as one pack, the input of
    Context::fromPack / api.Context.from_products / bench.py --pack."""
    import argparse

    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("directory")
    ap.add_argument("out")
    ap.add_argument("--no-check", action="store_true", help="skip the consistency checks (index-set nesting, factor residual)")
    a = ap.parse_args(argv)
    P = merge(a.directory, check=not a.no_check)
    pack.save_products(a.out, P)
    print(f"merged {P.nsub} ranks: n={P.n} N0={P.A0.shape[0]} symmetric={P.meta['symmetric']} -> {a.out}")


if __name__ == "__main__":
    main()
