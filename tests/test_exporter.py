"""the exporter a dune-ddm build calls on every rank (dune/ddm_b200/exporter.hh) and the merge of its rank files
into unique-DOF setup products (ddm_b200/setup/rankpacks.py).

A problem is split the way the reference holds it -- per rank an additive local matrix on the non-overlapping index set with
owner / copy attributes, the overlapping set, D_i, a CHOLMOD-layout factor, the basis vectors -- with scrambled global ids;
tests/cxx/export_check.cc puts each share into the reference's own container types (Dune::BCRSMatrix, BlockVector,
ParallelIndexSet) and calls exportRank; the merged products must equal the original ones and give the same PCG run.
Needs the Dune headers, so it runs where /root/reference is present."""
import os
import shutil
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
DUNE = [f"{REF}/extern/dune-common", f"{REF}/extern/dune-istl"]

pytestmark = pytest.mark.skipif(not (all(os.path.isdir(d) for d in DUNE) and shutil.which("g++")),
                                reason="needs the reference's dune-common / dune-istl headers and g++")


@pytest.fixture(scope="module")
def exporter(tmp_path_factory):
    d = tmp_path_factory.mktemp("cxx_export")
    stubs = d / "stubs"
    stubs.mkdir()
    for name in ("dune-common-config.hh", "dune-istl-config.hh"):
        (stubs / name).write_text("#pragma once\n")
    exe = str(d / "export_check")
    cmd = ["g++", "-std=c++20", "-O1", "-DHAVE_MPI=0", f"-I{stubs}", *[f"-I{p}" for p in DUNE], f"-I{ROOT}/include", f"-I{ROOT}/dune-ddm_b200",
           f"{ROOT}/tests/cxx/export_check.cc", *[f"{DUNE[0]}/dune/common/{f}.cc" for f in ("exceptions", "stdstreams", "debugallocator")], "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    return exe


def scalar_columns(F):
    """Columns of L from the supernodal panels: (rows, values) per column, diagonal first."""
    out = []
    for J in range(F.nsuper):
        f, l = int(F.super_[J]), int(F.super_[J + 1])
        rows = F.s[F.pi[J] : F.pi[J + 1]]
        P = F.x[F.px[J] : F.px[J + 1]].reshape((rows.size, l - f), order="F")
        for q in range(l - f):
            out.append((rows[q:].astype(np.int64), P[q:, q].copy()))
    return out


def cholmod_arrays(F, kind, itype):
    """The factor as CHOLMOD would hold it: 'super' (supernodal LL^T), 'll' (simplicial LL^T, packed), 'ldl' (simplicial
    LDL^T with slack between the columns, as after cholmod_updown / a non-packed factorisation)."""
    d = dict(L_Perm=F.perm.astype(np.int64), L_itype=np.array([itype], dtype=np.int64))
    if kind == "super":
        d.update(L_is_super=np.array([1], dtype=np.int64), L_is_ll=np.array([1], dtype=np.int64), L_super=F.super_.astype(np.int64),
                 L_pi=F.pi.astype(np.int64), L_px=F.px.astype(np.int64), L_s=F.s.astype(np.int64), L_x=F.x)
        return d
    cols = scalar_columns(F)
    slack = 0 if kind == "ll" else 2
    p, ii, xx, nz = [0], [], [], []
    for rows, vals in cols:
        if kind == "ldl":
            dj = vals[0] ** 2
            vals = np.concatenate(([dj], vals[1:] / vals[0]))
        ii.append(np.concatenate((rows, np.full(slack, -1, dtype=np.int64))))
        xx.append(np.concatenate((vals, np.full(slack, np.nan))))
        nz.append(rows.size)
        p.append(p[-1] + rows.size + slack)
    d.update(L_is_super=np.array([0], dtype=np.int64), L_is_ll=np.array([1 if kind == "ll" else 0], dtype=np.int64),
             L_p=np.array(p, dtype=np.int64), L_i=np.concatenate(ii), L_x=np.concatenate(xx), L_nz=np.array(nz, dtype=np.int64))
    return d


def rank_shares(P, factor_kinds, gid_of, ghosts=False):
    """Split unique-DOF products into what every rank of the reference holds.  ghosts: discontinuous Galerkin -- no DOF is
    shared between cells, the local index set carries the DOFs of the ghost cells as copies and every matrix row lives on the
    rank that owns it (make_additive, examples/pdelab_helper.hh:109-149)."""
    A = P.A.tocoo()
    taken = np.zeros(A.nnz, dtype=bool)
    shares = []
    block = int(P.meta.get("block", 1))
    for r in range(P.nsub):
        s = P.subs[r]
        nov = s if ghosts else s[: int(P.n_novlp[r])]
        inr = np.zeros(P.n, dtype=bool)
        inr[nov] = True
        if ghosts:
            mine = (P.owner[A.row] == r) & ~taken
            assert inr[A.col[mine]].all()
        else:
            mine = inr[A.row] & inr[A.col] & ~taken  # additive split: every entry lives on the lowest rank that sees both DOFs
        taken |= mine
        loc = np.full(P.n, -1, dtype=np.int64)
        loc[nov] = np.arange(nov.size)
        Al = sp.coo_matrix((A.data[mine], (loc[A.row[mine]], loc[A.col[mine]])), shape=(nov.size, nov.size)).tocsr()
        Al.sort_indices()
        Ai = P.A[s][:, s].tocsr()
        Ai.sort_indices()
        attr = lambda idx: np.where(P.owner[idx] == r, 1, 3).astype(np.uint8)  # owner / copy (overlap_extension.hh:261)
        d = dict(rank=np.array([r], dtype=np.int64), nranks=np.array([P.nsub], dtype=np.int64), A_block=np.array([block], dtype=np.int64),
                 A_rowptr=Al.indptr.astype(np.int64), A_col=Al.indices.astype(np.int32), A_val=Al.data,
                 Aovlp_rowptr=Ai.indptr.astype(np.int64), Aovlp_col=Ai.indices.astype(np.int32), Aovlp_val=Ai.data,
                 novlp_gid=gid_of(nov), novlp_attr=attr(nov), ovlp_gid=gid_of(s), ovlp_attr=attr(s), pou=P.pou[r],
                 Z=np.ascontiguousarray(P.Z[r].T), rhs=P.b[nov])
        kind = factor_kinds[r % len(factor_kinds)]
        if kind is not None:
            d.update(cholmod_arrays(P.factors[r], kind, itype=0 if r % 2 else 2))
        shares.append(d)
    assert taken.all()
    return shares


def export_and_merge(exe, P, tmp_path, factor_kinds, gid_of, ghosts=False):
    from ddm_b200.setup import pack, rankpacks

    out = tmp_path / "dump"
    out.mkdir()
    for r, d in enumerate(rank_shares(P, factor_kinds, gid_of, ghosts)):
        src = str(tmp_path / f"in{r}.ddmpack")
        pack.write(src, d)
        res = subprocess.run([exe, src, str(out)], capture_output=True, text=True, timeout=300)
        assert res.returncode == 0, res.stderr
    return rankpacks.merge(str(out))


def same_products(P, Q, ghosts=False, blocks=False):
    assert Q.n == P.n and Q.nsub == P.nsub
    if blocks:  # full 3x3 blocks carry explicit zeros the scalar pattern had pruned: same values, wider pattern
        D = (Q.A - P.A).tocsr()
        assert Q.A.nnz >= P.A.nnz and (D.nnz == 0 or abs(D).max() == 0.0)
    else:
        assert np.array_equal(Q.A.indptr, P.A.indptr) and np.array_equal(Q.A.indices, P.A.indices) and np.array_equal(Q.A.data, P.A.data)
    assert np.array_equal(Q.owner, P.owner) and np.array_equal(Q.b, P.b)
    assert np.array_equal(Q.n_novlp, [s.size for s in P.subs] if ghosts else P.n_novlp)
    for i in range(P.nsub):
        assert np.array_equal(Q.subs[i], P.subs[i]) and np.array_equal(Q.pou[i], P.pou[i]) and np.array_equal(Q.Z[i], P.Z[i])
    assert np.linalg.norm(Q.A0 - P.A0) <= 1e-12 * np.linalg.norm(P.A0)


def test_poisson_ranks_roundtrip(exporter, small3d, tmp_path):
    from ddm_b200.setup import factor
    from oracle import oracle

    P = small3d
    # the reference's global ids are neither ours nor contiguous: any strictly increasing relabelling must be undone
    Q = export_and_merge(exporter, P, tmp_path, ["super", "ll", "ldl"], gid_of=lambda idx: (7 * np.asarray(idx) + 3).astype(np.int64))
    same_products(P, Q)
    rng = np.random.default_rng(2)
    for r in range(P.nsub):
        F, G = P.factors[r], Q.factors[r]
        if r % 3 == 0:  # supernodal factors travel verbatim
            for a in ("perm", "super_", "pi", "px", "s", "x"):
                assert np.array_equal(getattr(F, a), getattr(G, a)), a
        else:  # simplicial ones are regrouped into fundamental supernodes: same L, other panels
            assert np.array_equal(F.perm, G.perm) and G.nsuper <= F.n
            d = rng.standard_normal(F.n)
            x, y = factor.solve_supernodal(F, d), factor.solve_supernodal(G, d)
            assert np.linalg.norm(x - y) <= 1e-12 * np.linalg.norm(x)
    x1, it1, c1, h1 = oracle.Oracle(P, nthreads=2).pcg(P.b, reduction=1e-8, maxit=200)
    x2, it2, c2, h2 = oracle.Oracle(Q, nthreads=2).pcg(Q.b, reduction=1e-8, maxit=200)
    assert c1 and c2 and it1 == it2 and np.linalg.norm(x1 - x2) <= 1e-10 * np.linalg.norm(x1)


def test_nonsymmetric_ranks_refactorise(exporter, cd_dg, tmp_path):
    """No CHOLMOD factor on a non-symmetric problem (the reference uses UMFPack): the merge factorises the exported A_ovlp."""
    from oracle import oracle

    P = cd_dg
    Q = export_and_merge(exporter, P, tmp_path, [None], gid_of=lambda idx: np.asarray(idx, dtype=np.int64) * 2, ghosts=True)
    same_products(P, Q, ghosts=True)
    assert Q.meta["symmetric"] is False and all(hasattr(F, "xu") for F in Q.factors)
    x1, it1, c1, _ = oracle.Oracle(P, nthreads=2).gmres(P.b, reduction=1e-8, restart=30, maxit=300, restricted=True)
    x2, it2, c2, _ = oracle.Oracle(Q, nthreads=2).gmres(Q.b, reduction=1e-8, restart=30, maxit=300, restricted=True)
    assert c1 and c2 and abs(it1 - it2) <= 1 and np.linalg.norm(x1 - x2) <= 1e-6 * np.linalg.norm(x1)


def test_merge_rejects_inconsistent_dumps(exporter, small3d, tmp_path):
    from ddm_b200.setup import pack, rankpacks

    P = small3d
    shares = rank_shares(P, ["super"], gid_of=lambda idx: np.asarray(idx, dtype=np.int64))
    out = tmp_path / "dump"
    out.mkdir()
    for r, d in enumerate(shares[:-1]):  # one rank's file is missing
        src = str(tmp_path / f"in{r}.ddmpack")
        pack.write(src, d)
        assert subprocess.run([exporter, src, str(out)], capture_output=True, timeout=300).returncode == 0
    with pytest.raises(ValueError):
        rankpacks.merge(str(out))
    # a share whose local matrix does not match its index set is refused by the exporter itself
    bad = dict(shares[-1])
    bad["A_rowptr"] = np.concatenate((bad["A_rowptr"], bad["A_rowptr"][-1:]))
    src = str(tmp_path / "bad.ddmpack")
    pack.write(src, bad)
    res = subprocess.run([exporter, src, str(out)], capture_output=True, text=True, timeout=300)
    assert res.returncode == 1 and "differ in size" in res.stderr


def test_elasticity_blocks_roundtrip(exporter, elast, tmp_path):
    """BCRSMatrix<FieldMatrix<double,3,3>> / BlockVector<FieldVector<double,3>> (examples/linearelasticity.cc): the exporter
    flattens node-major blocks, one index-set entry per node."""
    from oracle import oracle

    P = elast
    assert P.meta.get("block") == 3
    for s in P.subs:  # node-major: the three displacement components of a node travel together
        assert s.size % 3 == 0 and np.array_equal(s.reshape(-1, 3), 3 * (s[::3] // 3)[:, None] + np.arange(3))
    Q = export_and_merge(exporter, P, tmp_path, ["super", "ldl"], gid_of=lambda idx: np.asarray(idx, dtype=np.int64) + 3 * 1000)
    same_products(P, Q, blocks=True)
    assert Q.meta["block"] == 3
    x1, it1, c1, _ = oracle.Oracle(P, nthreads=2).pcg(P.b, reduction=1e-8, maxit=300)
    x2, it2, c2, _ = oracle.Oracle(Q, nthreads=2).pcg(Q.b, reduction=1e-8, maxit=300)
    assert c1 and c2 and it1 == it2 and np.linalg.norm(x1 - x2) <= 1e-9 * np.linalg.norm(x1)
