"""Setup restatement against the reference's own known-answer literals (tests/golden/, extracted by
scripts/make_golden.py from /root/reference/tests/test_galerkin_coarse_matrix.cc:21-67,152-176,219-283
and /root/reference/tests/test_build_matrix.cc:21-81)."""
import json
import os

import numpy as np
import scipy.io
import scipy.sparse as sp

from ddm_b200.setup import coarse, decomp

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _case():
    A = sp.csr_matrix(scipy.io.mmread(os.path.join(G, "test_matrix_9x9.mtx"), spmatrix=True))
    A0 = scipy.io.mmread(os.path.join(G, "coarse_matrix_4x4.mtx"), spmatrix=True).toarray()
    meta = json.load(open(os.path.join(G, "galerkin_case.json")))
    novlp = [np.array(v, dtype=np.int64) for v in meta["novlp"]]
    return A, A0, novlp, meta


def test_overlap_6_layers_reproduces_the_full_matrix():
    # test_galerkin_coarse_matrix.cc:198-212: after 6 layers rank 0 holds the whole 9x9 matrix
    A, _, novlp, _ = _case()
    subs = decomp.grow_overlap(A, novlp, 6)
    s0 = subs[0]
    assert s0.tolist() == list(range(9))  # novlp first, then layers in discovery order
    Aov = A[s0][:, s0].toarray()
    assert np.linalg.norm(Aov - A.toarray()) <= 1e-16


def test_overlap_1_index_lists_keep_novlp_first_then_new_layer():
    # overlap_extension.hh:100-104,260-262: new indices are appended after the rank's own
    A, _, novlp, _ = _case()
    subs = decomp.grow_overlap(A, novlp, 1)
    assert [s.tolist() for s in subs] == [[0, 1, 2, 3], [2, 3, 4, 1, 5], [4, 5, 6, 3, 7], [6, 7, 8, 5]]


def test_owner_is_lowest_rank():
    # pdelab_helper.hh:44-82 / the attributes of test_galerkin_coarse_matrix.cc:152-176
    _, _, novlp, meta = _case()
    owner = np.full(9, 4)
    for r in range(3, -1, -1):
        owner[novlp[r]] = r
    for r, (idx, att) in enumerate(zip(novlp, meta["attributes"])):
        assert [("owner" if owner[g] == r else "copy") for g in idx] == att


def test_handwritten_pou_sums_to_one():
    # test_galerkin_coarse_matrix.cc:219-263
    A, _, novlp, meta = _case()
    subs = decomp.grow_overlap(A, novlp, 1)
    tot = np.zeros(9)
    for s, w in zip(subs, meta["pou_overlap1"]):
        np.add.at(tot, s, np.asarray(w, dtype=np.float64))
    assert np.all(np.abs(tot - 1.0) <= 2e-16)


def test_coarse_matrix_literal():
    # test_galerkin_coarse_matrix.cc:265-283: A0 for the template vector 1, overlap 1, tol 1e-12
    A, A0_expected, novlp, meta = _case()
    subs = decomp.grow_overlap(A, novlp, 1)
    Z = [np.asarray(w, dtype=np.float64)[:, None] * np.ones((len(w), 1)) for w in meta["pou_overlap1"]]
    A0, off, _ = coarse.galerkin_matrix(A, subs, Z)
    assert off.tolist() == [0, 1, 2, 3, 4]  # exclusive scan of k_i in rank order, galerkin_preconditioner.hh:246-256
    assert np.linalg.norm(A0 - A0_expected) <= 1e-12
    # sparsity = neighbour-block structure (helpers.hh:204-253 drops exact zeros)
    assert ((A0 != 0) == (A0_expected != 0)).all()


def test_gather_rows_layout():
    # test_build_matrix.cc:21-81: ranks contribute 2/3 rows alternately, A[i][j] = i + j; the gathered
    # matrix lists the ranks' rows in rank order -- the ordering galerkin_matrix uses for A0 blocks.
    for nranks in (1, 4, 16, 32):
        counts = [2 if r % 2 == 0 else 3 for r in range(nranks)]
        off = np.concatenate(([0], np.cumsum(counts)))
        n = int(off[-1])
        rows = [np.add.outer(np.arange(off[r], off[r + 1]), np.arange(n)).astype(float) for r in range(nranks)]
        M = np.vstack(rows)
        i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
        assert (M == i + j).all()
