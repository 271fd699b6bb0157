"""The rank-local setup of the Q1 brick family (setup/local.py: every process builds only its own GPU's region of the mesh)
against the one-process setup of the whole problem followed by partition.shard: same index sets, halo plan, operator rows,
partition of unity, subdomain matrices and coarse bases.  Integer products must be identical (north_star: "index sets,
overlap maps ... bit-exact"), floating-point products agree to round-off of the assembly order."""
import numpy as np
import pytest


def shapes(world, ncell=8, p=2):
    g = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[world]
    return tuple(ncell * k for k in g), tuple(p * k for k in g), g


@pytest.mark.parametrize("world,coarse", [(1, "pou"), (2, "pou"), (4, "pou"), (8, "pou"), (2, "geneo")])
def test_local_products_equal_sharded_global_products(world, coarse):
    from ddm_b200.partition import shard
    from ddm_b200.setup import local, problem

    shape, pg, g = shapes(world)
    P = problem.poisson_q1_cube(shape if world > 1 else shape[0], pg, group=g, overlap=2, coarse=coarse, nev=4, workers=1)
    for r in range(world):
        S = shard(P, r, world)
        L = local.poisson_q1_cube_local(shape, pg, g, r, overlap=2, coarse=coarse, nev=4, workers=1, numeric="host", eigensolver="host")
        T = L.shard
        assert L.n_global == P.n and L.nsub_global == P.nsub
        for name in ("owned", "ghosts", "peers", "send_ptr", "send_idx", "recv_ptr", "rowptr", "col", "sub_ptr", "sub_dofs"):
            assert np.array_equal(getattr(S, name), getattr(T, name)), (r, name)
        assert T.sub_range == S.sub_range
        scale = np.abs(S.val).max()
        assert np.abs(S.val - T.val).max() <= 1e-14 * scale
        assert np.abs(L.b_own - P.b[S.owned]).max() <= 1e-13 * max(np.abs(P.b).max(), 1.0)
        a, b = S.sub_range
        for j, i in enumerate(range(a, b)):
            assert np.abs(L.pou[j] - P.pou[i]).max() <= 1e-15
            Fg, Fl = P.factors[i], L.factors[j]
            assert np.array_equal(Fg.perm, Fl.perm) and np.array_equal(Fg.s, Fl.s) and np.array_equal(Fg.px, Fl.px)
            assert np.abs(Fg.x - Fl.x).max() <= 1e-11 * np.abs(Fg.x).max()
            Zg, Zl = P.Z[i], L.Z[j]
            assert Zg.shape == Zl.shape
            if coarse == "pou":
                assert np.abs(Zg - Zl).max() <= 1e-14
            else:  # eigenvectors: same subspace (signs and rotations inside clusters are free)
                Q, _ = np.linalg.qr(Zg)
                assert np.linalg.norm(Zl - Q @ (Q.T @ Zl)) <= 1e-6 * np.linalg.norm(Zl)


def test_congruent_bricks_share_one_symbolic_analysis():
    from ddm_b200.setup import local

    L = local.poisson_q1_cube_local((16, 16, 16), (4, 4, 4), (1, 1, 1), 0, overlap=1, coarse="none", workers=1, numeric="device")
    ans = {id(F.an) for F in L.factors}
    assert len(L.factors) == 64 and len(ans) == 27  # per axis: first, inner, last brick; the 8 inner bricks differ by a translation only
    for F in L.factors:
        assert F.A.shape[0] == F.an.n


def test_brick_coloring_covers_every_nonzero_galerkin_block():
    """ddm_set_coarse_coloring's contract for the brick family: for every subdomain m and colour c at most one subdomain of colour
    c has a non-zero Galerkin block with m, and owner_of names it (local.brick_coloring); checked against the dense A0 of the
    one-process setup.  The generic colouring from the coupling graph (coarse.coupling_coloring) must satisfy the same."""
    from ddm_b200.setup import coarse, local, problem

    pg, g = (4, 4, 2), (2, 2, 1)
    P = problem.poisson_q1_cube((20, 20, 10), pg, group=g, overlap=2, coarse="pou", workers=1)
    off = np.concatenate(([0], np.cumsum([z.shape[1] for z in P.Z])))
    for color, nc, owner in (local.brick_coloring(pg, g), coarse.coupling_coloring(P.A, P.subs)):
        assert color.min() >= 0 and color.max() < nc and owner.shape == (P.nsub, nc)
        for m in range(P.nsub):
            assert owner[m, color[m]] == m
            for i in range(P.nsub):
                if np.abs(P.A0[off[m] : off[m + 1], off[i] : off[i + 1]]).max() > 0:
                    assert owner[m, color[i]] == i, (m, i)
    assert local.brick_coloring(pg, g)[1] == 27
