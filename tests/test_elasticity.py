"""Config C3 on the CPU side: the P1 elasticity element matrices, the node-major (3x3 block) layout and its sharding."""
import numpy as np


def test_element_matrices_have_the_rigid_body_kernel():
    from ddm_b200.setup import elasticity as el
    from ddm_b200.setup import mesh

    coords, cells = mesh.structured_simplices((2, 2, 2), lengths=(1.0, 0.5, 2.0))
    lam, mu = np.full(cells.shape[0], 3.0), np.full(cells.shape[0], 2.0)
    Ke, vol = el.p1_elasticity_element_matrices(coords, cells, lam, mu)
    assert np.isclose(vol.sum(), 1.0)
    assert np.allclose(Ke, np.transpose(Ke, (0, 2, 1)))
    x = coords[cells]  # (m, 4, 3)
    modes = [np.tile(np.eye(3)[c], (cells.shape[0], 4, 1)) for c in range(3)]
    for a, b in ((0, 1), (0, 2), (1, 2)):  # infinitesimal rotations
        r = np.zeros_like(x)
        r[..., a], r[..., b] = -x[..., b], x[..., a]
        modes.append(r)
    for m in modes:
        assert np.abs(np.einsum("mij,mj->mi", Ke, m.reshape(cells.shape[0], 12))).max() <= 1e-12 * np.abs(Ke).max()
    # uniaxial strain u = (x, 0, 0): energy = vol (lam + 2 mu)
    u = np.zeros_like(x)
    u[..., 0] = x[..., 0]
    uf = u.reshape(cells.shape[0], 12)
    assert np.isclose(np.einsum("mi,mij,mj->", uf, Ke, uf), 1.0 * (3.0 + 2 * 2.0))


def test_problem_layout_and_block_sharding(elast):
    import scipy.sparse as sp

    from ddm_b200 import partition

    P = elast
    assert P.meta["block"] == 3 and P.n % 3 == 0
    assert abs(P.A - P.A.T).max() <= 1e-12 * abs(P.A).max()
    B = P.A.tobsr((3, 3))
    assert 9 * B.indices.size < 1.25 * P.A.nnz  # node-major numbering: the scalar pattern is (nearly) full 3x3 blocks
    for i, s in enumerate(P.subs):  # subdomains hold whole nodes
        assert s.size % 3 == 0 and np.array_equal(np.unique(s // 3).repeat(3) * 3 + np.tile(np.arange(3), s.size // 3), np.sort(s))
    for world in (1, 2, 3):
        for r in range(world):
            S = partition.shard(P, r, world)
            assert S.n_own % 3 == 0 and S.n_ghost % 3 == 0
            assert not (S.owned[::3] % 3).any() and not (S.ghosts[::3] % 3).any()
            assert np.array_equal(S.owned[1::3], S.owned[::3] + 1) and np.array_equal(S.ghosts[2::3], S.ghosts[::3] + 2)


def test_oracle_solves_the_elasticity_problem(elast):
    from oracle import oracle

    P = elast
    O = oracle.Oracle(P, nthreads=2)
    x, it, conv, _ = O.pcg(P.b, reduction=1e-8, maxit=300)
    assert conv and it < 60
    x1, it1, conv1, _ = O.pcg(P.b, reduction=1e-8, maxit=300, with_coarse=False)
    assert conv1 and it1 > it  # the GenEO level pays off on the high-contrast bars
    assert np.linalg.norm(P.b - P.A @ x) <= 1.05e-8 * np.linalg.norm(P.b)
