"""tests/golden/c1_unitsquare.npz: the mesh-dependent setup products of config C1 exactly as BASELINE.json states it -- P1 Poisson on the
reference's data/unitsquare.msh (7 441 nodes), islands coefficient, 16 subdomains, overlap 2, GenEO nev 16 -- produced by the
host-side setup restatement on the CPU container: operator, right-hand side, index sets, partition of unity, GenEO basis.
The GPU box has no /root/reference, so these travel as a (compressed) fixture; subdomain factors and the Galerkin matrix follow
from them deterministically and are rebuilt by the loader (tests/test_zz_gpu_c1_mesh.py::load_c1), which keeps the file small.

  python scripts/make_c1_fixture.py        (needs /root/reference/data/unitsquare.msh)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))


def main():
    import numpy as np

    from ddm_b200.setup import problem

    P = problem.poisson_gmsh("/root/reference/data/unitsquare.msh", 2, 16, refine=0, overlap=2, nev=16, workers=1)
    out = os.path.join(ROOT, "tests", "golden", "c1_unitsquare.npz")
    sub_ptr = np.concatenate(([0], np.cumsum([s.size for s in P.subs])))
    np.savez_compressed(out, A_indptr=P.A.indptr, A_indices=P.A.indices, A_data=P.A.data, b=P.b, owner=P.owner, n_novlp=P.n_novlp,
                        sub_ptr=sub_ptr, sub_dofs=np.concatenate(P.subs), pou=np.concatenate(P.pou),
                        k=np.array([z.shape[1] for z in P.Z]), Z=np.concatenate([z.ravel(order="F") for z in P.Z]))
    print(out, os.path.getsize(out), "bytes; n =", P.n, "N0 =", P.A0.shape[0])


if __name__ == "__main__":
    main()
