"""tests/golden/exported_ranks.ddmpack: a 4-rank problem that went through the dune-ddm exporter path on the CPU container --
rank-wise split, exportRank on the reference's Dune containers (tests/cxx/export_check.cc), merge (ddm_b200/setup/rankpacks.py).
Ranks 0..3 carry a supernodal LL^T, a packed simplicial LL^T, a non-packed simplicial LDL^T and again a supernodal factor in
CHOLMOD's layout, so the merged pack holds regrouped fundamental supernodes next to verbatim panels.  The GPU box has no
/root/reference: tests/test_zz_gpu_exported.py loads this file.

  python scripts/make_exported_fixture.py        (needs /root/reference/extern and g++)
"""
import os
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dune-ddm_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

import numpy as np


def main():
    import test_exporter as T
    from ddm_b200.setup import pack, problem

    P = problem.unitcube_p1(7, 4, overlap=1, nev=4, workers=1)
    with tempfile.TemporaryDirectory() as d:
        d = Path(d)
        stubs = d / "stubs"
        stubs.mkdir()
        for name in ("dune-common-config.hh", "dune-istl-config.hh"):
            (stubs / name).write_text("#pragma once\n")
        exe = str(d / "export_check")
        cmd = ["g++", "-std=c++20", "-O1", "-DHAVE_MPI=0", f"-I{stubs}", *[f"-I{p}" for p in T.DUNE], f"-I{ROOT}/include", f"-I{ROOT}/dune-ddm_b200",
               f"{ROOT}/tests/cxx/export_check.cc", *[f"{T.DUNE[0]}/dune/common/{f}.cc" for f in ("exceptions", "stdstreams", "debugallocator")], "-o", exe]
        subprocess.run(cmd, check=True)
        Q = T.export_and_merge(exe, P, d, ["super", "ll", "ldl"], gid_of=lambda idx: (5 * np.asarray(idx) + 11).astype(np.int64))
    T.same_products(P, Q)
    out = os.path.join(ROOT, "tests", "golden", "exported_ranks.ddmpack")
    pack.save_products(out, Q)
    print(out, os.path.getsize(out), "bytes; n =", Q.n, "supernodes per rank:", [F.nsuper for F in Q.factors], "original:", [F.nsuper for F in P.factors])


if __name__ == "__main__":
    main()
