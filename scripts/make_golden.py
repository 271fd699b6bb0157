"""Extracts the reference's known-answer literals into tests/golden/ (run in the build container only).

Source: /root/reference/tests/test_galerkin_coarse_matrix.cc
  :21-48   9x9 tridiagonal test matrix (MatrixMarket literal)
  :51-67   4x4 coarse matrix A0 for the template vector 1 and overlap 1 (MatrixMarket literal)
  :152-176 the four ranks' non-overlapping index sets (global ids of local indices 0,1,2)
  :219-244 the hand-written partition of unity on the overlap-1 index lists
and /root/reference/tests/test_build_matrix.cc:21-81 (gatherMatrixFromRows layout: A[i][j] = i + j,
ranks contribute 2/3 rows alternately).
The GPU box has no /root/reference, so the literals are committed as small fixtures.
"""
import json
import os
import re

REF = "/root/reference/tests/test_galerkin_coarse_matrix.cc"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    src = open(REF).read()
    lits = re.findall(r'R"\((%%MatrixMarket.*?)\)"', src, flags=re.S)
    assert len(lits) == 2
    open(os.path.join(OUT, "test_matrix_9x9.mtx"), "w").write(lits[0] + "\n")
    open(os.path.join(OUT, "coarse_matrix_4x4.mtx"), "w").write(lits[1] + "\n")
    # index sets: paridxs.add(global, {local, attribute, public}) per rank
    ranks = re.split(r"helper\.rank\(\) == \d\)|else \{", src[src.index("paridxs.beginResize") : src.index("paridxs.endResize")])
    novlp, attrs = [], []
    for blk in ranks[1:]:
        adds = re.findall(r"paridxs\.add\((\d+), \{(\d+), Attribute::(\w+)", blk)
        loc = sorted((int(l), int(g), a) for g, l, a in adds)
        novlp.append([g for _, g, _ in loc])
        attrs.append([a for _, _, a in loc])
    # partition of unity: pou[k] = value per rank, after 'Dune::BlockVector<double> pou('
    ptxt = src[src.index("Dune::BlockVector<double> pou(") : src.index("auto pou_sum")]
    pblk = re.split(r"helper\.rank\(\) == \d\)|else \{", ptxt)[1:]
    pou = []
    for blk in pblk:
        vals = re.findall(r"pou\[(\d+)\] = ([^;]+);", blk)
        pou.append([eval(v.replace("1.", "1.0")) for _, v in sorted(vals, key=lambda t: int(t[0]))])
    json.dump(dict(novlp=novlp, attributes=attrs, pou_overlap1=pou, source=REF + ":21-67,152-176,219-244"),
              open(os.path.join(OUT, "galerkin_case.json"), "w"), indent=1)
    print(novlp, attrs, pou)


if __name__ == "__main__":
    main()
