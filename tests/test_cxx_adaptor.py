"""Dune::ddm_b200::Context::fromPack (dune-ddm_b200/dune/ddm_b200/adaptors.hh) without a GPU: tests/cxx/pack_upload_check.cc is
linked against recording stand-ins of the C-ABI upload entry points, and what the C++ adaptor hands to the library is compared
with the Python view of the same .ddmpack -- scalar CSR, the 3x3 block-CSR conversion of the elasticity operator (against
scipy's tobsr), Cholesky and LU factor arrays.  Needs the Dune headers, so it runs where /root/reference is present."""
import os
import shutil
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
DUNE = [f"{REF}/extern/dune-common", f"{REF}/extern/dune-istl"]

pytestmark = pytest.mark.skipif(not (all(os.path.isdir(d) for d in DUNE) and shutil.which("g++")),
                                reason="needs the reference's dune-common / dune-istl headers and g++")


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    d = tmp_path_factory.mktemp("cxx")
    stubs = d / "stubs"
    stubs.mkdir()
    for name in ("dune-common-config.hh", "dune-istl-config.hh"):
        (stubs / name).write_text("#pragma once\n")
    exe = str(d / "pack_upload_check")
    cmd = ["g++", "-std=c++20", "-O1", "-DHAVE_MPI=0", f"-I{stubs}", *[f"-I{p}" for p in DUNE], f"-I{ROOT}/include", f"-I{ROOT}/dune-ddm_b200",
           f"{ROOT}/tests/cxx/pack_upload_check.cc", *[f"{DUNE[0]}/dune/common/{f}.cc" for f in ("exceptions", "stdstreams", "debugallocator")], "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    return exe


def recorded(exe, P, tmp_path):
    from ddm_b200.setup import pack

    src, out = str(tmp_path / "p.ddmpack"), str(tmp_path / "p.dump")
    pack.save_products(src, P)
    r = subprocess.run([exe, src, out], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    raw, d, o = open(out, "rb").read(), {}, 0
    while o < len(raw):
        (nl,) = struct.unpack_from("<Q", raw, o)
        name = raw[o + 8 : o + 8 + nl].decode()
        (nb,) = struct.unpack_from("<Q", raw, o + 8 + nl)
        d[name] = raw[o + 16 + nl : o + 16 + nl + nb]
        o += 16 + nl + nb
    return d


def arr(d, k, dt):
    return np.frombuffer(d[k], dtype=dt)


def check_factors(d, P):
    for i, F in enumerate(P.factors):
        f = f"F{i}_"
        assert arr(d, f + "n", np.int64)[0] == P.subs[i].size
        assert np.array_equal(arr(d, f + "super", np.int32), F.super_) and np.array_equal(arr(d, f + "pi", np.int64), F.pi)
        assert np.array_equal(arr(d, f + "px", np.int64), F.px) and np.array_equal(arr(d, f + "s", np.int32), F.s)
        assert np.array_equal(arr(d, f + "x", np.float64), F.x) and np.array_equal(arr(d, f + "perm", np.int32), F.perm)
        xu = getattr(F, "xu", None)
        assert arr(d, f + "lu", np.int64)[0] == (xu is not None)
        if xu is not None:
            assert np.array_equal(arr(d, f + "xu", np.float64), xu)
    assert arr(d, "N0", np.int64)[0] == P.A0.shape[0] and np.array_equal(arr(d, "A0", np.float64), np.ascontiguousarray(P.A0).ravel())
    assert arr(d, "finalized", np.int64)[0] == 1


def test_scalar_cholesky_pack(checker, small3d, tmp_path):
    P = small3d
    d = recorded(checker, P, tmp_path)
    assert arr(d, "op_kind", np.int64)[0] == 1 and arr(d, "op_n_own", np.int64)[0] == P.n and arr(d, "op_n_ghost", np.int64)[0] == 0
    assert np.array_equal(arr(d, "op_rowptr", np.int64), P.A.indptr) and np.array_equal(arr(d, "op_col", np.int32), P.A.indices)
    assert np.array_equal(arr(d, "op_val", np.float64), P.A.data)
    assert np.array_equal(arr(d, "sub_dofs", np.int32), np.concatenate(P.subs)) and np.array_equal(arr(d, "pou", np.float64), np.concatenate(P.pou))
    check_factors(d, P)


def test_lu_pack(checker, cd_dg, tmp_path):
    d = recorded(checker, cd_dg, tmp_path)
    assert arr(d, "op_kind", np.int64)[0] == 1
    check_factors(d, cd_dg)
    assert all(arr(d, f"F{i}_lu", np.int64)[0] == 1 for i in range(len(cd_dg.factors)))


def test_block3_pack_matches_scipy_bsr(checker, elast, tmp_path):
    P = elast
    assert P.meta.get("block") == 3
    d = recorded(checker, P, tmp_path)
    nb = P.n // 3
    assert arr(d, "op_kind", np.int64)[0] == 3 and arr(d, "op_n_own", np.int64)[0] == nb
    B = P.A.tobsr((3, 3))
    B.sort_indices()
    rp, ci, va = arr(d, "op_rowptr", np.int64), arr(d, "op_col", np.int32), arr(d, "op_val", np.float64).reshape(-1, 3, 3)
    # the adaptor keeps every block the scalar pattern touches; scipy does the same (explicit zeros stay in the pattern)
    assert np.array_equal(rp, B.indptr) and np.array_equal(ci, B.indices) and np.array_equal(va, B.data)
    check_factors(d, P)


def _locate(raw, name):
    """Offsets of (element code, rank, first dimension) of array `name` in a .ddmpack byte string (layout: setup/pack.py)."""
    key = struct.pack("<i", len(name)) + name.encode()
    o = raw.index(key) + len(key)
    return o, o + 4, o + 8


def test_damaged_packs_are_refused(checker, small3d, tmp_path):
    """the C++ reader must not trust the file -- a truncated pack, an unknown element code, a dimension larger than the
    file and an array of the wrong element type each end in an error message, not in an out-of-bounds read."""
    from ddm_b200.setup import pack

    src = str(tmp_path / "good.ddmpack")
    pack.save_products(src, small3d)
    good = open(src, "rb").read()
    code_o, ndim_o, dim_o = _locate(good, "A_rowptr")
    assert struct.unpack_from("<i", good, code_o)[0] == 2 and struct.unpack_from("<i", good, ndim_o)[0] == 1

    def run(raw):
        bad = str(tmp_path / "bad.ddmpack")
        open(bad, "wb").write(raw)
        return subprocess.run([checker, bad, str(tmp_path / "bad.dump")], capture_output=True, text=True, timeout=120)

    cases = {
        "truncated": (good[: len(good) // 2], "ddmpack"),
        "unknown element code": (good[:code_o] + struct.pack("<i", 9) + good[code_o + 4 :], "unknown element type"),
        "dimension beyond the file": (good[:dim_o] + struct.pack("<q", 1 << 50) + good[dim_o + 8 :], "larger than the file"),
        "negative dimension": (good[:dim_o] + struct.pack("<q", -5) + good[dim_o + 8 :], "larger than the file"),
        "wrong element type": (good[:code_o] + struct.pack("<i", 0) + good[code_o + 4 :], "wrong element type"),  # f64 has the same width
        "foreign file": (b"NOTAPACK" + good[8:], "not a .ddmpack"),
    }
    for what, (raw, text) in cases.items():
        r = run(raw)
        assert r.returncode == 4, (what, r.returncode, r.stderr[-300:])
        assert text in r.stderr, (what, r.stderr[-300:])
    assert run(good).returncode == 0
