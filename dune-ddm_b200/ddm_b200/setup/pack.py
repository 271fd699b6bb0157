"""`.ddmpack`: flat binary container of the setup products ("checkpoint").

The C++ side (dune-ddm_b200/dune/ddm_b200/ddmpack.hh) reads the same file, so a Dune program, the oracle
and the Python harness all start from bit-identical inputs.  Layout (little endian):
    8 bytes  magic "DDMPACK1"
    int64    number of arrays
    per array: int32 name length, name bytes, int32 dtype (0 f64, 1 i32, 2 i64, 3 u8), int32 ndim,
               int64 dims[ndim], raw data in C order
"""
from __future__ import annotations

import struct

import numpy as np

MAGIC = b"DDMPACK1"
_CODES = {np.dtype(np.float64): 0, np.dtype(np.int32): 1, np.dtype(np.int64): 2, np.dtype(np.uint8): 3}
_TYPES = {v: k for k, v in _CODES.items()}


def write(path, arrays: dict):
    with open(path, "wb") as fh:
        fh.write(MAGIC)
        fh.write(struct.pack("<q", len(arrays)))
        for name, a in arrays.items():
            a = np.ascontiguousarray(a)
            if a.dtype == np.bool_:
                a = a.astype(np.uint8)
            code = _CODES[a.dtype]
            nb = name.encode()
            fh.write(struct.pack("<i", len(nb)))
            fh.write(nb)
            fh.write(struct.pack("<ii", code, a.ndim))
            fh.write(struct.pack(f"<{a.ndim}q", *a.shape))
            fh.write(a.tobytes())


def read(path) -> dict:
    out = {}
    with open(path, "rb") as fh:
        if fh.read(8) != MAGIC:
            raise ValueError(f"{path} is not a .ddmpack file")
        (count,) = struct.unpack("<q", fh.read(8))
        for _ in range(count):
            (ln,) = struct.unpack("<i", fh.read(4))
            name = fh.read(ln).decode()
            code, ndim = struct.unpack("<ii", fh.read(8))
            dims = struct.unpack(f"<{ndim}q", fh.read(8 * ndim))
            dt = _TYPES[code]
            n = int(np.prod(dims)) if ndim else 1
            out[name] = np.frombuffer(fh.read(n * dt.itemsize), dtype=dt).reshape(dims).copy()
    return out


def products_to_arrays(P) -> dict:
    """Flatten SetupProducts into the arrays the C ABI takes (whole problem, one rank)."""
    ns = P.nsub
    d = dict(
        n=np.array([P.n], dtype=np.int64),
        nsub=np.array([ns], dtype=np.int64),
        A_rowptr=P.A.indptr.astype(np.int64),
        A_col=P.A.indices.astype(np.int32),
        A_val=P.A.data.astype(np.float64),
        b=np.asarray(P.b, dtype=np.float64),
        owner=np.asarray(P.owner, dtype=np.int32),
        sub_ptr=np.concatenate(([0], np.cumsum([s.size for s in P.subs]))).astype(np.int64),
        sub_dofs=np.concatenate(P.subs).astype(np.int32),
        pou=np.concatenate(P.pou).astype(np.float64),
        n_novlp=np.asarray(P.n_novlp, dtype=np.int64),
        k=np.array([z.shape[1] for z in P.Z], dtype=np.int32),
        Z=np.concatenate([z.ravel(order="F") for z in P.Z]).astype(np.float64) if ns else np.zeros(0),
        A0=np.ascontiguousarray(P.A0, dtype=np.float64),
        block=np.array([int(P.meta.get("block", 1))], dtype=np.int64),
    )
    for i, F in enumerate(P.factors):
        d[f"F{i}_super"] = np.asarray(F.super_, dtype=np.int32)
        d[f"F{i}_pi"] = np.asarray(F.pi, dtype=np.int64)
        d[f"F{i}_px"] = np.asarray(F.px, dtype=np.int64)
        d[f"F{i}_s"] = np.asarray(F.s, dtype=np.int32)
        d[f"F{i}_x"] = np.asarray(F.x, dtype=np.float64)
        d[f"F{i}_perm"] = np.asarray(F.perm, dtype=np.int32)
        if hasattr(F, "xu"):  # LU factor of a non-symmetric subdomain matrix
            d[f"F{i}_xu"] = np.asarray(F.xu, dtype=np.float64)
    return d


def save_products(path, P):
    write(path, products_to_arrays(P))


def load_products(path):
    """Inverse of save_products (coords / dirichlet / meta are not stored: they are setup-only)."""
    import scipy.sparse as sp

    from . import factor
    from .problem import SetupProducts

    d = read(path)
    n, ns = int(d["n"][0]), int(d["nsub"][0])
    A = sp.csr_matrix((d["A_val"], d["A_col"], d["A_rowptr"]), shape=(n, n))
    sub_ptr, k = d["sub_ptr"], d["k"]
    subs = [d["sub_dofs"][sub_ptr[i] : sub_ptr[i + 1]].astype(np.int64) for i in range(ns)]
    pou = [d["pou"][sub_ptr[i] : sub_ptr[i + 1]] for i in range(ns)]
    Z, zo = [], 0
    for i in range(ns):
        ni = int(sub_ptr[i + 1] - sub_ptr[i])
        Z.append(np.asfortranarray(d["Z"][zo : zo + ni * int(k[i])].reshape((ni, int(k[i])), order="F")))
        zo += ni * int(k[i])
    factors = []
    for i in range(ns):
        kw = dict(n=subs[i].size, perm=d[f"F{i}_perm"], super_=d[f"F{i}_super"], pi=d[f"F{i}_pi"], px=d[f"F{i}_px"], s=d[f"F{i}_s"], x=d[f"F{i}_x"])
        factors.append(factor.SupernodalLU(xu=d[f"F{i}_xu"], **kw) if f"F{i}_xu" in d else factor.SupernodalFactor(**kw))
    return SetupProducts(A=A, b=d["b"], owner=d["owner"], subs=subs, pou=pou, factors=factors, Z=Z, A0=d["A0"],
                         n_novlp=d["n_novlp"], meta=dict(n=n, nsub=ns, N0=int(d["A0"].shape[0]), block=int(d["block"][0]) if "block" in d else 1))
