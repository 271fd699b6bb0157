"""Test infrastructure: the rank-local sequence of the multi-GPU path, replayed with numpy.

Mirrors, step for step, what libddm_b200.so does on one rank (halo_forward -> restrict -> subdomain solves
-> prolong -> halo_reverse_add; ddm_b200.cu) but with the oracle's CPU subdomain solve, so that the host-side
sharding logic (ddm_b200/partition.py: ownership, ghost numbering, halo plan, coarse offsets) can be checked
without a GPU, in one process or under torch.distributed/gloo.
"""
import numpy as np


class RankEmu:
    def __init__(self, P, S, O):
        self.P, self.S, self.O = P, S, O
        self.nloc = S.n_own + S.n_ghost

    # -- the two halo primitives; `exchange(sendbufs: {peer: array}, direction) -> {peer: array}` moves the data
    def halo_forward(self, v, exchange):
        S = self.S
        send = {int(p): v[S.send_idx[S.send_ptr[k] : S.send_ptr[k + 1]]].copy() for k, p in enumerate(S.peers)}
        recv = exchange(send, "fwd")
        for k, p in enumerate(S.peers):
            a, b = S.recv_ptr[k], S.recv_ptr[k + 1]
            if b > a:
                v[S.n_own + a : S.n_own + b] = recv[int(p)]

    def halo_reverse_add(self, v, exchange):
        S = self.S
        send = {int(p): v[S.n_own + S.recv_ptr[k] : S.n_own + S.recv_ptr[k + 1]].copy() for k, p in enumerate(S.peers)}
        recv = exchange(send, "rev")
        for k, p in enumerate(S.peers):  # fixed peer order
            idx = S.send_idx[S.send_ptr[k] : S.send_ptr[k + 1]]
            if idx.size:
                np.add.at(v, idx, recv[int(p)])

    def local(self, owned_values):
        v = np.zeros(self.nloc)
        v[: self.S.n_own] = owned_values
        return v

    def spmv(self, v, exchange):
        import scipy.sparse as sp

        S = self.S
        self.halo_forward(v, exchange)
        A = sp.csr_matrix((S.val, S.col, S.rowptr), shape=(S.n_own, self.nloc))
        return A @ v

    def twolevel(self, d, exchange, allreduce, restricted=False, with_coarse=True):
        """Additive two-level apply on this rank; returns the owned part of x."""
        S, P = self.S, self.P
        a, b = S.sub_range
        self.halo_forward(d, exchange)
        x = np.zeros(self.nloc)
        N0 = P.A0.shape[0]
        d0 = np.zeros(N0)
        koff = S.k_offset
        loc = []
        for j, i in enumerate(range(a, b)):
            idx = S.sub_dofs[S.sub_ptr[j] : S.sub_ptr[j + 1]]
            di = d[idx]
            loc.append((i, idx, di))
            k = P.Z[i].shape[1]
            d0[koff : koff + k] = P.Z[i].T @ di
            koff += k
        if with_coarse:
            d0 = allreduce(d0)
            y0 = np.linalg.solve(P.A0, d0)
        koff = S.k_offset
        for i, idx, di in loc:
            xi = self.O.sub_solve(i, di)
            if restricted:
                xi = xi * P.pou[i]
            if with_coarse:
                k = P.Z[i].shape[1]
                xi = xi + P.Z[i] @ y0[koff : koff + k]
                koff += k
            np.add.at(x, idx, xi)
        self.halo_reverse_add(x, exchange)
        return x[: S.n_own]
