"""TEST INFRASTRUCTURE - OracleEmulator for graphs too large for dense (N+M)^2 filters (~100k nodes).

Same arithmetic as oracle/emulator_oracle.py (the dense restatement of emulator.py); only the two places where the
reference multiplies by a dense graph matrix are re-expressed with scipy.sparse so that they fit in memory:
``filter @ (x @ W)`` of the joint GCN (emulator.py:220-223 with spektral's gcn_filter, restated on the CSR adjacency) and
the incidence products of post_proc (emulator.py:643-678).  graph_base=1, GCN only.
"""
import numpy as np
import scipy.sparse as sp

from .emulator_oracle import OracleEmulator, act_fn


def gcn_filter_sparse(rp, ci, val, n):
    """spektral.utils.gcn_filter on a CSR adjacency: D^-1/2 (A + I) D^-1/2, D = rowsum(A + I), inf -> 0 (float64)."""
    A = sp.csr_matrix((np.asarray(val, np.float64), np.asarray(ci), np.asarray(rp)), shape=(n, n)) + sp.identity(n, format="csr")
    with np.errstate(divide="ignore"):
        d = np.power(np.asarray(A.sum(1)).ravel(), -0.5)
    d[np.isinf(d)] = 0.0
    return (sp.diags(d) @ A @ sp.diags(d)).tocsr()


class SparseOracleEmulator(OracleEmulator):
    def __init__(self, conv, resnet, recurrent, args, dtype=np.float64):
        assert "GCN" in conv and getattr(args, "graph_base", 0) == 1, "sparse oracle: joint-graph GCN only"
        csr = args.adj_csr
        keep = {k: getattr(args, k, None) for k in ("adj", "edge_adj", "node_edge")}
        args.adj, args.edge_adj, args.node_edge = np.zeros((1, 1)), np.zeros((1, 1)), np.zeros((1, 1))
        try:
            super().__init__(conv, resnet, recurrent, args, dtype=dtype)
        finally:
            for k, v in keep.items():
                setattr(args, k, v)
        n = self.n_node + self.n_edge
        self._filt = gcn_filter_sparse(csr.row_ptr, csr.col_idx, csr.val if csr.val is not None else np.ones(len(csr.col_idx)), n).astype(self.dtype)
        e = np.asarray(args.edges)
        M, N = self.n_edge, self.n_node
        m = np.arange(M)
        ok = e[:, 0] != e[:, 1]
        self._pos = sp.csr_matrix((np.ones(ok.sum()), (e[ok, 0], m[ok])), shape=(N, M)).astype(self.dtype)   # clip(node_edge, 0, 1)
        self._neg = sp.csr_matrix((np.ones(ok.sum()), (e[ok, 1], m[ok])), shape=(N, M)).astype(self.dtype)   # |clip(node_edge, -1, 0)|
        self.node_edge = np.zeros((1, 1), np.float32)

    def _conv(self, x, filt, w):                                     # gcn_conv with the sparse filter, sample by sample
        h = np.matmul(x, w[0])
        out = np.stack([self._filt @ h[s] for s in range(h.shape[0])])
        return act_fn(self.activation)(out + w[1])

    def _mm(self, v, S):                                             # (..., rows of S) @ S
        return (v.reshape(-1, v.shape[-1]) @ S).reshape(v.shape[:-1] + (S.shape[1],))

    def post_proc(self, y, ey, a, b, tf=False):                      # emulator.py:643-678, incidence products sparse
        if self.tide:
            h = y[..., 0] * (1 - self.is_outfall) + b[..., -1]
            y = np.concatenate([h[..., None], y[..., 1:]], axis=-1)
        inoff = self._mm(self.normalize(y, "y", True)[..., 0] - self.hmin, self._pos)
        f = ey[..., -1]
        flow = (f * (f > 0) * (self.offset > 0) * (inoff > self.offset) + f * (f <= 0) * (self.offset > 0) + f * (self.offset == 0))[..., None]
        ey = np.concatenate([ey[..., :-1], flow], axis=-1)
        if self.act:
            ae = self.get_edge_action(a)
            ey = ey.copy()
            if not tf or np.min(self.pump) > 0:
                fl = self.pump * self._mm((y[..., 0] > 0.01).astype(self.dtype), self._pos)
                fl = fl * ((self.norm_e[0, :, 2] > 1e-3) / self.norm_e[0, :, 2])
                ey[..., -1] = ey[..., -1] * (fl == 0) + fl
            ey[..., -1:] *= ae
            if not self.edge_fusion:
                a_out, a_in = self.get_action(a[:, :self.seq_out, ...])
                thr = 0.0 if tf else 0.01
                fli = self.pump_in * (y[..., 0] > thr) / self.norm_y[0, :, 1]
                flo = self.pump_out * (y[..., 0] > thr) / self.norm_y[0, :, 2]
                y = y.copy()
                y[..., 1] = y[..., 1] * (fli == 0) + fli
                y[..., 2] = y[..., 2] * (flo == 0) + flo
                y[..., 2] *= a_out
                y[..., 1] *= a_in
        if self.edge_fusion:
            efl = self.normalize(ey, "e", True)[..., -1]
            up, dn = np.clip(efl, 0, np.inf), -np.clip(efl, -np.inf, 0)
            node_out = (self._mm(up, self._pos.T.tocsr()) + self._mm(dn, self._neg.T.tocsr()))[..., None]
            node_in = (self._mm(up, self._neg.T.tocsr()) + self._mm(dn, self._pos.T.tocsr()))[..., None]
            node_out = node_out * ((self.norm_y[0, :, 2:3] > 1e-3) / self.norm_y[0, :, 2:3])
            node_in = node_in * ((self.norm_y[0, :, 1:2] > 1e-3) / self.norm_y[0, :, 1:2])
            y = np.concatenate([y[..., :1], node_in, node_out, y[..., 1:]], axis=-1)
        return y, ey
