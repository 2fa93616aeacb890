"""NumPy restatement of ``surrogate/emulator.py::Emulator`` (forward only).

Oracle (test infrastructure, see ``oracle/__init__.py``; PARITY UNPINNED).
Dense formulation exactly as the reference writes it: dense ``N x N`` masked
softmax for GAT, dense ``filter @ (x @ W)`` for GCN, dense ``N x M`` incidence
matmuls for NodeEdge and the edge-fusion mass balance - so that the CSR / fused
CUDA product path is checked against an independent formulation.

Two precisions: ``dtype=np.float64`` (truth) and ``np.float32`` (reference-like).

Weights are passed as a list in Keras *creation order* (one entry per weighted
layer, each a list of arrays in the layer's ``weights`` order); ``build_spec``
enumerates that order by walking ``build_network`` (emulator.py:166-341).
"""
from types import SimpleNamespace

import numpy as np


# ----------------------------------------------------------------------------
# third-party layer semantics (keras 2.10 / spektral 1.3.1), restated
# ----------------------------------------------------------------------------
def act_fn(name):
    if name in (None, "linear"):
        return lambda v: v
    if name == "relu":
        return lambda v: np.maximum(v, 0)
    if name == "tanh":
        return np.tanh
    if name == "sigmoid":
        return lambda v: 1.0 / (1.0 + np.exp(-v))
    if name == "hard_sigmoid":                     # keras 2.x backend.hard_sigmoid
        return lambda v: np.clip(0.2 * v + 0.5, 0.0, 1.0)
    raise ValueError("activation %s" % name)


def dense(x, W, b, act):
    """keras.layers.Dense: act(x @ W + b) on the last axis."""
    return act_fn(act)(np.matmul(x, W) + b)


def gcn_filter(A):
    """spektral.utils.gcn_filter: D^-1/2 (A + I) D^-1/2, D = rowsum(A + I), inf -> 0.
    (emulator.py:133 GCNConv.preprocess; the reference's A already has 1 on the diagonal.)"""
    out = np.array(A, dtype=np.float64, copy=True)
    out[np.diag_indices_from(out)] += 1
    with np.errstate(divide="ignore"):
        d = np.power(out.sum(1), -0.5)
    d[np.isinf(d)] = 0.0
    return (d[:, None] * out) * d[None, :]


def gcn_conv(x, filt, W, b, act):
    """spektral GCNConv.call (mixed mode): act(filter @ (x @ W) + b)."""
    return act_fn(act)(np.matmul(filt, np.matmul(x, W)) + b)


def gat_conv(x, filt, W, a_self, a_neigh, b, act):
    """spektral GATConv._call_dense, heads=1, concat, add_self_loops=True, no dropout.
    W (C,1,E); a_self/a_neigh (E,1,1); filt (N,N)."""
    a = np.array(filt, dtype=x.dtype, copy=True)
    if a.ndim == 2:
        a[np.diag_indices_from(a)] = 1
    else:                                                      # per-sample adjacency (use_adj): tf.linalg.set_diag on every sample
        idx = np.arange(a.shape[-1])
        a[..., idx, idx] = 1
    h = np.matmul(x, W[:, 0, :])                              # (...,N,E)
    s_self = np.matmul(h, a_self[:, 0, :])                    # (...,N,1)
    s_neigh = np.matmul(h, a_neigh[:, 0, :])                  # (...,N,1)
    score = s_self + np.swapaxes(s_neigh, -1, -2)             # (...,N,N): [i,j] = self_i + neigh_j
    score = np.where(score > 0, score, x.dtype.type(0.2) * score)   # leaky_relu(alpha=0.2)
    score = score + np.where(a == 0, x.dtype.type(-10e9), x.dtype.type(0.0))
    score = score - score.max(axis=-1, keepdims=True)
    p = np.exp(score)
    p = p / p.sum(axis=-1, keepdims=True)
    return act_fn(act)(np.matmul(p, h) + b)


def conv1d_causal(x, W, b, dilation, act):
    """keras Conv1D(padding='causal', dilation_rate=d): y[t] = b + sum_j W[j] x[t-(k-1-j)d].
    x (R,T,C), W (k,C,H)."""
    k = W.shape[0]
    T = x.shape[1]
    y = np.zeros(x.shape[:2] + (W.shape[2],), dtype=x.dtype) + b
    for j in range(k):
        s = (k - 1 - j) * dilation
        if s >= T:
            continue
        y[:, s:, :] += np.matmul(x[:, :T - s, :], W[j])
    return act_fn(act)(y)


def _sig(v):
    return 1.0 / (1.0 + np.exp(-v))


def gru(x, K, R, b):
    """keras GRU v2 defaults (reset_after=True, return_sequences=True, zero init state)."""
    H = R.shape[0]
    h = np.zeros((x.shape[0], H), dtype=x.dtype)
    out = []
    mx_all = np.matmul(x, K) + b[0]
    for t in range(x.shape[1]):
        mx = mx_all[:, t]
        mh = np.matmul(h, R) + b[1]
        z = _sig(mx[:, :H] + mh[:, :H])
        r = _sig(mx[:, H:2 * H] + mh[:, H:2 * H])
        hh = np.tanh(mx[:, 2 * H:] + r * mh[:, 2 * H:])
        h = z * h + (1 - z) * hh
        out.append(h)
    return np.stack(out, axis=1)


def lstm(x, K, R, b):
    """keras LSTM v2 defaults (gate order i,f,c,o; return_sequences=True)."""
    H = R.shape[0]
    h = np.zeros((x.shape[0], H), dtype=x.dtype)
    c = np.zeros_like(h)
    out = []
    zx = np.matmul(x, K) + b
    for t in range(x.shape[1]):
        z = zx[:, t] + np.matmul(h, R)
        i, f, g, o = _sig(z[:, :H]), _sig(z[:, H:2 * H]), np.tanh(z[:, 2 * H:3 * H]), _sig(z[:, 3 * H:])
        c = f * c + i * g
        h = o * np.tanh(c)
        out.append(h)
    return np.stack(out, axis=1)


# ----------------------------------------------------------------------------
# the Emulator
# ----------------------------------------------------------------------------
class OracleEmulator:
    """Forward-only restatement of emulator.py:47-852 (GCN / GAT and the no-conv NN-cat variant)."""

    def __init__(self, conv=None, resnet=False, recurrent=None, args=None, dtype=np.float32):
        g = lambda k, d=None: getattr(args, k, d)
        self.dtype = np.dtype(dtype)
        # emulator.py:49-127
        self.n_node, self.n_in = g("state_shape", (40, 4))
        self.tide = g("tide", False)
        self.b_in = 2 if self.tide else 1
        self.act = g("act", False)
        self.act = self.act and self.act != "False"
        self.n_out = self.n_in - 1
        self.seq_in, self.seq_out = g("seq_in", 6), g("seq_out", 1)
        self.embed_size, self.hidden_dim = g("embed_size", 64), g("hidden_dim", 64)
        self.kernel_size = g("kernel_size", 3)
        self.n_sp_layer, self.n_tp_layer = g("n_sp_layer", 3), g("n_tp_layer", 2)
        self.activation = g("activation", "relu")
        self.if_flood = g("if_flood", 0)
        if self.if_flood:
            self.n_in += 1
        self.is_outfall = np.asarray(g("is_outfall", np.zeros(self.n_node)))
        self.epsilon = g("epsilon", -1.0)
        self.graph_base = g("graph_base", 0)
        self.edge_fusion = g("edge_fusion", False)
        self.edges = np.asarray(args.edges)
        self.n_edge, self.e_in = g("edge_state_shape", (40, 4))
        self.e_out = self.e_in - 1
        self.edge_adj = np.asarray(g("edge_adj", np.eye(self.n_edge)))
        self.ehmax = np.asarray(g("ehmax", np.full(self.n_edge, 0.5)))
        self.pump = np.asarray(g("pump", np.zeros(self.n_edge)))
        self.node_edge = np.asarray(args.node_edge, dtype=np.float32)
        if self.edge_fusion:
            self.n_out -= 2
        self.adj = np.asarray(g("adj", np.eye(self.n_node)))
        if self.act:
            self.act_edges = np.asarray(args.act_edges)
            self.use_adj = g("use_adj", False)
        self.area = np.asarray(g("area", np.zeros(self.n_node)))
        self.pump_in = np.asarray(g("pump_in", np.zeros(self.n_node)))
        self.pump_out = np.asarray(g("pump_out", np.zeros(self.n_node)))
        self.offset = np.asarray(g("offset", np.zeros(self.n_edge)))
        self.hmax = np.asarray(g("hmax", np.full(self.n_node, 1.5)))
        self.hmin = np.asarray(g("hmin", np.zeros(self.n_node)))
        self.conv = False if conv in ["None", "False", "NoneType", None, False] else conv
        self.resnet = resnet
        self.recurrent = recurrent
        self.roll = g("roll", 0)
        if self.conv:
            self._set_filters()
        else:
            self.kind = "nn"       # NN-cat: flattened (N*C) Dense stacks, emulator.py:180,197-211,231-233
        self.spec = self.build_spec()
        self.weights = None
        self.trace = None          # set to {} to record per-stage activations of the next model() call

    # emulator.py:129-152
    def _set_filters(self):
        if "GCN" in self.conv:
            self.kind = "gcn"
            self.filter = gcn_filter(self.adj)
            self.edge_filter = gcn_filter(self.edge_adj)
        elif "GAT" in self.conv:
            self.kind = "gat"
            self.filter = (self.adj > 0).astype(int)
            self.edge_filter = (self.edge_adj > 0).astype(int)
        else:
            raise AssertionError("Unknown Convolution layer %s" % str(self.conv))

    # ------------------------------------------------------------------ spec
    def _conv_shapes(self, cin):
        E = self.embed_size
        if self.kind == "gcn":
            return ("gcn_conv", [(cin, E), (E,)])
        return ("mixed_gat", [(cin, 1, E), (E, 1, 1), (E, 1, 1), (E,)])

    def _tem_shapes(self, cin):
        H, k, P = self.hidden_dim, self.kernel_size, self.n_tp_layer
        out = []
        for i in range(P):
            c = cin if i == 0 else H
            if self.recurrent == "Conv1D":
                out.append(("conv1d", [(k, c, H), (H,)]))
            elif self.recurrent == "GRU":
                out.append(("gru", [(c, 3 * H), (H, 3 * H), (2, 3 * H)]))
            elif self.recurrent == "LSTM":
                out.append(("lstm", [(c, 4 * H), (H, 4 * H), (4 * H,)]))
            else:
                raise ValueError("recurrent %s" % self.recurrent)
        return out

    def build_spec(self):
        """Creation order of weighted layers in build_network (emulator.py:195-336)."""
        E, H, S = self.embed_size, self.hidden_dim, self.n_sp_layer
        N, M = self.n_node, self.n_edge
        if not self.conv:
            return self._build_spec_nn()
        sp = [("dense", [(self.n_in, E), (E,)]),             # :198
              ("dense", [(self.b_in, E // 2), (E // 2,)]),   # :203
              ("dense", [(self.e_in, E), (E,)])]             # :206
        if self.act:
            sp.append(("dense", [(1, E // 2), (E // 2,)]))   # :212

        def spatial(cx, ce):
            for _ in range(S):
                if self.graph_base:
                    assert cx == ce
                    sp.append(self._conv_shapes(cx))         # :222 / :275
                else:
                    sp.append(("dense", [(ce, E // 2), (E // 2,)]))      # x_e :225
                    sp.append(("dense", [(cx, E // 2), (E // 2,)]))      # e_x :226
                    sp.append(("node_edge", [(N, M), (N, M)]))           # :227
                    sp.append(("node_edge", [(M, N), (M, N)]))           # :228
                    sp.append(self._conv_shapes(cx + E // 2))            # :229
                    sp.append(self._conv_shapes(ce + E // 2))            # :230
                cx = ce = E
        spatial(E, E)
        sp += self._tem_shapes(E)                            # x :247
        sp += self._tem_shapes(E)                            # e :254
        spatial(H + E // 2, H + (E // 2 if self.act else 0))
        sp += self._tem_shapes(E)                            # :302
        sp += self._tem_shapes(E)                            # :308
        sp.append(("dense_resx", [(H, E), (E,)]))            # :313
        sp.append(("dense", [(H, E), (E,)]))                 # :317
        sp.append(("dense", [(E, self.n_out), (self.n_out,)]))   # :324
        if self.if_flood:
            c = E
            for _ in range(self.if_flood):
                sp.append(("dense", [(c, E // 2), (E // 2,)]))   # :329
                c = E // 2
            sp.append(("dense", [(c, 1), (1,)]))             # :330
        sp.append(("dense", [(E, self.e_out), (self.e_out,)]))   # :336
        return sp

    def _build_spec_nn(self):
        """NN-cat creation order (emulator.py:195-336 with conv falsy): node / edge features flattened per time step."""
        E, H, S = self.embed_size, self.hidden_dim, self.n_sp_layer
        N, M = self.n_node, self.n_edge
        sp = [("dense", [(N * self.n_in, E), (E,)]),         # :197-198
              ("dense", [(N * self.b_in, E // 2), (E // 2,)]),   # :202-203
              ("dense", [(M * self.e_in, E), (E,)])]         # :205-206
        if self.act:
            sp.append(("dense", [(M, E // 2), (E // 2,)]))   # :211-212
        for _ in range(S):
            sp.append(("dense", [(2 * E, 2 * E), (2 * E,)]))     # :232
        sp += self._tem_shapes(E)                            # :247
        sp += self._tem_shapes(E)                            # :254
        c = 2 * H + E // 2 + (E // 2 if self.act else 0)
        for _ in range(S):
            sp.append(("dense", [(c, 2 * E), (2 * E,)]))     # :285
            c = 2 * E
        sp += self._tem_shapes(E)                            # :302
        sp += self._tem_shapes(E)                            # :308
        sp.append(("dense_resx", [(H, E), (E,)]))            # :313
        sp.append(("dense", [(H, E), (E,)]))                 # :317
        sp.append(("dense", [(E, self.n_out * N), (self.n_out * N,)]))   # :322-324
        if self.if_flood:
            c = E
            for _ in range(self.if_flood):
                sp.append(("dense", [(c, E // 2), (E // 2,)]))   # :329
                c = E // 2
            sp.append(("dense", [(c, N), (N,)]))             # :327,330
        sp.append(("dense", [(E, self.e_out * M), (self.e_out * M,)]))   # :335-336
        return sp

    def set_weights(self, weights):
        assert len(weights) == len(self.spec), (len(weights), len(self.spec))
        out = []
        for (kind, shapes), ws in zip(self.spec, weights):
            assert len(ws) == len(shapes), kind
            for w, s in zip(ws, shapes):
                assert tuple(w.shape) == tuple(s), (kind, w.shape, s)
            out.append([np.asarray(w, dtype=self.dtype) for w in ws])
        self.weights = out

    # --------------------------------------------------------------- forward
    def _conv(self, x, filt, w):
        if self.kind == "gcn":
            return gcn_conv(x, filt, w[0], w[1], self.activation)
        return gat_conv(x, filt, w[0], w[1], w[2], w[3], self.activation)

    def _temporal(self, x, it):
        for i in range(self.n_tp_layer):
            w = next(it)
            if self.recurrent == "Conv1D":
                x = conv1d_causal(x, w[0], w[1], 2 ** i, self.activation)
            elif self.recurrent == "GRU":
                x = gru(x, w[0], w[1], w[2])
            else:
                x = lstm(x, w[0], w[1], w[2])
        return x

    def get_adj_action(self, a):                                     # emulator.py:343-362 (numpy branch)
        """(B,T,n_act) -> (B,T,n,n): the adjacency with every controlled link's setting written into its (from,to) entry,
        then `adj.astype(int)` for GAT.  (GCN: GCNConv.preprocess on this 4-D array raises inside spektral's gcn_filter -
        np.diag_indices_from needs equal dimensions - so the reference's GCN + use_adj cannot run; not restated.)"""
        assert self.kind == "gat", "use_adj is only functional with GAT in the reference"
        def get_act(s):
            adj = self.adj.copy()
            adj[tuple(self.act_edges.T)] = s
            return adj
        adj = np.apply_along_axis(get_act, -1, np.asarray(a, dtype=np.float64))
        return adj.astype(int)

    def model(self, X, Bn, E_in, AE=None, A_dyn=None):
        """The Keras graph of build_network (emulator.py:195-340), normalised inputs."""
        dt = self.dtype
        X, Bn, E_in = [np.asarray(v, dtype=dt) for v in (X, Bn, E_in)]
        if not self.conv:
            return self._model_nn(X, Bn, E_in, AE)
        it = iter(self.weights)
        actv = act_fn(self.activation)
        N, M, Ti, To, H = self.n_node, self.n_edge, self.seq_in, self.seq_out, self.hidden_dim
        filt = np.asarray(self.filter, dtype=dt)
        efilt = np.asarray(self.edge_filter, dtype=dt)
        nabs = np.abs(self.node_edge).astype(dt)

        w = next(it); x = dense(X, w[0], w[1], "linear")
        res = x[:, -1:, ...]
        x = actv(x)
        w = next(it); b = dense(Bn, w[0], w[1], self.activation)
        w = next(it); e = dense(E_in, w[0], w[1], "linear")
        res_e = e[:, -1:, ...]
        e = actv(e)
        if self.act:
            w = next(it); ae = dense(np.asarray(AE, dtype=dt), w[0], w[1], self.activation)

        def spatial(x, e, filt=filt):
            for _ in range(self.n_sp_layer):
                if self.graph_base:
                    z = self._conv(np.concatenate([x, e], axis=-2), filt, next(it))
                    x, e = z[..., :N, :], z[..., N:, :]
                else:
                    w = next(it); x_e = dense(e, w[0], w[1], self.activation)
                    w = next(it); e_x = dense(x, w[0], w[1], self.activation)
                    w = next(it); x = np.concatenate([x, np.matmul(w[0] * nabs + w[1], x_e)], axis=-1)
                    w = next(it); e = np.concatenate([e, np.matmul(w[0] * nabs.T + w[1], e_x)], axis=-1)
                    x = self._conv(x, filt, next(it))
                    e = self._conv(e, efilt, next(it))
            return x, e

        B = X.shape[0]
        tr = self.trace if self.trace is not None else {}
        tr["embed"] = (x.copy(), e.copy())
        x = x.reshape((-1,) + x.shape[2:]); e = e.reshape((-1,) + e.shape[2:])
        x, e = spatial(x, e)
        x = x.reshape((B, Ti, N, -1)); e = e.reshape((B, Ti, M, -1))
        tr["sp1"] = (x.copy(), e.copy())
        x = np.transpose(x, (0, 2, 1, 3)).reshape((B * N, Ti, -1))
        x = self._temporal(x, it)[:, -To:, :]
        x = np.transpose(x.reshape((B, N, To, H)), (0, 2, 1, 3))
        e = np.transpose(e, (0, 2, 1, 3)).reshape((B * M, Ti, -1))
        e = self._temporal(e, it)[:, -To:, :]
        e = np.transpose(e.reshape((B, M, To, H)), (0, 2, 1, 3))
        tr["tp1"] = (x.copy(), e.copy())
        x = np.concatenate([x, b], axis=-1)
        if self.act:
            e = np.concatenate([e, ae], axis=-1)
        x = x.reshape((-1,) + x.shape[2:]); e = e.reshape((-1,) + e.shape[2:])
        if A_dyn is not None:                                        # emulator.py:268-269: A = reshape(A_in, (-1, n, n))
            x, e = spatial(x, e, np.asarray(A_dyn, dtype=dt).reshape((-1,) + A_dyn.shape[-2:]))
        else:
            x, e = spatial(x, e)
        x = x.reshape((B, To, N, -1)); e = e.reshape((B, To, M, -1))
        tr["sp2"] = (x.copy(), e.copy())
        x = np.transpose(x, (0, 2, 1, 3)).reshape((B * N, To, -1))
        x = self._temporal(x, it)
        x = np.transpose(x.reshape((B, N, To, H)), (0, 2, 1, 3))
        e = np.transpose(e, (0, 2, 1, 3)).reshape((B * M, To, -1))
        e = self._temporal(e, it)
        e = np.transpose(e.reshape((B, M, To, H)), (0, 2, 1, 3))

        tr["tp2"] = (x.copy(), e.copy())
        w = next(it); x_out = dense(x, w[0], w[1], "linear")
        x = actv(np.cumsum(x_out, axis=1, dtype=dt) + res) if self.resnet else actv(x_out)
        w = next(it); e_out = dense(e, w[0], w[1], "linear")
        e = actv(np.cumsum(e_out, axis=1, dtype=dt) + res_e) if self.resnet else actv(e_out)
        tr["res"] = (x.copy(), e.copy())
        w = next(it); out = dense(x, w[0], w[1], "hard_sigmoid")
        if self.if_flood:
            for _ in range(self.if_flood):
                w = next(it); x = dense(x, w[0], w[1], self.activation)
            w = next(it); flood = dense(x, w[0], w[1], "sigmoid")
            out = np.concatenate([out, flood], axis=-1)
        w = next(it); e_o = dense(e, w[0], w[1], "tanh")
        assert next(it, None) is None
        return out, e_o

    def _model_nn(self, X, Bn, E_in, AE=None):
        """NN-cat graph (emulator.py:195-340 with conv falsy): (B,T,N*C) rows through Dense / temporal stacks."""
        dt = self.dtype
        it = iter(self.weights)
        actv = act_fn(self.activation)
        N, M, To = self.n_node, self.n_edge, self.seq_out
        B = X.shape[0]
        flat = lambda v: v.reshape(v.shape[0], v.shape[1], -1)
        w = next(it); x = dense(flat(X), w[0], w[1], "linear")
        res = x[:, -1:, :]
        x = actv(x)
        w = next(it); b = dense(flat(Bn), w[0], w[1], self.activation)
        w = next(it); e = dense(flat(E_in), w[0], w[1], "linear")
        res_e = e[:, -1:, :]
        e = actv(e)
        if self.act:
            w = next(it); ae = dense(flat(np.asarray(AE, dtype=dt)), w[0], w[1], self.activation)
        tr = self.trace if self.trace is not None else {}

        def spatial(x, e):
            for _ in range(self.n_sp_layer):
                w = next(it); z = dense(np.concatenate([x, e], axis=-1), w[0], w[1], self.activation)   # :232
                x, e = np.split(z, 2, axis=-1)                                                         # :233
            return x, e

        x, e = spatial(x, e)
        tr["sp1"] = (x.copy(), e.copy())
        x = self._temporal(x, it)[:, -To:, :]
        e = self._temporal(e, it)[:, -To:, :]
        x = np.concatenate([x, b], axis=-1)
        if self.act:
            e = np.concatenate([e, ae], axis=-1)
        x, e = spatial(x, e)
        tr["sp2"] = (x.copy(), e.copy())
        x = self._temporal(x, it)
        e = self._temporal(e, it)
        tr["tp2"] = (x.copy(), e.copy())
        w = next(it); x_out = dense(x, w[0], w[1], "linear")
        x = actv(np.cumsum(x_out, axis=1, dtype=dt) + res) if self.resnet else actv(x_out)
        w = next(it); e_out = dense(e, w[0], w[1], "linear")
        e = actv(np.cumsum(e_out, axis=1, dtype=dt) + res_e) if self.resnet else actv(e_out)
        tr["res"] = (x.copy(), e.copy())
        w = next(it); out = dense(x, w[0], w[1], "hard_sigmoid").reshape(B, To, N, self.n_out)   # :324-325
        if self.if_flood:
            for _ in range(self.if_flood):
                w = next(it); x = dense(x, w[0], w[1], self.activation)
            w = next(it); flood = dense(x, w[0], w[1], "sigmoid").reshape(B, To, N, 1)
            out = np.concatenate([out, flood], axis=-1)
        w = next(it); e_o = dense(e, w[0], w[1], "tanh").reshape(B, To, M, self.e_out)          # :336-337
        assert next(it, None) is None
        return out, e_o

    # ----------------------------------------------------- helpers / surface
    def set_norm(self, norm_x, norm_b, norm_y, norm_r, norm_e):      # emulator.py:794-800
        self.norm_x, self.norm_b, self.norm_y, self.norm_r = norm_x, norm_b, norm_y, norm_r
        if norm_e is not None:
            self.norm_e = norm_e

    def normalize(self, dat, item, inverse=False):                   # emulator.py:803-810
        dim = dat.shape[-1]
        normal = getattr(self, "norm_%s" % item)
        maxi, mini = normal[0, ..., :dim], normal[1, ..., :dim]
        return dat * (maxi - mini) + mini if inverse else (dat - mini) / (maxi - mini)

    def act_edge_index(self):                                        # emulator.py:386-388
        ae = [np.where((self.edges == act_edge).all(1))[0] for act_edge in self.act_edges]
        ae = [i for e in ae for i in e]
        return sorted(list(set(ae)), key=ae.index)

    def get_edge_action(self, a, g=False):                           # emulator.py:385-398
        idx = self.act_edge_index()
        out = np.ones(a.shape[:-1] + (self.n_edge,), dtype=a.dtype)
        out[..., idx] = a
        return out[..., None]

    def get_action(self, a):                                         # emulator.py:364-383
        a_out = np.ones(a.shape[:-1] + (self.n_node,), dtype=a.dtype)
        a_in = np.ones_like(a_out)
        a_out[..., self.act_edges[:, 0]] = a
        a_in[..., self.act_edges[:, 1]] = a
        return a_out, a_in

    def post_proc(self, y, ey, a, b, tf=False):                      # emulator.py:643-678 (numpy) / 680-725 (tf=True)
        """tf=True follows post_proc_tf, which differs on pumped networks: the rated pump flow overrides the edge
        flow only when EVERY link is a pump (``self.pump.min()>0``, :698) and the node-side pump gates use ``>0``
        instead of ``>0.01`` (:710-711).  The inlet-offset gate (:688) is the same arithmetic in both."""
        ne = self.node_edge.astype(np.float64 if self.dtype == np.float64 else np.float32)
        if self.tide:
            h = y[..., 0] * (1 - self.is_outfall) + b[..., -1]
            y = np.concatenate([h[..., None], y[..., 1:]], axis=-1)
        inoff = np.matmul(self.normalize(y, "y", True)[..., 0] - self.hmin, np.clip(ne, 0, 1))
        f = ey[..., -1]
        flow = (f * (f > 0) * (self.offset > 0) * (inoff > self.offset)
                + f * (f <= 0) * (self.offset > 0) + f * (self.offset == 0))[..., None]
        ey = np.concatenate([ey[..., :-1], flow], axis=-1)
        if self.act:
            ae = self.get_edge_action(a)
            ey = ey.copy()
            if not tf or np.min(self.pump) > 0:
                fl = self.pump * np.matmul(y[..., 0] > 0.01, np.clip(ne, 0, 1))
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
            efl = self.normalize(ey, "e", True)[..., -1:]
            pos, neg = np.clip(ne, 0, 1), np.abs(np.clip(ne, -1, 0))
            node_out = np.matmul(pos, np.clip(efl, 0, np.inf)) + np.matmul(neg, -np.clip(efl, -np.inf, 0))
            node_in = np.matmul(neg, np.clip(efl, 0, np.inf)) + np.matmul(pos, -np.clip(efl, -np.inf, 0))
            node_out = node_out * ((self.norm_y[0, :, 2:3] > 1e-3) / self.norm_y[0, :, 2:3])
            node_in = node_in * ((self.norm_y[0, :, 1:2] > 1e-3) / self.norm_y[0, :, 1:2])
            y = np.concatenate([y[..., :1], node_in, node_out, y[..., 1:]], axis=-1)
        return y, ey

    def constrain(self, y, r, h0=None):                              # emulator.py:727-748
        h, q_us, q_ds = [y[..., i] for i in range(3)]
        r = np.squeeze(r, axis=-1)
        h = np.clip(h, self.hmin, self.hmax)
        q_w = np.clip(q_us + r - q_ds, 0, np.inf) * (1 - self.is_outfall)
        if self.if_flood:
            f = y[..., -1] > 0.5
            h = self.hmax * f + h * ~f
            y = np.stack([h, q_us, q_ds, y[..., -1]], axis=-1)
        else:
            y = np.stack([h, q_us, q_ds], axis=-1)
        if self.epsilon > 0:
            q_w = q_w * ((self.hmax - h) < self.epsilon)
        elif self.epsilon == 0:
            pass
        elif self.if_flood:
            q_w = q_w * f
        return q_w, y

    def has_pumps(self):
        return float(np.sum(self.pump_in) + np.sum(self.pump_out) + np.sum(self.pump)) > 0

    def predict(self, states, b, a=None, edge_state=None, tf=False):  # emulator.py:566-602 (tf=True: 604-641)
        dt = self.dtype
        x = np.asarray(states)[:, -self.seq_in:, ...].astype(dt)
        ex = np.asarray(edge_state)[:, -self.seq_in:, ...].astype(dt)
        b = np.asarray(b).astype(dt)
        assert b.shape[1] == self.seq_out
        ae = self.get_edge_action(np.asarray(a).astype(dt)) if self.act else None
        nb = self.normalize(b, "b").astype(dt)
        adj = self.get_adj_action(np.asarray(a)[:, :self.seq_out]) if self.act and getattr(self, "use_adj", False) else None   # emulator.py:573-574
        y, ey = self.model(self.normalize(x, "x"), nb, self.normalize(ex, "e"), ae, adj)
        y, ey = self.post_proc(y, ey, None if a is None else np.asarray(a).astype(dt), nb, tf=tf)
        y = self.normalize(y, "y", True)
        ey = self.normalize(ey, "e", True)
        ey = np.concatenate([np.clip(ey[..., 0], 0, self.ehmax)[..., None], ey[..., 1:]], axis=-1)
        if self.has_pumps():                                         # emulator.py:590-598
            ne = self.node_edge.astype(np.float64)
            ps = (self.area * np.matmul(np.clip(ne, 0, 1), self.pump)) > 0
            h, qin, qout = [y[..., i] for i in [0, 1, 2]]
            de = np.zeros_like(h)
            for t in range(de.shape[1]):
                de[:, t, :] = np.clip(x[:, -1, :, 0] if t == 0 else de[:, t - 1, :]
                                      + (qin - qout)[:, t, :] / (self.area + 1e-6), self.hmin, self.hmax)
            y = y.copy()
            y[..., 0] = h * (1 - ps) + de * ps
        q_w, y = self.constrain(y, b[..., :1], x[:, -1:, :, 0])
        y = np.concatenate([y, q_w[..., None]], axis=-1)
        return y.astype(dt), ey.astype(dt)

    def predict_tf(self, states, b, a=None, edge_state=None):        # emulator.py:604-641
        return self.predict(states, b, a, edge_state, tf=True)

    def simulate(self, states, runoff, a=None, edge_states=None):    # emulator.py:521-564
        """Per-index batch-1 loop in the reference; every window is independent, so the
        restatement evaluates them as one batch (F9: a=None is passed through)."""
        runoff = np.asarray(runoff)[:, :self.seq_out]
        return self.predict(states, runoff, a, edge_states)

    # ------------------------------------------------------ autoregressive loops
    def rollout_mpc(self, state, runoff, settings, edge_state):      # mpc.py:565-582
        To = self.seq_out
        state, edge_state = state[:, -self.seq_in:, ...], edge_state[:, -self.seq_in:, ...]
        ys, eys, perf = [], [], None
        for idx in range(runoff.shape[1] // To):
            ri = runoff[:, idx * To:(idx + 1) * To, ...]
            sett = settings[:, idx * To:(idx + 1) * To, :] if settings is not None else None
            if self.if_flood and idx > 0:
                f = (perf > 0).astype(state.dtype)
                state = np.concatenate([state[..., :-1], f, state[..., -1:]], axis=-1)
            y, ey = self.predict_tf(state, ri, sett, edge_state)          # mpc.py:575
            state, perf = np.concatenate([y[..., :-2], ri], axis=-1), y[..., -1:]
            ae = self.get_edge_action(sett) if self.act else np.ones(ey.shape[:-1] + (1,), ey.dtype)
            edge_state = np.concatenate([ey, ae.astype(ey.dtype)], axis=-1)
            ys.append(y); eys.append(ey)
        return np.concatenate(ys, axis=1), np.concatenate(eys, axis=1)

    def rollout_mbrl(self, x, b, settings, ex):                      # mbrl.py:337-346
        To = self.seq_out
        ys, eys = [], []
        for i in range(b.shape[1] // To):
            bi = b[:, i * To:(i + 1) * To, ...]
            sett = settings[:, i * To:(i + 1) * To, :] if settings is not None else None
            y, ey = self.predict_tf(x, bi, sett, ex)                     # mbrl.py:337
            if self.if_flood:
                x = np.concatenate([y[..., :-2], (y[..., -2:-1] > 0.5).astype(y.dtype), bi], axis=-1)
            else:
                x = np.concatenate([y[..., :-1], bi], axis=-1)
            ae = self.get_edge_action(sett) if self.act else np.ones(ey.shape[:-1] + (1,), ey.dtype)
            ex = np.concatenate([ey, ae.astype(ey.dtype)], axis=-1)
            ys.append(y); eys.append(ey)
        return np.concatenate(ys, axis=1), np.concatenate(eys, axis=1)

    def rollout_norm(self, x, a, b, ex):                             # emulator.py:400-438 (_model, roll>0)
        """Normalised-space curriculum loop: inputs already normalised, no constrain inside."""
        dt = self.dtype
        Ti, To = self.seq_in, self.seq_out
        x, b, ex = [np.asarray(v, dtype=dt) for v in (x, b, ex)]
        P, EP = [], []
        for i in range(self.roll):
            bi = b[:, i * To:(i + 1) * To, :]
            ai = a[:, i * To:(i + 1) * To, ...] if a is not None else None
            aei = self.get_edge_action(np.asarray(ai, dtype=dt)) if self.act else None
            y, ey = self.model(x[:, -Ti:], bi, ex[:, -Ti:], aei)
            y, ey = self.post_proc(y, ey, ai, bi, tf=True)               # emulator.py:413 post_proc_tf
            P.append(y); EP.append(ey)
            if self.if_flood:
                x_new = np.concatenate([y[..., :-1], (y[..., -1:] > 0.5).astype(dt), bi], axis=-1)
            else:
                x_new = np.concatenate([y, bi], axis=-1)
            x = np.concatenate([x[:, -(Ti - To):], x_new], axis=1) if Ti > To else x_new
            ae_new = aei if self.act else np.ones(ey.shape[:-1] + (1,), dt)
            ex_new = np.concatenate([ey, ae_new], axis=-1)
            ex = np.concatenate([ex[:, -(Ti - To):], ex_new], axis=1) if Ti > To else ex_new
        return np.clip(np.concatenate(P, axis=1), 0, 1).astype(dt), np.concatenate(EP, axis=1).astype(dt)


def astlingen_objective(preds, state, node_names, targets, gamma=None, rough=True):
    """envs/scenario/astlingen.py:75-99 (objective_pred_tf)."""
    q_w = preds[..., -1]
    q_in = np.concatenate([state[:, -1:, :, 1], preds[..., 1]], axis=1)
    idx = node_names.index
    obj = sum(q_w[..., idx(i)] * w for i, attr, w in targets if attr == "cumflooding")
    obj = obj + sum(q_in[:, 1:, idx(i)] * w for i, attr, w in targets if attr == "cuminflow" and "WWTP" in i)
    if rough:
        obj = obj + sum(np.abs(np.diff(q_in[..., idx(i)], axis=1)) * w
                        for i, attr, w in targets if attr == "cuminflow" and "WWTP" not in i)
    gamma = np.ones(preds.shape[1]) if gamma is None else np.asarray(gamma, dtype=np.float32)
    return (obj * gamma).sum(axis=-1)


def make_args(**kw):
    return SimpleNamespace(**kw)
