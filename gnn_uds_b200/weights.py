"""Parameter inventory of ``Emulator.build_network`` in Keras creation order.

The C library walks the same order (``LayoutBuilder`` in csrc/gnnuds.cu); this
module adds what the file formats need - Keras layer classes, auto-names and
weight names - and the default initialisers for synthetic ("random-init")
models.  Reference: surrogate/emulator.py:195-336 (creation order),
SURVEY.md 8(b) (.h5 naming), Appendix B (initialisers).
"""
import numpy as np


class LayerSpec:
    __slots__ = ("cls", "role", "weight_names", "shapes", "fixed_name")

    def __init__(self, cls, role, weight_names, shapes, fixed_name=None):
        self.cls, self.role, self.weight_names, self.shapes, self.fixed_name = cls, role, weight_names, shapes, fixed_name


def layer_specs(cfg):
    """cfg: object with the Emulator hyper-parameters (see emulator.Emulator._cfg)."""
    E, H, S, P, k = cfg.embed_size, cfg.hidden_dim, cfg.n_sp_layer, cfg.n_tp_layer, cfg.kernel_size
    N, M = cfg.n_node, cfg.n_edge
    out = []

    def dense(cin, cout, role, fixed=None):
        out.append(LayerSpec("dense", role, ["kernel:0", "bias:0"], [(cin, cout), (cout,)], fixed))

    def conv(cin, role):
        if cfg.conv_kind == "gat":
            out.append(LayerSpec("mixed_gat", role, ["kernel:0", "attn_kernel_self:0", "attn_kernel_neigh:0", "bias:0"],
                                 [(cin, 1, E), (E, 1, 1), (E, 1, 1), (E,)]))
        else:
            out.append(LayerSpec("gcn_conv", role, ["kernel:0", "bias:0"], [(cin, E), (E,)]))

    def temporal(cin, role):
        for i in range(P):
            c = cin if i == 0 else H
            if cfg.recurrent == "Conv1D":
                out.append(LayerSpec("conv1d", "%s%d" % (role, i), ["kernel:0", "bias:0"], [(k, c, H), (H,)]))
            elif cfg.recurrent == "GRU":
                out.append(LayerSpec("gru", "%s%d" % (role, i),
                                     ["gru_cell/kernel:0", "gru_cell/recurrent_kernel:0", "gru_cell/bias:0"],
                                     [(c, 3 * H), (H, 3 * H), (2, 3 * H)]))
            elif cfg.recurrent == "LSTM":
                out.append(LayerSpec("lstm", "%s%d" % (role, i),
                                     ["lstm_cell/kernel:0", "lstm_cell/recurrent_kernel:0", "lstm_cell/bias:0"],
                                     [(c, 4 * H), (H, 4 * H), (4 * H,)]))
            else:
                raise ValueError("recurrent %r is not on the CUDA path" % (cfg.recurrent,))

    def spatial(cx, ce, blk):
        for l in range(S):
            tag = "sp%d.%d." % (blk, l)
            if cfg.graph_base:
                conv(cx, tag + "joint")
            else:
                dense(ce, E // 2, tag + "x_e")
                dense(cx, E // 2, tag + "e_x")
                out.append(LayerSpec("node_edge", tag + "ne_n", ["weight:0", "bias:0"], [(N, M), (N, M)]))
                out.append(LayerSpec("node_edge", tag + "ne_m", ["weight:0", "bias:0"], [(M, N), (M, N)]))
                conv(cx + E // 2, tag + "conv_x")
                conv(ce + E // 2, tag + "conv_e")
            cx = ce = E

    if getattr(cfg, "conv_kind", None) == "nn":
        # NN-cat (emulator.py:180,197-211,231-233,284-286,322-337): node / edge features flattened per time step
        dense(N * cfg.n_in, E, "embed_x")
        dense(N * cfg.b_in, E // 2, "embed_b")
        dense(M * cfg.e_in, E, "embed_e")
        if cfg.act:
            dense(M, E // 2, "embed_a")
        for l in range(S):
            dense(2 * E, 2 * E, "sp1.%d.cat" % l)
        temporal(E, "tp1.x")
        temporal(E, "tp1.e")
        c = 2 * H + E // 2 + (E // 2 if cfg.act else 0)
        for l in range(S):
            dense(c, 2 * E, "sp2.%d.cat" % l)
            c = 2 * E
        temporal(E, "tp2.x")
        temporal(E, "tp2.e")
        dense(H, E, "res_x", fixed="dense_resx")
        dense(H, E, "res_e")
        dense(E, cfg.n_out * N, "out")
        if cfg.if_flood:
            c = E
            for i in range(cfg.if_flood):
                dense(c, E // 2, "flood%d" % i)
                c = E // 2
            dense(c, N, "flood_out")
        dense(E, cfg.e_out * M, "e_out")
        return out
    dense(cfg.n_in, E, "embed_x")
    dense(cfg.b_in, E // 2, "embed_b")
    dense(cfg.e_in, E, "embed_e")
    if cfg.act:
        dense(1, E // 2, "embed_a")
    spatial(E, E, 1)
    temporal(E, "tp1.x")
    temporal(E, "tp1.e")
    spatial(H + E // 2, H + (E // 2 if cfg.act else 0), 2)
    temporal(E, "tp2.x")
    temporal(E, "tp2.e")
    dense(H, E, "res_x", fixed="dense_resx")
    dense(H, E, "res_e")
    dense(E, cfg.n_out, "out")
    if cfg.if_flood:
        c = E
        for i in range(cfg.if_flood):
            dense(c, E // 2, "flood%d" % i)
            c = E // 2
        dense(c, 1, "flood_out")
    dense(E, cfg.e_out, "e_out")
    return out


def keras_names(specs, start=None):
    """Auto-names Keras gives in a fresh session (dense, dense_1, ...; 'dense_resx' literal)."""
    counters = dict(start or {})
    names = []
    for s in specs:
        if s.fixed_name:
            names.append(s.fixed_name)
            continue
        i = counters.get(s.cls, 0)
        names.append(s.cls if i == 0 else "%s_%d" % (s.cls, i))
        counters[s.cls] = i + 1
    return names


def _glorot(rng, shape):
    if len(shape) == 1:
        fan_in = fan_out = shape[0]
    elif len(shape) == 2:
        fan_in, fan_out = shape
    else:
        rec = int(np.prod(shape[:-2]))
        fan_in, fan_out = shape[-2] * rec, shape[-1] * rec
    lim = np.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def _orthogonal(rng, shape):
    a = rng.standard_normal((max(shape), min(shape)))
    q, r = np.linalg.qr(a)
    q = q * np.sign(np.diag(r))
    q = q if shape[0] >= shape[1] else q.T
    return q[:shape[0], :shape[1]].astype(np.float32)


def init_weights(specs, seed=0, node_edge_bias_std=0.0, bias_std=0.0):
    """Keras default initialisers restated: Glorot-uniform kernels,
    zero biases, orthogonal recurrent kernels (per gate block), NodeEdge w ~ N(0,0.05).
    ``node_edge_bias_std`` / ``bias_std`` > 0 perturb the zero biases so trained-like dense
    NodeEdge matrices and non-trivial biases are exercised."""
    rng = np.random.default_rng(seed)
    ws = []
    for s in specs:
        cur = []
        for name, shape in zip(s.weight_names, s.shapes):
            base = name.split("/")[-1]
            if s.cls == "node_edge":
                if base.startswith("weight"):
                    w = (rng.standard_normal(shape) * 0.05).astype(np.float32)
                else:
                    w = (rng.standard_normal(shape) * node_edge_bias_std).astype(np.float32)
            elif base.startswith("recurrent_kernel"):
                Hh = shape[0]
                w = np.concatenate([_orthogonal(rng, (Hh, Hh)) for _ in range(shape[1] // Hh)], axis=1)
            elif base.startswith("bias"):
                w = (rng.standard_normal(shape) * bias_std).astype(np.float32)
                if s.cls == "lstm" and bias_std == 0.0:
                    Hh = shape[0] // 4
                    w[Hh:2 * Hh] = 1.0          # unit_forget_bias
            else:
                w = _glorot(rng, shape)
            cur.append(np.ascontiguousarray(w, dtype=np.float32))
        ws.append(cur)
    return ws


def pack(weights):
    """-> (blob float32, offsets int64, sizes int64); every tensor 16-byte aligned."""
    offs, sizes, parts, pos = [], [], [], 0
    for layer in weights:
        for w in layer:
            w = np.ascontiguousarray(w, dtype=np.float32).ravel()
            offs.append(pos); sizes.append(w.size); parts.append(w)
            pad = (-w.size) % 4
            if pad:
                parts.append(np.zeros(pad, np.float32))
            pos += w.size + pad
    return np.concatenate(parts), np.asarray(offs, np.int64), np.asarray(sizes, np.int64)


def check_shapes(specs, weights):
    if len(specs) != len(weights):
        raise ValueError("expected %d weighted layers, got %d" % (len(specs), len(weights)))
    for s, ws in zip(specs, weights):
        if len(ws) != len(s.shapes):
            raise ValueError("layer %s (%s): expected %d tensors, got %d" % (s.role, s.cls, len(s.shapes), len(ws)))
        for w, shp, nm in zip(ws, s.shapes, s.weight_names):
            if tuple(w.shape) != tuple(shp):
                raise ValueError("layer %s (%s) weight %s: shape %s, expected %s" % (s.role, s.cls, nm, tuple(w.shape), tuple(shp)))
