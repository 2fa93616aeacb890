"""Drop-in ``Emulator`` for GNN-UDS running on hand-written sm_100a CUDA.

Same construct / load / predict surface as the reference class
``surrogate/emulator.py::Emulator`` (constructor emulator.py:48-127, ``load``
833-852, ``set_norm``/``normalize`` 794-810, ``predict`` 566-602, ``predict_tf``
604-641, ``simulate`` 521-564, ``get_edge_action`` 385-398), so that its callers
(``main.py --test``, ``mpc.py`` population evaluation, ``mbrl.py`` rollouts) can
switch by changing the import.  Forward only: ``fit_eval`` / ``fit_grad_norm`` /
``optimizer`` / ``model`` raise ``NotImplementedError``.

All arithmetic of the hot path runs in ``libgnnuds.so`` through ctypes
(``include/gnnuds.h``); there is no CPU fallback - without the built extension
or without a CUDA device construction fails.
"""
import ctypes as C
import os
import sys
from types import SimpleNamespace

import numpy as np

from . import _lib, graph as _graph, weights as _weights


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class _MSE:
    """keras.losses.MeanSquaredError()(y_true, y_pred) (used at main.py:368-373)."""

    def __call__(self, y_true, y_pred, sample_weight=None):
        d = (np.asarray(y_pred, np.float64) - np.asarray(y_true, np.float64)) ** 2
        d = d.mean(axis=-1)
        if sample_weight is not None:
            d = d * sample_weight
        return np.float32(d.mean())


class _BCE:
    """keras.losses.BinaryCrossentropy() on probabilities (epsilon 1e-7 clip)."""

    def __call__(self, y_true, y_pred, sample_weight=None):
        p = np.clip(np.asarray(y_pred, np.float64), 1e-7, 1 - 1e-7)
        t = np.asarray(y_true, np.float64)
        d = -(t * np.log(p) + (1 - t) * np.log(1 - p)).mean(axis=-1)
        if sample_weight is not None:
            d = d * sample_weight
        return np.float32(d.mean())


class Emulator:
    def __init__(self, conv=None, resnet=False, recurrent=None, args=None, device=0, precision="fp32",
                 max_chunk=0, workspace_mb=0):
        g = lambda k, d=None: getattr(args, k, d)
        self._ctor = (conv, resnet, recurrent, args, device, max_chunk, workspace_mb)   # for the fp32 twin of the reverse pass
        self._twin = None
        # ---- attribute parsing, emulator.py:49-127 ----
        self.n_node, self.n_in = g("state_shape", (40, 4))
        self.tide = bool(g("tide", False))
        self.b_in = 2 if self.tide else 1
        self.act = g("act", False)
        self.act = self.act and self.act != "False"
        self.n_out = self.n_in - 1
        self.seq_in, self.seq_out = g("seq_in", 6), g("seq_out", 1)
        self.embed_size, self.hidden_dim = g("embed_size", 64), g("hidden_dim", 64)
        self.kernel_size = g("kernel_size", 3)
        self.n_sp_layer, self.n_tp_layer = g("n_sp_layer", 3), g("n_tp_layer", 2)
        self.dropout = g("dropout", 0.0)          # inactive at inference (emulator.py:542,582,621)
        self.activation = g("activation", "relu")
        self.balance = g("balance", False)
        self.if_flood = int(g("if_flood", 0) or 0)
        if self.if_flood:
            self.n_in += 1
        self.is_outfall = np.asarray(g("is_outfall", np.zeros(self.n_node)))
        self.epsilon = float(g("epsilon", -1.0))
        self.graph_base = int(g("graph_base", 0) or 0)
        self.edge_fusion = bool(g("edge_fusion", False))
        if not hasattr(args, "edges"):
            raise AttributeError("args.edges is mandatory (emulator.py:76)")
        self.edges = np.asarray(args.edges)
        self.n_edge, self.e_in = g("edge_state_shape", (40, 4))
        self.e_out = self.e_in - 1
        self.ehmax = np.asarray(g("ehmax", np.full(self.n_edge, 0.5)), dtype=np.float64)
        self.pump = np.asarray(g("pump", np.zeros(self.n_edge)), dtype=np.float64)
        self.ewei = np.asarray(g("ewei", np.ones(self.n_edge)))
        if g("node_edge") is None and self.graph_base == 0 and not hasattr(args, "edges"):
            raise AttributeError("args.node_edge is mandatory (emulator.py:83)")
        self.node_edge = g("node_edge")
        if self.edge_fusion:
            self.n_out -= 2
        if self.act:
            self.act_edges = np.asarray(args.act_edges)
            self.use_adj = bool(g("use_adj", False))
            if self.use_adj and "GCN" in str(conv):
                raise NotImplementedError("use_adj with GCN: the reference hands the (B,T,N,N) adjacency of get_adj_action to "
                                          "spektral's gcn_filter, which raises for 4-D input (emulator.py:357-358)")
        self.area = np.asarray(g("area", np.zeros(self.n_node)), dtype=np.float64)
        self.pump_in = np.asarray(g("pump_in", np.zeros(self.n_node)), dtype=np.float64)
        self.pump_out = np.asarray(g("pump_out", np.zeros(self.n_node)), dtype=np.float64)
        self.offset = np.asarray(g("offset", np.zeros(self.n_edge)), dtype=np.float64)
        self.hmax = np.asarray(g("hmax", np.full(self.n_node, 1.5)), dtype=np.float64)
        self.hmin = np.asarray(g("hmin", np.zeros(self.n_node)), dtype=np.float64)
        self.nwei = np.asarray(g("nwei", np.ones(self.n_node)))
        # pumps / inlet offsets / pumped storage (emulator.py:590-598, 649-667) run on the CUDA path for graph_base == 0
        self.has_pumps = float(self.pump.sum() + self.pump_in.sum() + self.pump_out.sum()) > 0
        if (self.has_pumps or float(np.max(self.offset)) > 0) and self.graph_base:
            raise NotImplementedError("pumps / inlet offsets with graph_base != 0 are not on the CUDA path")
        self.conv = False if conv in ["None", "False", "NoneType", None, False] else conv
        if not self.conv:
            self.conv_kind = "nn"          # NN-cat: flattened Dense stacks (emulator.py:180,231-233); graph_base unused
            self.graph_base = 0
        elif "GCN" in self.conv:
            self.conv_kind = "gcn"
        elif "GAT" in self.conv:
            self.conv_kind = "gat"
        else:
            raise AssertionError("Unknown Convolution layer %s" % str(conv))   # emulator.py:151
        self.resnet = bool(resnet)
        self.recurrent = recurrent
        if recurrent not in ("Conv1D", "GRU", "LSTM"):
            raise NotImplementedError("recurrent=%r is not on the CUDA path" % (recurrent,))
        if self.activation not in _lib.ACT:
            raise NotImplementedError("activation %r" % (self.activation,))
        self.roll = int(g("roll", 0) or 0)
        self.model_dir = g("model_dir")
        self.mse = _MSE()
        if self.if_flood:
            self.bce = _BCE()
        self.n_act = int(len(self.act_edges)) if self.act else 0
        self.precision = precision
        self.device = int(device)

        # ---- filters: Emulator.get_conv (emulator.py:129-152) on CSR ----
        adj_csr = g("adj_csr")
        if adj_csr is None:
            adj_csr = self._csr_from_dense(np.asarray(g("adj", np.eye(self.n_node))))
        eadj_csr = g("edge_adj_csr")
        if eadj_csr is None and not self.graph_base:
            eadj_csr = self._csr_from_dense(np.asarray(g("edge_adj", np.eye(self.n_edge))))
        self.adj_csr, self.edge_adj_csr = adj_csr, eadj_csr
        if self.conv_kind == "gcn":
            self.filter_csr = _graph.gcn_normalise(adj_csr)
            self.edge_filter_csr = _graph.gcn_normalise(eadj_csr) if eadj_csr is not None else None
        else:
            self.filter_csr = _graph.gat_pattern(adj_csr)
            self.edge_filter_csr = _graph.gat_pattern(eadj_csr) if eadj_csr is not None else None

        # ---- native handle ----
        self._lib = _lib.load()
        cfg = _lib.Config()
        cfg.abi_version = _lib.ABI_VERSION
        cfg.n_node, cfg.n_edge = self.n_node, self.n_edge
        cfg.n_in, cfg.b_in, cfg.e_in, cfg.n_out, cfg.e_out = self.n_in, self.b_in, self.e_in, self.n_out, self.e_out
        cfg.seq_in, cfg.seq_out = self.seq_in, self.seq_out
        cfg.embed, cfg.hidden, cfg.kernel_size = self.embed_size, self.hidden_dim, self.kernel_size
        cfg.n_sp_layer, cfg.n_tp_layer = self.n_sp_layer, self.n_tp_layer
        cfg.conv = {"nn": _lib.CONV_NONE, "gcn": _lib.CONV_GCN, "gat": _lib.CONV_GAT}[self.conv_kind]
        cfg.recurrent = {"Conv1D": _lib.REC_CONV1D, "GRU": _lib.REC_GRU, "LSTM": _lib.REC_LSTM}[recurrent]
        cfg.activation = _lib.ACT[self.activation]
        cfg.resnet, cfg.if_flood, cfg.edge_fusion = int(self.resnet), self.if_flood, int(self.edge_fusion)
        cfg.act, cfg.graph_base, cfg.tide, cfg.n_act = int(bool(self.act)), self.graph_base, int(self.tide), self.n_act
        cfg.epsilon = self.epsilon
        cfg.precision = {"fp32": _lib.PREC_FP32, "bf16": _lib.PREC_BF16_TC, "fp16": _lib.PREC_BF16_TC, "tc": _lib.PREC_BF16_TC}[precision]   # "bf16" = historical name of the 16-bit tensor-core path (fp16 operands in the default build)
        cfg.device, cfg.max_chunk, cfg.workspace_mb = self.device, int(max_chunk), int(workspace_mb)
        self._cfg = cfg
        self._h = C.c_void_p()
        rc = self._lib.gnnuds_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            raise _lib.GnnUdsError("gnnuds_create failed (%d): %s" % (rc, self._lib.gnnuds_last_error(None).decode()))
        self._set_graph()
        if self.act and getattr(self, "use_adj", False) and self.conv_kind == "gat":
            self._set_adj_action()
        self.specs = _weights.layer_specs(self)
        nt = C.c_int32()
        _lib.check(self._h, self._lib.gnnuds_weight_layout(self._h, C.byref(nt), None))
        sizes = (C.c_int64 * nt.value)()
        _lib.check(self._h, self._lib.gnnuds_weight_layout(self._h, C.byref(nt), sizes))
        mine = [int(np.prod(shp)) for s in self.specs for shp in s.shapes]
        if mine != list(sizes):
            raise _lib.GnnUdsError("host/device weight layouts disagree")
        yc, ec = C.c_int32(), C.c_int32()
        self._lib.gnnuds_output_dims(self._h, C.byref(yc), C.byref(ec))
        self.y_channels, self.ey_channels = yc.value, ec.value
        self.weights = None

    # ------------------------------------------------------------------ plumbing
    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                self._lib.gnnuds_destroy(h)
            except Exception:
                pass
            self._h = None

    def _set_adj_action(self):
        """use_adj (emulator.py:268-271, 343-362): tell the library which entries of the node filter belong to controlled
        links and which of the others survive `adj.astype(int)` (get_adj_action casts the WEIGHTED adjacency)."""
        f, adj = self.filter_csr, self.adj_csr
        raw = {}
        for i in range(adj.n):
            for q in range(adj.row_ptr[i], adj.row_ptr[i + 1]):
                raw[(i, int(adj.col_idx[q]))] = float(adj.val[q])
        act_entry = np.zeros(f.nnz, np.int32)
        keep = np.zeros(f.nnz, np.uint8)
        pos = {}
        for i in range(f.n):
            for q in range(f.row_ptr[i], f.row_ptr[i + 1]):
                j = int(f.col_idx[q])
                pos[(i, j)] = q
                keep[q] = 1 if int(raw.get((i, j), 0.0)) != 0 else 0
        for k, (u, v) in enumerate(np.asarray(self.act_edges).reshape(-1, 2)):
            q = pos.get((int(u), int(v)))
            if q is None:
                raise ValueError("use_adj: controlled link (%d,%d) is not an entry of the adjacency pattern" % (u, v))
            act_entry[q] = k + 1                                   # the last controlled link writing an entry wins, as in get_act
        self._adj_act_entry, self._adj_keep = act_entry, keep
        _lib.check(self._h, self._lib.gnnuds_set_adj_action(self._h, _ptr(act_entry), _ptr(keep)))

    @staticmethod
    def _csr_from_dense(a):
        rows = [{int(j): float(a[i, j]) for j in np.nonzero(a[i])[0]} for i in range(a.shape[0])]
        return _graph.CSR(a.shape[0], rows)

    # dense views that callers of the reference read off the object
    @property
    def filter(self):
        d = self.filter_csr.to_dense()
        return d if self.conv_kind == "gcn" else d.astype(int)

    @property
    def edge_filter(self):
        d = self.edge_filter_csr.to_dense()
        return d if self.conv_kind == "gcn" else d.astype(int)

    def act_edge_index(self):
        """emulator.py:386-388: edge indices driven by each setting column (dedup, order kept)."""
        idx = [np.where((self.edges == e).all(1))[0] for e in self.act_edges]
        idx = [i for e in idx for i in e]
        return sorted(list(set(idx)), key=idx.index)

    def _set_graph(self):
        f, ef = self.filter_csr, self.edge_filter_csr
        gcn = self.conv_kind == "gcn"
        keep = dict(rp=_i32(f.row_ptr), ci=_i32(f.col_idx), val=_f32(f.val) if gcn else None,
                    edges=_i32(self.edges), hmin=_f32(self.hmin), hmax=_f32(self.hmax),
                    outf=_f32(self.is_outfall), ehmax=_f32(self.ehmax))
        if ef is not None:
            keep.update(erp=_i32(ef.row_ptr), eci=_i32(ef.col_idx), evl=_f32(ef.val) if gcn else None)
        g = _lib.Graph()
        pi, pf = C.POINTER(C.c_int32), C.POINTER(C.c_float)
        cast_i = lambda a: None if a is None else a.ctypes.data_as(pi)
        cast_f = lambda a: None if a is None else a.ctypes.data_as(pf)
        g.row_ptr, g.col_idx, g.val, g.n, g.nnz = cast_i(keep["rp"]), cast_i(keep["ci"]), cast_f(keep["val"]), f.n, f.nnz
        if ef is not None:
            g.e_row_ptr, g.e_col_idx, g.e_val, g.m, g.e_nnz = cast_i(keep["erp"]), cast_i(keep["eci"]), cast_f(keep["evl"]), ef.n, ef.nnz
        g.edges = cast_i(keep["edges"])
        if self.act:
            aoe = np.full(self.n_edge, -1, np.int32)
            for k, m in enumerate(self.act_edge_index()):
                aoe[m] = k
            aout, ain = np.full(self.n_node, -1, np.int32), np.full(self.n_node, -1, np.int32)
            for k, (u, v) in enumerate(self.act_edges):        # emulator.py:374-380 (last writer wins)
                aout[u] = k
                ain[v] = k
            keep.update(aoe=aoe, aout=aout, ain=ain)
            g.act_of_edge, g.act_out_of_node, g.act_in_of_node = cast_i(aoe), cast_i(aout), cast_i(ain)
        g.hmin, g.hmax, g.is_outfall, g.ehmax = cast_f(keep["hmin"]), cast_f(keep["hmax"]), cast_f(keep["outf"]), cast_f(keep["ehmax"])
        if self.has_pumps or float(np.max(self.offset)) > 0:
            keep.update(pump=_f32(self.pump), offset=_f32(self.offset), pump_in=_f32(self.pump_in),
                        pump_out=_f32(self.pump_out), area=_f32(self.area))
            g.pump, g.offset, g.pump_in, g.pump_out, g.area = (cast_f(keep[k]) for k in ("pump", "offset", "pump_in", "pump_out", "area"))
        _lib.check(self._h, self._lib.gnnuds_set_graph(self._h, C.byref(g)))

    # ------------------------------------------------------------------ weights / norms
    def set_weights(self, weights):
        """weights: list (creation order) of lists of arrays in Keras layout."""
        _weights.check_shapes(self.specs, weights)
        blob, offs, sizes = _weights.pack(weights)
        _lib.check(self._h, self._lib.gnnuds_set_weights(
            self._h, _ptr(blob), blob.size, offs.ctypes.data_as(C.POINTER(C.c_int64)),
            sizes.ctypes.data_as(C.POINTER(C.c_int64)), len(offs)))
        self.weights = [[np.array(w, dtype=np.float32) for w in l] for l in weights]

    def init_weights(self, seed=0, **kw):
        """Random-init weights with Keras' default initialisers (no trained weights ship)."""
        self.set_weights(_weights.init_weights(self.specs, seed, **kw))

    def set_norm(self, norm_x, norm_b, norm_y, norm_r, norm_e):        # emulator.py:794-800
        self.norm_x, self.norm_b, self.norm_y, self.norm_r = norm_x, norm_b, norm_y, norm_r
        if norm_e is not None:
            self.norm_e = norm_e
        self._push_norm()

    def _push_norm(self):
        if not all(hasattr(self, "norm_%s" % k) for k in "xbye"):
            return
        arrs = [_f32(getattr(self, "norm_%s" % k)) for k in "xbye"]
        for a, n, nm in zip(arrs, (self.n_node, self.n_node, self.n_node, self.n_edge), "xbye"):
            if a.ndim != 3 or a.shape[0] != 2 or a.shape[1] != n:
                raise ValueError("norm_%s must have shape (2,%d,c), got %s" % (nm, n, a.shape))
        _lib.check(self._h, self._lib.gnnuds_set_norm(
            self._h, _ptr(arrs[0]), arrs[0].shape[2], _ptr(arrs[1]), arrs[1].shape[2],
            _ptr(arrs[2]), arrs[2].shape[2], _ptr(arrs[3]), arrs[3].shape[2]))

    def normalize(self, dat, item, inverse=False):                      # emulator.py:803-810
        dim = dat.shape[-1]
        normal = getattr(self, "norm_%s" % item)
        maxi, mini = normal[0, ..., :dim], normal[1, ..., :dim]
        return dat * (maxi - mini) + mini if inverse else (dat - mini) / (maxi - mini)

    def load(self, model_dir=None, retrain=False):                      # emulator.py:833-852
        from . import h5lite
        model_dir = model_dir if model_dir is not None else self.model_dir
        if model_dir.endswith(".h5") or model_dir.endswith(".npz"):
            path, model_dir = model_dir, os.path.dirname(model_dir)
        else:
            path = os.path.join(model_dir, "model.h5")
            if not os.path.exists(path) and os.path.exists(os.path.join(model_dir, "model.npz")):
                path = os.path.join(model_dir, "model.npz")
        if path.endswith(".npz"):
            z = np.load(path)
            weights, i = [], 0
            for s in self.specs:
                weights.append([z["w%d" % (i + j)] for j in range(len(s.shapes))])
                i += len(s.shapes)
        else:
            weights = h5lite.load_keras_weights(path, self.specs)
        self.set_weights(weights)
        for item in "xbyre":
            p = os.path.join(model_dir, "norm_%s.npy" % item)
            if os.path.exists(p):
                setattr(self, "norm_%s" % item, np.load(p))
        self._push_norm()

    def save(self, model_dir=None):                                     # emulator.py:814-826
        from . import h5lite
        model_dir = model_dir if model_dir is not None else self.model_dir
        if model_dir.endswith(".h5"):
            path, model_dir = model_dir, os.path.dirname(model_dir)
        else:
            path = os.path.join(model_dir, "model.h5")
        os.makedirs(model_dir or ".", exist_ok=True)
        h5lite.save_keras_weights(path, self.specs, self.weights)
        for item in "xbyre":
            if hasattr(self, "norm_%s" % item):
                np.save(os.path.join(model_dir, "norm_%s.npy" % item), getattr(self, "norm_%s" % item))

    # ------------------------------------------------------------------ inference surface
    def get_edge_action(self, a, g=False):                              # emulator.py:385-398
        a = np.asarray(a)
        out = np.ones(a.shape[:-1] + (self.n_edge,), dtype=a.dtype)
        out[..., self.act_edge_index()] = a
        return out[..., None]

    def _run_predict(self, x, b, a, ex):
        B = x.shape[0]
        y, ey = self._out_arrays(B)
        _lib.check(self._h, self._lib.gnnuds_predict(self._h, _ptr(x), _ptr(b), _ptr(a), _ptr(ex), B,
                                                     _ptr(y), _ptr(ey), _lib.MEM_HOST))
        return y, ey

    def _out_arrays(self, B):
        """Result arrays of a predict call.  Fresh arrays as in the reference, except that a large pair handed out by the
        previous call is recycled once the caller has dropped every reference to it (views included): first-touch
        page faults of a fresh 0.9 GB result cost more than the copy that fills it."""
        shp = ((B, self.seq_out, self.n_node, self.y_channels), (B, self.seq_out, self.n_edge, self.ey_channels))
        c = getattr(self, "_out_cache", None)
        if c is not None and c[0].shape == shp[0] and c[1].shape == shp[1]:
            y, ey = c
            if sys.getrefcount(y) == 3 and sys.getrefcount(ey) == 3:   # the cache tuple, these locals, the call's argument
                return y, ey
        y, ey = np.empty(shp[0], np.float32), np.empty(shp[1], np.float32)
        self._out_cache = (y, ey) if y.nbytes + ey.nbytes >= (16 << 20) else None
        return y, ey

    def predict_ptrs(self, x, b, a, ex, B, y, ey, device=True):
        """Raw-pointer entry (ints): buffers already laid out as gnnuds_predict documents, on the
        handle's device (``device=True``, e.g. torch tensors' data_ptr()) or on the host."""
        _lib.check(self._h, self._lib.gnnuds_predict(self._h, x, b, a, ex, int(B), y, ey,
                                                     _lib.MEM_DEVICE if device else _lib.MEM_HOST))

    def _prep(self, states, b, a, edge_state):
        if edge_state is None:
            raise ValueError("edge_state is required (emulator.py:579 reads it unconditionally)")
        x = _f32(np.asarray(states)[:, -self.seq_in:, ...])
        ex = _f32(np.asarray(edge_state)[:, -self.seq_in:, ...])
        b = _f32(b)
        if x.shape[1:] != (self.seq_in, self.n_node, self.n_in):
            raise ValueError("states shape %s, expected (B,%d,%d,%d)" % (x.shape, self.seq_in, self.n_node, self.n_in))
        if ex.shape[1:] != (self.seq_in, self.n_edge, self.e_in) or ex.shape[0] != x.shape[0]:
            raise ValueError("edge_state shape %s" % (ex.shape,))
        if b.shape[0] != x.shape[0] or b.shape[2:] != (self.n_node, self.b_in):
            raise ValueError("b shape %s" % (b.shape,))
        if self.act:
            if a is None:
                raise ValueError("settings `a` required for an act model")
            a = _f32(a)
            if a.shape[0] != x.shape[0] or a.shape[2] != self.n_act:
                raise ValueError("a shape %s" % (a.shape,))
        else:
            a = None
        return x, b, a, ex

    def predict(self, states, b, a=None, edge_state=None):              # emulator.py:566-602
        x, b, a, ex = self._prep(states, b, a, edge_state)
        assert b.shape[1] == self.seq_out                               # emulator.py:570
        if a is not None:
            a = _f32(a[:, :self.seq_out])
        return self._run_predict(x, b, a, ex)

    def predict_tf(self, states, b, a=None, edge_state=None):           # emulator.py:604-641 (forward only)
        """Same pipeline with the reference's TF post-processing rules (post_proc_tf, emulator.py:680-725): on pumped
        networks the rated pump flow is applied to the links only when every link is a pump and the node-side pump
        gates use h > 0 - the variant the MPC / MBRL callers run (mpc.py:250,575; mbrl.py:337)."""
        _lib.check(self._h, self._lib.gnnuds_set_post_proc_variant(self._h, 1))
        try:
            return self.predict(states, b, a, edge_state)
        finally:
            self._lib.gnnuds_set_post_proc_variant(self._h, 0)

    def simulate(self, states, runoff, a=None, edge_states=None):       # emulator.py:521-564
        """All T' teacher-forced windows as one batch (the reference loops batch-1 per index).
        ``a=None`` is passed through when act is off (reference defect F9, emulator.py:545)."""
        runoff = np.asarray(runoff)[:, :self.seq_out]
        return self.predict(states, runoff, a, edge_states)

    def _replay_args(self, states, perfs, edge_states, settings):
        st = _f32(states)
        es = _f32(edge_states)
        T = st.shape[0]
        if st.shape[1:] != (self.n_node, 4) or es.shape != (T, self.n_edge, 4):
            raise ValueError("replay: states (T,%d,4) [h, cuminflow, outflow, lateral] and edge_states (T,%d,4) expected, got %s / %s"
                             % (self.n_node, self.n_edge, st.shape, es.shape))
        pf = None
        if self.if_flood:
            if perfs is None:
                raise ValueError("replay: an if_flood model needs the cumflooding series `perfs` (main.py:339)")
            pf = _f32(np.asarray(perfs).reshape(T, self.n_node))
        se = None
        if self.act:
            if settings is None:
                raise ValueError("replay: an act model needs the settings series (main.py:350)")
            se = _f32(np.asarray(settings).reshape(T, self.n_act))
        return st, pf, es, se, T

    def replay(self, states, perfs, edge_states, settings=None):           # main.py:327-361 + emulator.py:521-564
        """Teacher-forced replay of ONE recorded event: the T - seq_out windows of main.py:327-354 are built on the device
        from the raw series (states (T,N,4) [h, cuminflow, outflow, lateral], perfs (T,N) cumflooding, edge_states (T,M,4),
        settings (T,n_act)) and run as one batch.  Returns what ``emul.simulate(states_w, r, a, edge_states_w)`` returns:
        (T-seq_out, seq_out, N, y_channels), (T-seq_out, seq_out, M, ey_channels)."""
        st, pf, es, se, T = self._replay_args(states, perfs, edge_states, settings)
        W = max(T - self.seq_out, 0)
        y = np.empty((W, self.seq_out, self.n_node, self.y_channels), np.float32)
        ey = np.empty((W, self.seq_out, self.n_edge, self.ey_channels), np.float32)
        _lib.check(self._h, self._lib.gnnuds_replay(self._h, _ptr(st), _ptr(pf), _ptr(es), _ptr(se), T, _ptr(y), _ptr(ey), _lib.MEM_HOST))
        return y, ey

    def replay_many(self, events):
        """``replay`` of several recorded events in one call (gnnuds_replay_batch): ``events`` is a list of
        (states, perfs, edge_states[, settings]) tuples or dicts with those keys; returns the list of (y, ey) pairs
        ``replay`` returns event by event (bitwise).  Short events share a batch - a 600-window event alone leaves the
        persistent kernels a quarter of a wave per launch - and the results of one group of events are copied back while
        the next group computes."""
        args = []
        for e in events:
            if isinstance(e, dict):
                e = (e["states"], e.get("perfs"), e["edge_states"], e.get("settings"))
            e = tuple(e) + (None,) * (4 - len(e))
            args.append(self._replay_args(*e))
        n = len(args)
        if n == 0:
            return []
        VP = C.c_void_p * n
        out = []
        for st, pf, es, se, T in args:
            W = max(T - self.seq_out, 0)
            out.append((np.empty((W, self.seq_out, self.n_node, self.y_channels), np.float32),
                        np.empty((W, self.seq_out, self.n_edge, self.ey_channels), np.float32)))
        addr = lambda a: None if a is None else a.ctypes.data
        ps = VP(*[addr(a[0]) for a in args]); pp = VP(*[addr(a[1]) for a in args])
        pe = VP(*[addr(a[2]) for a in args]); pse = VP(*[addr(a[3]) for a in args])
        Ts = (C.c_int64 * n)(*[a[4] for a in args])
        py = VP(*[o[0].ctypes.data for o in out]); pey = VP(*[o[1].ctypes.data for o in out])
        _lib.check(self._h, self._lib.gnnuds_replay_batch(self._h, n, ps, pp, pe, pse, Ts, py, pey))
        return out

    def build_windows(self, states, perfs, edge_states, settings=None, first=0, count=None):
        """The replay windows themselves (device window builder, expand_seq semantics): x, b, a, ex of windows
        first..first+count-1, laid out as ``predict`` takes them."""
        st, pf, es, se, T = self._replay_args(states, perfs, edge_states, settings)
        W = max(T - self.seq_out, 0)
        count = W - first if count is None else count
        x = np.empty((count, self.seq_in, self.n_node, self.n_in), np.float32)
        b = np.empty((count, self.seq_out, self.n_node, self.b_in), np.float32)
        a = np.empty((count, self.seq_out, self.n_act), np.float32) if self.act else None
        ex = np.empty((count, self.seq_in, self.n_edge, self.e_in), np.float32)
        _lib.check(self._h, self._lib.gnnuds_build_windows(self._h, _ptr(st), _ptr(pf), _ptr(es), _ptr(se), T, int(first), int(count),
                                                           _ptr(x), _ptr(b), _ptr(a), _ptr(ex), _lib.MEM_HOST))
        return x, b, a, ex

    # ------------------------------------------------------------------ autoregressive loops
    def rollout(self, states, b, a=None, edge_state=None, rule="mpc"):
        """Chunked autoregressive rollout on the device.  rule: 'mpc' (mpc.py:565-582),
        'mbrl' (mbrl.py:337-346) or 'norm' (emulator.py:400-438, normalised inputs/outputs)."""
        code = {"mpc": _lib.ROLL_MPC, "mbrl": _lib.ROLL_MBRL, "norm": _lib.ROLL_NORM}[rule]
        x, b, a, ex = self._prep(states, b, a, edge_state)
        if b.shape[1] % self.seq_out:
            raise ValueError("horizon %d is not a multiple of seq_out" % b.shape[1])
        n_chunks = b.shape[1] // self.seq_out
        B = x.shape[0]
        nyc = self.y_channels if rule != "norm" else 3 + (1 if self.if_flood else 0)
        y = np.empty((B, b.shape[1], self.n_node, nyc), np.float32)
        ey = np.empty((B, b.shape[1], self.n_edge, self.ey_channels), np.float32)
        _lib.check(self._h, self._lib.gnnuds_rollout(self._h, code, _ptr(x), _ptr(b), _ptr(a), _ptr(ex), B, n_chunks,
                                                     _ptr(y), _ptr(ey), _lib.MEM_HOST))
        if rule == "norm":
            y = np.clip(y, 0, 1)                                        # emulator.py:437
        return y, ey

    def objective(self, preds, state, flood_w, wwtp_w, rough_w=None, gamma=None):
        """astlingen.objective_pred_tf (envs/scenario/astlingen.py:75-99) on the device."""
        y = _f32(preds)
        B, T = y.shape[0], y.shape[1]
        q0 = _f32(np.asarray(state)[:, -1, :, 1])
        obj = np.empty(B, np.float32)
        fw, ww = _f32(flood_w), _f32(wwtp_w)
        rw = None if rough_w is None else _f32(rough_w)
        gm = None if gamma is None else _f32(gamma)
        _lib.check(self._h, self._lib.gnnuds_objective(self._h, _ptr(y), _ptr(q0), B, T, _ptr(fw), _ptr(ww), _ptr(rw),
                                                       _ptr(gm), _ptr(obj), _lib.MEM_HOST))
        return obj

    def _grad_emulator(self):
        """The reverse pass runs on the fp32 path.  A tensor-core emulator keeps an fp32 twin with the same graph, weights
        and norms for it (created on first use; the weights are re-synchronised whenever they change)."""
        if self.precision == "fp32":
            return self
        if self._twin is None or self._twin_stamp != (id(self.weights), id(getattr(self, "norm_x", None))):
            conv, resnet, recurrent, args, device, max_chunk, workspace_mb = self._ctor
            if self._twin is None:
                self._twin = Emulator(conv, resnet, recurrent, args, device=device, precision="fp32", max_chunk=max_chunk,
                                      workspace_mb=workspace_mb)
            if self.weights is None or not hasattr(self, "norm_x"):
                raise RuntimeError("set_weights / load and set_norm before asking for gradients")
            self._twin.set_weights(self.weights)
            self._twin.set_norm(self.norm_x, self.norm_b, self.norm_y, self.norm_r, getattr(self, "norm_e", None))
            self._twin_stamp = (id(self.weights), id(self.norm_x))
        return self._twin

    def predict_vjp(self, states, b, a, edge_state, gy, gey, wrt=("a", "x", "ex")):
        """Vector-Jacobian product of ``predict_tf`` (what ``tf.GradientTape`` evaluates in mpc.py:596-607): cotangents
        gy (B,seq_out,N,ny) / gey (B,seq_out,M,e_out) of the two outputs -> dict with the forward results ``y``, ``ey``
        and the cotangents ``ga`` (B,seq_out,n_act), ``gx`` (B,seq_in,N,n_in), ``gex`` (B,seq_in,M,e_in) asked for in
        ``wrt``.  Runs on the fp32 path (gnnuds_predict_vjp); a tensor-core emulator delegates to its fp32 twin."""
        if self.precision != "fp32":
            return self._grad_emulator().predict_vjp(states, b, a, edge_state, gy, gey, wrt)
        x = _f32(np.asarray(states)[:, -self.seq_in:])
        ex = _f32(np.asarray(edge_state)[:, -self.seq_in:])
        b = _f32(b)
        a = _f32(a) if a is not None else None
        B = x.shape[0]
        gy, gey = _f32(gy), _f32(gey)
        ny = self.y_channels
        if gy.shape != (B, self.seq_out, self.n_node, ny) or gey.shape != (B, self.seq_out, self.n_edge, self.e_out):
            raise ValueError("cotangent shapes %s / %s do not match the outputs" % (gy.shape, gey.shape))
        out = {"y": np.empty(gy.shape, np.float32), "ey": np.empty(gey.shape, np.float32)}
        if "a" in wrt and self.act:
            out["ga"] = np.empty((B, self.seq_out, self.n_act), np.float32)
        if "x" in wrt:
            out["gx"] = np.empty(x.shape, np.float32)
        if "ex" in wrt:
            out["gex"] = np.empty(ex.shape, np.float32)
        _lib.check(self._h, self._lib.gnnuds_predict_vjp(self._h, _ptr(x), _ptr(b), _ptr(a), _ptr(ex), B, _ptr(gy), _ptr(gey),
                                                         _ptr(out["y"]), _ptr(out["ey"]), _ptr(out.get("ga")), _ptr(out.get("gx")),
                                                         _ptr(out.get("gex")), _lib.MEM_HOST))
        return out

    def grad_settings(self, states, b, a, edge_state, flood_w, wwtp_w, rough_w=None, gamma=None, rule="mpc"):
        """``mpc_problem_gr.gradient_fn`` (mpc.py:596-607) for the astlingen objective: (obj (B,), d obj / d a
        (B,T,n_act)) with T = b.shape[1] a multiple of seq_out - one ``predict_tf`` or the autoregressive chunk loop
        (mpc.py:565-582) differentiated end to end on the device (gnnuds_grad_settings, fp32 path; a tensor-core emulator
        delegates to its fp32 twin, so `obj` is the fp32 objective there)."""
        if self.precision != "fp32":
            return self._grad_emulator().grad_settings(states, b, a, edge_state, flood_w, wwtp_w, rough_w, gamma, rule)
        code = {"mpc": _lib.ROLL_MPC, "mbrl": _lib.ROLL_MBRL}[rule]
        x = _f32(np.asarray(states)[:, -self.seq_in:])
        ex = _f32(np.asarray(edge_state)[:, -self.seq_in:])
        b, a = _f32(b), _f32(a)
        B, T = b.shape[0], b.shape[1]
        if T % self.seq_out or a.shape != (B, T, self.n_act):
            raise ValueError("runoff %s / settings %s do not cover whole chunks of %d steps" % (b.shape, a.shape, self.seq_out))
        fw, ww = _f32(flood_w), _f32(wwtp_w)
        rw = None if rough_w is None else _f32(rough_w)
        gm = None if gamma is None else _f32(gamma)
        if gm is not None and gm.shape != (T,):
            raise ValueError("gamma must have one entry per horizon step")
        obj, ga = np.empty(B, np.float32), np.empty(a.shape, np.float32)
        _lib.check(self._h, self._lib.gnnuds_grad_settings(self._h, code, _ptr(x), _ptr(b), _ptr(a), _ptr(ex), B, T // self.seq_out,
                                                           _ptr(fw), _ptr(ww), _ptr(rw), _ptr(gm), _ptr(obj), _ptr(ga), _lib.MEM_HOST))
        return obj, ga

    def evaluate(self, state, runoff, edge_state, dec, r_step, horizon, flood_w, wwtp_w, rough_w=None, gamma=None,
                 rule="mpc", obj_ptr=None):
        """MPC population evaluation in one device-resident call (gnnuds_evaluate): ONE observed state
        (seq_in,N,n_in) / edge_state (seq_in,M,e_in) shared by the candidates, runoff (horizon,N,b_in) or
        (n_b,horizon,N,b_in), dec (pop,n_step,n_act) held r_step steps each (mpc.py:211-218), objective weights as in
        `objective`.  Returns the per-candidate objectives (pop,) as NumPy, or writes them to the device pointer
        `obj_ptr` (an int, e.g. a torch CUDA tensor's data_ptr()) and returns None."""
        code = {"mpc": _lib.ROLL_MPC, "mbrl": _lib.ROLL_MBRL}[rule]
        x0 = _f32(np.asarray(state)[-self.seq_in:])
        ex0 = _f32(np.asarray(edge_state)[-self.seq_in:])
        b = _f32(runoff)
        if b.ndim == 3:
            b = b[None]
        dec = _f32(dec)
        if x0.shape != (self.seq_in, self.n_node, self.n_in) or ex0.shape != (self.seq_in, self.n_edge, self.e_in):
            raise ValueError("state %s / edge_state %s: one observed window expected" % (x0.shape, ex0.shape))
        if horizon % self.seq_out or b.shape[1:] != (horizon, self.n_node, self.b_in):
            raise ValueError("runoff shape %s does not cover a horizon of %d steps" % (b.shape, horizon))
        if dec.ndim != 3 or dec.shape[2] != self.n_act:
            raise ValueError("dec shape %s, expected (pop, n_step, %d)" % (dec.shape, self.n_act))
        B = dec.shape[0]
        fw, ww = _f32(flood_w), _f32(wwtp_w)
        rw = None if rough_w is None else _f32(rough_w)
        gm = None if gamma is None else _f32(gamma)
        if gm is not None and gm.shape[0] < horizon:
            raise ValueError("gamma needs one weight per horizon step")
        obj = None if obj_ptr is not None else np.empty(B, np.float32)
        _lib.check(self._h, self._lib.gnnuds_evaluate(
            self._h, code, _ptr(x0), _ptr(ex0), _ptr(b), b.shape[0], _ptr(dec), dec.shape[1], int(r_step), B,
            horizon // self.seq_out, _ptr(fw), _ptr(ww), _ptr(rw), _ptr(gm), obj_ptr if obj_ptr is not None else _ptr(obj),
            _lib.MEM_HOST, _lib.MEM_DEVICE if obj_ptr is not None else _lib.MEM_HOST))
        return obj

    # ------------------------------------------------------------------ instrumentation
    def launch_count(self):
        return int(self._lib.gnnuds_launch_count(self._h))

    def set_profiling(self, level=1):
        """0 off, 1 time whole calls, 2 also time every rows-GEMM launch (CUDA events on the library stream)."""
        self._lib.gnnuds_set_profiling(self._h, int(level))

    def last_timing(self):
        t, g = C.c_float(), C.c_float()
        self._lib.gnnuds_last_timing(self._h, C.byref(t), C.byref(g))
        return t.value, g.value

    def measure_fp32_tflops(self):
        """fp32 FFMA peak of this device, measured now (TFLOP/s)."""
        t = C.c_float()
        _lib.check(self._h, self._lib.gnnuds_measure_simt_tflops(self._h, C.byref(t)))
        return float(t.value)

    def kernel_timing(self, family):
        """(ms, launches) of one kernel family in the last call (profiling level 2): 0 rows GEMM, 1 fused spatial, 2 gather."""
        t, n = C.c_float(), C.c_int64()
        _lib.check(self._h, self._lib.gnnuds_kernel_timing(self._h, int(family), C.byref(t), C.byref(n)))
        return t.value, n.value

    def debug_stage(self, stage, states, b, a=None, edge_state=None):
        """Test hook: run the forward up to `stage` and return the live (node, edge) streams."""
        x, b, a, ex = self._prep(states, b, a, edge_state)
        self._lib.gnnuds_debug_stop_after(self._h, int(stage))
        try:
            self._run_predict(x, b, a, ex)
            outs = []
            for which in (0, 1):
                rows, width = C.c_int64(), C.c_int32()
                _lib.check(self._h, self._lib.gnnuds_debug_fetch(self._h, which, None, 0, C.byref(rows), C.byref(width)))
                buf = np.empty((rows.value, width.value), np.float32)
                _lib.check(self._h, self._lib.gnnuds_debug_fetch(self._h, which, _ptr(buf), buf.size, C.byref(rows), C.byref(width)))
                outs.append(buf)
        finally:
            self._lib.gnnuds_debug_stop_after(self._h, 0)
        return outs

    # ------------------------------------------------------------------ not on this path
    def fit_eval(self, *a, **k):
        raise NotImplementedError("training (emulator.py:457-484) is out of scope of the forward rollout path")

    fit_grad_norm = fit_eval

    @property
    def model(self):
        raise NotImplementedError("there is no Keras model object behind this Emulator")

    @property
    def optimizer(self):
        raise NotImplementedError("there is no optimizer behind this Emulator")


def make_args(**kw):
    return SimpleNamespace(**kw)
