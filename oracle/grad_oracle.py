"""Gradient oracle: d objective / d settings through ``predict_tf`` (mpc.py:511-627: ``GradientTape`` through
``Emulator.predict_tf`` + ``objective_pred_tf``; ``gradient_fn`` / ``pred_fit``).

Oracle (test infrastructure, see ``oracle/__init__.py``; PARITY UNPINNED against TensorFlow).  The forward of
``oracle/emulator_oracle.py`` restated op for op in torch fp64 (dense formulation, same weights / norms / graph tensors of
an ``OracleEmulator`` instance), so that reverse-mode autograd - an implementation independent of the hand-written VJP
kernels of the product path - yields the gradients.  Sub-gradient conventions are those TensorFlow uses for the same ops
(``tf.clip_by_value`` / ``relu`` / ``abs`` / ``cast(x > c)``): torch.clamp passes the gradient on the closed interval,
comparisons carry no gradient.

Pinned here by: (1) forward equality with the NumPy oracle (1e-12), (2) central finite differences of the NumPy oracle
(``tests/test_grad_oracle.py``).
"""
import numpy as np
import torch

DT = torch.float64


def _t(v):
    return v if isinstance(v, torch.Tensor) else torch.as_tensor(np.asarray(v, dtype=np.float64), dtype=DT)


def _act(name):
    if name in (None, "linear"):
        return lambda v: v
    if name == "relu":
        return torch.relu
    if name == "tanh":
        return torch.tanh
    if name == "sigmoid":
        return torch.sigmoid
    if name == "hard_sigmoid":                      # keras 2.x backend.hard_sigmoid
        return lambda v: torch.clamp(0.2 * v + 0.5, 0.0, 1.0)
    raise ValueError(name)


def _dense(x, w, act):
    return _act(act)(torch.matmul(x, w[0]) + w[1])


class TorchEmulator:
    """Differentiable twin of ``OracleEmulator`` (conv models; emulator.py:195-340, 604-641, 680-748)."""

    def __init__(self, orc):
        self.o = orc
        self.w = [[_t(a) for a in ws] for ws in orc.weights]
        self.norm = {k: _t(getattr(orc, "norm_" + k)) for k in "xbye" if hasattr(orc, "norm_" + k)}

    # ---- layer rules (oracle/emulator_oracle.py, same formulas) ----
    def _conv(self, x, filt, w):
        o = self.o
        if o.kind == "gcn":
            return _act(o.activation)(torch.matmul(filt, torch.matmul(x, w[0])) + w[1])
        a = filt.clone()
        idx = torch.arange(a.shape[-1])
        a[..., idx, idx] = 1
        h = torch.matmul(x, w[0][:, 0, :])
        s_self = torch.matmul(h, w[1][:, 0, :])
        s_neigh = torch.matmul(h, w[2][:, 0, :])
        score = torch.nn.functional.leaky_relu(s_self + s_neigh.transpose(-1, -2), 0.2)
        score = score + torch.where(a == 0, torch.full_like(a, -10e9), torch.zeros_like(a))
        p = torch.softmax(score, dim=-1)
        return _act(o.activation)(torch.matmul(p, h) + w[3])

    def _conv1d(self, x, W, b, d):
        k, T = W.shape[0], x.shape[1]
        y = torch.zeros(x.shape[:2] + (W.shape[2],), dtype=DT) + b
        for j in range(k):
            s = (k - 1 - j) * d
            if s >= T:
                continue
            y = y + torch.nn.functional.pad(torch.matmul(x[:, :T - s, :], W[j]), (0, 0, s, 0))
        return _act(self.o.activation)(y)

    def _gru(self, x, K, R, b):
        H = R.shape[0]
        h = torch.zeros((x.shape[0], H), dtype=DT)
        out = []
        mx_all = torch.matmul(x, K) + b[0]
        for t in range(x.shape[1]):
            mx = mx_all[:, t]
            mh = torch.matmul(h, R) + b[1]
            z = torch.sigmoid(mx[:, :H] + mh[:, :H])
            r = torch.sigmoid(mx[:, H:2 * H] + mh[:, H:2 * H])
            hh = torch.tanh(mx[:, 2 * H:] + r * mh[:, 2 * H:])
            h = z * h + (1 - z) * hh
            out.append(h)
        return torch.stack(out, dim=1)

    def _lstm(self, x, K, R, b):
        H = R.shape[0]
        h = torch.zeros((x.shape[0], H), dtype=DT)
        c = torch.zeros_like(h)
        out = []
        zx = torch.matmul(x, K) + b
        for t in range(x.shape[1]):
            z = zx[:, t] + torch.matmul(h, R)
            i, f, g, o = torch.sigmoid(z[:, :H]), torch.sigmoid(z[:, H:2 * H]), torch.tanh(z[:, 2 * H:3 * H]), torch.sigmoid(z[:, 3 * H:])
            c = f * c + i * g
            h = o * torch.tanh(c)
            out.append(h)
        return torch.stack(out, dim=1)

    def _temporal(self, x, it):
        o = self.o
        for i in range(o.n_tp_layer):
            w = next(it)
            if o.recurrent == "Conv1D":
                x = self._conv1d(x, w[0], w[1], 2 ** i)
            elif o.recurrent == "GRU":
                x = self._gru(x, w[0], w[1], w[2])
            else:
                x = self._lstm(x, w[0], w[1], w[2])
        return x

    def model(self, X, Bn, E_in, AE=None, trace=None, A_dyn=None):
        """build_network (emulator.py:195-340) on normalised inputs; ``trace`` collects named intermediates
        (with retain_grad) for stage-wise debugging of a VJP."""
        o = self.o
        if not o.conv:
            return self._model_nn(X, Bn, E_in, AE)
        it = iter(self.w)
        actv = _act(o.activation)
        N, M, Ti, To, H = o.n_node, o.n_edge, o.seq_in, o.seq_out, o.hidden_dim
        filt, efilt = _t(o.filter), (_t(o.edge_filter) if not o.graph_base else None)
        nabs = _t(np.abs(o.node_edge))

        def keep(name, v):
            if trace is not None:
                if v.requires_grad:
                    v.retain_grad()
                trace[name] = v
            return v

        w = next(it); x = _dense(X, w, "linear")
        res = x[:, -1:, ...]
        x = actv(x)
        w = next(it); b = _dense(Bn, w, o.activation)
        w = next(it); e = _dense(E_in, w, "linear")
        res_e = e[:, -1:, ...]
        e = actv(e)
        if o.act:
            w = next(it); ae = keep("aemb", _dense(AE, w, o.activation))

        def spatial(x, e, blk, filt=filt):
            for l in range(o.n_sp_layer):
                if o.graph_base:
                    z = self._conv(torch.cat([x, e], dim=-2), filt, next(it))
                    x, e = z[..., :N, :], z[..., N:, :]
                else:
                    w = next(it); x_e = _dense(e, w, o.activation)
                    w = next(it); e_x = _dense(x, w, o.activation)
                    w = next(it); x = torch.cat([x, torch.matmul(w[0] * nabs + w[1], x_e)], dim=-1)
                    w = next(it); e = torch.cat([e, torch.matmul(w[0] * nabs.T + w[1], e_x)], dim=-1)
                    x = self._conv(x, filt, next(it))
                    e = self._conv(e, efilt, next(it))
                keep("sp%d_%d_x" % (blk, l), x); keep("sp%d_%d_e" % (blk, l), e)
            return x, e

        B = X.shape[0]
        x = x.reshape((-1,) + tuple(x.shape[2:])); e = e.reshape((-1,) + tuple(e.shape[2:]))
        x, e = spatial(x, e, 1)
        x = x.reshape((B, Ti, N, -1)); e = e.reshape((B, Ti, M, -1))
        x = self._temporal(x.permute(0, 2, 1, 3).reshape((B * N, Ti, -1)), it)[:, -To:, :]
        x = x.reshape((B, N, To, H)).permute(0, 2, 1, 3)
        e = self._temporal(e.permute(0, 2, 1, 3).reshape((B * M, Ti, -1)), it)[:, -To:, :]
        e = e.reshape((B, M, To, H)).permute(0, 2, 1, 3)
        keep("tp1_x", x); keep("tp1_e", e)
        x = torch.cat([x, b], dim=-1)
        if o.act:
            e = torch.cat([e, ae], dim=-1)
        x = x.reshape((-1,) + tuple(x.shape[2:])); e = e.reshape((-1,) + tuple(e.shape[2:]))
        if A_dyn is not None:                                        # use_adj (emulator.py:268-269): per-sample node adjacency
            x, e = spatial(x, e, 2, _t(A_dyn).reshape((-1,) + tuple(A_dyn.shape[-2:])))
        else:
            x, e = spatial(x, e, 2)
        x = x.reshape((B, To, N, -1)); e = e.reshape((B, To, M, -1))
        x = self._temporal(x.permute(0, 2, 1, 3).reshape((B * N, To, -1)), it)
        x = x.reshape((B, N, To, H)).permute(0, 2, 1, 3)
        e = self._temporal(e.permute(0, 2, 1, 3).reshape((B * M, To, -1)), it)
        e = e.reshape((B, M, To, H)).permute(0, 2, 1, 3)
        keep("tp2_x", x); keep("tp2_e", e)
        w = next(it); x_out = _dense(x, w, "linear")
        x = actv(torch.cumsum(x_out, dim=1) + res) if o.resnet else actv(x_out)
        w = next(it); e_out = _dense(e, w, "linear")
        e = actv(torch.cumsum(e_out, dim=1) + res_e) if o.resnet else actv(e_out)
        keep("res_x", x); keep("res_e", e)
        w = next(it); out = _dense(x, w, "hard_sigmoid")
        if o.if_flood:
            for _ in range(o.if_flood):
                w = next(it); x = _dense(x, w, o.activation)
            w = next(it); out = torch.cat([out, _dense(x, w, "sigmoid")], dim=-1)
        w = next(it); e_o = _dense(e, w, "tanh")
        assert next(it, None) is None
        return keep("out_n", out), keep("out_e", e_o)

    def _model_nn(self, X, Bn, E_in, AE=None):
        """NN-cat graph (emulator.py:195-340 with conv falsy), as OracleEmulator._model_nn."""
        o = self.o
        it = iter(self.w)
        actv = _act(o.activation)
        N, M, To = o.n_node, o.n_edge, o.seq_out
        B = X.shape[0]
        flat = lambda v: v.reshape(v.shape[0], v.shape[1], -1)
        w = next(it); x = _dense(flat(X), w, "linear")
        res = x[:, -1:, :]
        x = actv(x)
        w = next(it); b = _dense(flat(Bn), w, o.activation)
        w = next(it); e = _dense(flat(E_in), w, "linear")
        res_e = e[:, -1:, :]
        e = actv(e)
        if o.act:
            w = next(it); ae = _dense(flat(AE), w, o.activation)

        def spatial(x, e):
            for _ in range(o.n_sp_layer):
                w = next(it); z = _dense(torch.cat([x, e], dim=-1), w, o.activation)
                x, e = torch.chunk(z, 2, dim=-1)
            return x, e

        x, e = spatial(x, e)
        x = self._temporal(x, it)[:, -To:, :]
        e = self._temporal(e, it)[:, -To:, :]
        x = torch.cat([x, b], dim=-1)
        if o.act:
            e = torch.cat([e, ae], dim=-1)
        x, e = spatial(x, e)
        x = self._temporal(x, it)
        e = self._temporal(e, it)
        w = next(it); x_out = _dense(x, w, "linear")
        x = actv(torch.cumsum(x_out, dim=1) + res) if o.resnet else actv(x_out)
        w = next(it); e_out = _dense(e, w, "linear")
        e = actv(torch.cumsum(e_out, dim=1) + res_e) if o.resnet else actv(e_out)
        w = next(it); out = _dense(x, w, "hard_sigmoid").reshape(B, To, N, o.n_out)
        if o.if_flood:
            for _ in range(o.if_flood):
                w = next(it); x = _dense(x, w, o.activation)
            w = next(it); out = torch.cat([out, _dense(x, w, "sigmoid").reshape(B, To, N, 1)], dim=-1)
        w = next(it); e_o = _dense(e, w, "tanh").reshape(B, To, M, o.e_out)
        assert next(it, None) is None
        return out, e_o

    def normalize(self, dat, item, inverse=False):                   # emulator.py:803-810
        dim = dat.shape[-1]
        normal = self.norm[item]
        maxi, mini = normal[0, ..., :dim], normal[1, ..., :dim]
        return dat * (maxi - mini) + mini if inverse else (dat - mini) / (maxi - mini)

    def get_edge_action(self, a):                                    # emulator.py:385-398
        idx = self.o.act_edge_index()
        out = torch.ones(tuple(a.shape[:-1]) + (self.o.n_edge,), dtype=DT)
        out = out.index_copy(-1, torch.as_tensor(idx), a) if len(idx) == a.shape[-1] else out
        return out[..., None]

    def get_action(self, a):                                         # emulator.py:364-383
        o = self.o
        ones = torch.ones(tuple(a.shape[:-1]) + (o.n_node,), dtype=DT)
        a_out = ones.index_copy(-1, torch.as_tensor(o.act_edges[:, 0]), a)
        a_in = ones.index_copy(-1, torch.as_tensor(o.act_edges[:, 1]), a)
        return a_out, a_in

    def post_proc_tf(self, y, ey, a, b):                             # emulator.py:680-725
        o = self.o
        ne = _t(o.node_edge)
        pos, neg = torch.clamp(ne, 0, 1), torch.abs(torch.clamp(ne, -1, 0))
        if o.tide:
            h = y[..., 0] * _t(1 - o.is_outfall) + b[..., -1]
            y = torch.cat([h[..., None], y[..., 1:]], dim=-1)
        if np.max(o.offset) > 0:
            off = _t(o.offset)
            inoff = torch.matmul(self.normalize(y, "y", True)[..., 0] - _t(o.hmin), pos)
            f = ey[..., -1]
            flow = (f * (f > 0).to(DT) * (off > 0).to(DT) * (inoff > off).to(DT)
                    + f * (f <= 0).to(DT) * (off > 0).to(DT) + f * (off == 0).to(DT))[..., None]
            ey = torch.cat([ey[..., :-1], flow], dim=-1)
        if o.act:
            if np.min(o.pump) > 0:
                fl = _t(o.pump) * torch.matmul((y[..., 0] > 0.01).to(DT), pos)
                fl = fl * _t((o.norm_e[0, :, 2] > 1e-3) / o.norm_e[0, :, 2])
                flow = (ey[..., -1] * (fl == 0).to(DT) + fl)[..., None]
            else:
                flow = ey[..., -1:]
            ae = self.get_edge_action(a)
            ey = torch.cat([ey[..., :-1], flow * ae], dim=-1)
            if not o.edge_fusion:
                a_out, a_in = self.get_action(a[:, :o.seq_out, ...])
                fli = _t(o.pump_in) * (y[..., 0] > 0).to(DT) / _t(o.norm_y[0, :, 1])
                flo = _t(o.pump_out) * (y[..., 0] > 0).to(DT) / _t(o.norm_y[0, :, 2])
                inflow = y[..., 1] * (fli == 0).to(DT) + fli
                outflow = y[..., 2] * (flo == 0).to(DT) + flo
                y = torch.cat([torch.stack([y[..., 0], inflow * a_in, outflow * a_out], dim=-1), y[..., 3:]], dim=-1)
        if o.edge_fusion:
            efl = self.normalize(ey, "e", True)[..., -1:]
            node_out = torch.matmul(pos, torch.clamp(efl, min=0)) + torch.matmul(neg, -torch.clamp(efl, max=0))
            node_in = torch.matmul(neg, torch.clamp(efl, min=0)) + torch.matmul(pos, -torch.clamp(efl, max=0))
            node_out = node_out * _t((o.norm_y[0, :, 2:3] > 1e-3) / o.norm_y[0, :, 2:3])
            node_in = node_in * _t((o.norm_y[0, :, 1:2] > 1e-3) / o.norm_y[0, :, 1:2])
            y = torch.cat([y[..., :1], node_in, node_out, y[..., 1:]], dim=-1)
        return y, ey

    def constrain_tf(self, y, r):                                    # emulator.py:750-771
        o = self.o
        h, q_us, q_ds = [y[..., i] for i in range(3)]
        r = r[..., 0]
        h = torch.maximum(torch.minimum(h, _t(o.hmax)), _t(o.hmin))
        q_w = torch.clamp(q_us + r - q_ds, min=0) * _t(1 - o.is_outfall)
        if o.if_flood:
            f = (y[..., -1] > 0.5).to(DT)
            h = _t(o.hmax) * f + h * (1 - f)
            y = torch.stack([h, q_us, q_ds, y[..., -1]], dim=-1)
        else:
            y = torch.stack([h, q_us, q_ds], dim=-1)
        if o.epsilon > 0:
            q_w = q_w * ((_t(o.hmax) - h) < o.epsilon).to(DT)
        elif o.epsilon == 0:
            pass
        elif o.if_flood:
            q_w = q_w * f
        return q_w, y

    def predict_tf(self, states, b, a=None, edge_state=None, trace=None):   # emulator.py:604-641
        o = self.o
        x = _t(states)[:, -o.seq_in:, ...]
        ex = _t(edge_state)[:, -o.seq_in:, ...]
        b = _t(b)
        a = _t(a) if a is not None else None
        ae = self.get_edge_action(a) if o.act else None
        nb = self.normalize(b, "b")
        adj = None
        if o.act and getattr(o, "use_adj", False):                   # `adj.astype(int)`: a constant of the differentiation
            adj = o.get_adj_action(a.detach().numpy()[:, :o.seq_out]).astype(np.float64)
        y, ey = self.model(self.normalize(x, "x"), nb, self.normalize(ex, "e"), ae, trace, adj)
        y, ey = self.post_proc_tf(y, ey, a, nb)
        ey = self.normalize(ey, "e", True)
        ey = torch.cat([torch.maximum(torch.minimum(ey[..., 0], _t(o.ehmax)), torch.zeros(()).to(DT))[..., None], ey[..., 1:]], dim=-1)
        y = self.normalize(y, "y", True)
        if o.has_pumps():                                            # emulator.py:626-637 pumped-storage depth
            ps = _t((o.area * np.matmul(np.clip(o.node_edge.astype(np.float64), 0, 1), o.pump)) > 0)
            h, qin, qout = [y[..., i] for i in (0, 1, 2)]
            lo, hi, ar = _t(o.hmin), _t(o.hmax), _t(o.area + 1e-6)
            de = []
            for t in range(o.seq_out):
                v = x[:, -1, :, 0] if t == 0 else de[-1] + (qin - qout)[:, t, :] / ar
                de.append(torch.maximum(torch.minimum(v, hi), lo))
            de = torch.stack(de, dim=1)
            y = torch.cat([(h * (1 - ps) + de * ps)[..., None], y[..., 1:]], dim=-1)
        q_w, y = self.constrain_tf(y, b[..., :1])
        return torch.cat([y, q_w[..., None]], dim=-1), ey

    def rollout_mpc(self, state, runoff, settings, edge_state):      # mpc.py:565-582
        o = self.o
        To = o.seq_out
        state, edge_state, runoff, settings = _t(state)[:, -o.seq_in:], _t(edge_state)[:, -o.seq_in:], _t(runoff), _t(settings)
        ys, eys, perf = [], [], None
        for idx in range(runoff.shape[1] // To):
            ri = runoff[:, idx * To:(idx + 1) * To]
            sett = settings[:, idx * To:(idx + 1) * To]
            if o.if_flood and idx > 0:
                f = (perf > 0).to(DT)
                state = torch.cat([state[..., :-1], f, state[..., -1:]], dim=-1)
            y, ey = self.predict_tf(state, ri, sett, edge_state)
            state, perf = torch.cat([y[..., :-2], ri], dim=-1), y[..., -1:]
            edge_state = torch.cat([ey, self.get_edge_action(sett)], dim=-1)
            ys.append(y); eys.append(ey)
        return torch.cat(ys, dim=1), torch.cat(eys, dim=1)


def astlingen_objective_t(preds, q_in0, flood_w, wwtp_w, rough_w=None, gamma=None):
    """envs/scenario/astlingen.py:75-99 with the per-node weight vectors of ``gnnuds_objective``."""
    q_w = preds[..., -1]
    q_in = torch.cat([_t(q_in0)[:, None, :], preds[..., 1]], dim=1)
    obj = q_w * _t(flood_w) + q_in[:, 1:] * _t(wwtp_w)
    if rough_w is not None:
        obj = obj + torch.abs(q_in[:, 1:] - q_in[:, :-1]) * _t(rough_w)
    g = torch.ones(preds.shape[1], dtype=DT) if gamma is None else _t(gamma)
    return (obj.sum(dim=-1) * g).sum(dim=-1)


def grad_settings(orc, states, b, a, edge_state, flood_w, wwtp_w, rough_w=None, gamma=None, chunks=1, trace=None,
                  wrt_state=False):
    """(obj (B,), d sum(obj) / d a (B,T,n_act)[, d/d states, d/d edge_state]) - mpc.py:596-607 ``gradient_fn``: every
    candidate's objective depends on its own settings only, so the gradient of the sum is the per-candidate gradient."""
    te = TorchEmulator(orc)
    at = _t(a).clone().requires_grad_(True)
    st, et = _t(states).clone(), _t(edge_state).clone()
    if wrt_state:
        st.requires_grad_(True); et.requires_grad_(True)
    if chunks == 1:
        y, _ = te.predict_tf(st, b, at, et, trace)
    else:
        y, _ = te.rollout_mpc(st, b, at, et)
    obj = astlingen_objective_t(y, st[:, -1, :, 1].detach(), flood_w, wwtp_w, rough_w, gamma)
    obj.sum().backward()
    out = (obj.detach().numpy(), at.grad.numpy())
    if wrt_state:
        out += (st.grad.numpy(), et.grad.numpy())
    return out


def predict_vjp(orc, states, b, a, edge_state, gy, gey, trace=None):
    """VJP of predict_tf: (y, ey, ga, gx, gex) for the cotangents gy (B,To,N,ny), gey (B,To,M,e_out)."""
    te = TorchEmulator(orc)
    at = _t(a).clone().requires_grad_(True)
    st = _t(states).clone().requires_grad_(True)
    et = _t(edge_state).clone().requires_grad_(True)
    y, ey = te.predict_tf(st, b, at, et, trace)
    ((y * _t(gy)).sum() + (ey * _t(gey)).sum()).backward()
    return y.detach().numpy(), ey.detach().numpy(), at.grad.numpy(), st.grad.numpy(), et.grad.numpy()
