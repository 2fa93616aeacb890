"""Synthetic inputs for the Emulator hot path (no SWMM, no trained weights ship).

Follows SURVEY.md 8(d): physical units, ``np.random.default_rng(seed)``;
normalisation arrays shaped like ``DataGenerator.get_norm`` produces them
(dataloader.py:224-265: ``(2, N, C)`` with [0]=max, [1]=min=0, flood channel
max = 1, setting channel max = 1); node state channels ``[h, q_us, q_ds,
(flood flag), r]`` (dataloader.py:145-169), edge state ``[depth, volume,
flow_vol (signed), setting]``; runoff a Chicago-type pulse per scenario
(utils/utilities.py:107-125).
"""
import numpy as np

from . import graph as _graph


def make_norms(n_node, n_edge, if_flood, rng, tide=False, head_min=None, is_outfall=None, head_max=None):
    """head_min (scalar): head-based network (dataloader.py:231-232,246-253): the h channel is normalised between the
    global minimum head and the global maximum head, so norm_x/norm_y carry a NON-ZERO minimum in channel 0 and
    norm_b (tide) one at the outfalls."""
    q = rng.uniform(0.5, 5.0, size=(n_node, 4)).astype(np.float32)       # h, q_us, q_ds, r maxima
    one = np.ones((n_node, 1), np.float32)
    qw = np.full((n_node, 1), np.float32(q[:, 1].max()))
    if if_flood:
        nx = np.concatenate([q[:, :3], one, q[:, 3:4]], -1)
        ny = np.concatenate([q[:, :3], one, qw], -1)
    else:
        nx = q.copy()
        ny = np.concatenate([q[:, :3], qw], -1)
    nb = q[:, 3:4].copy()
    if tide:
        nb = np.concatenate([nb, q[:, :1]], -1)
    ne = rng.uniform(0.5, 5.0, size=(n_edge, 4)).astype(np.float32)
    ne[:, 3] = 1.0
    st = lambda a: np.stack([a, np.zeros_like(a)]).astype(np.float32)
    if head_min is not None:
        hm = np.float32(head_max + 1e-6)
        nx[:, 0] = hm; ny[:, 0] = hm
        out = dict(norm_x=st(nx), norm_y=st(ny), norm_r=st(q[:, 3:4]), norm_e=st(ne))
        out["norm_x"][1, :, 0] = np.float32(head_min)
        out["norm_y"][1, :, 0] = np.float32(head_min)
        if tide:
            nb[:, 1] = hm * np.asarray(is_outfall, np.float32) + 1e-6
        out["norm_b"] = st(nb)
        if tide:
            out["norm_b"][1, :, 1] = np.float32(head_min) * np.asarray(is_outfall, np.float32)
        return out
    return dict(norm_x=st(nx), norm_b=st(nb), norm_y=st(ny), norm_r=st(q[:, 3:4]), norm_e=st(ne))


def chicago_pulse(T, rng, size):
    """(size..., T) unit-peak hyetograph: rising/falling power-law limbs around a random peak."""
    r = rng.uniform(0.2, 0.8, size=size + (1,))
    t = (np.arange(T, dtype=np.float64) + 0.5) / T
    d = np.where(t < r, (r - t) / r, (t - r) / (1 - r))
    return (1.0 - d) ** 2.5


def make_inputs(emul, B, rng, horizon=None, shared_state=False):
    """states (B,Ti,N,n_in), b (B,Th,N,b_in), a (B,Th,n_act)|None, edge_state (B,Ti,M,e_in)."""
    N, M, Ti = emul.n_node, emul.n_edge, emul.seq_in
    Th = horizon or emul.seq_out
    nx, ne, nb = emul.norm_x[0], emul.norm_e[0], emul.norm_b[0]
    Bs = 1 if shared_state else B
    hmax = np.where(emul.hmax > 0, emul.hmax, 1.0)
    h = rng.uniform(0, 1, (Bs, Ti, N)) * hmax
    hmin = np.asarray(getattr(emul, "hmin", 0.0), np.float64)
    if np.max(hmin) > 0:                                   # head-based: h between the invert elevation and the rim
        h = hmin + rng.uniform(0, 1, (Bs, Ti, N)) * np.where(emul.hmax - hmin > 0, emul.hmax - hmin, 1.0)
    q_us = rng.uniform(0, 1, (Bs, Ti, N)) * nx[:, 1]
    q_ds = rng.uniform(0, 1, (Bs, Ti, N)) * nx[:, 2]
    pulse = chicago_pulse(Ti + Th, rng, (B, 1)) * rng.uniform(0, 1, (B, 1, 1))       # (B,1,Ti+Th)
    rain = np.transpose(np.repeat(pulse, N, axis=1), (0, 2, 1)) * nb[:, 0]          # (B,Ti+Th,N)
    r_hist = rain[:Bs, :Ti]
    cols = [h, q_us, q_ds]
    if emul.if_flood:
        cols.append((rng.uniform(0, 1, (Bs, Ti, N)) < 0.1).astype(np.float64))
    cols.append(r_hist)
    states = np.stack(cols, -1)
    b = rain[:, Ti:, :, None]
    if emul.tide:
        tide = rng.uniform(0, 1, (B, Th, N)) * nb[:, 1] * emul.is_outfall
        if np.max(hmin) > 0:                               # tide level as a head above the global minimum
            lo = float(emul.norm_b[1][:, 1].max())
            tide = (lo + rng.uniform(0, 1, (B, Th, N)) * (nb[:, 1] - lo)) * emul.is_outfall
        b = np.concatenate([b, tide[..., None]], -1)
    ehmax = np.where(emul.ehmax > 0, emul.ehmax, 1.0)
    depth = rng.uniform(0, 1, (Bs, Ti, M)) * ehmax
    vol = rng.uniform(0, 1, (Bs, Ti, M)) * ne[:, 1]
    flow = rng.uniform(-0.2, 1, (Bs, Ti, M)) * ne[:, 2]
    sett = np.ones((Bs, Ti, M))
    a = None
    if emul.act:
        a = rng.uniform(0.1, 1.0, (B, Th, emul.n_act)).astype(np.float32)
    edge_state = np.stack([depth, vol, flow, sett], -1)
    if shared_state:
        states = np.repeat(states, B, axis=0)
        edge_state = np.repeat(edge_state, B, axis=0)
    return states.astype(np.float32), b.astype(np.float32), a, edge_state.astype(np.float32)


def random_tree(n_node, rng, max_children=3):
    """Synthetic drainage tree: node 0 is the outfall, every other node drains to an
    earlier node with at most `max_children` inflows; returns a graph.Network."""
    parent = np.zeros(n_node, np.int64)
    slots = [0] * max_children          # open inflow slots (node ids, with multiplicity)
    slots = list(slots)
    for v in range(1, n_node):
        k = int(rng.integers(0, len(slots)))
        parent[v] = slots[k]
        slots[k] = slots[-1]
        slots.pop()
        slots.extend([v] * int(rng.integers(1, max_children + 1)))
        if not slots:
            slots.append(v)
    edges = np.stack([np.arange(1, n_node), parent[1:]], 1)       # from child (upstream) to parent
    lengths = rng.uniform(50, 500, n_node - 1)
    nodes = ["n%d" % i for i in range(n_node)]
    links = ["c%d" % i for i in range(n_node - 1)]
    is_out = np.zeros(n_node, int); is_out[0] = 1
    hmax = rng.uniform(1, 5, n_node); hmax[0] = 0.0
    return _graph.Network(nodes, links, edges, lengths, is_out, np.zeros(n_node, int), hmax,
                          np.zeros(n_node), rng.uniform(0.3, 2.0, n_node - 1), np.zeros(n_node))
