"""Event replay: the caller side of BASELINE.json configs[3] (``surrogate/main.py:306-382``, the test loop).

The reference loads (or simulates) one rainfall event, expands every series seq-fold into windows on the host
(``DataGenerator.expand_seq``, dataloader.py:93-95; main.py:327-354), calls ``emul.simulate`` window by window
(emulator.py:521-564), prints three losses and writes five ``.npy`` files (main.py:361-382).  Here the raw series go to
the device once and the windows are built there (``gnnuds_replay`` / ``replay_windows_kernel``); this module keeps the
rest of the loop: the event list, the ground-truth windows of the losses and writers (strided views, no copies), mode B
(a true autoregressive roll from the first full window, ``mpc.py:565-582`` feedback rule) and the sharding of whole
events over ranks.
"""
import os

import numpy as np
from numpy.lib.stride_tricks import sliding_window_view

from . import sharding

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def shunqing_events():
    """(names, sets, steps): the 148 events of envs/network/shunqing/events.csv; an event of `Duration` hours holds
    (Duration + 1) * 60 one-minute steps (103 020 steps in total)."""
    z = np.load(os.path.join(_DATA, "shunqing_events.npz"))
    return [str(s) for s in z["name"]], [str(s) for s in z["set"]], ((z["duration_h"] + 1) * 60).astype(np.int64)


def _windows(series, first, count, length):
    """Windows `first`..`first+count-1` of `length` steps, window w = steps w-length..w-1 of `series` (T, ...), zeros
    before the event start (expand_seq semantics) - as a strided view of a left-padded copy of the series."""
    pad = np.zeros((length,) + series.shape[1:], series.dtype)
    ext = np.concatenate([pad, series], 0)                         # ext[k] = series[k - length]
    v = sliding_window_view(ext, length, axis=0)                   # v[w] = ext[w..w+length-1] = series[w-length..w-1]
    v = np.moveaxis(v, -1, 1)
    return v[first:first + count]


def truth_windows(em, states, perfs, edge_states):
    """r, true, edge_true of main.py:331-365 for the losses and the .npy writers: the observed future of every window.
    states (T,N,4) [h, cuminflow, outflow, lateral], perfs (T,N[,1]) cumflooding, edge_states (T,M,4)."""
    Ti, To = em.seq_in, em.seq_out
    seq = max(Ti, To)
    T = states.shape[0]
    W = max(T - To, 0)
    st = np.array(states, copy=True)
    st[..., 1] = st[..., 1] - st[..., -1]                          # main.py:331
    pf = np.asarray(perfs).reshape(T, -1, 1)
    fut = _windows(st, To, W, seq)                                 # states[seq_out:]
    r = fut[..., -1:]
    true = fut[..., :-1]
    if em.tide:
        r = np.concatenate([r, fut[..., :1] * np.asarray(em.is_outfall).reshape(1, 1, -1, 1)], -1)
    pfw = _windows(pf, To, W, seq)
    if em.if_flood:
        true = np.concatenate([true, (pfw > 0).astype(st.dtype)], -1)
    true = np.concatenate([true, pfw], -1)[:, :To]                 # main.py:364-365: cumflooding appended, first seq_out steps
    edge_true = _windows(np.asarray(edge_states), To, W, seq)[..., :-1][:, :To]
    return r[:, :To], true, edge_true


def mse(a, b):
    return float(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))


def bce(p, y, eps=1e-7):
    p = np.clip(np.asarray(p, np.float64), eps, 1 - eps)
    y = np.asarray(y, np.float64)
    return float(np.mean(-(y * np.log(p) + (1 - y) * np.log(1 - p))))


def replay_event(em, states, perfs, edge_states, settings=None, result_dir=None, name="event", losses=True):
    """One event through ``Emulator.replay`` (teacher-forced windows, mode A) + the reference's test losses
    (main.py:367-375: node MSE and edge MSE in normalised space, flood BCE) and result files (main.py:377-382)."""
    pred, edge_pred = em.replay(states, perfs, edge_states, settings)
    out = {"pred": pred, "edge_pred": edge_pred}
    if losses or result_dir:
        r, true, edge_true = truth_windows(em, states, perfs, edge_states)
        out.update(r=r, true=true, edge_true=edge_true)
        if losses and pred.shape[0]:
            out["loss_node"] = mse(em.normalize(pred[..., :3], "y"), em.normalize(true[..., :3], "y"))
            if em.if_flood:
                out["loss_flood"] = bce(pred[..., -2:-1], true[..., -2:-1])
            out["loss_edge"] = mse(em.normalize(edge_pred, "e"), em.normalize(edge_true, "e"))
        if result_dir:
            os.makedirs(result_dir, exist_ok=True)
            for key, arr in (("runoff", r), ("true", true), ("pred", pred), ("edge_true", edge_true), ("edge_pred", edge_pred)):
                np.save(os.path.join(result_dir, "%s_%s.npy" % (name, key)), np.asarray(arr, np.float32))
    return out


def replay_event_ar(em, states, perfs, edge_states, settings=None):
    """Mode B: a TRUE autoregressive roll of one event from its first full window (steps 0..seq_in-1), the model fed with
    its own predictions chunk after chunk (mpc.py:565-582 feedback rule) and the recorded runoff / settings.
    Returns (pred (1, n_chunks*seq_out, N, y_channels), edge_pred, first_step): predictions of steps first_step.."""
    Ti, To = em.seq_in, em.seq_out
    if Ti != To:
        raise ValueError("the autoregressive roll feeds seq_out steps back as the next seq_in steps (mpc.py:573-579)")
    T = states.shape[0]
    n_chunks = (T - Ti) // To
    if n_chunks < 1:
        raise ValueError("event of %d steps is shorter than one window plus one chunk" % T)
    x, b, a, ex = em.build_windows(states, perfs, edge_states, settings, Ti, 1)    # window Ti: steps 0..Ti-1
    hor = n_chunks * To
    st = np.asarray(states, np.float32)
    bb = st[Ti:Ti + hor, :, -1:][None]
    if em.tide:
        bb = np.concatenate([bb, (st[Ti:Ti + hor, :, :1] * np.asarray(em.is_outfall, np.float32).reshape(1, -1, 1))[None]], -1)
    aa = None if settings is None or not em.act else np.asarray(settings, np.float32)[Ti:Ti + hor][None]
    y, ey = em.rollout(x, bb, aa, ex, rule="mpc")
    return y, ey, Ti


def replay_events(em, events, mode="windows"):
    """All events of this rank: `events` is a list of dicts with states / perfs / edge_states (/ settings); whole events are
    balanced over ranks by length (sharding.balance_events).  Returns {event index: result} for the events of this rank and
    the predicted node-timesteps of the rank (reference accounting: (T - seq_out) * seq_out * N per event in mode A)."""
    rank, ws = sharding.world()
    mine = sharding.balance_events([e["states"].shape[0] for e in events], ws)[rank]
    res, units = {}, 0
    if mode == "windows":                 # all events of the rank in one call: short events share a batch (gnnuds_replay_batch)
        for i, (p, ep) in zip(mine, em.replay_many([events[i] for i in mine])):
            res[i] = (p, ep)
            units += p.shape[0] * p.shape[1] * p.shape[2]
        return res, units
    for i in mine:
        e = events[i]
        p, ep, _ = replay_event_ar(em, e["states"], e.get("perfs"), e["edge_states"], e.get("settings"))
        res[i] = (p, ep)
        units += p.shape[0] * p.shape[1] * p.shape[2]
    return res, units


def synthetic_event(em, T, rng):
    """Recorded series of one synthetic event in physical units: states (T,N,4) [h, cuminflow, outflow, lateral],
    perfs (T,N) cumflooding (mostly zero), edge_states (T,M,4), settings (T,n_act) or None."""
    N, M = em.n_node, em.n_edge
    nx, ne = em.norm_x[0], em.norm_e[0]
    hmax = np.where(em.hmax > 0, em.hmax, 1.0)
    t = np.arange(T)[:, None]
    peak = rng.integers(T // 6 + 1, T // 2 + 2)
    hyeto = np.exp(-0.5 * ((t - peak) / (T / 12.0 + 1)) ** 2)                          # one storm pulse
    lateral = hyeto * rng.uniform(0.2, 1.0, (1, N)) * em.norm_b[0][:, 0]
    h = np.clip(hyeto * rng.uniform(0.1, 1.0, (1, N)) + 0.05 * rng.standard_normal((T, N)), 0, 1) * hmax
    outflow = hyeto * rng.uniform(0.1, 1.0, (1, N)) * nx[:, 2]
    cuminflow = lateral + hyeto * rng.uniform(0.1, 1.0, (1, N)) * nx[:, 1]
    states = np.stack([h, cuminflow, outflow, lateral], -1).astype(np.float32)
    perfs = (np.maximum(hyeto - 0.8, 0) * (rng.uniform(0, 1, (1, N)) < 0.15) * rng.uniform(0, 5, (1, N))).astype(np.float32)
    ehmax = np.where(em.ehmax > 0, em.ehmax, 1.0)
    depth = np.clip(hyeto * rng.uniform(0.1, 1.0, (1, M)), 0, 1) * ehmax
    vol = hyeto * rng.uniform(0.1, 1.0, (1, M)) * ne[:, 1]
    flow = hyeto * rng.uniform(-0.2, 1.0, (1, M)) * ne[:, 2]
    edge_states = np.stack([depth, vol, flow, np.ones((T, M))], -1).astype(np.float32)
    settings = rng.uniform(0.1, 1.0, (T, em.n_act)).astype(np.float32) if em.act else None
    if settings is not None:
        edge_states[:, em.act_edge_index(), 3] = settings
    return dict(states=states, perfs=perfs, edge_states=edge_states, settings=settings)
