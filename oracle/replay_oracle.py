"""TEST INFRASTRUCTURE - CPU restatement of the reference's replay window construction.

``expand_seq`` follows ``surrogate/dataloader.py:93-95`` and ``replay_windows`` follows ``surrogate/main.py:327-354``
statement by statement (NumPy, the reference's own formulation with explicit tiling and concatenation), so that the
device window builder (``replay_windows_kernel``) is checked against an independent construction.  Only tests/,
__graft_entry__.smoke() and bench.py's CPU legs import this package.
"""
import numpy as np


def expand_seq(dats, seq, zeros=True):                                   # dataloader.py:93-95
    dats = np.stack([np.concatenate([np.tile(np.zeros_like(s) if zeros else np.ones_like(s),
                                             (max(seq - idx, 0),) + tuple(1 for _ in s.shape)),
                                     dats[max(idx - seq, 0):idx]], axis=0) for idx, s in enumerate(dats)])
    return dats


def replay_windows(states, perfs, edge_states, settings, seq_in, seq_out, if_flood, tide=False, is_outfall=None, act=False):
    """main.py:327-354: returns (states_w, r, a, edge_states_w, true, edge_true) exactly as handed to emul.simulate
    (which then keeps states[:, -seq_in:] and r[:, :seq_out], emulator.py:531-533)."""
    seq = max(seq_in, seq_out)
    states, perfs = [expand_seq(dat, seq) for dat in [states, perfs]]
    edge_states = expand_seq(edge_states, seq)
    states[..., 1] = states[..., 1] - states[..., -1]
    r, true = states[seq_out:, ..., -1:], states[seq_out:, ..., :-1]
    if tide:
        t = states[seq_out:, ..., 0] * is_outfall
        r = np.concatenate([r, np.expand_dims(t, axis=-1)], axis=-1)
    if if_flood:
        f = (perfs > 0).astype(float)
        states = np.concatenate([states[..., :-1], f, states[..., -1:]], axis=-1)
        true = np.concatenate([true, f[seq_out:]], axis=-1)
    states = states[:-seq_out]
    edge_true = edge_states[seq_out:, ..., :-1]
    edge_states = edge_states[:-seq_out]
    edge_states = edge_states[:, -seq_in:, ...]
    edge_true = edge_true[:, :seq_out, ...]
    if act:
        a = expand_seq(settings, seq_out, zeros=False)
        a = a[seq_out:, :seq_out, ...]
    else:
        a = None
    true = np.concatenate([true, perfs[seq_out:, ...]], axis=-1)          # main.py:364-365
    true = true[:, :seq_out, ...]
    return states[:, -seq_in:], r[:, :seq_out], a, edge_states, true, edge_true
