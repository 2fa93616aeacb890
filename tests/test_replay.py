"""Event replay (BASELINE.json configs[3]; main.py:306-382): the device window builder against the NumPy restatement of
expand_seq / main.py:327-354 (bit-exact), replay against simulate on host-built windows (bitwise, fp32), the truth
windows of the losses, the event table, mode B against the rollout, sharding of whole events."""
import numpy as np
import pytest

import cases


def test_event_table():
    from gnn_uds_b200 import replay
    names, sets, steps = replay.shunqing_events()
    assert len(names) == 148 and sets.count("Test") == 25 and sets.count("Train") == 123
    assert int(steps.sum()) == 103020 and steps.min() == 7 * 60 and steps.max() == 25 * 60


def test_expand_seq_semantics():
    from oracle.replay_oracle import expand_seq
    d = np.arange(1, 7, dtype=np.float64).reshape(6, 1)
    w = expand_seq(d, 3)
    assert w.shape == (6, 3, 1)
    assert w[0].ravel().tolist() == [0, 0, 0] and w[2].ravel().tolist() == [0, 1, 2] and w[5].ravel().tolist() == [3, 4, 5]
    assert expand_seq(d, 3, zeros=False)[1].ravel().tolist() == [1, 1, 1]


class _Shim:
    def __init__(self, **kw):
        self.__dict__.update(kw)


@pytest.mark.parametrize("seq_in,seq_out,flood,tide", [(5, 5, True, False), (8, 4, False, True), (6, 6, True, True)])
def test_truth_windows_match_reference_construction(seq_in, seq_out, flood, tide):
    from gnn_uds_b200 import replay
    from oracle.replay_oracle import replay_windows
    rng = np.random.default_rng(3)
    T, N, M = 23, 7, 6
    states = rng.uniform(0, 2, (T, N, 4))
    perfs = (rng.uniform(0, 1, (T, N, 1)) > 0.7) * rng.uniform(0, 3, (T, N, 1))
    edge_states = rng.uniform(-1, 2, (T, M, 4))
    outfall = (rng.uniform(0, 1, N) > 0.6).astype(float)
    shim = _Shim(seq_in=seq_in, seq_out=seq_out, if_flood=flood, tide=tide, is_outfall=outfall)
    r, true, edge_true = replay.truth_windows(shim, states, perfs, edge_states)
    _, r0, _, _, true0, edge_true0 = replay_windows(states.copy(), perfs.copy(), edge_states.copy(), None, seq_in, seq_out, flood,
                                                    tide, outfall)
    assert np.array_equal(r, r0) and np.array_equal(true, true0) and np.array_equal(edge_true, edge_true0)


def test_balance_events_covers_all():
    from gnn_uds_b200 import replay, sharding
    _, _, steps = replay.shunqing_events()
    for ws in (1, 2, 4, 8):
        parts = sharding.balance_events(steps, ws)
        assert sorted(sum(parts, [])) == list(range(148))
        loads = [int(steps[p].sum()) for p in parts]
        assert max(loads) - min(loads) <= steps.max()


def _event(em, T, seed):
    from gnn_uds_b200 import replay
    return replay.synthetic_event(em, T, np.random.default_rng(seed))


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["c3_ast_act_5", "c2_shq_gat_conv_10", "ast_head_tide_offset"])
def test_window_builder_bit_exact(case):
    from oracle.replay_oracle import replay_windows
    em, _ = cases.setup_product(case)
    ev = _event(em, 47, 11)
    x, b, a, ex = em.build_windows(ev["states"], ev["perfs"], ev["edge_states"], ev["settings"])
    f32 = lambda t: None if t is None else t.astype(np.float32)
    x0, b0, a0, ex0, _, _ = replay_windows(f32(ev["states"]).copy(), f32(ev["perfs"])[..., None].copy(), f32(ev["edge_states"]).copy(),
                                           f32(ev["settings"]), em.seq_in, em.seq_out, bool(em.if_flood), bool(em.tide),
                                           np.asarray(em.is_outfall, np.float32), bool(em.act))
    assert x.shape == x0.shape and np.array_equal(x, x0.astype(np.float32))
    assert np.array_equal(b, b0.astype(np.float32))
    assert np.array_equal(ex, ex0.astype(np.float32))
    if em.act:
        assert np.array_equal(a, a0.astype(np.float32))
    # a sub-range gives the same windows
    x1, b1, a1, ex1 = em.build_windows(ev["states"], ev["perfs"], ev["edge_states"], ev["settings"], first=9, count=13)
    assert np.array_equal(x1, x[9:22]) and np.array_equal(b1, b[9:22]) and np.array_equal(ex1, ex[9:22])


@pytest.mark.gpu
@pytest.mark.parametrize("case,precision", [("c3_ast_act_5", "fp32"), ("c2_shq_gat_conv_10", "fp32"), ("c2_shq_gat_conv_10", "bf16")])
def test_replay_equals_simulate_on_reference_windows(case, precision):
    """replay (windows built on the device) == simulate on the windows the reference builds on the host; the oracle's
    forward on the same windows closes the loop to the reference arithmetic."""
    from gnn_uds_b200 import replay
    from oracle.replay_oracle import replay_windows
    em, _ = cases.setup_product(case, precision=precision)
    ev = _event(em, 33, 5)
    pred, edge_pred = em.replay(ev["states"], ev["perfs"], ev["edge_states"], ev["settings"])
    sw, r, a, ew, true, _ = replay_windows(ev["states"].copy(), ev["perfs"][..., None].copy(), ev["edge_states"].copy(), ev["settings"],
                                           em.seq_in, em.seq_out, bool(em.if_flood), bool(em.tide), np.asarray(em.is_outfall), bool(em.act))
    p2, e2 = em.simulate(sw, r, a, ew)
    assert pred.shape == p2.shape == (33 - em.seq_out, em.seq_out, em.n_node, em.y_channels)
    if precision == "fp32":
        assert np.array_equal(pred, p2) and np.array_equal(edge_pred, e2)
    else:
        assert np.abs(pred - p2).max() <= 1e-3 * max(1.0, np.abs(p2).max())
    if precision == "fp32":
        o, _ = cases.make_oracle(case, np.float64)
        y64, ey64 = o.predict(sw.astype(np.float64), r.astype(np.float64), None if a is None else a.astype(np.float64), ew.astype(np.float64))
        sc = lambda t: np.maximum(np.abs(t).reshape(-1, t.shape[-1]).max(0), 1.0)
        d = np.abs(pred - y64) / sc(y64)
        if em.if_flood:
            d[np.abs(y64[..., 3] - 0.5) < 1e-4] = 0
        assert d.max() < 2e-5 and (np.abs(edge_pred - ey64) / sc(ey64)).max() < 2e-5
    out = replay.replay_event(em, ev["states"], ev["perfs"], ev["edge_states"], ev["settings"])
    assert np.isfinite(out["loss_node"]) and np.isfinite(out["loss_edge"]) and out["true"].shape[:3] == pred.shape[:3]


@pytest.mark.gpu
def test_replay_mode_b_is_the_mpc_rollout(tmp_path):
    from gnn_uds_b200 import replay
    em, _ = cases.setup_product("c3_ast_act_5")
    ev = _event(em, 29, 8)
    y, ey, first = replay.replay_event_ar(em, ev["states"], ev["perfs"], ev["edge_states"], ev["settings"])
    n_chunks = (29 - em.seq_in) // em.seq_out
    assert first == em.seq_in and y.shape == (1, n_chunks * em.seq_out, em.n_node, em.y_channels)
    # the first chunk of the roll is the teacher-forced prediction of window seq_in
    p, ep = em.replay(ev["states"], ev["perfs"], ev["edge_states"], ev["settings"])
    assert np.array_equal(y[0, :em.seq_out], p[em.seq_in]) and np.array_equal(ey[0, :em.seq_out], ep[em.seq_in])
    out = replay.replay_event(em, ev["states"], ev["perfs"], ev["edge_states"], ev["settings"], result_dir=str(tmp_path), name="ev0")
    for key in ("runoff", "true", "pred", "edge_true", "edge_pred"):                  # main.py:377-382
        assert np.load(tmp_path / ("ev0_%s.npy" % key)).dtype == np.float32
    assert np.array_equal(np.load(tmp_path / "ev0_pred.npy"), out["pred"])


@pytest.mark.gpu
def test_replay_short_event_and_errors():
    em, _ = cases.setup_product("c3_ast_act_5")
    ev = _event(em, em.seq_out, 2)                                                     # T == seq_out: no window
    p, ep = em.replay(ev["states"], ev["perfs"], ev["edge_states"], ev["settings"])
    assert p.shape[0] == 0 and ep.shape[0] == 0
    ev = _event(em, 12, 2)
    with pytest.raises(ValueError):
        em.replay(ev["states"], None, ev["edge_states"], ev["settings"])              # flood model without cumflooding
    with pytest.raises(ValueError):
        em.replay(ev["states"][:, :, :3], ev["perfs"], ev["edge_states"], ev["settings"])


@pytest.mark.gpu
@pytest.mark.parametrize("case,precision,max_chunk", [("c3_ast_act_5", "fp32", 0), ("c3_ast_act_5", "fp32", 40),
                                                      ("c2_shq_gat_conv_10", "bf16", 64), ("c1_ast_gcn_gru", "fp32", 0)])
def test_replay_many_equals_event_by_event(case, precision, max_chunk):
    """gnnuds_replay_batch: events of mixed lengths (one without any window) cut into workspace chunks regardless of event
    boundaries - one chunk, several chunks, an event spanning two chunks - return bitwise what Emulator.replay
    returns event by event."""
    kw = {"max_chunk": max_chunk} if max_chunk else {}
    em, _ = cases.setup_product(case, precision=precision, **kw)
    lens = [em.seq_out + 7, em.seq_out, em.seq_out + 23, em.seq_out + 1, em.seq_out + 30, em.seq_out + 12]
    evs = [_event(em, t, 100 + i) for i, t in enumerate(lens)]
    many = em.replay_many(evs)
    assert len(many) == len(evs)
    for ev, (p, ep) in zip(evs, many):
        q, eq = em.replay(ev["states"], ev["perfs"], ev["edge_states"], ev["settings"])
        assert p.shape == q.shape and ep.shape == eq.shape
        if precision == "fp32" or em.n_node > 64:
            assert np.array_equal(p, q) and np.array_equal(ep, eq)
        else:      # small graphs on the tensor-core path: a sample's result depends on its slot in the MMA tile
            assert np.abs(p - q).max() <= 1e-3 * max(1.0, np.abs(q).max())
    assert em.replay_many([]) == []
    if max_chunk:                                           # one event longer than a chunk: spans two chunks
        big = _event(em, em.seq_out + max_chunk + 5, 7)
        (p, ep), = em.replay_many([big])
        q, eq = em.replay(big["states"], big["perfs"], big["edge_states"], big["settings"])
        assert p.shape == q.shape
        if precision == "fp32" or em.n_node > 64:
            assert np.array_equal(p, q) and np.array_equal(ep, eq)
