"""GPU tests of the reverse pass (gnnuds_predict_vjp / gnnuds_grad_settings; mpc.py:511-627): the hand-written VJP
kernels against the gradient oracle (torch fp64 autograd, oracle/grad_oracle.py, itself pinned against finite
differences of the NumPy oracle in tests/test_grad_oracle.py).  Tolerances (fp32 forward + fp32 VJP against fp64), written
below: 1e-4 of the largest entry for the settings gradient, 1e-3 for the state / edge-state gradients."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu

TOL = 1e-4          # settings gradients: of the largest gradient entry
TOL_STATE = 1e-3    # state / edge-state gradients: entries ~1e-3 come out of fp32 sums of O(1) cotangents (measured <= 3e-4)


def _weights(T):
    from gnn_uds_b200 import mpc
    net, _ = cases.load_net("astlingen")
    fw, ww, rw = mpc.target_weights(list(net.nodes))
    return fw, ww, rw, (0.9 ** np.arange(T)).astype(np.float32)


def _inputs(em, o, B, chunks, seed):
    from gnn_uds_b200 import synthetic
    rng = np.random.default_rng(seed)
    states, b, a, ex = synthetic.make_inputs(em, B, rng, horizon=em.seq_out * chunks)
    a = rng.uniform(0.15, 0.95, a.shape).astype(np.float32)
    return states, b, a, ex


def _rel(got, ref):
    return np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-12)


@pytest.mark.parametrize("case,chunks", [("c3_ast_act_5", 1), ("c3_ast_act_5", 3), ("ast_gcn_conv_act_nofusion", 1),
                                         ("ast_gcn_conv_act_nofusion", 2), ("ast_nn_conv_act_fusion", 1),
                                         ("ast_nn_conv_act_fusion", 2), ("ast_gcn_lstm_act_nofusion", 1),
                                         ("ast_gcn_lstm_act_nofusion", 2), ("ast_pump_act_fusion", 1), ("ast_pump_act_fusion", 2),
                                         ("ast_pump_act_nofusion", 1), ("ast_head_tide_offset", 1),
                                         ("ast_joint_gat_conv_act", 1), ("ast_gat_use_adj", 1), ("ast_gat_use_adj", 2),
                                         ("ast_edgebased_gcn_lstm_act", 1), ("ast_edgebased_gcn_lstm_act", 2)])
def test_grad_settings_matches_autograd(case, chunks):
    from oracle import grad_oracle as go
    em, _ = cases.setup_product(case)
    o, _ = cases.make_oracle(case, np.float64)
    states, b, a, ex = _inputs(em, o, 5, chunks, 31)
    fw, ww, rw, gm = _weights(em.seq_out * chunks)
    obj, ga = em.grad_settings(states, b, a, ex, fw, ww, rw, gm)
    ref_obj, ref_ga = go.grad_settings(o, states.astype(np.float64), b.astype(np.float64), a.astype(np.float64),
                                       ex.astype(np.float64), fw, ww, rw, gm, chunks=chunks)
    assert np.abs(ref_ga).max() > 1e-6
    print("%s chunks=%d: obj rel %.2e, grad rel %.2e (|grad| max %.3g)" % (case, chunks, _rel(obj, ref_obj), _rel(ga, ref_ga),
                                                                          np.abs(ref_ga).max()))
    assert np.allclose(obj, ref_obj, rtol=2e-4, atol=2e-4)
    assert _rel(ga, ref_ga) <= TOL
    # the objective returned with the gradient is the one the forward-only entry points return
    if chunks == 1:
        y, _ = em.predict_tf(states, b, a, ex)
    else:
        y, _ = em.rollout(states, b, a, ex, rule="mpc")
    assert np.array_equal(obj, em.objective(y, states, fw, ww, rw, gm))


@pytest.mark.parametrize("case", ["c3_ast_act_5", "ast_gcn_conv_eps0", "c2_shq_gat_conv_10", "ast_gcn_conv_act_nofusion",
                                  "ast_nn_conv_act_fusion", "c1_ast_gcn_gru", "ast_gat_gru_noflood", "ast_gcn_lstm_act_nofusion",
                                  "shq_nn_gru", "ast_pump_act_fusion", "ast_pump_act_nofusion", "ast_head_tide_offset",
                                  "ast_joint_gat_conv_act", "ast_joint_gcn_gru_act", "ast_gat_use_adj", "ast_edgebased_gcn_lstm_act",
                                  "c2_shq_gat_conv_60"])
def test_predict_vjp_matches_autograd(case):
    """Random cotangents on every output channel; gradients w.r.t. settings, states and edge states."""
    from oracle import grad_oracle as go
    em, inp = cases.setup_product(case)
    o, _ = cases.make_oracle(case, np.float64)
    states, b, a, ex = [None if v is None else np.asarray(v)[:2] for v in inp]
    rng = np.random.default_rng(7)
    y0, ey0 = em.predict_tf(states, b, a, ex)
    gy = rng.normal(size=y0.shape).astype(np.float32)
    gey = rng.normal(size=ey0.shape).astype(np.float32)
    out = em.predict_vjp(states, b, a, ex, gy, gey)
    f64 = lambda v: None if v is None else np.asarray(v, np.float64)
    if a is not None:
        ry, rey, rga, rgx, rgex = go.predict_vjp(o, f64(states), f64(b), f64(a), f64(ex), f64(gy), f64(gey))
    else:
        ry, rey, rga, rgx, rgex = _vjp_no_act(go, o, f64(states), f64(b), f64(ex), f64(gy), f64(gey))
    assert np.array_equal(out["y"], y0) and np.array_equal(out["ey"], ey0)
    msg = "%s: gx rel %.2e, gex rel %.2e" % (case, _rel(out["gx"], rgx), _rel(out["gex"], rgex))
    if a is not None:
        msg += ", ga rel %.2e" % _rel(out["ga"], rga)
        assert _rel(out["ga"], rga) <= TOL
    print(msg)
    assert _rel(out["gx"], rgx) <= TOL_STATE
    assert _rel(out["gex"], rgex) <= TOL_STATE


def _vjp_no_act(go, o, states, b, ex, gy, gey):
    import torch
    te = go.TorchEmulator(o)
    st = torch.as_tensor(states).clone().requires_grad_(True)
    et = torch.as_tensor(ex).clone().requires_grad_(True)
    y, ey = te.predict_tf(st, b, None, et)
    ((y * torch.as_tensor(gy)).sum() + (ey * torch.as_tensor(gey)).sum()).backward()
    return y.detach().numpy(), ey.detach().numpy(), None, st.grad.numpy(), et.grad.numpy()


def test_reverse_pass_argument_errors():
    from gnn_uds_b200._lib import GnnUdsError
    em, inp = cases.setup_product("c1_ast_gcn_gru")              # no control input: no settings gradient to ask for
    states, b, a, ex = inp
    fw, ww, rw, gm = _weights(em.seq_out)
    with pytest.raises(Exception):
        em.grad_settings(states, b, np.ones((states.shape[0], em.seq_out, 1), np.float32), ex, fw, ww, rw, gm)
    em2, inp2 = cases.setup_product("c3_ast_act_5")
    s2, b2, a2, e2 = inp2
    y, ey = em2.predict(s2, b2, a2, e2)
    with pytest.raises(ValueError):
        em2.predict_vjp(s2, b2, a2, e2, np.ones_like(y)[:, :-1], np.ones_like(ey))     # cotangent of the wrong shape
    with pytest.raises(ValueError):
        em2.grad_settings(s2, b2[:, :-1], a2, e2, fw, ww, rw, gm)                       # horizon not a whole number of chunks


def test_tensor_core_emulator_delegates_gradients_to_its_fp32_twin():
    """A tensor-core Emulator evaluates populations in 16-bit and differentiates on the fp32 path: grad_settings /
    predict_vjp go to a lazily created fp32 twin with the same weights and norms (the C ABI itself rejects a
    tensor-core handle)."""
    from gnn_uds_b200._lib import GnnUdsError
    import gnn_uds_b200._lib as L
    em2, inp2 = cases.setup_product("c3_ast_act_5", precision="bf16")
    em1, _ = cases.setup_product("c3_ast_act_5", precision="fp32")
    s2, b2, a2, e2 = inp2
    fw, ww, rw, gm = _weights(em2.seq_out)
    o2, g2 = em2.grad_settings(s2, b2, a2, e2, fw, ww, rw, gm)
    o1, g1 = em1.grad_settings(s2, b2, a2, e2, fw, ww, rw, gm)
    assert np.array_equal(o1, o2) and np.allclose(g1, g2, rtol=1e-5, atol=1e-7)
    # the raw entry point on the tensor-core handle: rejected, no silent fallback
    obj = np.empty(s2.shape[0], np.float32); ga = np.empty(a2.shape, np.float32)
    f = lambda t: np.ascontiguousarray(t, np.float32)
    x, bb, aa, ee = f(s2[:, -em2.seq_in:]), f(b2), f(a2), f(e2[:, -em2.seq_in:])
    rc = em2._lib.gnnuds_grad_settings(em2._h, L.ROLL_MPC, x.ctypes.data, bb.ctypes.data, aa.ctypes.data, ee.ctypes.data, x.shape[0], 1,
                                       fw.ctypes.data, ww.ctypes.data, rw.ctypes.data, gm.ctypes.data, obj.ctypes.data, ga.ctypes.data, L.MEM_HOST)
    assert rc == -1


def test_mpc_evaluator_gradient():
    """MpcEvaluator.gradient (mpc.py:596-607): decision vectors held 2 steps, 3 chunks of the 5->5 model, control horizon
    shorter than the evaluation horizon (the last decision is held to the end) - against finite differences of the
    evaluator's own objective in the bounded and the projected (sigmoid) parametrisation."""
    from gnn_uds_b200 import mpc, synthetic
    case = "c3_ast_act_5"
    em, _ = cases.setup_product(case)
    net, _ = cases.load_net("astlingen")
    T = em.seq_out * 3
    ev = mpc.MpcEvaluator(em, net.nodes, interval=1, eval_horizon=T, control_horizon=10, setting_duration=2,
                          gamma=0.95 ** np.arange(T))
    rng = np.random.default_rng(9)
    states, b, _, ex = synthetic.make_inputs(em, 1, rng, horizon=T)
    ev.load_state(states[0], b[0], ex[0])
    y = rng.uniform(0.2, 0.9, (3, ev.n_var))
    objs, g = ev.gradient(y)
    assert g.shape == y.shape and np.allclose(objs, ev.pred_emu(y.astype(np.float32)), rtol=1e-5, atol=1e-6)
    # directional derivative along a random direction, fp32 forward: a coarse step and a loose bound
    d = rng.normal(size=y.shape)
    eps = 2e-2
    fp = ev.pred_emu((y + eps * d).astype(np.float32)).astype(np.float64)
    fm = ev.pred_emu((y - eps * d).astype(np.float32)).astype(np.float64)
    fd = (fp - fm) / (2 * eps)
    an = (g * d).sum(axis=1)
    print("directional derivative: analytic %s, finite differences %s" % (an, fd))
    assert np.allclose(an, fd, rtol=5e-2, atol=5e-3)
    # projected parametrisation: chain rule of the sigmoid
    z = ev.project(y, inverse=True)
    objs2, gz = ev.gradient(z, bounded=False)
    yy = ev.project(z)
    assert np.allclose(gz, ev.gradient(yy)[1] * yy * (1 - yy), rtol=1e-5, atol=1e-7)
    hv = ev.hessp(y[0], d[0])
    assert hv.shape == y[0].shape and np.all(np.isfinite(hv))
    hvb = ev.hessp(y, d)                                     # a batch of (y, p) pairs: row i is the product of pair i
    assert hvb.shape == y.shape and np.allclose(hvb[0], hv, rtol=1e-4, atol=1e-4 * max(np.abs(hv).max(), 1e-12))


def test_grad_settings_graph_replay_and_eager_agree(monkeypatch):
    """The gradient body runs as a cached CUDA graph (a shape runs eagerly once, is captured on its second use and replayed afterwards);
    GNNUDS_GRAPHS=0 runs the same launches eagerly.  First call, replays with other inputs, and the eager path agree."""
    case = "c3_ast_act_5"
    em, _ = cases.setup_product(case)
    o, _ = cases.make_oracle(case, np.float64)
    fw, ww, rw, gm = _weights(em.seq_out * 2)
    s1, b1, a1, e1 = _inputs(em, o, 4, 2, 3)
    s2, b2, a2, e2 = _inputs(em, o, 4, 2, 4)
    o1, g1 = em.grad_settings(s1, b1, a1, e1, fw, ww, rw, gm)          # first use of the shape: eager
    o2, g2 = em.grad_settings(s2, b2, a2, e2, fw, ww, rw, gm)          # second use: capture + launch, other inputs
    o1r, g1r = em.grad_settings(s1, b1, a1, e1, fw, ww, rw, gm)        # replay, first inputs again
    o2r, g2r = em.grad_settings(s2, b2, a2, e2, fw, ww, rw, gm)        # replay
    assert np.array_equal(o2, o2r) and np.allclose(g2, g2r, rtol=1e-5, atol=1e-7)
    assert np.array_equal(o1, o1r) and np.allclose(g1, g1r, rtol=1e-5, atol=1e-7)
    assert not np.allclose(g1, g2)
    monkeypatch.setenv("GNNUDS_GRAPHS", "0")
    em0, _ = cases.setup_product(case)
    o1e, g1e = em0.grad_settings(s1, b1, a1, e1, fw, ww, rw, gm)
    o2e, g2e = em0.grad_settings(s2, b2, a2, e2, fw, ww, rw, gm)
    assert np.array_equal(o1, o1e) and np.array_equal(o2, o2e)
    assert np.allclose(g1, g1e, rtol=1e-5, atol=1e-7) and np.allclose(g2, g2e, rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("case,chunks", [("c3_ast_act_5", 3), ("ast_nn_conv_act_fusion", 2), ("ast_gcn_lstm_act_nofusion", 2),
                                         ("ast_pump_act_fusion", 2), ("ast_gat_use_adj", 2), ("ast_edgebased_gcn_lstm_act", 2)])
def test_grad_settings_kept_tapes_and_checkpointed_sweeps_agree(case, chunks, monkeypatch):
    """A multi-chunk roll keeps the tapes of all chunks when they fit GNNUDS_TAPE_MB (default 16 GiB: the mode the other tests
    of this file run) and otherwise re-runs every chunk with the tape on during the reverse sweep (GNNUDS_TAPE_MB=0 forces
    it).  Both sweeps against the gradient oracle and against each other, first call (eager) and graph replays."""
    from oracle import grad_oracle as go
    o, _ = cases.make_oracle(case, np.float64)
    em, _ = cases.setup_product(case)
    states, b, a, ex = _inputs(em, o, 5, chunks, 17)
    fw, ww, rw, gm = _weights(em.seq_out * chunks)
    ref_obj, ref_ga = go.grad_settings(o, states.astype(np.float64), b.astype(np.float64), a.astype(np.float64),
                                       ex.astype(np.float64), fw, ww, rw, gm, chunks=chunks)
    kept = [em.grad_settings(states, b, a, ex, fw, ww, rw, gm) for _ in range(3)]
    monkeypatch.setenv("GNNUDS_TAPE_MB", "0")
    em0, _ = cases.setup_product(case)
    ckpt = [em0.grad_settings(states, b, a, ex, fw, ww, rw, gm) for _ in range(3)]
    for obj, ga in kept + ckpt:
        assert np.array_equal(obj, kept[0][0])
        assert _rel(ga, ref_ga) <= TOL
        assert np.allclose(ga, kept[0][1], rtol=1e-5, atol=1e-6 * np.abs(ref_ga).max())
    assert np.allclose(kept[0][0], ref_obj, rtol=2e-4, atol=2e-4)


def test_grad_settings_blocks_of_decision_vectors():
    """More decision vectors than one internal pass holds (max_chunk=2, B=5: blocks of 2, 2 and 1 through the same arena and
    workspace, kept tapes and checkpointed): every vector's objective and gradient equal the single-block result."""
    case, chunks = "c3_ast_act_5", 3
    o, _ = cases.make_oracle(case, np.float64)
    em, _ = cases.setup_product(case)
    states, b, a, ex = _inputs(em, o, 5, chunks, 23)
    fw, ww, rw, gm = _weights(em.seq_out * chunks)
    obj, ga = em.grad_settings(states, b, a, ex, fw, ww, rw, gm)
    em2, _ = cases.setup_product(case, max_chunk=2)
    for _ in range(3):                                       # eager, captured, replayed
        obj2, ga2 = em2.grad_settings(states, b, a, ex, fw, ww, rw, gm)
        assert np.array_equal(obj, obj2)
        assert np.allclose(ga, ga2, rtol=1e-5, atol=1e-6 * np.abs(ga).max())


def test_device_pointer_entry_points_equal_host_ones():
    """gnnuds_grad_settings / gnnuds_predict_vjp with where = GNNUDS_MEM_DEVICE (caller-owned device buffers, allocated here
    through gnnuds_alloc_device / gnnuds_memcpy): the same results as the host-pointer calls, in place in the caller's
    buffers; B = 0 is a no-op."""
    import ctypes as C
    import gnn_uds_b200._lib as L
    em, _ = cases.setup_product("c3_ast_act_5")
    o, _ = cases.make_oracle("c3_ast_act_5", np.float64)
    chunks = 2
    states, b, a, ex = _inputs(em, o, 3, chunks, 17)
    fw, ww, rw, gm = _weights(em.seq_out * chunks)
    ref_obj, ref_ga = em.grad_settings(states, b, a, ex, fw, ww, rw, gm)
    ptrs = []

    def put(arr):
        arr = np.ascontiguousarray(arr, np.float32)
        p = C.c_void_p()
        L.check(em._h, em._lib.gnnuds_alloc_device(em._h, C.byref(p), max(arr.nbytes, 4)))
        L.check(em._h, em._lib.gnnuds_memcpy(em._h, p, arr.ctypes.data_as(C.c_void_p), arr.nbytes, 1))
        ptrs.append(p)
        return p

    def get(p, shape):
        out = np.empty(shape, np.float32)
        L.check(em._h, em._lib.gnnuds_memcpy(em._h, out.ctypes.data_as(C.c_void_p), p, out.nbytes, 2))
        return out

    B = states.shape[0]
    x0, e0 = states[:, -em.seq_in:], ex[:, -em.seq_in:]
    dx, db, da, de = put(x0), put(b), put(a), put(e0)
    dobj, dga = put(np.full(B, np.nan)), put(np.full(a.shape, np.nan))
    vp = lambda t: t.ctypes.data_as(C.c_void_p)
    for _ in range(3):                                              # eager, capture, replay
        L.check(em._h, em._lib.gnnuds_grad_settings(em._h, L.ROLL_MPC, dx, db, da, de, B, chunks, vp(fw), vp(ww), vp(rw), vp(gm), dobj, dga,
                                                    L.MEM_DEVICE))
        assert np.array_equal(get(dobj, (B,)), ref_obj)
        assert np.allclose(get(dga, a.shape), ref_ga, rtol=1e-5, atol=1e-7)
    assert em._lib.gnnuds_grad_settings(em._h, L.ROLL_MPC, dx, db, da, de, 0, chunks, vp(fw), vp(ww), vp(rw), vp(gm), dobj, dga, L.MEM_DEVICE) == 0
    # predict_vjp on device pointers (first chunk)
    b1, a1 = np.ascontiguousarray(b[:, :em.seq_out]), np.ascontiguousarray(a[:, :em.seq_out])
    y0, ey0 = em.predict_tf(x0, b1, a1, e0)
    rng = np.random.default_rng(2)
    gy, gey = rng.normal(size=y0.shape).astype(np.float32), rng.normal(size=ey0.shape).astype(np.float32)
    ref = em.predict_vjp(x0, b1, a1, e0, gy, gey)
    db1, da1, dgy, dgey = put(b1), put(a1), put(gy), put(gey)
    dy, dey, dga1, dgx, dgex = put(np.zeros_like(y0)), put(np.zeros_like(ey0)), put(np.zeros_like(a1)), put(np.zeros_like(x0)), put(np.zeros_like(e0))
    L.check(em._h, em._lib.gnnuds_predict_vjp(em._h, dx, db1, da1, de, B, dgy, dgey, dy, dey, dga1, dgx, dgex, L.MEM_DEVICE))
    assert np.array_equal(get(dy, y0.shape), y0) and np.array_equal(get(dey, ey0.shape), ey0)
    for p, name, shape in ((dga1, "ga", a1.shape), (dgx, "gx", x0.shape), (dgex, "gex", e0.shape)):
        assert np.allclose(get(p, shape), ref[name], rtol=1e-5, atol=1e-7), name
    for p in ptrs:
        em._lib.gnnuds_free_device(em._h, p)
