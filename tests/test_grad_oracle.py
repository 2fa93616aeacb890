"""The gradient oracle (oracle/grad_oracle.py: torch fp64 autograd through a restatement of predict_tf + the astlingen
objective, mpc.py:596-607) pinned on the CPU: its forward equals the NumPy oracle's, its gradients equal central finite
differences of the NumPy oracle."""
import numpy as np
import pytest

import cases

GRAD_CASES = ["c3_ast_act_5", "ast_gcn_lstm_act_nofusion", "ast_joint_gat_conv_act", "ast_head_tide_offset",
              "ast_nn_conv_act_fusion", "ast_pump_act_fusion", "ast_pump_act_nofusion", "ast_gat_use_adj",
              "ast_edgebased_gcn_lstm_act"]


def _weights(o, seed=3):
    from gnn_uds_b200 import mpc
    net, _ = cases.load_net("astlingen")
    fw, ww, rw = mpc.target_weights(list(net.nodes))
    return fw.astype(np.float64), ww.astype(np.float64), rw.astype(np.float64), 0.9 ** np.arange(o.seq_out)


@pytest.mark.parametrize("case", GRAD_CASES)
def test_torch_forward_equals_numpy_oracle(case):
    from oracle import grad_oracle as go
    o, inp = cases.make_oracle(case, np.float64)
    states, b, a, ex = [np.asarray(v, np.float64) if v is not None else None for v in inp]
    y, ey = o.predict_tf(states, b, a, ex)
    te = go.TorchEmulator(o)
    yt, eyt = te.predict_tf(states, b, a, ex)
    assert np.abs(yt.numpy() - y).max() <= 1e-11 * max(1.0, np.abs(y).max())
    assert np.abs(eyt.numpy() - ey).max() <= 1e-11 * max(1.0, np.abs(ey).max())


@pytest.mark.parametrize("case,chunks", [("c3_ast_act_5", 1), ("c3_ast_act_5", 2), ("ast_gcn_lstm_act_nofusion", 1),
                                         ("ast_nn_conv_act_fusion", 2), ("ast_pump_act_fusion", 1), ("ast_pump_act_nofusion", 1),
                                         ("ast_head_tide_offset", 1)])
def test_autograd_equals_finite_differences(case, chunks):
    from oracle import grad_oracle as go
    from oracle.emulator_oracle import astlingen_objective  # noqa: F401  (same formula, weight-vector form below)
    from gnn_uds_b200 import synthetic
    o, inp = cases.make_oracle(case, np.float64)
    shim = cases.SpecShim(o)
    shim.norm_x, shim.norm_b, shim.norm_e = o.norm_x, o.norm_b, o.norm_e
    rng = np.random.default_rng(11)
    B = 2
    states, b, a, ex = [np.asarray(v, np.float64) for v in synthetic.make_inputs(shim, B, rng, horizon=o.seq_out * chunks)]
    a = rng.uniform(0.15, 0.95, a.shape)
    fw, ww, rw, gm = _weights(o)
    gm = 0.9 ** np.arange(o.seq_out * chunks)

    def objective(av):
        if chunks == 1:
            y, _ = o.predict_tf(states, b, av, ex)
        else:
            y, _ = o.rollout_mpc(states, b, av, ex)
        q_in = np.concatenate([states[:, -1:, :, 1], y[..., 1]], axis=1)
        v = y[..., -1] * fw + q_in[:, 1:] * ww + np.abs(np.diff(q_in, axis=1)) * rw
        return (v.sum(-1) * gm).sum(-1)

    obj, ga = go.grad_settings(o, states, b, a, ex, fw, ww, rw, gm, chunks=chunks)
    assert np.allclose(obj, objective(a), rtol=1e-11, atol=1e-11)
    eps = 1e-6
    fd = np.zeros_like(a)
    for idx in np.ndindex(*a.shape[1:]):
        ap, am = a.copy(), a.copy()
        ap[(slice(None),) + idx] += eps
        am[(slice(None),) + idx] -= eps
        fd[(slice(None),) + idx] = (objective(ap) - objective(am)) / (2 * eps)
    scale = max(np.abs(fd).max(), 1e-12)
    assert scale > 1e-6, "degenerate case: the objective does not depend on the settings"
    err = np.abs(ga - fd).max() / scale
    print("%s chunks=%d: |grad| max %.3g, autograd vs FD rel err %.2e" % (case, chunks, scale, err))
    assert err <= 1e-5


@pytest.mark.parametrize("case", ["c3_ast_act_5", "ast_pump_act_fusion", "ast_head_tide_offset"])
def test_vjp_wrt_inputs_equals_finite_differences(case):
    """predict_vjp: random cotangents, gradient w.r.t. a few entries of the state / edge state / settings (the pumped case
    runs the pumped-storage depth scan, emulator.py:626-637, whose depth depends on the input state and the flows)."""
    from oracle import grad_oracle as go
    o, inp = cases.make_oracle(case, np.float64)
    states, b, a, ex = [np.asarray(v, np.float64)[:2] for v in inp]
    rng = np.random.default_rng(5)
    y0, ey0 = o.predict_tf(states, b, a, ex)
    gy, gey = rng.normal(size=y0.shape), rng.normal(size=ey0.shape)
    if o.if_flood:
        gy[..., -2] = 0      # flood probability feeds thresholds only after de-normalisation; keep it but unweighted
    _, _, ga, gx, gex = go.predict_vjp(o, states, b, a, ex, gy, gey)

    def f(s, aa, e):
        y, ey = o.predict_tf(s, b, aa, e)
        return (y * gy).sum() + (ey * gey).sum()

    eps = 1e-6
    for arr, g, which in ((states, gx, 0), (a, ga, 1), (ex, gex, 2)):
        flat = rng.choice(arr.size, size=12, replace=False)
        for k in flat:
            idx = np.unravel_index(k, arr.shape)
            if which == 0 and o.if_flood and idx[-1] == arr.shape[-1] - 2:
                continue                                   # the flood flag input is a 0/1 indicator
            p, m = arr.copy(), arr.copy()
            p[idx] += eps; m[idx] -= eps
            args = [states, a, ex]
            args[which] = p; fp = f(*args)
            args[which] = m; fm = f(*args)
            fd = (fp - fm) / (2 * eps)
            assert abs(fd - g[idx]) <= 1e-5 * max(1.0, abs(fd)), (which, idx, fd, g[idx])
