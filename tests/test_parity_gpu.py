"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle.

fp32 bar (north_star): within 1e-5 relative per step.  "Relative" is taken against the
per-channel output scale (max |truth| of that channel, floor 1): outputs are physical
depths/flows of O(1) and many entries are exactly 0 after clipping, so element-wise
relative error is undefined there.  The float64 oracle is the truth; the float32 oracle's
own distance from it is printed for reference."""
import os

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _scale(t):
    return np.maximum(np.abs(t).reshape(-1, t.shape[-1]).max(0), 1.0)


def _check(name, got, truth, ref32=None, mask=None, rtol=RTOL):
    err = np.abs(got.astype(np.float64) - truth) / _scale(truth)
    if mask is not None:
        err = np.where(mask, 0.0, err)
    worst = err.max()
    msg = "%s: max scaled err %.3g" % (name, worst)
    if ref32 is not None:
        msg += " (float32 oracle: %.3g)" % (np.abs(ref32.astype(np.float64) - truth) / _scale(truth)).max()
    print(msg)
    assert np.isfinite(got).all(), name
    assert worst <= rtol, msg


def _flag_mask(em, y64):
    """Elements whose value depends on a threshold the truth sits within 1e-4 of."""
    if not em.if_flood:
        return None
    near = np.abs(y64[..., 3] - 0.5) < 1e-4
    m = np.zeros(y64.shape, bool)
    m[..., 0] = near
    m[..., -1] = near
    return m


STAGES = [(1, "embed"), (2, "sp1"), (3, "tp1"), (4, "sp2"), (5, "tp2"), (6, "res")]


@pytest.mark.parametrize("case", ["c2_shq_gat_conv_10", "c1_ast_gcn_gru", "c3_ast_act_5"])
def test_stage_by_stage(case):
    em, inp = cases.setup_product(case)
    o, inp_o = cases.make_oracle(case, np.float64)
    o.trace = {}
    o.predict(*inp_o)
    for stage, key in STAGES:
        xs, es = em.debug_stage(stage, *inp)
        tx, te = o.trace[key]
        w = xs.shape[1]
        _check("%s/%s node" % (case, key), xs, tx.reshape(-1, tx.shape[-1])[:, :w], rtol=2e-5)
        _check("%s/%s edge" % (case, key), es, te.reshape(-1, te.shape[-1])[:, :es.shape[1]], rtol=2e-5)


@pytest.mark.parametrize("case", cases.GOLDEN_CASES + ["c2_shq_gat_conv_60"])
def test_predict_matches_oracle(case):
    em, inp = cases.setup_product(case)
    y, ey = em.predict(*inp)
    o, inp_o = cases.make_oracle(case, np.float64)
    y64, ey64 = o.predict(*inp_o)
    o32, _ = cases.make_oracle(case, np.float32)
    y32, ey32 = o32.predict(*inp_o)
    assert y.shape == y64.shape and ey.shape == ey64.shape
    _check(case + " y", y, y64, y32, _flag_mask(em, y64))
    _check(case + " ey", ey, ey64, ey32)
    if em.if_flood:
        flips = ((y[..., 3] > 0.5) != (y64[..., 3] > 0.5)).mean()
        print("flood flag flip rate", flips)
        assert flips <= 1e-3


@pytest.mark.parametrize("case", ["ast_pump_act_fusion", "ast_pump_act_nofusion", "c3_ast_act_5"])
@pytest.mark.parametrize("all_pumps", [False, True])
def test_predict_tf_follows_post_proc_tf(case, all_pumps, monkeypatch):
    """predict_tf runs post_proc_tf (emulator.py:680-725), which differs from post_proc on pumped networks: rated pump
    flows override link flows only if EVERY link is a pump (:698), node-side gates use h > 0 (:710-711)."""
    if all_pumps:
        if "pump" not in case:
            pytest.skip("pump-free case")
        orig = cases.build_args

        def patched(c):
            args, B, seed = orig(c)
            args.pump = np.where(args.pump > 0, args.pump, 0.07)      # every link a pump -> pump.min() > 0
            return args, B, seed
        monkeypatch.setattr(cases, "build_args", patched)
    em, inp = cases.setup_product(case)
    o, inp_o = cases.make_oracle(case, np.float64)
    y, ey = em.predict_tf(*inp)
    y64, ey64 = o.predict_tf(*inp_o)
    _check(case + " tf y", y, y64, mask=_flag_mask(em, y64))
    _check(case + " tf ey", ey, ey64)
    yn, eyn = em.predict(*inp)                                        # the handle is back on the numpy rules
    y64n, ey64n = o.predict(*inp_o)
    _check(case + " np y", yn, y64n, mask=_flag_mask(em, y64n))
    _check(case + " np ey", eyn, ey64n)
    if "pump" in case and not all_pumps:
        assert np.abs(ey64 - ey64n).max() > 1e-3                       # the two variants really differ here


@pytest.mark.parametrize("case", cases.GOLDEN_CASES)
def test_predict_matches_committed_golden(case):
    z = np.load(os.path.join(cases.GOLD, "fwd_%s.npz" % case))
    em, inp = cases.setup_product(case)
    y, ey = em.predict(*inp)
    _check(case + " y (golden)", y, z["y64"], z["y32"], _flag_mask(em, z["y64"]))
    _check(case + " ey (golden)", ey, z["ey64"], z["ey32"])


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("max_chunk", [4, 1])
def test_chunking_is_invisible(max_chunk, precision):
    """Host buffers streamed through 2 and 6 internal chunks (the pipelined staging of gnnuds_predict reuses a
    staging set from the third chunk on; on the tensor-core path a chunk of one scenario also ends in a partial
    sample group): scenarios are independent, so the result is bitwise that of the single pass."""
    em, inp = cases.setup_product("c3_ast_act_5", precision=precision)
    y, ey = em.predict(*inp)
    em2, _ = cases.setup_product("c3_ast_act_5", precision=precision, max_chunk=max_chunk)
    y2, ey2 = em2.predict(*inp)
    if precision == "fp32":
        assert np.array_equal(y, y2) and np.array_equal(ey, ey2)      # shards are independent -> bitwise
    else:
        # tensor-core path on a small graph: 4 samples share one 128-row MMA tile (block-diagonal NodeEdge operand), and
        # which slot of the tile a sample occupies depends on the chunking.  The tensor core's multi-term fp32
        # accumulation is not bitwise slot-invariant for fp16 operands (it was for bf16, whose products are shorter), so
        # the results agree to a few fp16 ulps of the intermediate activations instead of bitwise.
        sc = lambda t: np.maximum(np.abs(t).reshape(-1, t.shape[-1]).max(0), 1.0)
        same = (y[..., 3] > 0.5) == (y2[..., 3] > 0.5)
        assert same.mean() > 0.999
        assert (np.abs(y - y2)[same] / sc(y)).max() < 1e-3 and (np.abs(ey - ey2) / sc(ey)).max() < 1e-3


def test_simulate_equals_predict_windows():
    em, (states, b, a, ex) = cases.setup_product("c2_shq_gat_conv_10")
    runoff = np.concatenate([b, b[:, :3]], axis=1)                 # windows longer than seq_out
    y, ey = em.simulate(states, runoff, None, ex)                  # a=None passes through (F9)
    y2, ey2 = em.predict(states, b, None, ex)
    assert np.array_equal(y, y2) and np.array_equal(ey, ey2)


@pytest.mark.parametrize("rule", ["mpc", "mbrl"])
def test_rollout_matches_oracle(rule):
    case = "c3_ast_act_5"
    em, _ = cases.setup_product(case)
    o, _ = cases.make_oracle(case, np.float64)
    rng = np.random.default_rng(7)
    from gnn_uds_b200 import synthetic
    states, b, a, ex = synthetic.make_inputs(em, 5, rng, horizon=4 * em.seq_out)
    y, ey = em.rollout(states, b, a, ex, rule=rule)
    fn = o.rollout_mpc if rule == "mpc" else o.rollout_mbrl
    y64, ey64 = fn(states.astype(np.float64), b.astype(np.float64), a.astype(np.float64), ex.astype(np.float64))
    assert y.shape == y64.shape
    # feedback thresholds (q_w>0, prob>0.5) can flip on rounding; compare scenario-steps before any flip
    flag_g = (y[..., -1] > 0) if rule == "mpc" else (y[..., 3] > 0.5)
    flag_o = (y64[..., -1] > 0) if rule == "mpc" else (y64[..., 3] > 0.5)
    ok = np.ones(y.shape[0], bool)
    good = 0
    for k in range(4):
        sl = slice(k * em.seq_out, (k + 1) * em.seq_out)
        if ok.any():
            _check("%s chunk %d y" % (rule, k), y[ok, sl], y64[ok, sl], mask=None if not em.if_flood else _flag_mask(em, y64[ok, sl]), rtol=5e-5 * (k + 1))
            _check("%s chunk %d ey" % (rule, k), ey[ok, sl], ey64[ok, sl], rtol=5e-5 * (k + 1))
            good += int(ok.sum())
        near = (np.abs(y64[:, sl, :, 3] - 0.5) < 1e-4).any(axis=(1, 2)) | ((np.abs(y64[:, sl, :, -1]) < 1e-5) & (y64[:, sl, :, -1] > 0)).any(axis=(1, 2))
        ok &= (flag_g[:, sl] == flag_o[:, sl]).all(axis=(1, 2)) & ~near
    assert good >= 5 * 2        # at least the first two chunks of every scenario were comparable


def test_rollout_norm_rule():
    case = "c3_ast_act_5"
    em, _ = cases.setup_product(case)
    o, _ = cases.make_oracle(case, np.float64)
    o.roll = 3
    rng = np.random.default_rng(9)
    from gnn_uds_b200 import synthetic
    states, b, a, ex = synthetic.make_inputs(em, 4, rng, horizon=3 * em.seq_out)
    xn, bn, exn = em.normalize(states, "x").astype(np.float32), em.normalize(b, "b").astype(np.float32), em.normalize(ex, "e").astype(np.float32)
    y, ey = em.rollout(xn, bn, a, exn, rule="norm")
    y64, ey64 = o.rollout_norm(xn.astype(np.float64), a.astype(np.float64), bn.astype(np.float64), exn.astype(np.float64))
    # every chunk of the roll loop (emulator.py:416-423 feeds post_proc'ed, UNclipped predictions + the 0.5-thresholded
    # flood flag back): a scenario stays comparable until its flag (prob > 0.5) could flip on rounding
    T = em.seq_out
    ok = np.ones(y.shape[0], bool)
    good = 0
    for k in range(3):
        sl = slice(k * T, (k + 1) * T)
        assert ok.any()
        _check("norm chunk %d y" % k, y[ok, sl], y64[ok, sl], rtol=2e-5 * (k + 1))
        _check("norm chunk %d ey" % k, ey[ok, sl], ey64[ok, sl], rtol=2e-5 * (k + 1))
        good += int(ok.sum())
        near = (np.abs(y64[:, sl, :, -1] - 0.5) < 1e-4).any(axis=(1, 2))
        ok &= ((y[:, sl, :, -1] > 0.5) == (y64[:, sl, :, -1] > 0.5)).all(axis=(1, 2)) & ~near
    assert good >= 4 * 2 + 2, good      # the feedback step itself is verified on (nearly) every scenario


def test_simulate_c_abi_window_repack_and_error_path():
    """gnnuds_simulate straight through ctypes: runoff windows longer than seq_out (the C side repacks the first
    seq_out steps, emulator.py:532) equal predict on the cut windows; device pointers with seq_r > seq_out are refused."""
    import ctypes as C
    from gnn_uds_b200 import _lib
    em, (states, b, a, ex) = cases.setup_product("c2_shq_gat_conv_10")
    B, To = states.shape[0], em.seq_out
    seq_r = To + 3
    rng = np.random.default_rng(11)
    runoff = np.concatenate([b, rng.uniform(0, 1, (B, 3) + b.shape[2:]).astype(np.float32)], axis=1)
    y = np.empty((B, To, em.n_node, em.y_channels), np.float32)
    ey = np.empty((B, To, em.n_edge, em.ey_channels), np.float32)
    vp = lambda t: t.ctypes.data_as(C.c_void_p)
    rc = em._lib.gnnuds_simulate(em._h, vp(states), vp(runoff), seq_r, None, vp(ex), B, vp(y), vp(ey), _lib.MEM_HOST)
    assert rc == 0, em._lib.gnnuds_last_error(em._h)
    y2, ey2 = em.predict(states, b, None, ex)
    assert np.array_equal(y, y2) and np.array_equal(ey, ey2)
    rc = em._lib.gnnuds_simulate(em._h, vp(states), vp(runoff), seq_r, None, vp(ex), B, vp(y), vp(ey), _lib.MEM_DEVICE)
    assert rc == _lib.E_INVALID if hasattr(_lib, "E_INVALID") else rc == -1
    assert b"host buffers" in em._lib.gnnuds_last_error(em._h)
    rc = em._lib.gnnuds_simulate(em._h, vp(states), vp(runoff), To - 1, None, vp(ex), B, vp(y), vp(ey), _lib.MEM_HOST)
    assert rc == -1


def test_objective_matches_oracle():
    from oracle.emulator_oracle import astlingen_objective
    em, inp = cases.setup_product("c3_ast_act_5")
    y, ey = em.predict(*inp)
    net, _ = cases.load_net("astlingen")
    targets = [("T%d" % i, "cumflooding", 2 if i == 6 else 1) for i in range(1, 7)] + \
              [("CSO7", "cumflooding", 2), ("CSO8", "cumflooding", 1), ("CSO9", "cumflooding", 2), ("CSO10", "cumflooding", 1),
               ("Out_to_WWTP", "cuminflow", -0.1)] + [(n, "cuminflow", 0.01) for n in ("J19", "J17", "J14", "J6", "J2", "J9")]
    fw, ww, rw = np.zeros(30), np.zeros(30), np.zeros(30)
    for n, attr, w in targets:
        i = net.nodes.index(n)
        if attr == "cumflooding":
            fw[i] += w
        elif "WWTP" in n:
            ww[i] += w
        else:
            rw[i] += w
    gamma = 0.95 ** np.arange(em.seq_out)
    for rough in (True, False):
        obj = em.objective(y, inp[0], fw, ww, rw if rough else None, gamma)
        ref = astlingen_objective(y.astype(np.float64), inp[0].astype(np.float64), net.nodes, targets, gamma, rough)
        assert np.allclose(obj, ref, rtol=1e-5, atol=1e-5), (obj, ref)


def test_large_batch_properties():
    """Full-size style checks the domain offers: batch permutation equivariance, replication
    (8192 MPC candidates share one state), determinism."""
    em, _ = cases.setup_product("c3_ast_act_5")
    from gnn_uds_b200 import synthetic
    rng = np.random.default_rng(3)
    states, b, a, ex = synthetic.make_inputs(em, 2048, rng, shared_state=True)
    y, ey = em.predict(states, b, a, ex)
    y2, ey2 = em.predict(states, b, a, ex)
    assert np.array_equal(y, y2)
    perm = rng.permutation(2048)
    yp, eyp = em.predict(states[perm], b[perm], a[perm], ex[perm])
    assert np.array_equal(yp, y[perm]) and np.array_equal(eyp, ey[perm])
    a2 = a.copy(); a2[1] = a[0]; b2 = b.copy(); b2[1] = b[0]
    y3, _ = em.predict(states, b2, a2, ex)
    assert np.array_equal(y3[0], y3[1])
    # physical invariants of constrain (emulator.py:727-748)
    assert (y[..., 0] >= -1e-6).all() and (y[..., 0] <= em.hmax + 1e-5).all() and (y[..., -1] >= 0).all()
    assert (y[..., em.is_outfall.astype(bool), -1] == 0).all()
    assert (ey[..., 0] >= 0).all() and (ey[..., 0] <= em.ehmax + 1e-6).all()


@pytest.mark.parametrize("case", ["ast_gat_use_adj", "ast_gat_use_adj_weighted"])
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_use_adj_dynamic_mask(case, precision):
    """use_adj (emulator.py:268-271, 343-362): the block-2 node attention runs over the per-sample `adj.astype(int)` mask.
    Besides the oracle comparison, switching the dynamic mask off must change the result by what it changes in the
    oracle (the settings of this case alternate between 1.0 = link kept and fractions = link masked)."""
    em, inp = cases.setup_product(case, precision=precision)
    y, ey = em.predict(*inp)
    o, inp_o = cases.make_oracle(case, np.float64)
    y64, ey64 = o.predict(*inp_o)
    o.use_adj = False
    y64s, _ = o.predict(*inp_o)
    sc = lambda t: np.maximum(np.abs(t).reshape(-1, t.shape[-1]).max(0), 1.0)
    tol = 1e-5 if precision == "fp32" else 5e-3
    d = np.abs(y - y64) / sc(y64)
    if em.if_flood:
        d[np.abs(y64[..., 3] - 0.5) < (1e-4 if precision == "fp32" else 5e-3)] = 0
    assert d.max() < tol and (np.abs(ey - ey64) / sc(ey64)).max() < tol
    if precision == "fp32":
        effect = np.abs(y64 - y64s).max()
        assert effect > 20 * np.abs(y - y64).max(), "the dynamic mask must matter more than the fp32 error (%.3g vs %.3g)" % (effect, np.abs(y - y64).max())
        assert np.abs((y - y64s)).max() > 0.5 * effect


def test_use_adj_gcn_rejected():
    import copy
    args, B, seed = cases.build_args("ast_gcn_lstm_act_nofusion")
    args = copy.copy(args)
    args.use_adj = True
    from gnn_uds_b200.emulator import Emulator
    with pytest.raises(NotImplementedError):
        Emulator(args.conv, args.resnet, args.recurrent, args)
