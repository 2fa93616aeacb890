"""Edge cases of the reference-facing calls (Emulator.predict / rollout, emulator.py:566-602, mpc.py:565-582) on the
CUDA path: empty and single-scenario batches, ragged chunking, longer input histories, argument errors, and the
NN-cat variant through chunking and the autoregressive loop.  Bit-exact where the property is structural
(slicing, chunking on the fp32 path), oracle tolerance elsewhere.
"""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _scaled_err(a, ref):
    sc = np.maximum(np.abs(ref).reshape(-1, ref.shape[-1]).max(0), 1.0)
    return float((np.abs(a - ref) / sc).max())


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_empty_batch(precision):
    em, inp = cases.setup_product("c3_ast_act_5", precision=precision)
    empty = [None if t is None else t[:0] for t in inp]
    y, ey = em.predict(*empty)
    assert y.shape == (0, em.seq_out, em.n_node, em.y_channels) and ey.shape == (0, em.seq_out, em.n_edge, em.ey_channels)
    # the handle is still usable afterwards
    y1, _ = em.predict(*inp)
    assert np.isfinite(y1).all()


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("case", ["c2_shq_gat_conv_10", "c3_ast_act_5", "shq_nn_gru"])
def test_single_scenario_equals_its_row_in_the_batch(case, precision):
    em, inp = cases.setup_product(case, precision=precision)
    y, ey = em.predict(*inp)
    one = [None if t is None else t[1:2] for t in inp]
    y1, ey1 = em.predict(*one)
    if precision == "fp32":
        assert np.array_equal(y1, y[1:2]) and np.array_equal(ey1, ey[1:2])
    else:   # small graphs share an MMA tile between samples: slot-dependent fp32 accumulation order
        assert _scaled_err(y1[..., :3], y[1:2][..., :3]) < 1e-3 and _scaled_err(ey1, ey[1:2]) < 1e-3


def test_ragged_chunks_fp32_bitwise():
    em, inp = cases.setup_product("ast_nn_conv_act_fusion")            # B = 5
    y, ey = em.predict(*inp)
    for mc in (2, 3):
        em2, _ = cases.setup_product("ast_nn_conv_act_fusion", max_chunk=mc)
        y2, ey2 = em2.predict(*inp)
        assert np.array_equal(y, y2) and np.array_equal(ey, ey2), "max_chunk=%d" % mc


def test_longer_history_uses_the_last_seq_in_steps():
    em, inp = cases.setup_product("ast_gat_gru_noflood")               # 6 -> 3
    states, b, a, ex = inp
    pad_s = np.concatenate([np.full_like(states[:, :2], 7.0), states], axis=1)
    pad_e = np.concatenate([np.full_like(ex[:, :2], -3.0), ex], axis=1)
    y, ey = em.predict(states, b, a, ex)
    y2, ey2 = em.predict(pad_s, b, a, pad_e)                           # emulator.py:567,569: [:, -seq_in:]
    assert np.array_equal(y, y2) and np.array_equal(ey, ey2)


def test_argument_errors():
    em, inp = cases.setup_product("c3_ast_act_5")
    states, b, a, ex = inp
    with pytest.raises(ValueError):
        em.predict(states, b, None, ex)                                 # act model without settings
    with pytest.raises(ValueError):
        em.predict(states, b, a, None)                                  # edge_state is mandatory (emulator.py:579)
    with pytest.raises(AssertionError):
        em.predict(states, b[:, :-1], a, ex)                            # emulator.py:570
    with pytest.raises(ValueError):
        em.predict(states[:, :, :-1], b, a, ex)                         # wrong node count
    with pytest.raises(ValueError):
        em.predict(states, b, a[..., :-1], ex)                          # wrong number of settings
    with pytest.raises(ValueError):
        em.rollout(states, np.concatenate([b, b[:, :1]], 1), np.concatenate([a, a[:, :1]], 1), ex)   # horizon % seq_out


def test_nn_cat_rollout_matches_oracle():
    """mpc.py:565-582 chunked autoregression with the no-conv model: 3 chunks of 6 steps."""
    case = "ast_nn_conv_act_fusion"
    from gnn_uds_b200 import synthetic
    em, _ = cases.setup_product(case)
    o, _ = cases.make_oracle(case, np.float64)
    states, b, a, ex = synthetic.make_inputs(em, 6, np.random.default_rng(7), horizon=3 * em.seq_out)
    y, ey = em.rollout(states, b, a, ex, rule="mpc")
    y64, ey64 = o.rollout_mpc(states.astype(np.float64), b.astype(np.float64), a.astype(np.float64), ex.astype(np.float64))
    assert y.shape == y64.shape and ey.shape == ey64.shape
    # the fed-back thresholds (q_w > 0, prob > 0.5) may flip on rounding: compare the scenarios without a flip
    same = ((y[..., 3] > 0.5) == (y64[..., 3] > 0.5)).all(axis=(1, 2)) & ((y[..., -1] > 0) == (y64[..., -1] > 0)).all(axis=(1, 2))
    assert same.sum() >= 4
    assert _scaled_err(y[same], y64[same]) < 2e-4                      # fp32 through 3 fed-back chunks
    assert _scaled_err(ey[same], ey64[same]) < 2e-4
