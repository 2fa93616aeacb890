"""CPU guard of the oracle itself: the NumPy restatement (oracle/emulator_oracle.py) must keep reproducing the committed
golden vectors (tests/golden/fwd_<case>.npz, made by tools/make_golden.py --forward), so a change of the oracle cannot
silently move the target the CUDA path is compared with.  Also checks two structural properties of the NN-cat variant
(emulator.py:180,231-233) that do not need the reference to run.
"""
import os

import numpy as np
import pytest

import cases


@pytest.mark.parametrize("case", cases.GOLDEN_CASES)
def test_oracle_reproduces_committed_golden(case):
    z = np.load(os.path.join(cases.GOLD, "fwd_%s.npz" % case))
    o, inp = cases.make_oracle(case, np.float64)
    y, ey = o.predict(*inp)
    assert y.shape == z["y64"].shape and ey.shape == z["ey64"].shape
    # float64 on another BLAS: far below the 1e-5 parity tolerance of the CUDA fp32 path
    sy = np.maximum(np.abs(z["y64"]).reshape(-1, y.shape[-1]).max(0), 1.0)
    se = np.maximum(np.abs(z["ey64"]).reshape(-1, ey.shape[-1]).max(0), 1.0)
    flip = np.zeros(y.shape, bool)
    if o.if_flood:                                   # a flood probability within 1e-9 of 0.5 may flip h / q_w
        near = np.abs(z["y64"][..., 3] - 0.5) < 1e-9
        flip[..., 0] = near
        flip[..., -1] = near
    assert (np.abs(y - z["y64"]) / sy)[~flip].max() < 1e-9
    assert (np.abs(ey - z["ey64"]) / se).max() < 1e-9


def test_nn_cat_weight_order_and_shapes():
    """Creation order of the no-conv model (emulator.py:195-336 with conv falsy): flattened embeddings, Dense(2E)
    spatial layers over concat([x, e]), heads of width n_out*N / N / e_out*M."""
    o, _ = cases.make_oracle("ast_nn_conv_act_fusion", np.float32)
    N, M, E, H = o.n_node, o.n_edge, o.embed_size, o.hidden_dim
    shapes = [tuple(w.shape) for layer in o.weights for w in layer]
    assert shapes[0] == (N * o.n_in, E) and shapes[2] == (N * o.b_in, E // 2)
    assert shapes[4] == (M * o.e_in, E) and shapes[6] == (M, E // 2)
    assert shapes[8] == (2 * E, 2 * E)
    assert (2 * H + E, 2 * E) in shapes              # block 2, first layer: [x(H) | b(E/2) | e(H) | ae(E/2)]
    assert shapes[-2] == (E, o.e_out * M) and (E // 2, N) in shapes and (E, o.n_out * N) in shapes


def test_nn_cat_rows_are_independent_of_batch():
    """Every (b) scenario is an independent row set: predicting a sub-batch gives the same rows."""
    o, inp = cases.make_oracle("shq_nn_gru", np.float64)
    y, ey = o.predict(*inp)
    sub = [None if t is None else t[1:2] for t in inp]
    y1, ey1 = o.predict(*sub)
    assert np.allclose(y[1:2], y1, rtol=0, atol=1e-12) and np.allclose(ey[1:2], ey1, rtol=0, atol=1e-12)
