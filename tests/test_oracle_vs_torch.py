"""Pins the oracle's restated Keras/spektral layer rules (the dagger rules) against an INDEPENDENT
library: torch.nn (CPU, float64).  The reference's arithmetic lives in keras 2.10 / spektral 1.3.1, which cannot be
installed here; torch implements the same published cells, so agreement to 1e-12 pins
  * GRU  reset_after=True, gate order z,r,h  (torch: r,z,n with b_hn inside the reset product)
  * LSTM gate order i,f,c,o                  (torch: i,f,g,o)
  * Conv1D causal + dilation (left zero pad d(k-1), tap j reads x[t-(k-1-j)d])  (torch conv1d is a correlation too)
  * Dense, leaky_relu(0.2) masked softmax attention of GATConv, gcn_filter, sigmoid / tanh
(emulator.py:154-163 builds the temporal layers; 18-25, 129-152 the conv layers.)"""
import numpy as np
import pytest
import torch

from oracle import emulator_oracle as O

torch.set_default_dtype(torch.float64)
RNG = np.random.default_rng(7)


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.mark.parametrize("C,H,T", [(128, 64, 10), (24, 16, 7)])
def test_gru_matches_torch(C, H, T):
    x = RNG.standard_normal((5, T, C))
    K, R, b = RNG.standard_normal((C, 3 * H)) * 0.2, RNG.standard_normal((H, 3 * H)) * 0.2, RNG.standard_normal((2, 3 * H)) * 0.1
    perm = np.r_[H:2 * H, 0:H, 2 * H:3 * H]                      # keras z,r,h -> torch r,z,n
    m = torch.nn.GRU(C, H, batch_first=True)
    with torch.no_grad():
        m.weight_ih_l0.copy_(_t(K[:, perm].T)); m.weight_hh_l0.copy_(_t(R[:, perm].T))
        m.bias_ih_l0.copy_(_t(b[0][perm])); m.bias_hh_l0.copy_(_t(b[1][perm]))
        ref = m(_t(x))[0].numpy()
    assert np.abs(O.gru(x, K, R, b) - ref).max() < 1e-12


@pytest.mark.parametrize("C,H,T", [(128, 64, 10), (24, 16, 7)])
def test_lstm_matches_torch(C, H, T):
    x = RNG.standard_normal((5, T, C))
    K, R, b = RNG.standard_normal((C, 4 * H)) * 0.2, RNG.standard_normal((H, 4 * H)) * 0.2, RNG.standard_normal(4 * H) * 0.1
    m = torch.nn.LSTM(C, H, batch_first=True)                    # same i,f,g(c),o order
    with torch.no_grad():
        m.weight_ih_l0.copy_(_t(K.T)); m.weight_hh_l0.copy_(_t(R.T))
        m.bias_ih_l0.copy_(_t(b)); m.bias_hh_l0.zero_()
        ref = m(_t(x))[0].numpy()
    assert np.abs(O.lstm(x, K, R, b) - ref).max() < 1e-12


@pytest.mark.parametrize("k,d,T", [(3, 1, 10), (3, 2, 10), (3, 4, 10), (2, 8, 6), (4, 1, 5)])
def test_conv1d_causal_matches_torch(k, d, T):
    C, H = 12, 8
    x = RNG.standard_normal((6, T, C))
    W, b = RNG.standard_normal((k, C, H)), RNG.standard_normal(H)
    xt = torch.nn.functional.pad(_t(x).transpose(1, 2), (d * (k - 1), 0))          # (R,C,T+pad), left zero pad
    ref = torch.nn.functional.conv1d(xt, _t(W).permute(2, 1, 0).contiguous(), _t(b), dilation=d)
    ref = torch.relu(ref).transpose(1, 2).numpy()
    assert np.abs(O.conv1d_causal(x, W, b, d, "relu") - ref).max() < 1e-12


def test_dense_and_activations_match_torch():
    x, W, b = RNG.standard_normal((4, 9, 20)), RNG.standard_normal((20, 6)), RNG.standard_normal(6)
    lin = torch.nn.functional.linear(_t(x), _t(W.T), _t(b))
    assert np.abs(O.dense(x, W, b, "linear") - lin.numpy()).max() < 1e-12
    assert np.abs(O.dense(x, W, b, "tanh") - torch.tanh(lin).numpy()).max() < 1e-12
    assert np.abs(O.dense(x, W, b, "sigmoid") - torch.sigmoid(lin).numpy()).max() < 1e-12
    assert np.abs(O.dense(x, W, b, "relu") - torch.relu(lin).numpy()).max() < 1e-12
    # keras 2.x hard_sigmoid = clip(0.2 v + 0.5, 0, 1)  (NOT torch's hardsigmoid, which is v/6 + 0.5)
    hs = torch.clamp(0.2 * lin + 0.5, 0.0, 1.0).numpy()
    assert np.abs(O.dense(x, W, b, "hard_sigmoid") - hs).max() < 1e-12


def test_gat_conv_matches_edge_list_formulation():
    """Dense masked softmax (how spektral writes it) against a per-edge scatter formulation in torch."""
    N, C, E = 11, 10, 8
    A = (RNG.random((N, N)) < 0.3).astype(np.float64)
    A = np.maximum(A, A.T)
    np.fill_diagonal(A, 0)                                       # GATConv forces the self loops itself
    x = RNG.standard_normal((3, N, C))
    W, a_s, a_n, b = RNG.standard_normal((C, 1, E)), RNG.standard_normal((E, 1, 1)), RNG.standard_normal((E, 1, 1)), RNG.standard_normal(E)
    out = O.gat_conv(x, A, W, a_s, a_n, b, "relu")
    h = _t(x) @ _t(W[:, 0, :])
    ss, sn = (h @ _t(a_s[:, 0, :]))[..., 0], (h @ _t(a_n[:, 0, :]))[..., 0]
    Ad = A.copy(); np.fill_diagonal(Ad, 1)
    ref = torch.zeros(3, N, E)
    for i in range(N):
        nb = np.nonzero(Ad[i])[0]
        sc = torch.nn.functional.leaky_relu(ss[:, i:i + 1] + sn[:, nb], 0.2)
        al = torch.softmax(sc, dim=-1)
        ref[:, i] = (al[..., None] * h[:, nb]).sum(1)
    ref = torch.relu(ref + _t(b)).numpy()
    assert np.abs(out - ref).max() < 1e-12


def test_gcn_filter_matches_torch():
    N = 9
    A = (RNG.random((N, N)) < 0.4).astype(np.float64)
    A = np.maximum(A, A.T)
    np.fill_diagonal(A, 1)                                       # the reference's adjacency carries 1 on the diagonal
    At = _t(A) + torch.eye(N)
    dinv = At.sum(1).pow(-0.5)
    ref = (dinv[:, None] * At * dinv[None, :]).numpy()
    assert np.abs(O.gcn_filter(A) - ref).max() < 1e-14
    x, W, b = RNG.standard_normal((2, N, 5)), RNG.standard_normal((5, 4)), RNG.standard_normal(4)
    ref2 = torch.relu(_t(ref) @ (_t(x) @ _t(W)) + _t(b)).numpy()
    assert np.abs(O.gcn_conv(x, O.gcn_filter(A), W, b, "relu") - ref2).max() < 1e-12
