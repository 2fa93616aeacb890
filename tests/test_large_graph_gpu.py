"""GPU parity of the large-graph path: graph_base=1 joint node+edge graph on a synthetic
drainage tree whose rows do not fit one CTA's shared memory, so the CSR aggregation runs through
agg_large_kernel (one warp per row, 128-bit gathers from HBM/L2)."""
from types import SimpleNamespace

import numpy as np
import pytest

import cases
from gnn_uds_b200 import graph as G, synthetic, weights as Wt

pytestmark = pytest.mark.gpu


def _tree_model(conv, n_node, precision, seed=3):
    from gnn_uds_b200.emulator import Emulator
    from oracle.emulator_oracle import OracleEmulator
    net = synthetic.random_tree(n_node, np.random.default_rng(seed))
    act_links = [net.links[i] for i in (3, n_node // 2, n_node - 5)]
    ga = G.get_args(net, False, 0, 1, 1, act_links=act_links)
    d = dict(vars(ga))
    d.update(conv=conv, recurrent="GRU", resnet=True, seq_in=4, seq_out=4, embed_size=64, hidden_dim=64, n_sp_layer=2,
             n_tp_layer=1, graph_base=1, act="conti", if_flood=1, model_dir="/nonexistent")
    args = SimpleNamespace(**d)
    em = Emulator(args.conv, args.resnet, args.recurrent, args, precision=precision)
    o = OracleEmulator(args.conv, args.resnet, args.recurrent, args, dtype=np.float64)
    ws = Wt.init_weights(Wt.layer_specs(em), seed, bias_std=0.05)
    rng = np.random.default_rng(seed)
    n = synthetic.make_norms(em.n_node, em.n_edge, em.if_flood, rng)
    for m in (em, o):
        m.set_weights(ws)
        m.set_norm(n["norm_x"], n["norm_b"], n["norm_y"], n["norm_r"], n["norm_e"])
    inp = synthetic.make_inputs(em, 2, rng)
    return em, o, inp


def _scale(t):
    return np.maximum(np.abs(t).reshape(-1, t.shape[-1]).max(0), 1.0)


@pytest.mark.parametrize("conv", ["GCN", "GAT"])
def test_joint_tree_fp32(conv):
    em, o, inp = _tree_model(conv, 500, "fp32")
    y, ey = em.predict(*inp)
    y64, ey64 = o.predict(*[None if t is None else t.astype(np.float64) for t in inp])
    near = np.abs(y64[..., 3] - 0.5) < 1e-4
    dy = np.abs(y - y64) / _scale(y64)
    dy[near] = 0
    de = np.abs(ey - ey64) / _scale(ey64)
    print(conv, "tree fp32: y", dy.max(), "ey", de.max())
    assert dy.max() <= 2e-5 and de.max() <= 2e-5


def test_joint_tree_bf16():
    em, o, inp = _tree_model("GCN", 500, "bf16")
    y, ey = em.predict(*inp)
    y64, ey64 = o.predict(*[None if t is None else t.astype(np.float64) for t in inp])
    de = np.abs(ey - ey64) / _scale(ey64)
    rm = np.sqrt(((ey - ey64) ** 2).reshape(-1, 3).mean(0)) / _scale(ey64)
    print("tree bf16: ey max", de.max(), "rmse", rm)
    assert de.max() <= 2e-2 and (rm <= 3e-3).all()   # max over 6k elements of a 9-layer bf16 net; RMSE is the bar


def test_joint_tree_100k_nodes_vs_sparse_oracle():
    """C5-sized graph (100 000 nodes + 99 999 links = 199 999 joint rows): the dense (N+M)^2 oracle cannot run there, so
    the float64 truth comes from oracle/sparse_oracle.py (same arithmetic, graph matrices as scipy.sparse; checked
    against the dense oracle to 1e-15 on a 300-node tree).  fp32 path <= 2e-5, tensor-core path RMSE <= 3e-3."""
    from gnn_uds_b200.emulator import Emulator
    from oracle.sparse_oracle import SparseOracleEmulator
    n_node, seed = 100000, 7
    net = synthetic.random_tree(n_node, np.random.default_rng(seed))
    act_links = [net.links[i] for i in (3, n_node // 2, n_node - 5)]
    ga = G.get_args(net, False, 0, 1, 1, act_links=act_links, dense=False)
    d = dict(vars(ga))
    d.update(conv="GCN", recurrent="GRU", resnet=True, seq_in=3, seq_out=3, embed_size=64, hidden_dim=64, n_sp_layer=2,
             n_tp_layer=1, graph_base=1, act="conti", if_flood=1, model_dir="/nonexistent")
    args = SimpleNamespace(**d)
    o = SparseOracleEmulator(args.conv, args.resnet, args.recurrent, args, dtype=np.float64)
    rng = np.random.default_rng(seed)
    n = synthetic.make_norms(n_node, n_node - 1, 1, rng)
    res = {}
    for prec in ("fp32", "bf16"):
        em = Emulator(args.conv, args.resnet, args.recurrent, args, precision=prec)
        ws = Wt.init_weights(Wt.layer_specs(em), seed, bias_std=0.05)
        em.set_weights(ws)
        em.set_norm(n["norm_x"], n["norm_b"], n["norm_y"], n["norm_r"], n["norm_e"])
        if prec == "fp32":
            o.set_weights(ws)
            o.set_norm(n["norm_x"], n["norm_b"], n["norm_y"], n["norm_r"], n["norm_e"])
            inp = synthetic.make_inputs(em, 1, np.random.default_rng(seed + 1))
            y64, ey64 = o.predict(*[None if t is None else t.astype(np.float64) for t in inp])
        res[prec] = em.predict(*inp)
        del em
    y, ey = res["fp32"]
    near = np.abs(y64[..., 3] - 0.5) < 1e-4
    dy = np.abs(y - y64) / _scale(y64)
    dy[near] = 0
    de = np.abs(ey - ey64) / _scale(ey64)
    print("100k tree fp32: y", dy.max(), "ey", de.max())
    assert dy.max() <= 2e-5 and de.max() <= 2e-5
    y16, ey16 = res["bf16"]
    rm = np.sqrt(((ey16 - ey64) ** 2).reshape(-1, 3).mean(0)) / _scale(ey64)
    print("100k tree tensor-core: ey max", (np.abs(ey16 - ey64) / _scale(ey64)).max(), "rmse", rm)
    assert (rm <= 3e-3).all()
