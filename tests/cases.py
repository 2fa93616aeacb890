"""Shared test cases: build the same model for the product (gnn_uds_b200) and the oracle
from the committed network fixtures (tests/golden/*_net.npz), with seeded synthetic
weights, norms and inputs.  Nothing here reads /root/reference."""
import os
import sys
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from gnn_uds_b200 import graph as G, synthetic, weights as Wt   # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

# name -> hyper-parameters.  C1/C2/C3 follow SURVEY 8 (config sizes), scaled-down batch.
CASES = {
    # C1: astlingen GCN+GRU 10->10 (utils/config.yaml defaults)
    "c1_ast_gcn_gru": dict(net="astlingen", conv="GCNconv", recurrent="GRU", resnet=False, seq_in=10, seq_out=10,
                           embed_size=128, hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1, B=4, seed=43),
    # C2 quick variant: shunqing GAT+Conv1D+edge_fusion+resnet+flood, 10->10
    "c2_shq_gat_conv_10": dict(net="shunqing", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=10, seq_out=10,
                               embed_size=128, hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=1200.0, order=1,
                               edge_fusion=True, if_flood=3, B=3, seed=44),
    # C2 full horizon, tiny batch
    "c2_shq_gat_conv_60": dict(net="shunqing", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=60, seq_out=60,
                               embed_size=128, hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=1200.0, order=1,
                               edge_fusion=True, if_flood=3, B=1, seed=44),
    # C3b: astlingen act+edge_fusion+flood MPC evaluator, 5->5
    "c3_ast_act_5": dict(net="astlingen", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=5, seq_out=5,
                         embed_size=128, hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=1.0, order=1,
                         edge_fusion=True, if_flood=3, act="conti", B=6, seed=45),
    # variants for coverage
    "ast_gat_gru_noflood": dict(net="astlingen", conv="GAT", recurrent="GRU", resnet=False, seq_in=6, seq_out=3,
                                embed_size=64, hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1, B=3, seed=46),
    "ast_gcn_lstm_act_nofusion": dict(net="astlingen", conv="GCN", recurrent="LSTM", resnet=True, seq_in=4, seq_out=4,
                                      embed_size=128, hidden_dim=64, n_sp_layer=1, n_tp_layer=1, length=0, order=2,
                                      act="conti", if_flood=1, epsilon=0.1, B=2, seed=47),
    "ast_gcn_conv_eps0": dict(net="astlingen", conv="GCN", recurrent="Conv1D", resnet=True, seq_in=8, seq_out=4,
                              embed_size=128, hidden_dim=64, n_sp_layer=2, n_tp_layer=3, length=501, order=1,
                              edge_fusion=True, if_flood=2, epsilon=0.0, tide=True, B=2, seed=48),
    # reverse-pass coverage (tests/test_grad_gpu.py): GCN + Conv1D, act WITHOUT edge fusion (node-head flows times
    # a_in / a_out, emulator.py:708-715), no flood head, no resnet, seq_in == seq_out for the chunked roll
    "ast_gcn_conv_act_nofusion": dict(net="astlingen", conv="GCN", recurrent="Conv1D", resnet=False, seq_in=4, seq_out=4,
                                      embed_size=64, hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1,
                                      act="conti", if_flood=1, B=3, seed=60),
    # graph_base=1: one conv over the joint node+edge graph (emulator.py:220-223); needs act so that both
    # streams enter block 2 with H+E/2 channels.  C5 is this architecture on a 100k-node tree.
    "ast_joint_gcn_gru_act": dict(net="astlingen", conv="GCN", recurrent="GRU", resnet=True, seq_in=6, seq_out=6,
                                  embed_size=128, hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1,
                                  graph_base=1, act="conti", if_flood=2, B=3, seed=49),
    "ast_joint_gat_conv_act": dict(net="astlingen", conv="GAT", recurrent="Conv1D", resnet=False, seq_in=5, seq_out=5,
                                   embed_size=64, hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1,
                                   graph_base=1, act="conti", edge_fusion=True, B=2, seed=50),
    # pumped / offset networks (chaohu/hague features, emulator.py:590-598, 649-667) on the astlingen table: the four
    # controlled links become rated pumps fed by the tanks T2,T3,T4,T6 (storage area > 0 -> pumped-storage depth scan),
    # a few conduits get inlet offsets
    "ast_pump_act_fusion": dict(net="astlingen", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=6, seq_out=6,
                                embed_size=128, hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=1.0, order=1,
                                edge_fusion=True, if_flood=2, act="conti", epsilon=0.1, pumps=True, B=4, seed=51),
    "ast_pump_act_nofusion": dict(net="astlingen", conv="GCN", recurrent="GRU", resnet=False, seq_in=5, seq_out=5,
                                  embed_size=128, hidden_dim=64, n_sp_layer=1, n_tp_layer=1, length=0, order=1,
                                  act="conti", pumps=True, B=3, seed=52),
    # hague-style head-based network (base.py:289-292, dataloader.py:231-253, hague.py:115-136): h is a head, hmin =
    # invert elevation > 0, norm_x / norm_y / norm_b carry NON-ZERO minima, tide boundary, inlet offsets
    "ast_head_tide_offset": dict(net="astlingen", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=6, seq_out=6,
                                 embed_size=128, hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=1.0, order=1,
                                 edge_fusion=True, if_flood=2, act="conti", epsilon=0.1, tide=True, head=True,
                                 offsets=True, B=4, seed=56),
    "ast_head_gcn_gru_nofusion": dict(net="astlingen", conv="GCN", recurrent="GRU", resnet=False, seq_in=5, seq_out=5,
                                      embed_size=128, hidden_dim=64, n_sp_layer=1, n_tp_layer=1, length=0, order=1,
                                      head=True, B=3, seed=57),
    # use_adj (emulator.py:268-271, 343-362): per-sample attention mask from the settings; unit adjacency (length 0) and
    # Gaussian-weighted adjacency (length 1.0, where `adj.astype(int)` drops every off-diagonal entry)
    "ast_gat_use_adj": dict(net="astlingen", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=5, seq_out=5, embed_size=128,
                            hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1, edge_fusion=True, if_flood=2,
                            act="conti", use_adj=True, B=5, seed=58),
    "ast_gat_use_adj_weighted": dict(net="astlingen", conv="GAT", recurrent="GRU", resnet=False, seq_in=4, seq_out=4, embed_size=64,
                                     hidden_dim=64, n_sp_layer=2, n_tp_layer=1, length=501, order=1, act="conti", use_adj=True,
                                     B=3, seed=59),
    # graph_base=2: the same joint conv over the link-based graph of base.py:500-532 (asymmetric rows, order 2)
    "ast_edgebased_gcn_lstm_act": dict(net="astlingen", conv="GCN", recurrent="LSTM", resnet=True, seq_in=4, seq_out=4,
                                       embed_size=64, hidden_dim=64, n_sp_layer=2, n_tp_layer=1, length=0, order=2,
                                       graph_base=2, act="conti", edge_fusion=True, if_flood=1, B=2, seed=55),
    # NN-cat (conv falsy, emulator.py:180,231-233: flattened Dense stacks; 25 of the 86 saved model configs)
    "ast_nn_conv_act_fusion": dict(net="astlingen", conv="False", recurrent="Conv1D", resnet=True, seq_in=6, seq_out=6,
                                   embed_size=128, hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=0, order=1,
                                   edge_fusion=True, if_flood=2, act="conti", B=5, seed=53),
    "shq_nn_gru": dict(net="shunqing", conv=None, recurrent="GRU", resnet=False, seq_in=8, seq_out=4,
                       embed_size=64, hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1, B=3, seed=54),
}
GOLDEN_CASES = ["c1_ast_gcn_gru", "c2_shq_gat_conv_10", "c3_ast_act_5", "ast_gat_gru_noflood",
                "ast_gcn_lstm_act_nofusion", "ast_gcn_conv_eps0", "ast_joint_gcn_gru_act", "ast_joint_gat_conv_act",
                "ast_pump_act_fusion", "ast_pump_act_nofusion", "ast_nn_conv_act_fusion", "shq_nn_gru",
                "ast_edgebased_gcn_lstm_act", "ast_head_tide_offset", "ast_head_gcn_gru_nofusion",
                "ast_gat_use_adj", "ast_gat_use_adj_weighted"]


def load_net(name):
    z = np.load(os.path.join(GOLD, name + "_net.npz"))
    net = G.Network([str(s) for s in z["nodes"]], [str(s) for s in z["links"]], z["edges"], z["lengths"],
                    z["is_outfall"], z["is_storage"], z["hmax"], z["elevation"], z["ehmax"], z["area"])
    act_links = [str(s) for s in z["act_links"]] if "act_links" in z.files else None
    return net, act_links


def build_args(case):
    """Namespace in the reference's format (what env.get_args + parser.yaml give main.py:153-158)."""
    hp = dict(CASES[case])
    net, act_links = load_net(hp.pop("net"))
    length, order = hp.pop("length"), hp.pop("order")
    B, seed = hp.pop("B"), hp.pop("seed")
    head = hp.pop("head", False)
    ga = G.get_args(net, False, length, order, hp.get("graph_base", 0), act_links=act_links if hp.get("act") else None,
                    head=head)
    d = dict(vars(ga))
    if hp.pop("offsets", False):
        rng = np.random.default_rng(2000 + seed)
        off = np.zeros(len(net.links))
        off[::5] = rng.uniform(0.2, 1.0, size=len(off[::5]))
        d["offset"] = off
    if hp.pop("pumps", False):
        # synthetic pump table in the format of chaohu.get_args (chaohu.py:203-226): rated flows on the controlled
        # links, their in/out-flow seen from the end nodes, inlet offsets on every 7th conduit
        rng = np.random.default_rng(1000 + seed)
        M, N = len(net.links), len(net.nodes)
        pump, offset, pump_in, pump_out = np.zeros(M), np.zeros(M), np.zeros(N), np.zeros(N)
        for (u, v) in np.asarray(ga.act_edges):
            m = int(np.where((np.asarray(ga.edges) == (u, v)).all(1))[0][0])
            pump[m] = rng.uniform(0.05, 0.5)
            pump_out[u] += pump[m]
            pump_in[v] += pump[m]
        offset[::7] = rng.uniform(0.2, 1.0, size=len(offset[::7])) * (pump[::7] == 0)
        d.update(pump=pump, offset=offset, pump_in=pump_in, pump_out=pump_out)
    d.update(hp)
    d["model_dir"] = "/nonexistent"
    args = SimpleNamespace(**d)
    return args, B, seed


def make_product(case, device=0, precision="fp32", **kw):
    from gnn_uds_b200.emulator import Emulator
    args, B, seed = build_args(case)
    em = Emulator(args.conv, args.resnet, args.recurrent, args, device=device, precision=precision, **kw)
    return em, args, B, seed


def model_data(case, specs_from):
    """Seeded weights + norms + inputs for a case. specs_from: object exposing the hyper-parameters
    weights.layer_specs needs (product Emulator or the shim below)."""
    args, B, seed = build_args(case)
    specs = Wt.layer_specs(specs_from)
    ws = Wt.init_weights(specs, seed, node_edge_bias_std=0.05, bias_std=0.05)
    rng = np.random.default_rng(seed)
    hmin = np.asarray(getattr(args, "hmin", 0.0), np.float64)
    head_kw = {}
    if np.max(hmin) > 0:
        head_kw = dict(head_min=float(hmin.min()) - 0.5, head_max=float(np.max(args.hmax)), is_outfall=args.is_outfall)
    norms = synthetic.make_norms(specs_from.n_node, specs_from.n_edge, specs_from.if_flood, rng,
                                 tide=getattr(specs_from, "tide", False), **head_kw)
    return ws, norms, rng, B


class SpecShim:
    """Hyper-parameter view of an OracleEmulator for weights.layer_specs / synthetic.make_inputs."""

    def __init__(self, o):
        self.embed_size, self.hidden_dim, self.n_sp_layer, self.n_tp_layer = o.embed_size, o.hidden_dim, o.n_sp_layer, o.n_tp_layer
        self.kernel_size, self.n_node, self.n_edge = o.kernel_size, o.n_node, o.n_edge
        self.conv_kind, self.recurrent, self.graph_base = o.kind, o.recurrent, o.graph_base
        self.n_in, self.b_in, self.e_in, self.n_out, self.e_out = o.n_in, o.b_in, o.e_in, o.n_out, o.e_out
        self.act, self.if_flood, self.tide = o.act, o.if_flood, o.tide
        self.seq_in, self.seq_out = o.seq_in, o.seq_out
        self.hmax, self.ehmax, self.is_outfall = o.hmax, o.ehmax, o.is_outfall
        self.hmin = o.hmin
        self.n_act = len(o.act_edges) if o.act else 0


def _tweak_inputs(case, inputs):
    """use_adj cases: half of the settings are exactly 1 (fully open: the link stays in the attention mask), the others
    stay fractional (trunc -> 0: the link is masked out), so both branches of `adj.astype(int)` are exercised."""
    if not CASES[case].get("use_adj"):
        return inputs
    states, b, a, ex = inputs
    a = np.array(a, copy=True)
    idx = np.indices(a.shape).sum(0) % 2 == 0
    a[idx] = 1.0
    return states, b, a, ex


def make_oracle(case, dtype=np.float64):
    from oracle.emulator_oracle import OracleEmulator
    args, B, seed = build_args(case)
    o = OracleEmulator(args.conv, args.resnet, args.recurrent, args, dtype=dtype)
    shim = SpecShim(o)
    ws, norms, rng, B = model_data(case, shim)
    o.set_weights(ws)
    o.set_norm(norms["norm_x"], norms["norm_b"], norms["norm_y"], norms["norm_r"], norms["norm_e"])
    shim.norm_x, shim.norm_b, shim.norm_e = norms["norm_x"], norms["norm_b"], norms["norm_e"]
    inputs = _tweak_inputs(case, synthetic.make_inputs(shim, B, rng))
    return o, inputs


def oracle_run(case):
    """Outputs stored in tests/golden/fwd_<case>.npz."""
    o64, inp = make_oracle(case, np.float64)
    y64, ey64 = o64.predict(*_order(inp))
    o32, _ = make_oracle(case, np.float32)
    y32, ey32 = o32.predict(*_order(inp))
    return dict(y64=y64.astype(np.float64), ey64=ey64.astype(np.float64), y32=y32, ey32=ey32)


def _order(inp):
    states, b, a, edge_state = inp
    return states, b, a, edge_state


def setup_product(case, device=0, precision="fp32", **kw):
    """Product Emulator with the case's seeded weights/norms + the case inputs."""
    em, args, B, seed = make_product(case, device, precision, **kw)
    ws, norms, rng, B = model_data(case, em)
    em.set_weights(ws)
    em.set_norm(norms["norm_x"], norms["norm_b"], norms["norm_y"], norms["norm_r"], norms["norm_e"])
    inputs = _tweak_inputs(case, synthetic.make_inputs(em, B, rng))
    return em, inputs
