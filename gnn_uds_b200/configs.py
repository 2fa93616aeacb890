"""Named workloads of BASELINE.json (SURVEY.md 8, config sizes C1..C5)."""
import os
from types import SimpleNamespace

import numpy as np

from . import graph as G, synthetic

_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")   # network tables (tools/make_golden.py)

WORKLOADS = {
    # configs[0]: astlingen GCN+GRU 10->10, B=64 (the reference's CPU-runnable case)
    "C1": dict(net="astlingen", conv="GCNconv", recurrent="GRU", resnet=False, seq_in=10, seq_out=10, embed_size=128,
               hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1, B=64, seed=42),
    # configs[1]: shunqing GAT+Conv1D edge_fusion resnet flood, 60->60, B=4096 (model/shunqing/r1_edgef_flood_gat_0)
    "C2": dict(net="shunqing", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=60, seq_out=60, embed_size=128,
               hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=1200.0, order=1, edge_fusion=True, if_flood=3,
               B=4096, seed=43),
    "C2q": dict(net="shunqing", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=10, seq_out=10, embed_size=128,
                hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=1200.0, order=1, edge_fusion=True, if_flood=3,
                B=4096, seed=43),
    # NN-cat twin of C2 (conv falsy: flattened Dense stacks, emulator.py:231-233; surrogate/test.bat evaluates GAT and
    # NN-cat side by side)
    "C2nn": dict(net="shunqing", conv="False", recurrent="Conv1D", resnet=True, seq_in=60, seq_out=60, embed_size=128,
                 hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=1200.0, order=1, edge_fusion=True, if_flood=3,
                 B=4096, seed=43),
    # LSTM twin of C2 (6 of the saved model configs use recurrent: LSTM)
    "C2lstm": dict(net="shunqing", conv="GAT", recurrent="LSTM", resnet=True, seq_in=60, seq_out=60, embed_size=128,
                   hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=1200.0, order=1, edge_fusion=True, if_flood=3,
                   B=1024, seed=43),
    # configs[2]: astlingen act+edge_fusion+flood MPC evaluator; (a) 60->60 5 layers, (b) 12 AR chunks of 5->5
    "C3a": dict(net="astlingen", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=60, seq_out=60, embed_size=128,
                hidden_dim=64, n_sp_layer=5, n_tp_layer=5, length=1.0, order=1, edge_fusion=True, if_flood=5,
                act="conti", B=8192, seed=44),
    "C3b": dict(net="astlingen", conv="GAT", recurrent="Conv1D", resnet=True, seq_in=5, seq_out=5, embed_size=128,
                hidden_dim=64, n_sp_layer=3, n_tp_layer=3, length=1.0, order=1, edge_fusion=True, if_flood=3,
                act="conti", B=8192, seed=44, chunks=12),
    # configs[4]: synthetic drainage tree, graph_base=1 joint graph (NodeEdge is impossible at this size),
    # GCN+GRU, 10->10; act (4 controlled links) so both streams enter block 2 with H+E/2 channels as the
    # reference's concat on axis -2 needs (emulator.py:273).  HBM-bound CSR gather path.
    "C5": dict(net="tree:100000", conv="GCN", recurrent="GRU", resnet=True, seq_in=10, seq_out=10, embed_size=128,
               hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1, graph_base=1, act="conti", B=256, seed=46),
    "C5s": dict(net="tree:2000", conv="GCN", recurrent="GRU", resnet=True, seq_in=10, seq_out=10, embed_size=128,
                hidden_dim=64, n_sp_layer=2, n_tp_layer=2, length=0, order=1, graph_base=1, act="conti", B=8, seed=46),
}


def load_net(name):
    if name.startswith("tree:"):
        n = int(name.split(":")[1])
        net = synthetic.random_tree(n, np.random.default_rng(1000 + n))
        step = max(1, (n - 1) // 4)
        return net, [net.links[i] for i in range(step // 2, n - 1, step)][:4]
    z = np.load(os.path.join(_GOLD, name + "_net.npz"))
    net = G.Network([str(s) for s in z["nodes"]], [str(s) for s in z["links"]], z["edges"], z["lengths"],
                    z["is_outfall"], z["is_storage"], z["hmax"], z["elevation"], z["ehmax"], z["area"])
    return net, ([str(s) for s in z["act_links"]] if "act_links" in z.files else None)


def make_args(name, inp_path=None):
    """Namespace as main.py:153-158 / mpc.py:801-813 would hand to Emulator (graph from the committed
    network table, or from a SWMM .inp when given)."""
    hp = dict(WORKLOADS[name])
    netname = hp.pop("net")
    net, act_links = load_net(netname) if inp_path is None else (G.Network.from_inp(inp_path), None)
    length, order = hp.pop("length"), hp.pop("order")
    ga = G.get_args(net, False, length, order, hp.get("graph_base", 0), act_links=act_links if hp.get("act") else None,
                    dense=net.n_node <= 4096)
    d = dict(vars(ga)); d.update(hp); d["model_dir"] = None; d["env_name"] = netname
    return SimpleNamespace(**d)


def make_emulator(name, device=0, precision="fp32", **kw):
    from .emulator import Emulator
    args = make_args(name)
    em = Emulator(args.conv, args.resnet, args.recurrent, args, device=device, precision=precision, **kw)
    em.init_weights(args.seed, node_edge_bias_std=0.05, bias_std=0.05)
    rng = np.random.default_rng(args.seed)
    n = synthetic.make_norms(em.n_node, em.n_edge, em.if_flood, rng, tide=em.tide)
    em.set_norm(n["norm_x"], n["norm_b"], n["norm_y"], n["norm_r"], n["norm_e"])
    return em, args, rng
