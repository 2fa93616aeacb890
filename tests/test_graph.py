"""G1-G6: product CSR graph construction vs the networkx/dense oracle (bit-exact patterns)."""
import os

import numpy as np
import pytest

import cases
from gnn_uds_b200 import graph as G

GOLD = cases.GOLD
VARIANTS = {"astlingen": [(0, 1), (0, 2), (1.0, 1), (501, 1)], "shunqing": [(0, 1), (0, 2), (1200.0, 1), (1000.0, 1)]}


def _chk(csr, z, key):
    assert np.array_equal(csr.row_ptr, z[key + "_rp"]), key
    assert np.array_equal(csr.col_idx, z[key + "_ci"]), key
    assert np.array_equal(csr.val, z[key + "_val"]), key       # fp64, same formula -> identical bits


@pytest.mark.parametrize("name", ["astlingen", "shunqing"])
def test_csr_matches_networkx_golden(name):
    net, _ = cases.load_net(name)
    for length, order in VARIANTS[name]:
        z = np.load(os.path.join(GOLD, "graph_%s_l%g_o%d.npz" % (name, length, order)))
        _chk(G.node_adjacency(net, False, length, order), z, "adj")
        _chk(G.edge_adjacency(net, False, length, order), z, "edge_adj")
        if "joint_rp" in z.files:
            _chk(G.joint_adjacency(net, False, 0, 1), z, "joint")


def test_known_sizes():
    # README.md:37-38,44-45 of the reference: shunqing 113 nodes / 131 conduits, astlingen 30 / 29
    s, _ = cases.load_net("shunqing")
    a, al = cases.load_net("astlingen")
    assert (s.n_node, s.n_edge) == (113, 131) and (a.n_node, a.n_edge) == (30, 29)
    assert a.nodes[23] == "Out_to_WWTP" and a.is_outfall[23] == 1 and a.nodes[24:] == ["T%d" % i for i in range(1, 7)]
    assert al == ["V2", "V3", "V4", "V6"]
    assert a.edge_list(al).tolist() == [[25, 12], [26, 9], [27, 19], [29, 22]]
    assert G.node_adjacency(s).nnz == 375 and G.edge_adjacency(s).nnz == 615
    assert G.node_adjacency(s, length=1200.0).nnz == 763 and G.edge_adjacency(s, length=1200.0).max_row() == 16


def test_live_oracle_dense_equals_csr():
    from oracle import graph_oracle as go
    net, _ = cases.load_net("astlingen")
    for length, order, directed in [(0, 1, False), (0, 3, False), (300.0, 1, False), (0, 1, True), (0, 2, True), (400.0, 1, True)]:
        L = net.lengths if length else None
        assert np.array_equal(G.node_adjacency(net, directed, length, order).to_dense(), go.get_adj(net.edges, L, directed, length, order))
        assert np.array_equal(G.edge_adjacency(net, directed, length, order).to_dense(), go.get_edge_adj(net.edges, L, directed, length, order))
        assert np.array_equal(G.joint_adjacency(net, directed, length, order).to_dense(), go.get_node_based_adj(net.edges, L, directed, length, order))
        assert np.array_equal(G.edge_based_adjacency(net, directed, length, order).to_dense(), go.get_edge_based_adj(net.edges, L, directed, length, order))
    assert np.array_equal(G.incidence(net), go.get_node_edge(net.n_node, net.edges))


def test_edge_based_adjacency_known_answer():
    """graph_base=2 (base.py:500-532) on the path 0-1-2: links get the ids 3 and 4, are tied to each other through
    node 1 and to their end nodes; order=1 rows are {self} + neighbours."""
    from oracle import graph_oracle as go
    edges = np.array([[0, 1], [1, 2]])
    want = np.zeros((5, 5))
    for r, cols in {0: [0, 3], 1: [1, 3, 4], 2: [2, 4], 3: [3, 0, 1, 4], 4: [4, 1, 2, 3]}.items():
        want[r, cols] = 1
    assert np.array_equal(go.get_edge_based_adj(edges), want)
    net = G.Network(["a", "b", "c"], ["l0", "l1"], edges, np.array([10.0, 30.0]), np.zeros(3), np.zeros(3), np.ones(3),
                    np.zeros(3), np.ones(2), np.zeros(3))
    assert np.array_equal(G.edge_based_adjacency(net).to_dense(), want)
    # order=0: identity; length mode: exp(-(l/std)^2) of half-length ties (l0-a: 5, l0-l1: 20, l1-b: 15)
    assert np.array_equal(G.edge_based_adjacency(net, order=0).to_dense(), np.eye(5))
    d = G.edge_based_adjacency(net, length=16.0).to_dense()
    std = np.std([10.0, 30.0])
    assert np.isclose(d[3, 0], np.exp(-(5.0 / (std + 1e-5)) ** 2)) and np.isclose(d[4, 1], np.exp(-(15.0 / (std + 1e-5)) ** 2))
    assert d[3, 4] == 0 and d[0, 1] == np.exp(-(10.0 / (std + 1e-5)) ** 2)      # l0-l1 is 20 > cutoff; a-l0-b is 5+5
    assert np.array_equal(d, go.get_edge_based_adj(edges, np.array([10.0, 30.0]), False, 16.0, 1))


def test_gcn_filter_and_gat_pattern():
    from oracle.emulator_oracle import gcn_filter
    net, _ = cases.load_net("shunqing")
    for length in (0, 1200.0):
        csr = G.node_adjacency(net, False, length, 1)
        dense = gcn_filter(csr.to_dense())
        mine = G.gcn_normalise(csr).to_dense()
        assert np.array_equal((dense != 0), (mine != 0))
        assert np.array_equal(dense.astype(np.float32), mine.astype(np.float32)) or np.allclose(dense, mine, rtol=1e-15, atol=0)
        pat = G.gat_pattern(csr).to_dense()
        ref = (csr.to_dense() > 0).astype(int)
        ref[np.diag_indices_from(ref)] = 1
        assert np.array_equal(pat.astype(int), ref)


@pytest.mark.skipif(not os.path.isdir("/root/reference/surrogate/envs/network"), reason="reference tree absent")
def test_parser_vs_oracle_parser_all_networks():
    from oracle import graph_oracle as go
    base = "/root/reference/surrogate/envs/network"
    for name in ("astlingen", "shunqing", "chaohu", "hague", "RedChicoSur"):
        p = os.path.join(base, name, name + ".inp")
        mine, ref = G.Network.from_inp(p), go.parse_network(p)
        assert mine.nodes == ref["nodes"]
        assert np.array_equal(mine.edges, ref["edges"]) and np.array_equal(mine.lengths, ref["lengths"])
        assert np.array_equal(mine.hmax, ref["hmax"]) and np.array_equal(mine.is_outfall, ref["is_outfall"])
        assert np.array_equal(mine.ehmax[:len(ref["ehmax"])], ref["ehmax"][:len(mine.ehmax)])


def test_random_tree_is_a_tree():
    from gnn_uds_b200 import synthetic
    net = synthetic.random_tree(2000, np.random.default_rng(1))
    assert net.n_edge == net.n_node - 1
    assert (net.edges[:, 1] < net.edges[:, 0]).all()
    j = G.joint_adjacency(net)
    assert j.n == 2 * net.n_node - 1 and j.nnz == 6 * net.n_edge + j.n


# ---- pumped / tidal scenarios: chaohu (multigraph line graph, pump tables), hague (offsets, functional storage) ----
def _scenario(name):
    return np.load(os.path.join(GOLD, "scenario_%s.npz" % name))


@pytest.mark.parametrize("key,length,order", [("eadj_l0_o1", 0, 1), ("eadj_l0_o2", 0, 2), ("eadj_l200_o1", 200.0, 1)])
def test_chaohu_multigraph_edge_adjacency(key, length, order):
    """chaohu.py:248-282: parallel pumps between the same two nodes stay distinct line-graph vertices (bit-exact CSR
    against the networkx MultiGraph restatement run on the reference's chaohu.inp)."""
    z = _scenario("chaohu")
    M = len(z["links"])
    net = G.Network([str(s) for s in z["nodes"]], [str(s) for s in z["links"]], z["edges"], z["lengths"], np.zeros(len(z["nodes"])),
                    np.zeros(len(z["nodes"])), np.zeros(len(z["nodes"])), np.zeros(len(z["nodes"])), np.zeros(M), np.zeros(len(z["nodes"])))
    pairs = [tuple(e) for e in z["edges"].tolist()]
    assert len(set(pairs)) < len(pairs), "chaohu has parallel links"
    csr = G.edge_adjacency(net, False, length, order, multigraph=True)
    assert np.array_equal(csr.row_ptr, z[key + "_rp"]) and np.array_equal(csr.col_idx, z[key + "_ci"])
    assert np.array_equal(csr.val, z[key + "_val"])
    try:                                                                  # the base-class version merges the twins: the merged
        simple = G.edge_adjacency(net, False, length, order)              # link is no line-graph vertex (networkx raises too)
        assert simple.nnz != csr.nnz
    except KeyError:
        pass


@pytest.mark.parametrize("name", ["chaohu", "hague"])
def test_scenario_tables_from_inp(name):
    """get_args(scenario=...) derives area / pump / pump_in / pump_out / offset from the .inp as chaohu.py:203-208 and
    hague.py:127-132 do (goldens made by tools/make_golden.py --scenarios through the oracle's restatement)."""
    path = os.path.join("/root/reference/surrogate/envs/network", name, name + ".inp")
    if not os.path.exists(path):
        pytest.skip("reference networks not present on this machine")
    z = _scenario(name)
    a = G.get_args(path, False, 0, 1, 0, scenario=name)
    assert [str(s) for s in z["nodes"]] == a.nodes and [str(s) for s in z["links"]] == a.links
    for k in ("area", "pump", "pump_in", "pump_out"):
        assert np.array_equal(np.asarray(getattr(a, k), np.float64), z["tab_" + k]), k
    if name == "hague":
        assert np.array_equal(np.asarray(a.offset, np.float64), z["tab_offset"]) and a.offset.max() > 0
        assert a.tide is True
    else:
        assert (a.pump > 0).sum() == 7 and a.area.max() > 0                # chaohu: 7 pumps, two storage tanks
    key = "eadj_l0_o1"
    assert np.array_equal(a.edge_adj_csr.row_ptr, z[key + "_rp"]) and np.array_equal(a.edge_adj_csr.col_idx, z[key + "_ci"])
