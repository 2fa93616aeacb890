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
