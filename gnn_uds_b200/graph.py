"""Graph tensors for the Emulator, built sparse (CSR) from a SWMM ``.inp``.

Replaces, for the hot path only, the reference's
``envs/scenario/base.py::get_args`` family (base.py:277-498) and
``Emulator.get_conv`` (emulator.py:129-152).  The reference materialises dense
``N x N`` / ``M x M`` float64 matrices through swmm_api + networkx; here the
same traversals run on plain adjacency lists and emit CSR directly, so the
100k-node path never allocates ``N^2``.  Dense views (``.adj``, ``.edge_adj``,
``.node_edge``) are materialised lazily for small networks because callers of
the reference read them off ``args``.

Section order follows swmm_api.input_file.section_lists (NODE_SECTIONS =
JUNCTIONS, OUTFALLS, DIVIDERS, STORAGE; LINK_SECTIONS = CONDUITS, PUMPS,
ORIFICES, WEIRS, OUTLETS), file order inside a section.
"""
import heapq
import math
from itertools import combinations
from types import SimpleNamespace

import numpy as np

NODE_SECTIONS = ("JUNCTIONS", "OUTFALLS", "DIVIDERS", "STORAGE")
LINK_SECTIONS = ("CONDUITS", "PUMPS", "ORIFICES", "WEIRS", "OUTLETS")


# ----------------------------------------------------------------------------
# .inp reading (only the sections get_args touches)
# ----------------------------------------------------------------------------
def _num(tok, default=0.0):
    try:
        return float(tok)
    except (TypeError, ValueError):
        return default


def read_inp(path):
    """-> {SECTION: [[tokens]...]} with ';' comments stripped."""
    sections, cur = {}, None
    with open(path, "r", errors="replace") as fh:
        for raw in fh:
            line = raw.split(";", 1)[0].strip()
            if not line:
                continue
            if line[0] == "[":
                cur = line.strip("[] \t").upper()
                sections.setdefault(cur, [])
            elif cur is not None:
                sections[cur].append(line.split())
    return sections


class Network:
    """Node/link tables of one drainage network in the reference's ordering."""

    def __init__(self, nodes, links, edges, lengths, is_outfall, is_storage,
                 hmax, elevation, ehmax, area):
        self.nodes, self.links = list(nodes), list(links)
        self.edges = np.asarray(edges, dtype=np.int64).reshape(-1, 2)
        self.lengths = np.asarray(lengths, dtype=np.float64)
        self.is_outfall = np.asarray(is_outfall)
        self.is_storage = np.asarray(is_storage)
        self.hmax = np.asarray(hmax, dtype=np.float64)
        self.elevation = np.asarray(elevation, dtype=np.float64)
        self.ehmax = np.asarray(ehmax, dtype=np.float64)
        self.area = np.asarray(area, dtype=np.float64)

    @property
    def n_node(self):
        return len(self.nodes)

    @property
    def n_edge(self):
        return len(self.edges)

    @classmethod
    def from_inp(cls, path):
        s = read_inp(path)
        nodes, kind, hmax, elev, curve_of = [], [], [], [], []
        for sec in NODE_SECTIONS:
            for t in s.get(sec, ()):
                nodes.append(t[0]); kind.append(sec); elev.append(_num(t[1] if len(t) > 1 else 0))
                if sec == "JUNCTIONS":     # Name Elev MaxDepth InitDepth SurDepth Aponded
                    hmax.append(_num(t[2] if len(t) > 2 else 0) + _num(t[4] if len(t) > 4 else 0))
                    curve_of.append(None)
                elif sec == "STORAGE":     # Name Elev MaxDepth InitDepth Type Curve ...
                    hmax.append(_num(t[2] if len(t) > 2 else 0))
                    curve_of.append(t[5] if len(t) > 5 and t[4].upper() == "TABULAR" else None)
                else:                      # base.py:287 getattr(node,'MaxDepth',0)
                    hmax.append(0.0); curve_of.append(None)
        pos = {n: i for i, n in enumerate(nodes)}
        links, edges, lengths = [], [], []
        for sec in LINK_SECTIONS:
            for t in s.get(sec, ()):
                if t[1] in pos and t[2] in pos:          # base.py:359
                    links.append(t[0]); edges.append((pos[t[1]], pos[t[2]]))
                    lengths.append(_num(t[3]) if sec == "CONDUITS" else 0.0)   # base.py:361
        geom1 = {t[0]: _num(t[2]) for t in s.get("XSECTIONS", ()) if len(t) > 2}
        first_pt = {}
        for t in s.get("CURVES", ()):
            vals = t[1:]
            if vals and _num(vals[0], None) is None:
                vals = vals[1:]
            if t[0] not in first_pt and len(vals) >= 2:
                first_pt[t[0]] = _num(vals[1])
        net = cls(nodes, links, edges, lengths,
                  [1 if k == "OUTFALLS" else 0 for k in kind],
                  [1 if k == "STORAGE" else 0 for k in kind],
                  hmax, elev, [geom1.get(l, 0.0) for l in links],
                  [first_pt.get(c, 0.0) if c else 0.0 for c in curve_of])
        # pumped / tidal scenarios (chaohu.py:203-208, hague.py:127-132): rated pump flow = first point of the pump curve
        # * 60/1000, inlet offset = Orifice/Outlet Offset or Conduit InOffset, FUNCTIONAL storage area = first coefficient
        pump, offset, kept = [], [], set(links)
        col = {"CONDUITS": 5, "ORIFICES": 4, "OUTLETS": 3}
        for sec in LINK_SECTIONS:
            for t in s.get(sec, ()):
                if t[0] in kept and t[1] in pos and t[2] in pos:
                    pump.append(first_pt.get(t[3], 0.0) * 60 / 1000 if sec == "PUMPS" else 0.0)
                    offset.append(_num(t[col[sec]]) if sec in col and len(t) > col[sec] else 0.0)
        net.pump, net.offset = np.asarray(pump), np.asarray(offset)
        net.area_functional = np.asarray([_num(t[5]) if len(t) > 5 and t[4].upper() == "FUNCTIONAL" else 0.0
                                          for t in s.get("STORAGE", ())])
        return net

    def area_functional_nodes(self):
        """hague.py:127: area per node = first coefficient of the FUNCTIONAL storage curve (0 for non-storage nodes)."""
        out, k = np.zeros(self.n_node), 0
        for i, st in enumerate(self.is_storage):
            if st:
                out[i] = self.area_functional[k]; k += 1
        return out

    def edge_list(self, link_names):
        """base.py:349-365 with ``links=...``: (from,to) rows for the named links."""
        idx = {l: i for i, l in enumerate(self.links)}
        return np.array([tuple(self.edges[idx[k]]) for k in link_names], dtype=np.int64)


# ----------------------------------------------------------------------------
# adjacency-list graphs with networkx-compatible iteration order
# ----------------------------------------------------------------------------
class _AdjList:
    """Insertion-ordered adjacency (what nx.Graph / nx.DiGraph iterate in)."""

    def __init__(self, directed=False):
        self.directed = directed
        self.adj = {}       # node -> {nbr: attr dict}
        self.pred = {}      # directed only

    def add_edge(self, u, v, **attr):
        for n in (u, v):
            if n not in self.adj:
                self.adj[n] = {}
                self.pred[n] = {}
        d = self.adj[u].get(v, {})
        d.update(attr)
        self.adj[u][v] = d
        if self.directed:
            self.pred[v][u] = d
        else:
            self.adj[v][u] = d

    def limited_dfs(self, source, depth_limit):
        """Node set of nx.dfs_preorder_nodes(G, source, depth_limit) (iterative,
        neighbour-order dependent: a node first reached at the depth limit is not
        expanded and is not revisited through a shorter path)."""
        if source not in self.adj:
            raise KeyError("node %r not in graph" % (source,))
        visited = {source}
        order = [source]
        stack = [(source, iter(self.adj[source]))]
        depth_now = 1
        while stack:
            _, children = stack[-1]
            for child in children:
                if child not in visited:
                    visited.add(child)
                    order.append(child)
                    if depth_now < depth_limit:
                        stack.append((child, iter(self.adj[child])))
                        depth_now += 1
                        break
            else:
                stack.pop()
                depth_now -= 1
        return order

    def dijkstra(self, source, cutoff):
        """{node: distance <= cutoff} as nx.single_source_dijkstra_path_length."""
        if source not in self.adj:
            raise KeyError("node %r not in graph" % (source,))
        dist, seen = {}, {source: 0.0}
        heap, c = [(0.0, 0, source)], 1
        while heap:
            d, _, v = heapq.heappop(heap)
            if v in dist:
                continue
            dist[v] = d
            for u, attr in self.adj[v].items():
                vu = d + attr.get("length", 1)
                if vu > cutoff:
                    continue
                if u in dist:
                    continue
                if u not in seen or vu < seen[u]:
                    seen[u] = vu
                    heapq.heappush(heap, (vu, c, u)); c += 1
        return dist


class CSR:
    """Row-compressed pattern (+ optional fp64 values) of one adjacency."""

    def __init__(self, n, rows):
        # rows: list of {col: val}
        self.n = n
        self.row_ptr = np.zeros(n + 1, dtype=np.int32)
        cols, vals = [], []
        for i, r in enumerate(rows):
            ks = sorted(k for k, v in r.items() if v != 0)
            cols.extend(ks); vals.extend(r[k] for k in ks)
            self.row_ptr[i + 1] = len(cols)
        self.col_idx = np.asarray(cols, dtype=np.int32)
        self.val = np.asarray(vals, dtype=np.float64)

    @property
    def nnz(self):
        return int(self.row_ptr[-1])

    def max_row(self):
        return int(np.diff(self.row_ptr).max()) if self.n else 0

    def to_dense(self):
        out = np.zeros((self.n, self.n))
        for i in range(self.n):
            s, e = self.row_ptr[i], self.row_ptr[i + 1]
            out[i, self.col_idx[s:e]] = self.val[s:e]
        return out

    def with_values(self, val):
        c = object.__new__(CSR)
        c.n, c.row_ptr, c.col_idx, c.val = self.n, self.row_ptr, self.col_idx, np.asarray(val, dtype=np.float64)
        return c


def _traverse(G, n_ele, length, order, l_std, symmetric):
    rows = [dict() for _ in range(n_ele)]
    for n in range(n_ele):
        if length:
            for a, l in G.dijkstra(n, length).items():
                rows[n][a] = float(np.exp(-(np.float64(l) / (l_std + 1e-5)) ** 2))   # base.py:380, numpy scalar math
        else:
            for a in (G.limited_dfs(n, order) if order > 0 else [n]):
                rows[n][a] = 1.0
                if symmetric:
                    rows[a][n] = 1.0
    return rows


def node_adjacency(net, directed=False, length=0, order=1):
    """base.py:367-391 get_adj -> CSR."""
    G = _AdjList(directed)
    l_std = np.std(net.lengths) if length else 0.0
    for (u, v), l in zip(net.edges, net.lengths):
        if length:
            G.add_edge(int(u), int(v), length=float(l))
        else:
            G.add_edge(int(u), int(v))
    n_node = int(net.edges.max()) + 1
    return CSR(n_node, _traverse(G, n_node, length, order, l_std, symmetric=(not directed) and not length))


def _multi_edge_adjacency(net, directed, length, order):
    """chaohu.py:248-282: get_edge_adj over a MULTIgraph - parallel links between the same two nodes (chaohu's twin pumps)
    stay distinct line-graph vertices.  Incident edges of a node are enumerated as nx.MultiGraph.edges(n, data=True) does:
    neighbours in first-insertion order, parallel edges in key order; a self loop once."""
    l_std = np.std(net.lengths) if length else 0.0
    adj, pred = {}, {}                     # node -> {nbr: [edge ids]}

    def touch(n):
        if n not in adj:
            adj[n] = {}; pred[n] = {}
    for i, (u, v) in enumerate(net.edges):
        u, v = int(u), int(v)
        touch(u); touch(v)
        adj[u].setdefault(v, []).append(i)
        if directed:
            pred[v].setdefault(u, []).append(i)
        elif u != v:
            adj[v].setdefault(u, []).append(i)
    EX = _AdjList(directed)

    def link(a, b):
        if length:
            EX.add_edge(a, b, length=(float(net.lengths[a]) + float(net.lengths[b])) / 2)
        else:
            EX.add_edge(a, b)
    for n in list(adj):
        if directed:
            for a in [e for es in pred[n].values() for e in es]:
                for b in [e for es in adj[n].values() for e in es]:
                    link(a, b)
        else:
            for a, b in combinations([e for es in adj[n].values() for e in es], 2):
                link(a, b)
    return CSR(net.n_edge, _traverse(EX, net.n_edge, length, order, l_std, symmetric=False))


def edge_adjacency(net, directed=False, length=0, order=1, multigraph=False):
    """base.py:393-429 get_edge_adj (line graph) -> CSR; multigraph=True follows the chaohu override."""
    if multigraph:
        return _multi_edge_adjacency(net, directed, length, order)
    X = _AdjList(directed)
    l_std = np.std(net.lengths) if length else 0.0
    for i, (u, v) in enumerate(net.edges):
        if length:
            X.add_edge(int(u), int(v), edge=i, length=float(net.lengths[i]))
        else:
            X.add_edge(int(u), int(v), edge=i)
    EX = _AdjList(directed)

    def link(u, v):
        if length:
            EX.add_edge(u["edge"], v["edge"], length=(u["length"] + v["length"]) / 2)
        else:
            EX.add_edge(u["edge"], v["edge"])
    for n in list(X.adj):
        if directed:
            for a in X.pred[n]:
                for d in X.adj[n]:
                    link(X.adj[a][n], X.adj[n][d])
        else:
            for b, d in combinations(list(X.adj[n]), 2):
                link(X.adj[n][b], X.adj[n][d])
    return CSR(net.n_edge, _traverse(EX, net.n_edge, length, order, l_std, symmetric=False))


def joint_adjacency(net, directed=False, length=0, order=1):
    """base.py:471-498 get_node_based_adj (graph_base=1): nodes then links as one graph."""
    n_node, n_edge = int(net.edges.max()) + 1, net.n_edge
    G = _AdjList(directed)
    l_std = np.std(net.lengths) if length else 0.0
    for i, ((u, v), l) in enumerate(zip(net.edges, net.lengths)):
        kw = dict(length=float(l) / 2) if length else {}
        G.add_edge(int(u), int(v), **kw)
        G.add_edge(int(u), n_node + i, **kw)
        G.add_edge(n_node + i, int(v), **kw)
    n = n_node + n_edge
    return CSR(n, _traverse(G, n, length, order, l_std, symmetric=(not directed) and not length))


def edge_based_adjacency(net, directed=False, length=0, order=1):
    """base.py:500-532 get_edge_based_adj (graph_base=2): line graph of the links (ids n_node+i) plus a tie from every
    link to its end nodes; one-way fill (the reference does not mirror these rows)."""
    n_node, n_edge = int(net.edges.max()) + 1, net.n_edge
    l_std = np.std(net.lengths) if length else 0.0
    X = _AdjList(directed)
    for i, (u, v) in enumerate(net.edges):
        X.add_edge(int(u), int(v), edge=n_node + i, length=float(net.lengths[i]) if length else 0)
    EX = _AdjList(directed)

    def tie(a, b, l):
        if length:
            EX.add_edge(a, b, length=l)
        else:
            EX.add_edge(a, b)
    for n in list(X.adj):
        if directed:
            for a in X.pred[n]:
                for d in X.adj[n]:
                    u, v = X.adj[a][n], X.adj[n][d]
                    tie(u["edge"], v["edge"], (u["length"] + v["length"]) / 2)
            for a in X.pred[n]:
                tie(X.adj[a][n]["edge"], n, X.adj[a][n]["length"] / 2)
            for d in X.adj[n]:
                tie(n, X.adj[n][d]["edge"], X.adj[n][d]["length"] / 2)
        else:
            for b, d in combinations(list(X.adj[n]), 2):
                u, v = X.adj[n][b], X.adj[n][d]
                tie(u["edge"], v["edge"], (u["length"] + v["length"]) / 2)
            for b in list(X.adj[n]):
                tie(X.adj[n][b]["edge"], n, X.adj[n][b]["length"] / 2)
    n = n_node + n_edge
    return CSR(n, _traverse(EX, n, length, order, l_std, symmetric=False))


def gcn_normalise(csr):
    """spektral.utils.gcn_filter on CSR: D^-1/2 (A + I) D^-1/2, D = rowsum(A + I).
    The reference's A already carries 1 on the diagonal, so the diagonal becomes 2."""
    rows = []
    for i in range(csr.n):
        s, e = csr.row_ptr[i], csr.row_ptr[i + 1]
        r = dict(zip(csr.col_idx[s:e].tolist(), csr.val[s:e].tolist()))
        r[i] = r.get(i, 0.0) + 1.0
        rows.append(r)
    deg = np.array([math.fsum(r.values()) for r in rows])
    with np.errstate(divide="ignore"):
        d = np.power(deg, -0.5)
    d[np.isinf(d)] = 0.0
    for i, r in enumerate(rows):
        for j in r:
            r[j] = (d[i] * r[j]) * d[j]
    return CSR(csr.n, rows)


def gat_pattern(csr):
    """(adj > 0) with the diagonal forced (GATConv add_self_loops -> set_diag(a,1))."""
    rows = []
    for i in range(csr.n):
        s, e = csr.row_ptr[i], csr.row_ptr[i + 1]
        r = {int(c): 1.0 for c, v in zip(csr.col_idx[s:e], csr.val[s:e]) if v > 0}
        r[i] = 1.0
        rows.append(r)
    return CSR(csr.n, rows)


def incidence(net):
    """base.py:432-439 get_node_edge as a dense signed matrix (small nets only)."""
    ne = np.zeros((net.n_node, net.n_edge))
    for i, (u, v) in enumerate(net.edges):
        ne[u, i] += 1
        ne[v, i] += -1
    return ne


def get_args(inp_path, directed=False, length=0, order=1, graph_base=0,
             act_links=None, head=False, dense=True, scenario=None, **extra):
    """Drop-in for ``env.get_args(directed,length,order,graph_base)`` restricted to the
    keys ``Emulator`` reads (emulator.py:49-127).  Returns a namespace; ``adj_csr`` /
    ``edge_adj_csr`` are what this package's Emulator consumes, the dense ``adj`` /
    ``edge_adj`` / ``node_edge`` are there for callers that read them (``dense=False``
    skips them, required for 100k-node graphs)."""
    net = Network.from_inp(inp_path) if isinstance(inp_path, str) else inp_path
    a = SimpleNamespace(**extra)
    a.network = net
    a.nodes, a.links = net.nodes, net.links
    a.edges = net.edges
    a.is_outfall, a.is_storage = net.is_outfall, net.is_storage
    a.hmin = net.elevation.copy() if head else np.zeros_like(net.hmax)
    a.hmax = net.hmax + a.hmin if head else net.hmax
    a.ehmax, a.area = net.ehmax, net.area
    a.state_shape = (net.n_node, 4)
    a.edge_state_shape = (net.n_edge, 4)
    a.graph_base = graph_base
    if graph_base == 1:
        a.adj_csr = joint_adjacency(net, directed, length, order)
    elif graph_base == 2:
        a.adj_csr = edge_based_adjacency(net, directed, length, order)
    else:
        a.adj_csr = node_adjacency(net, directed, length, order)
    a.edge_adj_csr = edge_adjacency(net, directed, length, order, multigraph=(scenario == "chaohu"))
    if scenario in ("chaohu", "hague"):
        # chaohu.py:199-208 / hague.py:115-132: storage areas, rated pump flows, inlet offsets and the node-side pump tables
        a.area = net.area_functional_nodes() if scenario == "hague" else net.area
        a.pump = net.pump
        if scenario == "hague":
            a.offset = net.offset
        a.pump_out = np.zeros(net.n_node); a.pump_in = np.zeros(net.n_node)
        for m, (u, v) in enumerate(net.edges):
            if u != v:                       # a self loop cancels in node_edge (+1 - 1)
                a.pump_out[u] += net.pump[m]; a.pump_in[v] += net.pump[m]
        if scenario == "hague":
            a.tide, a.head = True, head
    if dense:
        a.adj = a.adj_csr.to_dense()
        a.edge_adj = a.edge_adj_csr.to_dense()
        a.node_edge = incidence(net)
    else:
        a.node_edge = None
    if act_links is not None:
        a.act_edges = net.edge_list(act_links)
    return a
