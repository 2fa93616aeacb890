"""Dense graph-tensor restatement of the reference's ``get_args`` family.

Oracle (test infrastructure, see ``oracle/__init__.py``).  Follows
``/root/reference/surrogate/envs/scenario/base.py:277-498`` with networkx, the
way the reference does it (dense ``N x N`` float64 matrices, ``nx.dfs_preorder_nodes``
and ``nx.single_source_dijkstra_path_length``).  The product builds CSR directly
with its own traversal (``gnn_uds_b200/graph.py``); tests compare the two.

swmm-api (0.2.0.18.3, absent here) is replaced by a minimal section reader that
keeps only what ``get_args`` touches: section order NODE_SECTIONS =
JUNCTIONS, OUTFALLS, DIVIDERS, STORAGE and LINK_SECTIONS = CONDUITS, PUMPS,
ORIFICES, WEIRS, OUTLETS (swmm_api.input_file.section_lists), file order inside
a section.
"""
from itertools import combinations

import numpy as np
import networkx as nx

NODE_SECTIONS = ["JUNCTIONS", "OUTFALLS", "DIVIDERS", "STORAGE"]
LINK_SECTIONS = ["CONDUITS", "PUMPS", "ORIFICES", "WEIRS", "OUTLETS"]


def read_sections(path):
    """Split an .inp into {SECTION: [token lists]} (comments/blank lines dropped)."""
    secs, cur = {}, None
    with open(path, "r", errors="replace") as f:
        for raw in f:
            line = raw.split(";")[0].strip()
            if not line:
                continue
            if line.startswith("["):
                cur = line.strip("[]").upper()
                secs.setdefault(cur, [])
            elif cur is not None:
                secs[cur].append(line.split())
    return secs


def _f(tok, default=0.0):
    try:
        return float(tok)
    except (TypeError, ValueError):
        return default


def parse_network(path):
    """Nodes/links in the reference's order with the attributes get_args reads.

    base.py:280-294 (is_outfall, hmax = MaxDepth + SurDepth, hmin),
    base.py:335-365 (get_features / get_edge_list), base.py:318 (ehmax = XSECTIONS.Geom1),
    astlingen.py:147-149 (area = first point of the storage curve).
    """
    s = read_sections(path)
    nodes, sec_of = [], []
    hmax, elev, curve = [], [], []
    for sec in NODE_SECTIONS:
        for t in s.get(sec, []):
            nodes.append(t[0])
            sec_of.append(sec)
            elev.append(_f(t[1]))
            if sec == "JUNCTIONS":
                # Name Elevation MaxDepth InitDepth SurDepth Aponded
                md = _f(t[2]) if len(t) > 2 else 0.0
                sd = _f(t[4]) if len(t) > 4 else 0.0
                hmax.append(md + sd)
                curve.append(None)
            elif sec == "STORAGE":
                # Name Elev MaxDepth InitDepth Type Curve ...; swmm-api 0.2.0.18.3
                # Storage has no SurDepth attribute -> getattr(...,0)
                hmax.append(_f(t[2]))
                curve.append(t[5] if len(t) > 5 and t[4].upper() == "TABULAR" else None)
            else:
                hmax.append(0.0)     # Outfall/Divider: no MaxDepth attribute in this swmm-api
                curve.append(None)
    links, frm, to, length = [], [], [], []
    for sec in LINK_SECTIONS:
        for t in s.get(sec, []):
            links.append(t[0])
            frm.append(t[1])
            to.append(t[2])
            length.append(_f(t[3]) if sec == "CONDUITS" else 0.0)
    xs = {t[0]: _f(t[2]) for t in s.get("XSECTIONS", [])}
    curves = {}
    for t in s.get("CURVES", []):
        name = t[0]
        vals = t[1:]
        if vals and _f(vals[0], None) is None:
            vals = vals[1:]           # curve type keyword on the first row
        pts = [(_f(vals[i]), _f(vals[i + 1])) for i in range(0, len(vals) - 1, 2)]
        curves.setdefault(name, []).extend(pts)
    idx = {n: i for i, n in enumerate(nodes)}
    edges, lengths, kept = [], [], []
    for k, (u, v, l) in enumerate(zip(frm, to, length)):
        if u in idx and v in idx:        # base.py:359
            edges.append((idx[u], idx[v]))
            lengths.append(l)
            kept.append(k)
    area = np.array([curves[c][0][1] if (c is not None and c in curves) else 0.0 for c in curve])
    return dict(
        nodes=nodes, links=links,
        edges=np.array(edges, dtype=np.int64), lengths=np.array(lengths, dtype=np.float64),
        is_outfall=np.array([1 if x == "OUTFALLS" else 0 for x in sec_of]),
        is_storage=np.array([1 if x == "STORAGE" else 0 for x in sec_of]),
        hmax=np.array(hmax, dtype=np.float64), elevation=np.array(elev, dtype=np.float64),
        ehmax=np.array([xs.get(l, 0.0) for l in links], dtype=np.float64),
        area=area,
    )


def get_adj(edges, lengths=None, directed=False, length=0, order=1):
    """base.py:367-391."""
    X = nx.DiGraph() if directed else nx.Graph()
    n_node = int(edges.max()) + 1
    adj = np.zeros((n_node, n_node))
    if length:
        l_std = np.std(lengths)
        for (u, v), l in zip(edges, lengths):
            X.add_edge(int(u), int(v), length=l)
        for n in range(n_node):
            pl = nx.single_source_dijkstra_path_length(X, n, weight="length", cutoff=length)
            for a, l in pl.items():
                adj[n, a] = np.exp(-(l / (l_std + 1e-5)) ** 2)
    else:
        for u, v in edges:
            X.add_edge(int(u), int(v))
        for n in range(n_node):
            for a in list(nx.dfs_preorder_nodes(X, n, order) if order > 0 else [n]):
                adj[n, a] = 1
                if not directed:
                    adj[a, n] = 1
    return adj


def get_edge_adj(edges, lengths=None, directed=False, length=0, order=1):
    """base.py:393-429 (line graph, then the same traversal)."""
    X = nx.DiGraph() if directed else nx.Graph()
    if length:
        l_std = np.std(lengths)
    for i, (u, v) in enumerate(edges):
        X.add_edge(int(u), int(v), edge=i)
        if length:
            X[int(u)][int(v)].update(length=lengths[i])
    EX = nx.DiGraph() if directed else nx.Graph()
    for n in X.nodes():
        if directed:
            for a, b in X.in_edges(n):
                for c, d in X.out_edges(n):
                    u, v = X[a][b], X[c][d]
                    EX.add_edge(u["edge"], v["edge"])
                    if length:
                        EX[u["edge"]][v["edge"]].update(length=(u["length"] + v["length"]) / 2)
        else:
            for (a, b), (c, d) in combinations(X.edges(n), 2):
                u, v = X[a][b], X[c][d]
                EX.add_edge(u["edge"], v["edge"])
                if length:
                    EX[u["edge"]][v["edge"]].update(length=(u["length"] + v["length"]) / 2)
    n_edge = edges.shape[0]
    adj = np.zeros((n_edge, n_edge))
    for n in range(n_edge):
        if length:
            pl = nx.single_source_dijkstra_path_length(EX, n, weight="length", cutoff=length)
            for a, l in pl.items():
                adj[n, a] = np.exp(-(l / (l_std + 1e-5)) ** 2)
        else:
            for a in list(nx.dfs_preorder_nodes(EX, n, order) if order > 0 else [n]):
                adj[n, a] = 1
    return adj


def get_node_edge(n_node, edges):
    """base.py:432-439 (signed incidence)."""
    ne = np.zeros((n_node, len(edges)))
    for i, (u, v) in enumerate(edges):
        ne[u, i] += 1
        ne[v, i] += -1
    return ne


def get_node_based_adj(edges, lengths=None, directed=False, length=0, order=1):
    """base.py:471-498 (joint node+edge graph, graph_base=1)."""
    n_node, n_edge = int(edges.max()) + 1, edges.shape[0]
    adj = np.zeros((n_node + n_edge, n_node + n_edge))
    X = nx.DiGraph() if directed else nx.Graph()
    if length:
        l_std = np.std(lengths)
        for i, ((u, v), l) in enumerate(zip(edges, lengths)):
            X.add_edge(int(u), int(v), length=l / 2)
            X.add_edge(int(u), n_node + i, length=l / 2)
            X.add_edge(n_node + i, int(v), length=l / 2)
        for n in range(n_node + n_edge):
            pl = nx.single_source_dijkstra_path_length(X, n, weight="length", cutoff=length)
            for a, l in pl.items():
                adj[n, a] = np.exp(-(l / (l_std + 1e-5)) ** 2)
    else:
        for i, (u, v) in enumerate(edges):
            X.add_edge(int(u), int(v))
            X.add_edge(int(u), n_node + i)
            X.add_edge(n_node + i, int(v))
        for n in range(n_node + n_edge):
            for a in list(nx.dfs_preorder_nodes(X, n, order) if order > 0 else [n]):
                adj[n, a] = 1
                if not directed:
                    adj[a, n] = 1
    return adj


def get_edge_based_adj(edges, lengths=None, directed=False, length=0, order=1):
    """base.py:500-532 (graph_base=2): the line graph of the links (ids n_node+i), every link also tied to its end
    nodes; rows are filled one way only (no symmetric mirror, unlike get_node_based_adj)."""
    n_node, n_edge = int(edges.max()) + 1, edges.shape[0]
    l_std = np.std(lengths) if length else 0.0
    X = nx.DiGraph() if directed else nx.Graph()
    for i, (u, v) in enumerate(edges):
        X.add_edge(int(u), int(v), edge=n_node + i, length=lengths[i] if length else 0)
    EX = nx.DiGraph() if directed else nx.Graph()
    half = lambda d: d["length"] / 2 if length else 0
    for n in X.nodes():
        if directed:
            pairs = [(X[a][b], X[c][d]) for (a, b) in X.in_edges(n) for (c, d) in X.out_edges(n)]
        else:
            pairs = [(X[a][b], X[c][d]) for (a, b), (c, d) in combinations(X.edges(n), 2)]
        for u, v in pairs:
            EX.add_edge(u["edge"], v["edge"], length=(u["length"] + v["length"]) / 2 if length else 0)
        if directed:
            for a, b in X.in_edges(n):
                EX.add_edge(X[a][b]["edge"], n, length=half(X[a][b]))
            for c, d in X.out_edges(n):
                EX.add_edge(n, X[c][d]["edge"], length=half(X[c][d]))
        else:
            for a, b in X.edges(n):
                EX.add_edge(X[a][b]["edge"], n, length=half(X[a][b]))
    adj = np.zeros((n_node + n_edge, n_node + n_edge))
    for n in range(n_node + n_edge):
        if length:
            for a, l in nx.single_source_dijkstra_path_length(EX, n, weight="length", cutoff=length).items():
                adj[n, a] = np.exp(-(l / (l_std + 1e-5)) ** 2)
        else:
            for a in (list(nx.dfs_preorder_nodes(EX, n, order)) if order > 0 else [n]):
                adj[n, a] = 1
    return adj


def get_args(path, directed=False, length=0, order=1, graph_base=0, act_links=None, head=False):
    """Restates base.py:277-332 (+ astlingen.py:136-150 for area/act_edges)."""
    net = parse_network(path)
    n_node, n_edge = len(net["nodes"]), len(net["edges"])
    args = dict(net)
    args["hmin"] = net["elevation"].copy() if head else np.zeros_like(net["hmax"])
    if head:
        args["hmax"] = net["hmax"] + args["hmin"]
    args["state_shape"] = (n_node, 4)
    args["edge_state_shape"] = (n_edge, 4)
    L = net["lengths"] if length else None
    if graph_base == 1:
        args["adj"] = get_node_based_adj(net["edges"], L, directed, length, order)
    elif graph_base == 2:
        args["adj"] = get_edge_based_adj(net["edges"], L, directed, length, order)
    else:
        args["adj"] = get_adj(net["edges"], L, directed, length, order)
    args["edge_adj"] = get_edge_adj(net["edges"], L, directed, length, order)
    args["node_edge"] = get_node_edge(n_node, net["edges"])
    if act_links is not None:
        # get_edge_list(links=...) base.py:352-365
        frm_to = []
        for k in act_links:
            i = net["links"].index(k)
            frm_to.append(tuple(net["edges"][i]))
        args["act_edges"] = np.array(frm_to)
    return args


def get_edge_adj_multigraph(edges, lengths=None, directed=False, length=0, order=1):
    """envs/scenario/chaohu.py:248-282: the line graph of a MULTIgraph (chaohu has parallel pumps between the same two
    nodes, which nx.Graph would merge), then the same traversal."""
    X = nx.MultiDiGraph() if directed else nx.MultiGraph()
    if length:
        l_std = np.std(lengths)
    for i, (u, v) in enumerate(edges):
        if length:
            X.add_edge(int(u), int(v), edge=i, length=lengths[i])
        else:
            X.add_edge(int(u), int(v), edge=i)
    EX = nx.DiGraph() if directed else nx.Graph()
    for n in X.nodes():
        if directed:
            for _, _, u in X.in_edges(n, data=True):
                for _, _, v in X.out_edges(n, data=True):
                    EX.add_edge(u["edge"], v["edge"])
                    if length:
                        EX[u["edge"]][v["edge"]].update(length=(u["length"] + v["length"]) / 2)
        else:
            for (_, _, u), (_, _, v) in combinations(X.edges(n, data=True), 2):
                EX.add_edge(u["edge"], v["edge"])
                if length:
                    EX[u["edge"]][v["edge"]].update(length=(u["length"] + v["length"]) / 2)
    n_edge = edges.shape[0]
    adj = np.zeros((n_edge, n_edge))
    for n in range(n_edge):
        if length:
            p_l = nx.single_source_dijkstra_path_length(EX, n, weight="length", cutoff=length)
            for a, l in p_l.items():
                adj[n, a] = np.exp(-(l / (l_std + 1e-5)) ** 2)
        else:
            for a in list(nx.dfs_preorder_nodes(EX, n, order)):
                adj[n, a] = 1
    return adj


def scenario_tables(path, scenario):
    """The extra get_args vectors of the pumped / tidal scenarios (restated on the raw .inp sections; swmm-api attribute
    names in brackets): chaohu.py:203-208 - area = first point of the storage curve [CURVES[node.Curve].points[0][1]],
    pump = rated flow of the pump curve's first point * 60/1000 [CURVES[link.Curve].points[0][1]], pump_in / pump_out
    = rows of |clip(node_edge)| * pump; hague.py:127-132 - area = first coefficient of the FUNCTIONAL storage curve
    [node.Curve[0]], pump as above, offset = Orifice/Outlet.Offset + Conduit.InOffset."""
    s = read_sections(path)
    net = parse_network(path)
    curves = {}
    for t in s.get("CURVES", []):
        vals = t[1:]
        if vals and _f(vals[0], None) is None:
            vals = vals[1:]
        curves.setdefault(t[0], []).extend([(_f(vals[i]), _f(vals[i + 1])) for i in range(0, len(vals) - 1, 2)])
    area, pump, offset = [], [], []
    for sec in NODE_SECTIONS:
        for t in s.get(sec, []):
            if sec != "STORAGE":
                area.append(0.0)
            elif scenario == "hague":
                area.append(_f(t[5]))                                  # FUNCTIONAL A B C: Curve = [A, B, C]
            else:
                area.append(curves[t[5]][0][1])                        # TABULAR curve name
    for sec in LINK_SECTIONS:
        for t in s.get(sec, []):
            pump.append(curves[t[3]][0][1] * 60 / 1000 if sec == "PUMPS" else 0.0)
            offset.append({"CONDUITS": lambda: _f(t[5]) if len(t) > 5 else 0.0, "ORIFICES": lambda: _f(t[4]) if len(t) > 4 else 0.0,
                           "OUTLETS": lambda: _f(t[3]) if len(t) > 3 else 0.0}.get(sec, lambda: 0.0)())
    area, pump, offset = np.array(area), np.array(pump), np.array(offset)
    ne = get_node_edge(len(net["nodes"]), net["edges"])
    return dict(area=area, pump=pump, offset=offset, pump_in=(-np.clip(ne, -1, 0) * pump).sum(axis=1),
                pump_out=(np.clip(ne, 0, 1) * pump).sum(axis=1))
