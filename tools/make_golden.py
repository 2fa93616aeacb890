"""Generate the committed fixtures under tests/golden/ from the reference's own network
files (read-only under /root/reference; absent on the GPU box, hence the fixtures).

  python tools/make_golden.py            # network tables + graph tensors
  python tools/make_golden.py --forward  # also oracle forward vectors (needs numpy only)

Network tables and dense graph tensors come from oracle/graph_oracle.py, which restates
envs/scenario/base.py:277-498 with networkx exactly as the reference does.  Forward vectors
come from oracle/emulator_oracle.py (float64 truth and float32) on seeded synthetic weights
and inputs; tests rebuild the same weights/inputs from the seeds through the product's
generators and compare.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import graph_oracle as go   # noqa: E402

REF = "/root/reference/surrogate/envs/network"
OUT = os.path.join(ROOT, "tests", "golden")
AST_ACT_LINKS = ["V2", "V3", "V4", "V6"]       # envs/config/astlingen.yaml:114-163 action_space keys


def events_table():
    """envs/network/shunqing/events.csv -> gnn_uds_b200/data/shunqing_events.npz (name, Train/Test, duration in hours):
    the event list of BASELINE.json configs[3] (replay of all 148 synthetic rainfalls)."""
    import csv
    rows = list(csv.reader(open(os.path.join(REF, "shunqing", "events.csv"))))[1:]
    out = os.path.join(ROOT, "gnn_uds_b200", "data", "shunqing_events.npz")
    np.savez_compressed(out, name=np.array([r[0] for r in rows]), set=np.array([r[1] for r in rows]),
                        duration_h=np.array([float(r[2]) for r in rows]))
    print("wrote", out, len(rows), "events")


def scenario_goldens():
    """chaohu (multigraph line graph + pump tables) and hague (offsets, functional storage areas) from the reference .inp
    files through the networkx restatements of oracle/graph_oracle.py."""
    for name, variants in (("chaohu", [(0, 1), (0, 2), (200.0, 1)]), ("hague", [(0, 1)])):
        path = os.path.join(REF, name, name + ".inp")
        net = go.parse_network(path)
        d = dict(nodes=np.array(net["nodes"]), links=np.array(net["links"]), edges=net["edges"], lengths=net["lengths"])
        d.update({"tab_" + k: v for k, v in go.scenario_tables(path, name).items()})
        for length, order in variants:
            L = net["lengths"] if length else None
            m = go.get_edge_adj_multigraph(net["edges"], L, False, length, order) if name == "chaohu" else go.get_edge_adj(net["edges"], L, False, length, order)
            rp, ci, vv = dense_to_csr(m)
            key = "eadj_l%g_o%d" % (length, order)
            d[key + "_rp"], d[key + "_ci"], d[key + "_val"] = rp, ci, vv
        np.savez_compressed(os.path.join(OUT, "scenario_%s.npz" % name), **d)
        print("wrote scenario_%s.npz" % name, len(net["nodes"]), "nodes", len(net["links"]), "links")


def dense_to_csr(a):
    rp, ci, vv = [0], [], []
    for i in range(a.shape[0]):
        nz = np.nonzero(a[i])[0]
        ci.extend(nz.tolist()); vv.extend(a[i, nz].tolist()); rp.append(len(ci))
    return np.array(rp, np.int32), np.array(ci, np.int32), np.array(vv, np.float64)


def networks():
    for name in ("astlingen", "shunqing"):
        net = go.parse_network(os.path.join(REF, name, name + ".inp"))
        extra = {}
        if name == "astlingen":
            extra["act_edges"] = np.array([net["edges"][net["links"].index(k)] for k in AST_ACT_LINKS])
            extra["act_links"] = np.array(AST_ACT_LINKS)
        np.savez_compressed(os.path.join(OUT, name + "_net.npz"),
                            nodes=np.array(net["nodes"]), links=np.array(net["links"]), edges=net["edges"],
                            lengths=net["lengths"], is_outfall=net["is_outfall"], is_storage=net["is_storage"],
                            hmax=net["hmax"], elevation=net["elevation"], ehmax=net["ehmax"], area=net["area"], **extra)
        import shutil
        os.makedirs(os.path.join(ROOT, "gnn_uds_b200", "data"), exist_ok=True)
        shutil.copy(os.path.join(OUT, name + "_net.npz"), os.path.join(ROOT, "gnn_uds_b200", "data", name + "_net.npz"))
        variants = {"astlingen": [(0, 1), (0, 2), (1.0, 1), (501, 1)],
                    "shunqing": [(0, 1), (0, 2), (1200.0, 1), (1000.0, 1)]}[name]
        for length, order in variants:
            L = net["lengths"] if length else None
            adj = go.get_adj(net["edges"], L, False, length, order)
            eadj = go.get_edge_adj(net["edges"], L, False, length, order)
            d = {}
            for key, m in (("adj", adj), ("edge_adj", eadj)):
                rp, ci, vv = dense_to_csr(m)
                d[key + "_rp"], d[key + "_ci"], d[key + "_val"] = rp, ci, vv
            if (length, order) == (0, 1):
                rp, ci, vv = dense_to_csr(go.get_node_based_adj(net["edges"], None, False, 0, 1))
                d["joint_rp"], d["joint_ci"], d["joint_val"] = rp, ci, vv
            np.savez_compressed(os.path.join(OUT, "graph_%s_l%g_o%d.npz" % (name, length, order)), **d)
            print(name, length, order, "adj nnz", int((adj != 0).sum()), "max row", int((adj != 0).sum(1).max()),
                  "edge_adj nnz", int((eadj != 0).sum()), "max row", int((eadj != 0).sum(1).max()))
        print(name, "N", len(net["nodes"]), "M", len(net["edges"]), "len mean/std", net["lengths"].mean(), net["lengths"].std())


def forward_vectors():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import cases   # tests/cases.py
    for case in cases.GOLDEN_CASES:
        d = cases.oracle_run(case)
        np.savez_compressed(os.path.join(OUT, "fwd_%s.npz" % case), **d)
        print(case, {k: v.shape for k, v in d.items()})


if __name__ == "__main__" and "--scenarios" in sys.argv:
    scenario_goldens()
    sys.exit(0)

if __name__ == "__main__" and "--events" in sys.argv:
    events_table()
    sys.exit(0)

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--forward", action="store_true")
    a = ap.parse_args()
    os.makedirs(OUT, exist_ok=True)
    if os.path.isdir(REF):
        networks()
    else:
        print("reference networks not present; skipping network fixtures")
    if a.forward:
        forward_vectors()
