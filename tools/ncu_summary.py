#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel.

  python tools/ncu_summary.py gpurun_out/launches.csv > profiles/rNN_launches.md

Per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.
"""
import csv
import re
import sys
from collections import OrderedDict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ns = v * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        rows.append((name, ns, r["Grid Size"], r["Block Size"]))
    agg = OrderedDict()
    for name, ns, grid, block in rows:
        a = agg.setdefault(name, [0, 0.0, grid, block])
        a[0] += 1
        a[1] += ns
    tot = sum(a[1] for a in agg.values()) or 1.0
    print("| kernel | launches | total ms | share | avg us | example grid x block |")
    print("|---|---:|---:|---:|---:|---|")
    for name, (n, ns, grid, block) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %d | %.3f | %.1f%% | %.1f | %s x %s |" % (name, n, ns / 1e6, 100 * ns / tot, ns / n / 1e3, grid, block))
    print("\n%d launches, %.3f ms summed (ncu-serialised, cold cache)" % (len(rows), tot / 1e6))


if __name__ == "__main__":
    main(sys.argv[1])
