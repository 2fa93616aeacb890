"""Per-region (W SASS lines) instruction and stall-sample histogram of one kernel instance.
  python tools/ncu_regions.py src.csv block [W] [min_samples]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
blocks, cur = [], None
for r in rows:
    if len(r) > 5 and r[0] == "Address":
        cur = {"hdr": r, "data": []}
        blocks.append(cur)
    elif cur is not None and len(r) > 10:
        cur["data"].append(r)
b = blocks[int(sys.argv[2])]
W = int(sys.argv[3]) if len(sys.argv) > 3 else 40
ms = int(sys.argv[4]) if len(sys.argv) > 4 else 250
hdr, data = b["hdr"], b["data"]
H = {h: i for i, h in enumerate(hdr)}
ia, isrc, ismp = H["Instructions Executed"], H["Source"], H["# Samples"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("samples", sum(int(r[ismp]) for r in data), "inst", sum(int(r[ia]) for r in data))
for i in range(0, len(data), W):
    ch = data[i:i + W]
    n = sum(int(r[ia]) for r in ch)
    s = sum(int(r[ismp]) for r in ch)
    if s > ms:
        agg = {}
        for r in ch:
            for st in stalls:
                v = int(r[H[st]])
                if v:
                    agg[st[6:]] = agg.get(st[6:], 0) + v
        top = sorted(agg.items(), key=lambda kv: -kv[1])[:4]
        hot = max(ch, key=lambda r: int(r[ismp]))
        print("%5d inst %10d samp %6d  %-46s hot: %s (%s)" % (i, n, s, top, hot[isrc].strip()[:48], hot[ismp]))
