"""Hot SASS lines of one kernel instance from `ncu -i rep --page source --csv` output.
  python tools/ncu_hot.py src.csv [block_index]"""
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
b = blocks[int(sys.argv[2]) if len(sys.argv) > 2 else 0]
hdr, data = b["hdr"], b["data"]
ia, isrc, ismp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
tot = sum(int(r[ia]) for r in data)
smp = sum(int(r[ismp]) for r in data)
print("blocks", len(blocks), "total inst", tot, "samples", smp, "sass lines", len(data))
for i, r in enumerate(data):
    n, s = int(r[ia]), int(r[ismp])
    if n > tot * 0.006 or s > smp * 0.01:
        print("%5d %-90s %12d %5.1f%% %7d %5.1f%%" % (i, r[isrc].strip()[:90], n, 100.0 * n / tot, s, 100.0 * s / max(smp, 1)))
