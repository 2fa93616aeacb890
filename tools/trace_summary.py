"""Mean offsets of the sp_conv clock64 stamps relative to graph_start (GNNUDS_TRACE_FILE output)."""
import sys

import numpy as np

NAMES = ["tma_issue", "in_full", "ne_issued", "ne_done", "nx_ready", "conv_done", "h_full", "graph_start", "softmax_done",
         "gather_done", "slowest", "warp15"]
for path in sys.argv[1:]:
    rows = [[int(v) for v in l.split()] for l in open(path) if not l.startswith("#")]
    a = np.array(rows)[10:40]
    print(path, "period %.0f" % np.diff(a[:, 8]).mean(),
          " ".join("%s=%.0f" % (n, (a[:, 1 + k] - a[:, 8]).mean()) for k, n in enumerate(NAMES) if k >= 5))
