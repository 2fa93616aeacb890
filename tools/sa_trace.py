"""Summary of a sp_agg_kernel clock64 trace (GNNUDS_TRACE_FILE): median cycles of every phase over samples 8..60."""
import sys
import numpy as np

lines = open(sys.argv[1]).read().splitlines()
print(lines[0][:110])
t = np.array([[int(v) for v in l.split()[1:]] for l in lines[1:] if l.strip()], dtype=np.int64)
t = t[8:60]
per = np.median(np.diff(t[:, 0]))
seq = [("A tmem", 0, 6), ("A store", 6, 7), ("A fence+arrive", 7, 1), ("C ne drain", 1, 3), ("B wait+bar", 3, 8), ("B zero+bar", 8, 9), ("B softmax", 9, 2),
       ("D wait agg", 2, 4), ("D tmem", 4, 10), ("D stage", 10, 11), ("D fence+bar", 11, 12), ("D tma issue", 12, 5)]
print("period %.0f cycles" % per)
print("  " + "  ".join("%s %.0f" % (n, np.median(t[:, b] - t[:, a])) for n, a, b in seq))
print("  next conv_done after out_done: %.0f" % np.median(t[1:, 0] - t[:-1, 5]))
print("  MMA thread: conv issued at %+.0f, agg issue at %+.0f (alpha_ready %+.0f), agg issued %+.0f, agg_done seen %+.0f  (relative to conv_done_seen)" % (
    np.median(t[:, 13] - t[:, 0]), np.median(t[:, 14] - t[:, 0]), np.median(t[:, 2] - t[:, 0]), np.median(t[:, 15] - t[:, 0]), np.median(t[:, 4] - t[:, 0])))
if t.shape[1] > 18:
    print("  conv[i] issue took %.0f, NE[i+1] issue %+.0f .. %+.0f (took %.0f)" % (np.median(t[:, 13] - t[:, 18]), np.median(t[1:, 16] - t[:-1, 0]), np.median(t[1:, 17] - t[:-1, 0]), np.median(t[:, 17] - t[:, 16])))
