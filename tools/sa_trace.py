"""Summary of a sp_agg_kernel clock64 trace (GNNUDS_TRACE_FILE): median cycles between consecutive worker stamps of a
sample (in the order they occur) over samples 8..60, and the MMA thread's issue windows."""
import sys
import numpy as np

lines = open(sys.argv[1]).read().splitlines()
print(lines[0][:110])
t = np.array([[int(v) for v in l.split()[1:]] for l in lines[1:] if l.strip()], dtype=np.int64)
t = t[8:60]
names = {0: "conv_done_seen", 6: "A:tmem", 7: "A:stored", 1: "h_ready", 3: "ne_drained", 8: "B:bar1", 9: "B:zeroed", 2: "alpha_ready",
         4: "agg_done_seen", 10: "D:tmem", 11: "D:staged", 12: "D:bar", 5: "out_done"}
per = np.median(np.diff(t[:, 0]))
off = {k: np.median(t[:, k] - t[:, 0]) for k in names}
order = sorted(names, key=lambda k: off[k])
print("period %.0f cycles" % per)
print("  " + "  ".join("%s +%.0f" % (names[k], off[k] - off[order[i - 1]] if i else 0) for i, k in enumerate(order)))
print("  next conv_done after out_done: %.0f" % np.median(t[1:, 0] - t[:-1, 5]))
print("  MMA thread (relative to conv_done_seen[i]): conv[i] issued %+.0f, agg issue %+.0f .. %+.0f (alpha_ready %+.0f, agg_done seen %+.0f)" % (
    np.median(t[:, 13] - t[:, 0]), np.median(t[:, 14] - t[:, 0]), np.median(t[:, 15] - t[:, 0]), off[2], off[4]))
if t.shape[1] > 18:
    print("  conv[i+1] issue %+.0f .. %+.0f, NE issue windows (by sample index) %+.0f .. %+.0f (took %.0f)" % (
        np.median(t[1:, 18] - t[:-1, 0]), np.median(t[1:, 13] - t[:-1, 0]), np.median(t[:, 16] - t[:, 0]), np.median(t[:, 17] - t[:, 0]),
        np.median(t[:, 17] - t[:, 16])))
