"""Latency of a 12-chunk autoregressive roll of ONE scenario (Emulator.rollout, host buffers) - the regime of the mode-B
event replay and of a single MPC candidate.  python tools/roll_latency.py [precision]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gnn_uds_b200 import configs, synthetic  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
em, args, rng = configs.make_emulator("C3b", precision=prec)
T = em.seq_out * 12
for B in (1, 16):
    states, b, a, ex = synthetic.make_inputs(em, B, np.random.default_rng(3), horizon=T)
    for _ in range(5):
        y, ey = em.rollout(states, b, a, ex, rule="mpc")
    ts = []
    for _ in range(30):
        t0 = time.perf_counter(); y, ey = em.rollout(states, b, a, ex, rule="mpc"); ts.append(time.perf_counter() - t0)
    print("%s graphs=%s B=%d: median %.2f ms, min %.2f ms, checksum %.6f" % (prec, os.environ.get("GNNUDS_GRAPHS", "1"), B,
                                                                       1e3 * np.median(ts), 1e3 * min(ts), float(y.sum())))
