"""One Emulator.predict of a named workload at a small batch (for ncu captures).
  python tools/one_forward.py [workload] [batch] [precision]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gnn_uds_b200 import configs, synthetic  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
prec = sys.argv[3] if len(sys.argv) > 3 else "bf16"
em, args, rng = configs.make_emulator(wl, precision=prec)
inp = synthetic.make_inputs(em, min(B, 64), rng)
rep = (B + 63) // 64
inp = [None if t is None else np.ascontiguousarray(np.concatenate([t] * rep, 0)[:B]) for t in inp]
y, ey = em.predict(*inp)
print("ok", y.shape, float(np.abs(y).mean()), em.launch_count())
