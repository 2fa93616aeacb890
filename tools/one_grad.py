"""Two MpcEvaluator.gradient calls of the C3grad workload at a small population (for ncu launch lists; run with
GNNUDS_GRAPHS=0 so every kernel is a stream launch).  python tools/one_grad.py [pop]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gnn_uds_b200 import configs, synthetic, mpc  # noqa: E402

pop = int(sys.argv[1]) if len(sys.argv) > 1 else 1
em, args, rng = configs.make_emulator("C3b", precision="fp32")
net, _ = configs.load_net("astlingen")
chunks = configs.WORKLOADS["C3b"]["chunks"]
T = em.seq_out * chunks
ev = mpc.MpcEvaluator(em, net.nodes, interval=1, eval_horizon=T, control_horizon=T, setting_duration=5, gamma=0.99 ** np.arange(T))
rng = np.random.default_rng(args.seed)
states, b, _, ex = synthetic.make_inputs(em, 1, rng, horizon=T)
ev.load_state(states[0], b[0], ex[0])
y = mpc.latin_hypercube(pop, np.full(ev.n_var, 0.05), np.ones(ev.n_var), rng)
l0 = em.launch_count()
objs, g = ev.gradient(y)
l1 = em.launch_count()
objs, g = ev.gradient(y)
print("ok", objs[:2], float(np.abs(g).max()), "launches per gradient", l1 - l0)
