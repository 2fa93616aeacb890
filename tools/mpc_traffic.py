"""Three MPC population evaluations (gnnuds_evaluate) of `pop` candidates x 12 chunks for an ncu pass that sums the DRAM
traffic of the last one:  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv ...
  python tools/mpc_traffic.py [pop]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gnn_uds_b200 import configs, mpc, synthetic  # noqa: E402

pop = int(sys.argv[1]) if len(sys.argv) > 1 else 256
em, args, rng = configs.make_emulator("C3b", precision="bf16")
net, _ = configs.load_net("astlingen")
T = em.seq_out * configs.WORKLOADS["C3b"]["chunks"]
ev = mpc.MpcEvaluator(em, net.nodes, interval=1, eval_horizon=T, control_horizon=T, setting_duration=5, gamma=0.99 ** np.arange(T))
rng = np.random.default_rng(1)
states, b, _, ex = synthetic.make_inputs(em, 1, rng, horizon=T)
ev.load_state(states[0], b[0], ex[0])
x = mpc.latin_hypercube(pop, np.full(ev.n_var, 0.05), np.ones(ev.n_var), rng)
l0 = em.launch_count()
for i in range(3):
    F = ev.pred_emu_device(x)
    print("eval", i, "launches so far", em.launch_count() - l0, float(F.sum()))
