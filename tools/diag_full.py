"""Where does the tensor-core path deviate most on picked C2 scenarios? (diagnostic)"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gnn_uds_b200 import configs, synthetic
from oracle.emulator_oracle import OracleEmulator
em, args, rng = configs.make_emulator("C2", precision="bf16")
base = synthetic.make_inputs(em, 64, rng)
pick = [3, 41]
sub = [None if t is None else np.ascontiguousarray(t[pick]) for t in base]
y, ey = em.predict(*sub)
em32, _, _ = configs.make_emulator("C2", precision="fp32")
y32, ey32 = em32.predict(*sub)
o = OracleEmulator(args.conv, args.resnet, args.recurrent, args, dtype=np.float64)
o.set_weights(em.weights); o.set_norm(em.norm_x, em.norm_b, em.norm_y, em.norm_r, em.norm_e)
y64, ey64 = o.predict(*sub)
sc = np.maximum(np.abs(ey64).reshape(-1, 3).max(0), 1.0)
for name, g in (("bf16", ey), ("fp32", ey32)):
    d = np.abs(g - ey64) / sc
    idx = np.unravel_index(np.argmax(d), d.shape)
    print(name, "max", d.max(), "at", idx, "got", g[idx], "truth", ey64[idx], "scale", sc, "rmse", np.sqrt((d ** 2).mean((0, 1, 2))))
    top = np.argsort(d.reshape(-1))[-8:]
    for t in top:
        i = np.unravel_index(t, d.shape); print("   ", i, d[i], g[i], ey64[i])
print("ehmax", em.ehmax[:5], "norm_e max", em.norm_e[0][:3])
