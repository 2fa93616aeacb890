"""GPU diagnostic for csrc/umma.cuh descriptors: both LBO/SBO conventions, both operand views."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from test_umma_gpu import run, bf16_round

em, _ = cases.setup_product("c3_ast_act_5")
rng = np.random.default_rng(0)
for swap in (0,):
    for (Ca, Nout, shift, Ra) in [(16, 16, 0, 128), (64, 64, 0, 128), (192, 128, 0, 136), (128, 64, 5, 144)]:
        act = rng.standard_normal((Ra, Ca)).astype(np.float32)
        mat = rng.standard_normal((Nout, Ca)).astype(np.float32)
        got = run(em, 0, act, mat, shift, swap)
        ref = bf16_round(act[shift:shift + 128]).astype(np.float64) @ bf16_round(mat).astype(np.float64).T
        print("swap", swap, "mode0", (Ca, Nout, shift, Ra), "err", np.abs(got - ref).max() / np.abs(ref).max(), flush=True)
    for (Km, Ca, Ra) in [(16, 16, 128), (128, 64, 128), (144, 128, 150)]:
        act = rng.standard_normal((Ra, Ca)).astype(np.float32)
        mat = rng.standard_normal((128, Km)).astype(np.float32)
        got = run(em, 1, act, mat, 0, swap)
        ref = bf16_round(mat).astype(np.float64) @ bf16_round(act[:Km]).astype(np.float64)
        print("swap", swap, "mode1", (Km, Ca, Ra), "err", np.abs(got - ref).max() / np.abs(ref).max(), flush=True)
