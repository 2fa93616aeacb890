"""Diagnostic: tensor-core (bf16) path against the float64 oracle, stage by stage and end to end.
Run on the GPU box:  timeout 300 python tools/tc_diag.py [case ...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402

STAGES = [(1, "embed"), (2, "sp1"), (3, "tp1"), (4, "sp2"), (5, "tp2"), (6, "res")]


def sc(t):
    return np.maximum(np.abs(t).reshape(-1, t.shape[-1]).max(0), 1.0)


def main():
    names = sys.argv[1:] or ["c2_shq_gat_conv_10", "c1_ast_gcn_gru", "c3_ast_act_5", "ast_gat_gru_noflood",
                             "ast_gcn_lstm_act_nofusion", "ast_gcn_conv_eps0"]
    for case in names:
        print("==", case, flush=True)
        em, inp = cases.setup_product(case, precision="bf16")
        o, inp_o = cases.make_oracle(case, np.float64)
        o.trace = {}
        y64, ey64 = o.predict(*inp_o)
        for stage, key in STAGES:
            xs, es = em.debug_stage(stage, *inp)
            tx, te = o.trace[key]
            for nm, got, tr in (("node", xs, tx), ("edge", es, te)):
                tr = tr.reshape(-1, tr.shape[-1])[:, :got.shape[1]]
                d = np.abs(got - tr)
                print("  %-5s %-4s max|d| %.4g  rms d %.4g  rms truth %.4g  finite %s" % (
                    key, nm, d.max(), np.sqrt((d ** 2).mean()), np.sqrt((tr ** 2).mean()), np.isfinite(got).all()), flush=True)
        y, ey = em.predict(*inp)
        dy = np.abs(y - y64) / sc(y64)
        dey = np.abs(ey - ey64) / sc(ey64)
        print("  predict: y max scaled err per channel", dy.reshape(-1, y.shape[-1]).max(0))
        print("           ey max scaled err per channel", dey.reshape(-1, ey.shape[-1]).max(0))
        rmse = np.sqrt(((y - y64) ** 2).reshape(-1, y.shape[-1]).mean(0))
        print("           y rmse per channel", rmse, " launches", em.launch_count(), flush=True)
        if em.if_flood:
            print("           flood flag flips", ((y[..., 3] > 0.5) != (y64[..., 3] > 0.5)).mean())


if __name__ == "__main__":
    main()
