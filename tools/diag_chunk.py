import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, cases
em, inp = cases.setup_product("c3_ast_act_5", precision="bf16")
y, ey = em.predict(*inp)
for mc in (4, 2, 1):
    em2, _ = cases.setup_product("c3_ast_act_5", precision="bf16", max_chunk=mc)
    y2, ey2 = em2.predict(*inp)
    d = np.abs(y - y2); de = np.abs(ey - ey2)
    print("max_chunk", mc, "nan", np.isnan(y2).sum(), np.isnan(ey2).sum(), "maxdiff", np.nanmax(d), np.nanmax(de), "where", np.argwhere(d > 0)[:6].tolist(), np.argwhere(de > 0)[:6].tolist())
