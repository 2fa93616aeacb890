"""MPC population evaluator: the caller-side of the hot path (R4/R5 in SURVEY 8a).

Mirrors ``surrogate/mpc.py::mpc_problem`` (``load_state`` :186, ``pred_emu`` :210-233,
``pred_emu_tf`` :235-259, ``_evaluate`` :261-271) and the chunked autoregressive loop of
``mpc_problem_gr.predict`` (:565-582) without pymoo/TF: ``x`` is the (pop, n_step*n_act) decision
matrix a pymoo ``Problem._evaluate`` receives, the result is ``out['F']`` (objective + 1e-6).
Candidates are sharded over ranks (``sharding.shard_range``); the per-candidate objectives are
all-gathered once.  All arithmetic (rollout + objective reduction) runs in libgnnuds.so.
"""
import numpy as np

from . import sharding

# envs/config/astlingen.yaml:165-237 performance_targets (node, attribute, weight)
ASTLINGEN_TARGETS = (
    [("T%d" % i, "cumflooding", 2 if i == 6 else 1) for i in range(1, 7)]
    + [("CSO7", "cumflooding", 2), ("CSO8", "cumflooding", 1), ("CSO9", "cumflooding", 2), ("CSO10", "cumflooding", 1),
       ("Out_to_WWTP", "cuminflow", -0.1)]
    + [(n, "cuminflow", 0.01) for n in ("J19", "J17", "J14", "J6", "J2", "J9")])


def target_weights(node_names, targets=ASTLINGEN_TARGETS):
    """Per-node weight vectors of astlingen.objective_pred(_tf) (envs/scenario/astlingen.py:57-99):
    flooding (on q_w), WWTP inflow (on q_in[1:]) and roughness (on |diff q_in|)."""
    n = len(node_names)
    fw, ww, rw = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.float32)
    for name, attr, w in targets:
        i = node_names.index(name)
        if attr == "cumflooding":
            fw[i] += w
        elif "WWTP" in name:
            ww[i] += w
        else:
            rw[i] += w
    return fw, ww, rw


class MpcEvaluator:
    """args: interval, eval_horizon, control_horizon, setting_duration (minutes, as ``utils/mpc.yaml``),
    optional stochastic, gamma, rough.  emul: a gnn_uds_b200 Emulator with act + if_flood."""

    def __init__(self, emul, node_names, interval=1, eval_horizon=60, control_horizon=60, setting_duration=5,
                 stochastic=0, gamma=None, rough=True, targets=ASTLINGEN_TARGETS):
        self.emul = emul
        self.step = interval
        self.eval_hrz = eval_horizon
        self.n_step = control_horizon // setting_duration          # mpc.py:154
        self.r_step = setting_duration // interval                 # mpc.py:155
        self.n_act = emul.n_act
        self.n_var = self.n_act * self.n_step
        self.stochastic = int(stochastic or 0)
        self.gamma, self.rough = gamma, rough
        self.fw, self.ww, self.rw = target_weights(list(node_names), targets)

    def load_state(self, state, runoff, edge_state):               # mpc.py:186-187
        self.state, self.runoff, self.edge_state = state, runoff, edge_state

    def expand_settings(self, y):
        """mpc.py:211-218: (pop, n_step*n_act) -> (pop[*stochastic], eval_hrz/step, n_act)."""
        settings = np.asarray(y, np.float32).reshape((-1, self.n_step, self.n_act))
        settings = np.repeat(settings, self.r_step, axis=1)
        T = self.eval_hrz // self.step
        if settings.shape[1] < T:
            settings = np.concatenate([settings, np.repeat(settings[:, -1:, :], T - settings.shape[1], axis=1)], axis=1)
        if self.stochastic:
            settings = np.repeat(settings, self.stochastic, axis=0)
        return settings

    def pred_emu(self, y):
        """Objectives of ALL candidates on every rank (this rank evaluates its slice only)."""
        settings = self.expand_settings(y)
        pop = settings.shape[0] // max(self.stochastic, 1)
        rank, ws = sharding.world()
        lo, hi = sharding.shard_range(pop, rank, ws)
        k = max(self.stochastic, 1)
        sett = settings[lo * k: hi * k]
        nb = sett.shape[0]
        if self.stochastic:                                         # runoff (stochastic, T, N, b_in): tiled per candidate
            runoff = np.tile(self.runoff, (hi - lo,) + (1,) * (self.runoff.ndim - 1))
        else:
            runoff = np.repeat(np.expand_dims(self.runoff, 0), nb, axis=0)
        state = np.repeat(np.expand_dims(self.state, 0), nb, axis=0)
        edge_state = np.repeat(np.expand_dims(self.edge_state, 0), nb, axis=0)
        T = settings.shape[1]
        em = self.emul
        if nb == 0:
            objs = np.zeros(0, np.float32)
        else:
            if T == em.seq_out:                                     # one forward (mpc.py:226 / :250)
                preds = em.predict(state, runoff, sett, edge_state)[0]
            else:                                                   # chunked AR loop (mpc.py:565-582)
                preds = em.rollout(state, runoff, sett, edge_state, rule="mpc")[0]
            gamma = None if self.gamma is None else np.asarray(self.gamma, np.float32)
            objs = em.objective(preds, state, self.fw, self.ww, self.rw if self.rough else None, gamma)
            if self.stochastic:
                objs = objs.reshape(hi - lo, self.stochastic).mean(axis=1)
        return sharding.gather_objectives(objs, pop, device=em.device)

    def pred_emu_device(self, y):
        """pred_emu with nothing but the decision matrix and the objectives crossing PCIe: the shared state / runoff are
        broadcast on the device, the chunk loop and the objective reduction run in gnnuds_evaluate, and (NCCL) the
        per-candidate objectives go from the kernel's output buffer straight into the all-gather; ONE device->host
        copy of the gathered vector at the end.  Bitwise equal to pred_emu on the fp32 path."""
        dec = np.asarray(y, np.float32).reshape((-1, self.n_step, self.n_act))
        pop = dec.shape[0]
        rank, ws = sharding.world()
        lo, hi = sharding.shard_range(pop, rank, ws)
        k = max(self.stochastic, 1)
        mine = np.repeat(dec[lo:hi], k, axis=0) if self.stochastic else dec[lo:hi]
        T = self.eval_hrz // self.step
        em = self.emul
        runoff = np.asarray(self.runoff, np.float32)                # (T,N,b_in) or (stochastic,T,N,b_in)
        gamma = None if self.gamma is None else np.asarray(self.gamma, np.float32)
        rw = self.rw if self.rough else None
        if ws > 1 and sharding.backend() == "nccl" and not self.stochastic:
            import torch
            per = (pop + ws - 1) // ws
            buf = torch.zeros(per, dtype=torch.float32, device=torch.device("cuda", em.device))
            if hi > lo:
                em.evaluate(self.state, runoff, self.edge_state, mine, self.r_step, T, self.fw, self.ww, rw, gamma,
                            obj_ptr=buf.data_ptr())
            return sharding.gather_objectives_device(buf, pop)
        objs = np.zeros(0, np.float32)
        if hi > lo:
            objs = em.evaluate(self.state, runoff, self.edge_state, mine, self.r_step, T, self.fw, self.ww, rw, gamma)
            if self.stochastic:
                objs = objs.reshape(hi - lo, self.stochastic).mean(axis=1)
        return sharding.gather_objectives(objs, pop, device=em.device)

    # ---- gradient MPC (mpc.py:511-627, mpc_problem_gr) ----
    def collapse_settings_grad(self, ga):
        """Transpose of ``expand_settings`` (mpc.py:211-218 / :550-557): d/d settings (pop, T, n_act) -> d/d y
        (pop, n_step*n_act).  Every decision is held r_step steps and the last one to the end of the horizon, so its
        gradient is the sum over the steps it is held."""
        ga = np.asarray(ga, np.float64)
        pop, T = ga.shape[0], ga.shape[1]
        gy = np.zeros((pop, self.n_step, self.n_act))
        for i in range(self.n_step):
            lo = i * self.r_step
            hi = T if i == self.n_step - 1 else min((i + 1) * self.r_step, T)
            if lo < T:
                gy[:, i] = ga[:, lo:hi].sum(axis=1)
        return gy.reshape(pop, self.n_var)

    def project(self, y, xl=0.0, xu=1.0, inverse=False):            # mpc.py:619-625
        y = np.asarray(y, np.float64)
        if inverse:
            y = (y - xl) / (xu - xl)
            return np.log(y / (1 - y + 1e-6))
        return 1.0 / (1.0 + np.exp(-y)) * (xu - xl) + xl

    def gradient(self, y, bounded=True, xl=0.0, xu=1.0):
        """``gradient_fn`` (mpc.py:596-607): (objs (pop,), d obj / d y (pop, n_var)) for decision vectors y
        (n_var,) or (pop, n_var); with ``bounded=False`` y lives in the unconstrained space of ``project``
        (sigmoid, mpc.py:601-602).  The rollout, the objective and the whole reverse pass run in
        ``gnnuds_grad_settings``; only the chain rule of the setting expansion / projection is done here."""
        if self.stochastic:
            raise NotImplementedError("gradient with stochastic runoff scenarios")
        y = np.atleast_2d(np.asarray(y, np.float64))
        yy = y if bounded else self.project(y, xl, xu)
        sett = self.expand_settings(yy.astype(np.float32))
        pop = sett.shape[0]
        state = np.repeat(np.expand_dims(self.state, 0), pop, axis=0)
        edge_state = np.repeat(np.expand_dims(self.edge_state, 0), pop, axis=0)
        runoff = np.repeat(np.expand_dims(self.runoff, 0), pop, axis=0)
        gamma = None if self.gamma is None else np.asarray(self.gamma, np.float32)
        objs, ga = self.emul.grad_settings(state, runoff, sett, edge_state, self.fw, self.ww,
                                           self.rw if self.rough else None, gamma)
        gy = self.collapse_settings_grad(ga)
        if not bounded:
            sg = (yy - xl) / (xu - xl)
            gy = gy * sg * (1 - sg) * (xu - xl)
        return objs, gy

    def hessp(self, y, p, eps=1e-3, **kw):
        """``hessp_fn`` (mpc.py:616-624) by central differences of ``gradient`` along p (the reference differentiates the
        gradient a second time with a nested tape; the rollout is piecewise linear in long stretches - relu,
        clipped flows - so a difference quotient of the exact gradient is what a second tape returns up to O(eps^2))."""
        y, p = np.asarray(y, np.float64), np.asarray(p, np.float64)
        y2, p2 = np.atleast_2d(y), np.atleast_2d(p)
        n = y2.shape[0]
        _, g = self.gradient(np.concatenate([y2 + eps * p2, y2 - eps * p2], axis=0), **kw)    # both points in ONE reverse pass
        return ((g[:n] - g[n:]) / (2 * eps)).reshape(np.shape(y))

    def _evaluate(self, x, out, *args, **kwargs):                   # mpc.py:261-264
        out["F"] = self.pred_emu_device(x) + 1e-6

    def pred(self, x):                                              # mpc.py:273-276
        return self.pred_emu_device(x) + 1e-6


def latin_hypercube(pop, lo, hi, rng):
    """pymoo is absent: candidates by Latin-hypercube sampling in [lo, hi] per variable."""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    n_var = lo.shape[0]
    u = (rng.permuted(np.tile(np.arange(pop), (n_var, 1)), axis=1).T + rng.uniform(0, 1, (pop, n_var))) / pop
    return (lo + u * (hi - lo)).astype(np.float32)
