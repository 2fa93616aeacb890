"""CPU tests of the multi-rank host logic: shard ranges, event balancing, the objective all-gather
(gloo, world_size 2) and the MPC settings expansion (mpc.py:211-218)."""
import os
import socket
import sys

import numpy as np
import pytest

from gnn_uds_b200 import sharding, mpc


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 8, 8192, 8193):
        for ws in (1, 2, 3, 8):
            got = []
            for r in range(ws):
                lo, hi = sharding.shard_range(n, r, ws)
                assert 0 <= lo <= hi <= n
                got.extend(range(lo, hi))
            assert got == list(range(n))


def test_balance_events():
    rng = np.random.default_rng(0)
    lengths = rng.integers(420, 1500, 148)
    parts = sharding.balance_events(lengths, 8)
    assert sorted(i for p in parts for i in p) == list(range(148))
    loads = [int(lengths[p].sum()) for p in parts]
    assert max(loads) - min(loads) <= lengths.max()


def test_expand_settings_matches_reference_rule():
    class E:
        n_act, seq_out, device = 4, 5, 0
    ev = mpc.MpcEvaluator(E(), ["n%d" % i for i in range(3)], interval=1, eval_horizon=60, control_horizon=30,
                          setting_duration=5, targets=[("n0", "cumflooding", 1)])
    assert ev.n_step == 6 and ev.r_step == 5 and ev.n_var == 24
    y = np.arange(2 * 24, dtype=np.float32).reshape(2, 24)
    s = ev.expand_settings(y)
    assert s.shape == (2, 60, 4)
    ref = np.repeat(y.reshape(2, 6, 4), 5, axis=1)
    ref = np.concatenate([ref, np.repeat(ref[:, -1:], 30, axis=1)], axis=1)
    assert np.array_equal(s, ref)
    ev.stochastic = 3
    assert ev.expand_settings(y).shape == (6, 60, 4)


def test_latin_hypercube_strata():
    rng = np.random.default_rng(1)
    x = mpc.latin_hypercube(64, np.zeros(5), np.ones(5), rng)
    assert x.shape == (64, 5)
    for j in range(5):
        assert sorted(np.floor(x[:, j] * 64).astype(int).tolist()) == list(range(64))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, ws, port, n_total, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    lo, hi = sharding.shard_range(n_total, rank, ws)
    local = np.arange(lo, hi, dtype=np.float32) * 0.5 + 1.0
    out = sharding.gather_objectives(local, n_total)
    q.put((rank, out))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [10, 11])
def test_gather_objectives_gloo_world2(n_total):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = dict(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
    want = np.arange(n_total, dtype=np.float32) * 0.5 + 1.0
    for r in range(2):
        assert np.array_equal(res[r], want)


def test_collapse_settings_grad_is_transpose_of_expand():
    """MpcEvaluator.collapse_settings_grad (gradient MPC, mpc.py:550-557): <expand(y), g> == <y, collapse(g)>."""
    import numpy as np
    from types import SimpleNamespace
    from gnn_uds_b200 import mpc
    em = SimpleNamespace(n_act=3)
    for T, ch, sd in ((15, 10, 2), (10, 10, 5), (12, 12, 3), (7, 4, 2)):
        ev = mpc.MpcEvaluator(em, ["a"], interval=1, eval_horizon=T, control_horizon=ch, setting_duration=sd, targets=())
        rng = np.random.default_rng(T)
        y = rng.normal(size=(4, ev.n_var))
        g = rng.normal(size=(4, T, 3))
        s = ev.expand_settings(y)[:, :T]
        assert s.shape == (4, T, 3)
        assert np.allclose((s * g).sum(), (y * ev.collapse_settings_grad(g)).sum(), rtol=1e-5)
