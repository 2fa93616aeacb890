"""Sharding of scenarios / MPC candidates / events over ranks (one process per GPU).

Every scenario of a rollout is independent (batch axis of ``predict``; ``mpc.py:237-240`` tiles one
state over the population; ``main.py:306`` loops events), so ranks take contiguous slices and there
is no inter-GPU traffic while a rollout runs.  The only collective is one all-gather of the
per-candidate objectives (4 KiB per rank at 8192 candidates over 8 GPUs): NCCL over NVLink on the
GPU box, gloo in the CPU tests.  torch.distributed is plumbing here; no kernel depends on it.
"""
import numpy as np


def world():
    """(rank, world_size) of the default process group, (0, 1) when not initialised."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except Exception:
        pass
    return 0, 1


def backend():
    """Backend name of the default process group ('nccl', 'gloo'), None when not initialised."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_backend()
    except Exception:
        pass
    return None


def shard_range(n, rank, world_size):
    """Contiguous slice [lo, hi) of n items for `rank`: ceil(n/G) per rank."""
    per = (n + world_size - 1) // world_size
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def balance_events(lengths, world_size):
    """Whole events to ranks, balanced by length (longest first onto the lightest rank).
    Returns a list of index lists, one per rank (C4: replay of the 148 shunqing events)."""
    order = np.argsort(-np.asarray(lengths), kind="stable")
    load = [0] * world_size
    out = [[] for _ in range(world_size)]
    for i in order:
        r = int(np.argmin(load))
        out[r].append(int(i))
        load[r] += int(lengths[i])
    return [sorted(o) for o in out]


def gather_objectives(local_obj, n_total, device=None):
    """All-gather the per-candidate objectives of every rank's slice -> (n_total,) float32 on every rank.
    `local_obj` is this rank's slice in shard_range order (NumPy).  Single process: returned as is."""
    import torch
    import torch.distributed as dist
    rank, ws = world()
    local_obj = np.ascontiguousarray(local_obj, dtype=np.float32)
    if ws == 1:
        assert local_obj.shape[0] == n_total
        return local_obj
    per = (n_total + ws - 1) // ws
    use_cuda = dist.get_backend() == "nccl"
    dev = torch.device("cuda", torch.cuda.current_device() if device is None else device) if use_cuda else torch.device("cpu")
    buf = torch.zeros(per, dtype=torch.float32, device=dev)
    buf[: local_obj.shape[0]] = torch.from_numpy(local_obj).to(dev)
    out = torch.empty(per * ws, dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(out, buf)
    out = out.cpu().numpy().reshape(ws, per)
    parts = []
    for r in range(ws):
        lo, hi = shard_range(n_total, r, ws)
        parts.append(out[r, : hi - lo])
    return np.concatenate(parts)


def gather_objectives_device(local_buf, n_total):
    """Same all-gather for objectives that already sit in a padded device tensor of ceil(n_total/ws) floats (written by
    gnnuds_evaluate): the shards stay on the device through the collective, ONE device->host copy at the end."""
    import torch
    import torch.distributed as dist
    rank, ws = world()
    per = (n_total + ws - 1) // ws
    assert local_buf.numel() == per
    out = torch.empty(per * ws, dtype=torch.float32, device=local_buf.device)
    dist.all_gather_into_tensor(out, local_buf)
    out = out.cpu().numpy().reshape(ws, per)
    return np.concatenate([out[r, : shard_range(n_total, r, ws)[1] - shard_range(n_total, r, ws)[0]] for r in range(ws)])
