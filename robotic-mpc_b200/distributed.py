"""Sharding of independent problem instances over the GPUs of one box and the only collective of the path.

Instances share nothing: each rank (one process per GPU, ``torchrun``) solves a contiguous chunk of
every batch with no data-path communication; at the end the per-instance summary rows are all-gathered
(NCCL over NVLink on GPUs; the same code runs on ``gloo`` for the CPU tests)."""
from __future__ import annotations

import numpy as np


def world():
    """(rank, world_size) of the default process group, (0, 1) when torch.distributed is not initialised."""
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except ImportError:
        pass
    return 0, 1


def shard_bounds(n, world_size):
    """Contiguous, balanced split of n items: list of (begin, end) per rank; the first n % world ranks get one more."""
    base, extra = divmod(int(n), int(world_size))
    out, b = [], 0
    for r in range(world_size):
        e = b + base + (1 if r < extra else 0)
        out.append((b, e))
        b = e
    return out


def all_gather_rows(local, n_total, device=None):
    """All-gather ``local`` ([n_local, k] float64 numpy or torch) from every rank into one [n_total, k] numpy array in
    rank order.  Chunks may differ by one row (see ``shard_bounds``): rows are padded to the largest chunk."""
    import torch
    import torch.distributed as dist

    rank, ws = world()
    t = torch.as_tensor(local, dtype=torch.float64)
    if ws == 1:
        return t.cpu().numpy()
    if device is None:
        device = t.device if t.is_cuda else (torch.device("cuda", torch.cuda.current_device())
                                             if dist.get_backend() == "nccl" else torch.device("cpu"))
    bounds = shard_bounds(n_total, ws)
    nmax = max(e - b for b, e in bounds)
    k = t.shape[1]
    pad = torch.zeros((nmax, k), dtype=torch.float64, device=device)
    pad[: t.shape[0]] = t.to(device)
    out = torch.empty((ws * nmax, k), dtype=torch.float64, device=device)
    dist.all_gather_into_tensor(out, pad)
    out = out.cpu().numpy().reshape(ws, nmax, k)
    return np.concatenate([out[r, : e - b] for r, (b, e) in enumerate(bounds)], axis=0)


def reduce_group_stats(values, groups, n_groups, device=None):
    """Per-group count / mean / variance / min / max of a per-instance scalar over ALL ranks (tracking-error
    statistics per horizon for the horizon sweep).  ``values`` [n_local] float, ``groups`` [n_local] int in [0, n_groups).
    Three small all-reduces (SUM of [count, sum, sum of squares], MIN, MAX); with one rank it is a local reduction.
    Returns a dict of [n_groups] numpy arrays; empty groups get count 0 and NaN statistics."""
    import torch
    import torch.distributed as dist

    rank, ws = world()
    v = torch.as_tensor(np.asarray(values, dtype=np.float64))
    g = torch.as_tensor(np.asarray(groups, dtype=np.int64))
    if ws > 1:
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
        v, g = v.to(device), g.to(device)
    acc = torch.zeros((3, n_groups), dtype=torch.float64, device=v.device)
    acc[0].index_add_(0, g, torch.ones_like(v))
    acc[1].index_add_(0, g, v)
    acc[2].index_add_(0, g, v * v)
    lo = torch.full((n_groups,), float("inf"), dtype=torch.float64, device=v.device)
    hi = torch.full((n_groups,), float("-inf"), dtype=torch.float64, device=v.device)
    lo.scatter_reduce_(0, g, v, reduce="amin")
    hi.scatter_reduce_(0, g, v, reduce="amax")
    if ws > 1:
        dist.all_reduce(acc, op=dist.ReduceOp.SUM)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    acc, lo, hi = acc.cpu().numpy(), lo.cpu().numpy(), hi.cpu().numpy()
    n = acc[0]
    with np.errstate(invalid="ignore", divide="ignore"):
        mean = acc[1] / n
        var = np.maximum(acc[2] / n - mean * mean, 0.0)
    empty = n == 0
    for a in (mean, var, lo, hi):
        a[empty] = np.nan
    return {"count": n.astype(np.int64), "mean": mean, "var": var, "min": lo, "max": hi}
