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
