"""Host-side batch split for multi-GPU runs: messages are independent, so a batch is cut
into contiguous message ranges of (nearly) equal byte weight, one per device, with no
collective on the data path.  torch.distributed is used only for the
barrier / max-over-ranks plumbing of measurements."""
from __future__ import annotations

import numpy as np


def split_by_bytes(lengths: np.ndarray, parts: int) -> np.ndarray:
    """Boundaries b[0..parts] such that shard r is messages [b[r], b[r+1]) and every shard's
    byte total is as close as a contiguous cut allows to total/parts.  Message order is kept,
    so results concatenate back in rank order."""
    n = len(lengths)
    w = lengths.astype(np.uint64) + np.uint64(16)  # per-message fixed cost keeps empty bodies spread out
    csum = np.concatenate([[0], np.cumsum(w)]).astype(np.float64)
    total = csum[-1]
    b = np.zeros(parts + 1, dtype=np.int64)
    b[parts] = n
    for r in range(1, parts):
        b[r] = int(np.searchsorted(csum, total * r / parts, side="left"))
        b[r] = min(max(b[r], b[r - 1]), n)
    return b


def shard_for_rank(lengths: np.ndarray, rank: int, world: int):
    b = split_by_bytes(lengths, world)
    return int(b[rank]), int(b[rank + 1])


def max_over_ranks(value: float, device=None) -> float:
    """Max of a per-rank scalar (e.g. device milliseconds); identity when not distributed."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_throughput(units_this_rank: float, seconds_this_rank: float, device=None) -> float:
    """Whole-job rate: units processed by all ranks / the slowest rank's time."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return units_this_rank / seconds_this_rank
    u = torch.tensor([units_this_rank], dtype=torch.float64, device=device)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(u.item()) / max_over_ranks(seconds_this_rank, device)
