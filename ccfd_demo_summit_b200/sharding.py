"""Row sharding for multi-GPU scoring.

Rows are independent, so a batch is cut into contiguous row ranges, one per rank, and every rank
scores its range on its own GPU with a full forest replica.  There is no exchange step on the
data path; the optional gather below only exists for callers that want the whole result on every
rank (16 bytes per row over NCCL/NVLink, or gloo on CPU-only test runs).
"""
from __future__ import annotations

import numpy as np

ALIGN_ROWS = 32  # shard starts are multiples of 32 rows: keeps each shard's first row 16-byte aligned


def shard_bounds(n: int, world: int, align: int = ALIGN_ROWS) -> np.ndarray:
    """world+1 row offsets: contiguous, covering [0, n), sizes differing by at most `align`."""
    if n < 0 or world < 1:
        raise ValueError("n >= 0 and world >= 1 required")
    blocks = -(-n // align)
    cuts = [min(n, (blocks * r // world) * align) for r in range(world)] + [n]
    return np.asarray(cuts, dtype=np.int64)


def shard_range(n: int, world: int, rank: int, align: int = ALIGN_ROWS) -> tuple[int, int]:
    b = shard_bounds(n, world, align)
    return int(b[rank]), int(b[rank + 1])


def gather_rows(local, n: int, world: int, group=None):
    """All-gather per-rank (rows, k) tensors of a shard_bounds partition into one (n, k) tensor."""
    import torch
    import torch.distributed as dist

    if world == 1:
        return local
    b = shard_bounds(n, world)
    sizes = (b[1:] - b[:-1]).tolist()
    pad = max(sizes)
    buf = local.new_zeros((pad,) + tuple(local.shape[1:]))
    buf[: local.shape[0]] = local
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)], dim=0)
