"""Multi-GPU plumbing: NMPC instances / vehicles / trajectories are independent (the reference steps its
controllers one after the other with no interaction, simulation/racing.py:230-241), so a batch is cut into
contiguous per-rank shards and there is NO data-path collective.  torch.distributed only carries the final
result gather (about 32 bytes per instance) and the timing reductions of bench.py.

Works on any backend: "nccl" (one process per GPU, NVLink) in production, "gloo" in the CPU tests.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) slice of n units owned by `rank`; sizes differ by at most one, earlier ranks larger."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world {rank}/{world_size}")
    q, r = divmod(int(n), world_size)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def shard(t, rank: int = None, world_size: int = None):
    """Slice the leading (instance) dimension of a tensor / ndarray for this rank."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    lo, hi = shard_range(t.shape[0], rank, world_size)
    return t[lo:hi]


def gather_results(local: Dict[str, torch.Tensor], n_total: int) -> Dict[str, torch.Tensor]:
    """All ranks contribute their shard of per-instance results (first actions, objective, status, iterations);
    every rank gets the full [n_total, ...] arrays back in instance order.  One all_gather per key on padded
    shards (shards differ by at most one row)."""
    rank, w = world()
    if w == 1:
        return dict(local)
    out = {}
    q, r = divmod(int(n_total), w)
    pad_rows = q + (1 if r else 0)
    for k, t in local.items():
        lo, hi = shard_range(n_total, rank, w)
        if t.shape[0] != hi - lo:
            raise ValueError(f"{k}: shard has {t.shape[0]} rows, expected {hi - lo}")
        buf = torch.zeros((pad_rows,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        buf[: t.shape[0]] = t
        parts = [torch.empty_like(buf) for _ in range(w)]
        dist.all_gather(parts, buf)
        rows = []
        for i, p in enumerate(parts):
            a, b = shard_range(n_total, i, w)
            rows.append(p[: b - a])
        out[k] = torch.cat(rows, 0)
    return out
