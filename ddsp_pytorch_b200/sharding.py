"""Voice (batch) sharding helpers.  Every op of the path is independent across voices, so the only
multi-GPU plumbing is: which contiguous voice range a rank owns, the global index of its first voice (the Philox
counter of the noise synthesiser is keyed by it, so any sharding draws the same noise), and max-over-ranks timing."""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def voice_shard(n_voices: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous, balanced [lo, hi) range of voices for `rank`; the first n_voices % world ranks get one extra."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(n_voices, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def max_over_ranks(value: float, device=None) -> float:
    """Slowest rank's time (the number a multi-GPU measurement reports)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_voices(local: torch.Tensor, n_voices: int) -> torch.Tensor:
    """All ranks' voice shards concatenated in rank order (shards may differ by one voice)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    sizes = [voice_shard(n_voices, world, r) for r in range(world)]
    pad = max(hi - lo for lo, hi in sizes)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, sizes)], dim=0)
