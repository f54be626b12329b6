"""Voice (batch) sharding helpers.  Every op of the path is independent across voices, so the only
multi-GPU plumbing is: which contiguous voice range a rank owns, the global index of its first voice (the Philox
counter of the noise synthesiser is keyed by it, so any sharding draws the same noise), and max-over-ranks timing.

Equal voice COUNTS are not equal WORK: the oscillator kernels skip harmonics above Nyquist, so an 80 Hz voice at 48 kHz
carries 256 active harmonics and a 1 kHz voice 24.  `voice_work` is the kernels' cost model per voice and
`balanced_voice_shards` cuts the voice axis into contiguous ranges of (min-max optimal) equal work; the ranges stay
contiguous, so `voice_offset = lo` keeps every voice's Philox key and the results do not change bit for bit."""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def voice_shard(n_voices: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous, balanced [lo, hi) range of voices for `rank`; the first n_voices % world ranks get one extra."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(n_voices, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def voice_work(f0: torch.Tensor, n_harmonics: int, sample_rate: int, hop: int, per_sample_cost: float = 55.0) -> torch.Tensor:
    """f0 [B,T] or [B,T,1] (Hz) -> [B] float64 relative cost of one voice through oscillator + noise, forward + backward.

    Oscillator: harmonic k of frame t is active when k*f0 <= sample_rate//2 (harmonic_oscillator.py:31); a hop segment
    evaluates the larger active count of its two frames for each of its samples (half segments at both ends).  Everything
    that does not depend on pitch -- the noise kernels (M + M*NB/hop flop per sample each way), prepare/finalize, and the
    oscillator kernels' own per-segment work (seeds, reductions, stores: about 37 harmonics' worth) -- is `per_sample_cost`
    harmonic-sample equivalents per audio sample.  Fitted on B200 at cfg 4 by least squares over eleven shards of the 8-rank
    job timed alone on one GPU (t = a * harmonic-samples + b * voices: a = 2.89e-13 s, b = 3.06 us per voice => 55;
    residuals within 1.2 %; profiles/summary_r02.md section 11)."""
    f = f0.detach().reshape(f0.shape[0], f0.shape[1]).to(torch.float32).cpu()
    B, T = f.shape
    nyq = float(sample_rate // 2)
    kact = torch.clamp(torch.floor(nyq / f.clamp_min(1e-3)), max=float(n_harmonics))
    # floor(nyq/f0) can be off by one against the kernels' k*f0 <= nyq test in fp32; settle it the way they do
    kact = kact + ((kact + 1) * f <= nyq).float() * (kact < n_harmonics).float() - (kact * f > nyq).float()
    kact = kact.clamp(0, n_harmonics).double()
    left = torch.cat([kact[:, :1], kact], dim=1)
    right = torch.cat([kact, kact[:, -1:]], dim=1)
    n = torch.full((T + 1,), float(hop), dtype=torch.float64)
    n[0] = n[-1] = hop / 2
    return (torch.maximum(left, right) * n).sum(dim=1) + per_sample_cost * T * hop


def balanced_voice_shards(work: Sequence[float], world: int) -> List[Tuple[int, int]]:
    """Contiguous [lo, hi) voice ranges, one per rank, minimising the largest rank's total work (linear-partition
    problem: bisection on the bottleneck with a greedy feasibility sweep).  Every rank gets at least one voice when there
    are at least `world` voices."""
    w = [float(x) for x in work]
    n = len(w)
    if world < 1:
        raise ValueError("world must be positive")
    if n < world:
        return [(min(r, n), min(r + 1, n)) for r in range(world)]

    def cuts(limit: float):
        out, lo, acc = [], 0, 0.0
        for i, x in enumerate(w):
            remaining_ranks = world - len(out) - 1              # ranks still to be opened after the current one
            if i > lo and (acc + x > limit or n - i <= remaining_ranks):
                out.append((lo, i))
                lo, acc = i, 0.0
            acc += x
        out.append((lo, n))
        return out

    lo_b, hi_b = max(w), sum(w)
    for _ in range(60):
        mid = 0.5 * (lo_b + hi_b)
        if len(cuts(mid)) <= world:
            hi_b = mid
        else:
            lo_b = mid
    ranges = cuts(hi_b)
    while len(ranges) < world:                                   # fewer cuts than ranks: split the heaviest splittable range
        j = max((r for r in range(len(ranges)) if ranges[r][1] - ranges[r][0] > 1),
                key=lambda r: sum(w[ranges[r][0]:ranges[r][1]]))
        a, b = ranges[j]
        half, acc, k = 0.5 * sum(w[a:b]), 0.0, a
        while k < b - 1 and acc + w[k] <= half:
            acc += w[k]
            k += 1
        k = max(k, a + 1)
        ranges[j:j + 1] = [(a, k), (k, b)]
    return ranges


def max_over_ranks(value: float, device=None) -> float:
    """Slowest rank's time (the number a multi-GPU measurement reports)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_voices(local: torch.Tensor, n_voices: int, ranges: Sequence[Tuple[int, int]] = None) -> torch.Tensor:
    """All ranks' voice shards concatenated in rank order (`ranges`: the shards in use; default the equal-count split)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    sizes = list(ranges) if ranges is not None else [voice_shard(n_voices, world, r) for r in range(world)]
    pad = max(hi - lo for lo, hi in sizes)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, sizes)], dim=0)
