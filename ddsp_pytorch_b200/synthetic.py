"""Seeded synthetic controls for benches and examples (SURVEY.md §8d shapes).  Host-side input generation only: voice
``b`` depends on ``seed`` and its GLOBAL index ``voice_offset + b``, so every sharding of a batch sees the same voices.
Draws the same values as the test oracle's generator (tests/test_sharding_gloo.py compares them)."""
from __future__ import annotations

import math

import torch


def _msig(z: torch.Tensor) -> torch.Tensor:            # decoder.py:110-116: the range the Controller's outputs live in
    return torch.sigmoid(z).pow(2.3026) * 2.0 + 1e-7


def synthetic_controls(B: int, T: int, K: int, NB: int, sample_rate: int, hop: int, seed: int, voice_offset: int = 0,
                       dtype=torch.float32):
    """-> f0 [B,T,1] (Hz: 80..1000 base with 5 Hz vibrato and jitter), c [B,T,K], a [B,T,1], H [B,T,NB] on the CPU."""
    f0 = torch.empty(B, T, 1, dtype=torch.float64)
    c = torch.empty(B, T, K, dtype=torch.float64)
    a = torch.empty(B, T, 1, dtype=torch.float64)
    H = torch.empty(B, T, NB, dtype=torch.float64)
    t = torch.arange(T, dtype=torch.float64) * hop / sample_rate
    for b in range(B):
        g = torch.Generator().manual_seed(seed * 1_000_003 + voice_offset + b)
        base = 80.0 * (1000.0 / 80.0) ** torch.rand(1, generator=g, dtype=torch.float64)
        f = base * (1.0 + 0.01 * torch.sin(2.0 * math.pi * 5.0 * t)) + torch.randn(T, generator=g, dtype=torch.float64)
        f0[b, :, 0] = f.clamp(20.0, sample_rate / 2 - 1.0)
        c[b] = _msig(torch.randn(T, K, generator=g, dtype=torch.float64))
        a[b] = _msig(torch.randn(T, 1, generator=g, dtype=torch.float64))
        H[b] = _msig(torch.randn(T, NB, generator=g, dtype=torch.float64))
    return f0.to(dtype), c.to(dtype), a.to(dtype), H.to(dtype)


def synthetic_f0(B: int, T: int, sample_rate: int, hop: int, seed: int, voice_offset: int = 0, dtype=torch.float32):
    """Only the f0 tracks of `synthetic_controls` (same draws: the pitch is the first thing each voice's generator
    produces) -> [B,T,1].  Cheap enough for every rank to evaluate for ALL voices when it balances the shards."""
    f0 = torch.empty(B, T, 1, dtype=torch.float64)
    t = torch.arange(T, dtype=torch.float64) * hop / sample_rate
    for b in range(B):
        g = torch.Generator().manual_seed(seed * 1_000_003 + voice_offset + b)
        base = 80.0 * (1000.0 / 80.0) ** torch.rand(1, generator=g, dtype=torch.float64)
        f = base * (1.0 + 0.01 * torch.sin(2.0 * math.pi * 5.0 * t)) + torch.randn(T, generator=g, dtype=torch.float64)
        f0[b, :, 0] = f.clamp(20.0, sample_rate / 2 - 1.0)
    return f0.to(dtype)
