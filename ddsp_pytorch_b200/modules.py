"""nn.Module mirror of the reference's hot-path classes: same constructor arguments, forward signatures,
attributes and state-dict keys, with the arithmetic done by libddsp_b200.so."""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops


class OscillatorBank(nn.Module):
    """Drop-in for model/ddsp/harmonic_oscillator.py:7-75."""

    def __init__(self, conf):
        super().__init__()
        self.n_harmonics = conf.n_harmonics
        self.sample_rate = conf.sample_rate
        self.hop_size = conf.hop_length
        # state-dict compatibility (harmonic_oscillator.py:15-22): both are int64, non-trainable Parameters
        self.harmonics = nn.Parameter(torch.arange(1, self.n_harmonics + 1), requires_grad=False)
        self.last_phases = nn.Parameter(torch.zeros(self.n_harmonics, dtype=torch.int64), requires_grad=False)

    def forward(self, x):
        c = x["c"]
        if c.shape[-1] != self.n_harmonics:
            raise RuntimeError(f"OscillatorBank: x['c'] has {c.shape[-1]} harmonics, module has {self.n_harmonics}")
        return ops.oscillator_bank(x["f0"], c, x["a"], self.sample_rate, self.hop_size)


class FilteredNoise(nn.Module):
    """Drop-in for model/ddsp/filtered_noise.py:35-53.

    The reference draws ``torch.rand`` on the CPU and copies it over every call (:44-48); here the noise is a
    Philox stream generated inside the kernel, keyed by ``seed`` and a call counter, so every forward sees
    fresh noise, backward regenerates the same block, and nothing crosses PCIe.  ``voice_offset`` is the
    global index of this rank's first voice, which makes the draw independent of batch sharding.
    """

    def __init__(self, conf, seed=None, voice_offset: int = 0):
        super().__init__()
        self.block_size = conf.hop_length
        self.seed = int(torch.initial_seed() if seed is None else seed) & (2 ** 63 - 1)
        self.voice_offset = int(voice_offset)
        self.calls = 0

    def forward(self, x, noise=None):
        offset = self.calls
        self.calls += 1
        return ops.filtered_noise(x["H"], self.block_size, self.seed, offset, self.voice_offset, noise=noise)
