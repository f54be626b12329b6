"""LoudnessEncoder -- drop-in for model/autoencoder/encoder.py:131-156 on the CUDA path.
The reference builds its A-weighting table with librosa (encoder.py:138), which is not installable here; the table is
computed from the IEC 61672 formula librosa implements (min_db = -80)."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from . import ops


def a_weighting_db(freqs, min_db: float = -80.0) -> np.ndarray:
    f2 = np.asarray(freqs, dtype=np.float64) ** 2
    k = np.square(np.array([12194.217, 20.598997, 107.65265, 737.86223]))
    with np.errstate(divide="ignore"):
        gain = (np.log10(k[0]) + 2 * np.log10(f2) - np.log10(f2 + k[0]) - np.log10(f2 + k[1])
                - 0.5 * (np.log10(f2 + k[2]) + np.log10(f2 + k[3])))
    return np.maximum(min_db, 2.0 + 20.0 * gain)


class LoudnessEncoder(nn.Module):
    def __init__(self, conf):
        super().__init__()
        self.n_fft, self.hop_length, self.sample_rate = conf.n_fft, conf.hop_length, conf.sample_rate
        freqs = np.linspace(0, float(self.sample_rate) / 2, int(1 + self.n_fft // 2), endpoint=True, dtype="float32")
        self.a_weight = nn.Parameter(torch.from_numpy(a_weighting_db(freqs).astype("float32")), requires_grad=False)

    @torch.no_grad()
    def forward(self, signal):
        return ops.loudness(signal, self.a_weight, self.n_fft, self.hop_length)
