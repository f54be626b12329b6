"""Controller and Decoder around the synthesis path.

The Controller (decoder.py:41-116) is NOT part of the hand-written hot path: its GEMMs and GRU go through torch
(cuBLAS / cuDNN).  It is restated here only so that a training step and a real-time step can be driven without the
reference tree; parameter names and shapes follow the reference exactly, so a Lightning checkpoint of the reference
(`rt/utils.py:7-24` strips the ``model.`` prefix) loads with ``strict=True`` and vice versa.  Everything downstream
of the controller -- oscillator bank, filtered noise, reverb -- is this package's CUDA path.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops
from .modules import FilteredNoise, OscillatorBank, Reverb


def modified_sigmoid(x: torch.Tensor) -> torch.Tensor:
    """2 * sigmoid(x)^ln(10) + 1e-7   (decoder.py:110-116)."""
    return 2.0 * torch.sigmoid(x).pow(2.3026) + 1e-7


class MLP(nn.Module):
    """n_layer x (Linear -> LayerNorm -> LeakyReLU); children are named mlp_layer1.. as in decoder.py:9-38."""

    def __init__(self, n_input: int, n_units: int, n_layer: int):
        super().__init__()
        self.n_layer = n_layer
        width = n_input
        for i in range(1, n_layer + 1):
            self.add_module(f"mlp_layer{i}", nn.Sequential(nn.Linear(width, n_units), nn.LayerNorm(n_units), nn.LeakyReLU()))
            width = n_units

    def forward(self, x):
        for i in range(1, self.n_layer + 1):
            x = getattr(self, f"mlp_layer{i}")(x)
        return x


class Controller(nn.Module):
    """decoder.py:41-116.  The tail -- three Linear heads + modified_sigmoid -- runs as one fused kernel
    (`ops.controller_heads`); the parameters stay the three nn.Linear modules of the reference, so state
    dicts are unchanged.  There is no CPU fallback: CPU tensors raise.  `fused_tail=False` is an explicit opt-out that
    evaluates the tail as the reference writes it (three nn.Linear + modified_sigmoid in torch), for comparisons."""

    def __init__(self, conf, fused_tail: bool = True):
        super().__init__()
        self.config = conf
        self.fused_tail = bool(fused_tail)
        u, layers = conf.decoder_mlp_units, conf.decoder_mlp_layers
        self.mlp_f0 = MLP(1, u, layers)
        self.mlp_loudness = MLP(1, u, layers)
        self.num_mlp = 2
        self.gru = nn.GRU(input_size=2 * u, hidden_size=conf.decoder_gru_units, num_layers=conf.decoder_gru_layers,
                          batch_first=True)
        self.mlp_gru = MLP(conf.decoder_gru_units + 2 * u, u, layers)
        self.dense_harmonic = nn.Linear(u, conf.n_harmonics)
        self.dense_loudness = nn.Linear(u, 1)
        self.dense_filter = nn.Linear(u, conf.n_noise_filters)

    def forward(self, batch, hidden=None):
        lf0 = self.mlp_f0(batch["normalized_cents"])
        lld = self.mlp_loudness(batch["loudness"])
        x = torch.cat((lf0, lld), dim=-1)
        x, h = self.gru(x) if hidden is None else self.gru(x, hidden)
        x = self.mlp_gru(torch.cat((x, lf0, lld), dim=-1))
        if self.fused_tail:
            c, a, H = ops.controller_heads(x, self.dense_harmonic.weight, self.dense_harmonic.bias, self.dense_loudness.weight,
                                           self.dense_loudness.bias, self.dense_filter.weight, self.dense_filter.bias)
        else:
            c, a, H = (modified_sigmoid(head(x)) for head in (self.dense_harmonic, self.dense_loudness, self.dense_filter))
        controls = dict(f0=batch["f0"], c=c, hidden=h, H=H, a=a)
        if hidden is not None:
            return controls, hidden                     # the reference returns the INPUT hidden state here (decoder.py:106-107)
        return controls


class Decoder(nn.Module):
    """decoder.py:119-147 with the synthesis modules of this package."""

    def __init__(self, conf, noise_seed=None, voice_offset: int = 0, device_counter: bool = False, fused_tail: bool = True):
        super().__init__()
        self.controller = Controller(conf, fused_tail=fused_tail)
        self.harmonics = OscillatorBank(conf)
        self.noise = FilteredNoise(conf, seed=noise_seed, voice_offset=voice_offset, device_counter=device_counter)
        self.reverb = Reverb(conf)

    def forward(self, z):
        ctrl = self.controller(z)
        # `harmonics + noise` (decoder.py:132) without the separate add: the oscillator kernel accumulates into the noise
        dry = self.harmonics(ctrl, out=self.noise(ctrl))
        return self.reverb(dry)

    def forward_live(self, z, hidden):
        ctrl, hidden = self.controller(z, hidden)
        audio = self.reverb.live_forward(self.harmonics.live(ctrl) + self.noise(ctrl))
        return audio.cpu().squeeze(0).numpy(), hidden
