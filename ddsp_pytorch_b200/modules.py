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

    def forward(self, x, out=None):
        """`out` (an extension over harmonic_oscillator.py:57-62): a [B,T*hop] fp32 CUDA buffer the result is ADDED into and
        returned as -- the Decoder passes the filtered noise here, which removes the separate `harmonics + noise` pass."""
        c = x["c"]
        if c.shape[-1] != self.n_harmonics:
            raise RuntimeError(f"OscillatorBank: x['c'] has {c.shape[-1]} harmonics, module has {self.n_harmonics}")
        return ops.oscillator_bank(x["f0"], c, x["a"], self.sample_rate, self.hop_size, out=out)

    # ---- the reference's public helper methods (harmonic_oscillator.py:24-55; no caller outside the class) -------------
    # forward() above does NOT go through them -- the kernels never build a [B,N,K] tensor.  They are kept for code that
    # inspects the intermediate quantities, stated op for op as the reference states them on top of ops.rescale, so they
    # materialise what the reference materialises and carry the reference's own fp32 phase drift.
    def rescale(self, x):
        """[B,T,C] -> [B,T*hop,C]  (:52-55)."""
        return ops.rescale(x, self.hop_size)

    def prepare_harmonics(self, f0, harm_amps):
        """(:24-37) -> (radians per sample of every harmonic, upsampled [B,N,K]; Nyquist-masked, normalised amplitudes [B,T,K])."""
        harmonics = self.harmonics.to(f0.device).view(1, 1, -1) * f0              # int64 x fp32 -> fp32 Hz (:26-29)
        mask = harmonics > self.sample_rate // 2                                   # (:31)
        harm_amps = harm_amps.masked_fill(mask, 0.)
        harm_amps = harm_amps / harm_amps.sum(-1, keepdim=True)                    # (:33), out of place: the caller's tensor is kept
        harmonics = harmonics * (2 * torch.pi) / self.sample_rate                  # (:34-35)
        return self.rescale(harmonics), harm_amps

    @staticmethod
    def generate_phases(harmonics):
        """(:39-43) cumulative sum over samples, wrapped to [0, 2 pi)."""
        return torch.cumsum(harmonics, dim=1) % (2 * torch.pi)

    def generate_signal(self, harm_amps, loudness, phases):
        """(:45-50) sum_k a_up * c_up[k] * sin(phase[k])."""
        return torch.sum(self.rescale(loudness) * self.rescale(harm_amps) * torch.sin(phases), dim=2)


class FilteredNoise(nn.Module):
    """Drop-in for model/ddsp/filtered_noise.py:35-53.

    The reference draws ``torch.rand`` on the CPU and copies it over every call (:44-48); here the noise is a
    Philox stream generated inside the kernel, keyed by ``seed`` and a call counter, so every forward sees
    fresh noise, backward regenerates the same block, and nothing crosses PCIe.  ``voice_offset`` is the
    global index of this rank's first voice, which makes the draw independent of batch sharding.
    """

    def __init__(self, conf, seed=None, voice_offset: int = 0, device_counter: bool = False):
        super().__init__()
        self.block_size = conf.hop_length
        self.seed = int(torch.initial_seed() if seed is None else seed) & (2 ** 63 - 1)
        self.voice_offset = int(voice_offset)
        self.calls = 0
        # device_counter: keep the call counter on the device and advance it there, so that a CUDA graph that captured
        # this forward (and its backward) draws fresh noise on every replay.  Each forward keeps a snapshot of the counter
        # for its own backward, so several forwards may precede a backward.
        self.device_counter = bool(device_counter)
        if self.device_counter:
            self.register_buffer("_calls_dev", torch.zeros(1, dtype=torch.int32), persistent=False)

    def forward(self, x, noise=None):
        if self.device_counter and noise is None:
            self._calls_dev.add_(1)
            return ops.filtered_noise(x["H"], self.block_size, self.seed, 0, self.voice_offset, offset_tensor=self._calls_dev)
        offset = self.calls
        self.calls += 1
        return ops.filtered_noise(x["H"], self.block_size, self.seed, offset, self.voice_offset, noise=noise)


def _live(self, x):
    """OscillatorBank.live (harmonic_oscillator.py:64-75): carries per-harmonic phases across calls (voice 0)."""
    y, new_last = ops.oscillator_live(x["f0"], x["c"], x["a"], self.last_phases.data, self.sample_rate, self.hop_size)
    self.last_phases.data = new_last            # int64 zeros before the first call, fp32 afterwards -- as the reference
    return y


OscillatorBank.live = _live


def fft_convolve(signal: torch.Tensor, kernel: torch.Tensor) -> torch.Tensor:
    """filtered_noise.py:25-32: causal convolution along the last axis, truncated to its length n.  A single kernel
    shared by a [B,n] signal (how Reverb.forward calls it) goes through the partitioned FFT convolution; any other
    broadcastable pair of shapes (how FilteredNoise calls it: [B,T,n] x [B,T,n]) through the per-row kernel."""
    if signal.shape[-1] != kernel.shape[-1]:
        raise RuntimeError(f"fft_convolve: last dimensions differ ({signal.shape[-1]} vs {kernel.shape[-1]})")
    if signal.dim() == 2 and kernel.numel() == kernel.shape[-1]:
        return ops.causal_conv(signal, kernel)
    shape = torch.broadcast_shapes(signal.shape, kernel.shape)
    return ops.row_conv(signal.expand(shape).contiguous(), kernel.expand(shape).contiguous())


amp_to_impulse_response = ops.amp_to_impulse_response


class Reverb(nn.Module):
    """Drop-in for model/ddsp/reverb.py:8-49 (same parameters, buffers-as-Parameters and state-dict keys)."""

    def __init__(self, conf, initial_wet=0, initial_decay=5):
        super().__init__()
        self.length = conf.sample_rate
        self.sampling_rate = conf.sample_rate
        self.noise = nn.Parameter(torch.rand(self.length) * 2 - 1)
        self.decay = nn.Parameter(torch.tensor(float(initial_decay)))
        self.wet = nn.Parameter(torch.tensor(float(initial_wet)))
        t = (torch.arange(self.length) / self.sampling_rate).reshape(1, -1)
        self.t = nn.Parameter(t, requires_grad=False)
        self.buffer = nn.Parameter(torch.zeros(1, self.length), requires_grad=False)

    def build_impulse(self):
        return ops.reverb_impulse(self.noise, self.decay, self.wet, self.t)

    def forward(self, x):
        # reverb.py:34 pads (or crops) the impulse to len(x); taps at lags >= len(x) cannot reach an output,
        # so the kernel takes the impulse as it is
        return ops.causal_conv(x, self.build_impulse())

    def live_forward(self, x):
        with torch.no_grad():
            return ops.reverb_live(self.buffer.data, x, self.build_impulse())


class _HannWindowHolder(nn.Module):
    """Keeps the ``stft.window`` buffer of torchaudio's Spectrogram in the state dict (mss_loss.py:23)."""

    def __init__(self, n_fft):
        super().__init__()
        self.register_buffer("window", torch.hann_window(n_fft))


class SSSLoss(nn.Module):
    """Drop-in for loss/mss_loss.py:11-33."""

    def __init__(self, n_fft, alpha=1.0, overlap=0.75, eps=1e-7):
        super().__init__()
        self.n_fft = n_fft
        self.alpha = alpha
        self.eps = eps
        self.overlap = overlap
        self.hop_length = int(n_fft * (1 - overlap))
        self.stft = _HannWindowHolder(n_fft)

    def forward(self, x_pred, x_true):
        return ops.mss_loss(x_pred, x_true, [self.n_fft], self.alpha, self.overlap, self.eps, windows=[self.stft.window])


class MSSLoss(nn.Module):
    """Drop-in for loss/mss_loss.py:36-68: all scales are evaluated by one call into the library."""

    def __init__(self, n_ffts: list, alpha=1.0, overlap=0.75, eps=1e-7, use_reverb=True):
        super().__init__()
        self.losses = nn.ModuleList([SSSLoss(n_fft, alpha, overlap, eps) for n_fft in n_ffts])
        self.n_ffts, self.alpha, self.overlap, self.eps = list(n_ffts), alpha, overlap, eps
        self.signal_key = "audio_reverb" if use_reverb else "audio_synth"

    def forward(self, x_pred, x_true):
        if isinstance(x_pred, dict):
            x_pred = x_pred[self.signal_key]
        if isinstance(x_true, dict):
            x_true = x_true["audio"]
        return ops.mss_loss(x_pred, x_true, self.n_ffts, self.alpha, self.overlap, self.eps,
                            windows=[m.stft.window for m in self.losses])
