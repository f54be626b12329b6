"""CPU oracle for the DDSP synthesis hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

This file restates, op for op, what kureta/ddsp-pytorch computes on the path
``model/ddsp/{harmonic_oscillator,filtered_noise,reverb}.py`` + ``loss/mss_loss.py``.
It is written as dtype-generic torch functions (run it in fp64 for ground truth,
in fp32 to see the reference's own drift) and uses the same torch library ops the
reference calls (``F.interpolate``, ``cumsum``, ``sin``, ``rfft``/``irfft``,
``stft``), so that timing it on host cores is a fair stand-in for the reference's
CPU path (``bench.py --impl reference`` and the ``cpu_baseline`` leg).

Who may import this: ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
cpu_baseline / ``--impl reference`` legs.  Nothing under ``ddsp_pytorch_b200/``
imports it; the product path fails loudly when the CUDA library is missing.

Parity pinning: the reference ships no tests or golden vectors, so
the oracle is pinned by *executing the reference itself* in the build container:
``tests/golden/make_golden.py`` imports ``/root/reference`` and stores its outputs
(fp64 and fp32) as fixtures under ``tests/golden/``; ``tests/test_oracle.py``
checks every function here against those fixtures and against the closed-form
known answers of SURVEY.md §4.

All ``file:line`` citations are relative to the reference repository root.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

TWO_PI = 2.0 * math.pi


# ----------------------------------------------------------------------------------------
# Oscillator bank  (model/ddsp/harmonic_oscillator.py)
# ----------------------------------------------------------------------------------------
def upsample_linear(x: torch.Tensor, hop: int) -> torch.Tensor:
    """[B,T,C] -> [B,T*hop,C], half-pixel linear interpolation with edge clamp.

    Follows ``OscillatorBank.rescale`` (harmonic_oscillator.py:52-55): channels-first
    ``F.interpolate(scale_factor=hop, mode='linear')`` (align_corners=False).
    """
    y = F.interpolate(x.transpose(1, 2), scale_factor=hop, mode="linear")
    return y.transpose(1, 2)


def upsample_linear_explicit(x: torch.Tensor, hop: int) -> torch.Tensor:
    """Same map as :func:`upsample_linear`, spelled out with index arithmetic.

    out[i] = (1-w)*x[i0] + w*x[i1],  src = max((i+0.5)/hop - 0.5, 0),
    i0 = floor(src), i1 = min(i0+1, T-1), w = src - i0.   
    Kept to document the semantics the CUDA kernels implement in closed form.
    """
    B, T, C = x.shape
    i = torch.arange(T * hop, dtype=torch.float64)
    src = torch.clamp((i + 0.5) / hop - 0.5, min=0.0)
    i0 = torch.floor(src).long()
    i1 = torch.clamp(i0 + 1, max=T - 1)
    w = (src - i0.double()).to(x.dtype).view(1, -1, 1)
    return (1 - w) * x[:, i0, :] + w * x[:, i1, :]


def osc_prepare(f0: torch.Tensor, c: torch.Tensor, n_harmonics: int, sample_rate: int,
                hop: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Frame-rate harmonic set-up; returns (omega_up [B,N,K] rad/sample, c_norm [B,T,K]).

    harmonic_oscillator.py:24-37 -- k*f0 (integer k times float f0), strict mask
    ``k*f0 > sample_rate // 2`` (:31), masked amplitudes renormalised to sum 1 (:32-33;
    an all-masked frame divides 0/0 and yields NaN, reproduced here), conversion to
    radians/sample (:34-35) and upsampling (:36).
    """
    k = torch.arange(1, n_harmonics + 1, dtype=f0.dtype, device=f0.device)
    fk = k.view(1, 1, -1) * f0                                  # :26-29
    above = fk > (sample_rate // 2)                             # :31
    c_masked = torch.where(above, torch.zeros_like(c), c)       # :32
    c_norm = c_masked / c_masked.sum(dim=-1, keepdim=True)      # :33
    omega = fk * TWO_PI / sample_rate                           # :34-35
    return upsample_linear(omega, hop), c_norm                  # :36


def osc_phases(omega_up: torch.Tensor) -> torch.Tensor:
    """Running phase per harmonic, wrapped to [0, 2pi)  (harmonic_oscillator.py:39-43)."""
    return torch.remainder(torch.cumsum(omega_up, dim=1), TWO_PI)


def osc_signal(c_norm: torch.Tensor, a: torch.Tensor, phases: torch.Tensor, hop: int) -> torch.Tensor:
    """y[b,n] = sum_k a_up * c_up * sin(phi)  (harmonic_oscillator.py:45-50)."""
    a_up = upsample_linear(a, hop)
    c_up = upsample_linear(c_norm, hop)
    return (a_up * c_up * torch.sin(phases)).sum(dim=2)


def oscillator_bank(f0: torch.Tensor, c: torch.Tensor, a: torch.Tensor, n_harmonics: int,
                    sample_rate: int, hop: int) -> torch.Tensor:
    """``OscillatorBank.forward`` (harmonic_oscillator.py:57-62): f0,a [B,T,1], c [B,T,K] -> [B,T*hop]."""
    omega_up, c_norm = osc_prepare(f0, c, n_harmonics, sample_rate, hop)
    return osc_signal(c_norm, a, osc_phases(omega_up), hop)


def oscillator_bank_live(f0, c, a, last_phases, n_harmonics, sample_rate, hop):
    """``OscillatorBank.live`` (harmonic_oscillator.py:64-75).

    ``last_phases`` [K] is added to the first sample's phase increment of batch row 0
    only (:70); the new state is the wrapped phase of the last sample of row 0 (:72).
    Returns (signal [B,T*hop], new_last_phases [K]).
    """
    omega_up, c_norm = osc_prepare(f0, c, n_harmonics, sample_rate, hop)
    omega_up = omega_up.clone()
    omega_up[0, 0, :] += last_phases.to(omega_up.dtype)
    phases = osc_phases(omega_up)
    return osc_signal(c_norm, a, phases, hop), phases[0, -1, :].clone()


# ----------------------------------------------------------------------------------------
# Filtered noise  (model/ddsp/filtered_noise.py)
# ----------------------------------------------------------------------------------------
def amp_to_ir(H: torch.Tensor, target: int) -> torch.Tensor:
    """Zero-phase windowed FIR from band magnitudes: [...,NB] -> [...,target].

    filtered_noise.py:7-22 -- irfft of the real spectrum (length M = 2(NB-1)), rotate by
    M/2, periodic Hann(M), zero-pad to ``target`` and rotate back by M/2.
    """
    r = torch.fft.irfft(torch.complex(H, torch.zeros_like(H)))           # :8-10
    M = r.shape[-1]
    r = torch.roll(r, M // 2, dims=-1)                                   # :14
    r = r * torch.hann_window(M, dtype=r.dtype, device=r.device)         # :15-17
    r = F.pad(r, (0, int(target) - M))                                   # :19
    return torch.roll(r, -(M // 2), dims=-1)                             # :20


def amp_to_ir_closed_form(H: torch.Tensor, target: int) -> torch.Tensor:
    """Closed form of :func:`amp_to_ir` (SURVEY.md §4 KAT): what the noise kernel builds.

    out[0:M/2] = r[0:M/2]*w[M/2:M], out[target-M/2:target] = r[M/2:M]*w[0:M/2], zero between,
    r = irfft(H), w = periodic Hann(M).
    """
    r = torch.fft.irfft(torch.complex(H, torch.zeros_like(H)))
    M = r.shape[-1]
    w = torch.hann_window(M, dtype=r.dtype, device=r.device)
    out = torch.zeros(*H.shape[:-1], target, dtype=r.dtype, device=r.device)
    out[..., : M // 2] = r[..., : M // 2] * w[M // 2:]
    out[..., target - M // 2:] = r[..., M // 2:] * w[: M // 2]
    return out


def causal_conv(signal: torch.Tensor, kernel: torch.Tensor) -> torch.Tensor:
    """Causal linear convolution truncated to len(signal), last dim.

    ``fft_convolve`` (filtered_noise.py:25-32): signal right-padded and kernel LEFT-padded
    to twice the length, spectra multiplied, upper half of the inverse transform kept.
    Equals ``np.convolve(s, k)[:len(s)]``.
    """
    n = signal.shape[-1]
    s2 = F.pad(signal, (0, n))
    k2 = F.pad(kernel, (kernel.shape[-1], 0))
    full = torch.fft.irfft(torch.fft.rfft(s2) * torch.fft.rfft(k2))
    return full[..., full.shape[-1] // 2:]


def filtered_noise(H: torch.Tensor, hop: int, noise: torch.Tensor) -> torch.Tensor:
    """``FilteredNoise.forward`` with the uniform noise supplied by the caller.

    filtered_noise.py:40-53.  ``noise`` [B,T,hop] is already mapped to [-1,1) (the reference
    draws ``torch.rand(...)*2-1`` on the CPU generator, :44-48).  Each frame is filtered on
    its own, the tail is dropped and frames are concatenated -- no overlap-add (:50-51).
    """
    ir = amp_to_ir(H, hop)
    y = causal_conv(noise.to(ir.dtype), ir)
    return y.reshape(y.shape[0], -1)


# ----------------------------------------------------------------------------------------
# Reverb  (model/ddsp/reverb.py)
# ----------------------------------------------------------------------------------------
def reverb_ir(noise: torch.Tensor, decay: torch.Tensor, wet: torch.Tensor, sample_rate: int) -> torch.Tensor:
    """``Reverb.build_impulse`` (reverb.py:24-29): noise[L] * exp(-softplus(-decay)*t*500) * sigmoid(wet), ir[0]=1."""
    L = noise.shape[-1]
    t = (torch.arange(L, device=noise.device) / sample_rate).to(noise.dtype).view(1, -1)   # reverb.py:18-20
    env = torch.exp(-F.softplus(-decay) * t * 500.0)
    ir = noise.view(1, -1) * env * torch.sigmoid(wet)
    ir = ir.clone()
    ir[:, 0] = 1.0
    return ir


def reverb_apply(x: torch.Tensor, ir: torch.Tensor) -> torch.Tensor:
    """``Reverb.forward`` (reverb.py:31-38): pad (or crop, when N < L) the IR to N, causal conv."""
    n = x.shape[1]
    ir_n = F.pad(ir, (0, n - ir.shape[-1]))
    return causal_conv(x, ir_n)


def reverb_live(buffer: torch.Tensor, x: torch.Tensor, ir: torch.Tensor):
    """``Reverb.live_forward`` (reverb.py:40-49): slide the 1-s buffer, convolve, keep the last len(x).

    Returns (y [1,len(x)], new_buffer [1,L]).
    """
    n = x.shape[1]
    new_buf = torch.cat([buffer[:, n:], x.to(buffer.dtype)], dim=1)
    y = causal_conv(new_buf, ir)
    return y[:, -n:], new_buf


# ----------------------------------------------------------------------------------------
# Multi-scale spectral loss  (loss/mss_loss.py)
# ----------------------------------------------------------------------------------------
def power_spectrogram(x: torch.Tensor, n_fft: int, hop: int) -> torch.Tensor:
    """torchaudio ``Spectrogram(n_fft, hop_length=hop)`` defaults (mss_loss.py:23):
    centre reflect-pad n_fft/2, periodic Hann, onesided, unnormalised, power=2 -> [B, n_fft/2+1, 1+N//hop]."""
    # the window is a float32 buffer of the torchaudio module; ``module.double()`` keeps its fp32-rounded values
    win = torch.hann_window(n_fft, dtype=torch.float32, device=x.device).to(x.dtype)
    X = torch.stft(x, n_fft, hop_length=hop, win_length=n_fft, window=win, center=True,
                   pad_mode="reflect", normalized=False, onesided=True, return_complex=True)
    return X.real ** 2 + X.imag ** 2


def sss_loss(x_pred: torch.Tensor, x_true: torch.Tensor, n_fft: int, alpha: float = 1.0,
             overlap: float = 0.75, eps: float = 1e-7) -> torch.Tensor:
    """``SSSLoss.forward`` (mss_loss.py:25-33): mean|Sp-St| + alpha*mean|log2(St+eps)-log2(Sp+eps)|."""
    hop = int(n_fft * (1 - overlap))                                      # :21
    s_true = power_spectrogram(x_true, n_fft, hop)
    s_pred = power_spectrogram(x_pred, n_fft, hop)
    lin = (s_pred - s_true).abs().mean()                                  # :29
    log = ((s_true + eps).log2() - (s_pred + eps).log2()).abs().mean()    # :30
    return lin + alpha * log


def mss_loss(x_pred: torch.Tensor, x_true: torch.Tensor, n_ffts: Sequence[int], alpha: float = 1.0,
             overlap: float = 0.75, eps: float = 1e-7) -> torch.Tensor:
    """``MSSLoss.forward`` (mss_loss.py:57-68): sum of single-scale losses."""
    total = x_pred.new_zeros(())
    for n in n_ffts:
        total = total + sss_loss(x_pred, x_true, n, alpha, overlap, eps)
    return total


# ----------------------------------------------------------------------------------------
# Whole path as Decoder wires it (model/autoencoder/decoder.py:127-135), controller excluded
# ----------------------------------------------------------------------------------------
def synth_path(f0, c, a, H, noise, rev_noise, rev_decay, rev_wet, n_harmonics, sample_rate, hop,
               use_reverb: bool = True) -> torch.Tensor:
    """harmonics + filtered noise (+ reverb): decoder.py:129-133."""
    y = oscillator_bank(f0, c, a, n_harmonics, sample_rate, hop) + filtered_noise(H, hop, noise)
    if use_reverb:
        y = reverb_apply(y, reverb_ir(rev_noise, rev_decay, rev_wet, sample_rate))
    return y


# ----------------------------------------------------------------------------------------
# Synthetic inputs -- shared by tests, bench and the reference arm
# ----------------------------------------------------------------------------------------
def modified_sigmoid(x: torch.Tensor) -> torch.Tensor:
    """Range of the controller heads: 2*sigmoid(x)**2.3026 + 1e-7 (decoder.py:110-116)."""
    return 2.0 * torch.sigmoid(x) ** 2.3026 + 1e-7


def synthetic_controls(B: int, T: int, K: int, NB: int, sample_rate: int, hop: int, seed: int,
                       voice_offset: int = 0, dtype=torch.float32):
    """Seeded f0/c/a/H of SURVEY.md §8d.  Voice ``b`` is generated from ``seed`` and its GLOBAL
    index ``voice_offset + b`` so any batch sharding reproduces the same voices."""
    f0 = torch.empty(B, T, 1, dtype=torch.float64)
    c = torch.empty(B, T, K, dtype=torch.float64)
    a = torch.empty(B, T, 1, dtype=torch.float64)
    H = torch.empty(B, T, NB, dtype=torch.float64)
    t = torch.arange(T, dtype=torch.float64) * hop / sample_rate
    for b in range(B):
        g = torch.Generator().manual_seed(seed * 1_000_003 + voice_offset + b)
        base = 80.0 * (1000.0 / 80.0) ** torch.rand(1, generator=g, dtype=torch.float64)
        f = base * (1.0 + 0.01 * torch.sin(TWO_PI * 5.0 * t)) + torch.randn(T, generator=g, dtype=torch.float64)
        f0[b, :, 0] = f.clamp(20.0, sample_rate / 2 - 1.0)
        c[b] = modified_sigmoid(torch.randn(T, K, generator=g, dtype=torch.float64))
        a[b] = modified_sigmoid(torch.randn(T, 1, generator=g, dtype=torch.float64))
        H[b] = modified_sigmoid(torch.randn(T, NB, generator=g, dtype=torch.float64))
    return f0.to(dtype), c.to(dtype), a.to(dtype), H.to(dtype)


# ----------------------------------------------------------------------------------------
# Loudness encoder  (model/autoencoder/encoder.py:131-156)
# ----------------------------------------------------------------------------------------
def a_weighting(freqs, min_db: float = -80.0):
    """A-weighting in dB at `freqs` [Hz] -- the IEC 61672 curve as librosa.A_weighting states it (the reference calls
    librosa, encoder.py:138; librosa is not installable here, so this restatement is pinned by the standard's tabulated
    values only: 0 dB at 1 kHz, -19.1 dB at 100 Hz, +1.2 dB at 2.5 kHz, -2.5 dB at 10 kHz -- tests/test_oracle.py)."""
    import numpy as np
    f_sq = np.asarray(freqs, dtype=np.float64) ** 2.0
    c = np.array([12194.217, 20.598997, 107.65265, 737.86223]) ** 2.0
    with np.errstate(divide="ignore"):
        w = 2.0 + 20.0 * (np.log10(c[0]) + 2.0 * np.log10(f_sq) - np.log10(f_sq + c[0]) - np.log10(f_sq + c[1])
                          - 0.5 * np.log10(f_sq + c[2]) - 0.5 * np.log10(f_sq + c[3]))
    return np.maximum(min_db, w)


def loudness_a_weight(n_fft: int, sample_rate: int) -> torch.Tensor:
    """The `a_weight` parameter of LoudnessEncoder (encoder.py:137-139): fp32 bin frequencies, weights in fp32."""
    import numpy as np
    freqs = np.linspace(0, float(sample_rate) / 2, int(1 + n_fft // 2), endpoint=True, dtype="float32")
    return torch.from_numpy(a_weighting(freqs).astype("float32"))


def loudness_encoder(signal: torch.Tensor, a_weight: torch.Tensor, n_fft: int, hop: int) -> torch.Tensor:
    """``LoudnessEncoder.forward`` (encoder.py:141-156): rectangular-window STFT, center=False."""
    stft = torch.stft(signal, n_fft=n_fft, hop_length=hop, center=False, return_complex=True).permute(0, 2, 1)
    db = torch.log10(torch.abs(stft) + 1e-20) * 20
    db = db + a_weight.to(db.dtype)
    db = db / 90 + 1
    return torch.mean(db, dim=-1, keepdim=True)


# ----------------------------------------------------------------------------------------------
# Controller tail -- model/autoencoder/decoder.py:96-100 with modified_sigmoid :110-116
# ----------------------------------------------------------------------------------------------
def controller_heads(x, w_c, b_c, w_a, b_a, w_h, b_h):
    """The three Linear heads on the last MLP's output and their activation (decoder.py:74-76, 96-100):
    c = msig(x W_c^T + b_c), a = msig(x W_a^T + b_a), H = msig(x W_h^T + b_h), msig = `modified_sigmoid` above
    (decoder.py:110-116: sigmoid -> pow(2.3026) -> mul(2) -> add(1e-7)); nn.Linear weight layout [out,in]."""
    lin = torch.nn.functional.linear
    return modified_sigmoid(lin(x, w_c, b_c)), modified_sigmoid(lin(x, w_a, b_a)), modified_sigmoid(lin(x, w_h, b_h))
