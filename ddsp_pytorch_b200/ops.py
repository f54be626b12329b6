"""autograd.Functions over the C ABI.  PyTorch supplies device memory, streams and autograd plumbing only;
all arithmetic happens in libddsp_b200.so.  Inputs may arrive in fp16/bf16 under autocast (train/train.py:50
runs ``precision=16``) and are cast to fp32 at this boundary."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _f32(t: torch.Tensor) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError("ddsp_pytorch_b200 ops need CUDA tensors (there is no CPU path)")
    return t.detach().to(torch.float32).contiguous()


def _bytes(n: int, device) -> torch.Tensor:
    return torch.empty(max(int(n), 1), dtype=torch.uint8, device=device)


class _Oscillator(torch.autograd.Function):
    @staticmethod
    def forward(ctx, f0, c, a, sample_rate: int, hop: int, out: Optional[torch.Tensor]):
        f0_, c_, a_ = _f32(f0), _f32(c), _f32(a)
        B, T, K = c_.shape
        if f0_.numel() != B * T or a_.numel() != B * T:
            raise RuntimeError(f"oscillator: f0 {tuple(f0.shape)} / a {tuple(a.shape)} do not match c {tuple(c.shape)}")
        dev = c_.device
        need_grad = c.requires_grad or a.requires_grad
        with torch.cuda.device(dev):
            lib = _lib.load()
            accumulate = out is not None
            y = out if accumulate else torch.empty(B, T * hop, dtype=torch.float32, device=dev)
            z = torch.empty(B, T * hop, dtype=torch.float32, device=dev) if need_grad else None
            state = _bytes(lib.ddsp_osc_state_bytes(B, T, K), dev)
            _lib.call("ddsp_osc_forward", f0_.data_ptr(), c_.data_ptr(), a_.data_ptr(), y.data_ptr(),
                      z.data_ptr() if z is not None else None, state.data_ptr(), int(accumulate), B, T, K, hop,
                      sample_rate, _stream())
        if need_grad:
            ctx.save_for_backward(f0_, a_, z, state)
            ctx.dims = (B, T, K, hop, sample_rate)
            ctx.shapes = (c.shape, a.shape, c.dtype, a.dtype)
        if accumulate:
            ctx.mark_dirty(out)
        return y

    @staticmethod
    def backward(ctx, g):
        f0_, a_, z, state = ctx.saved_tensors
        B, T, K, hop, sample_rate = ctx.dims
        c_shape, a_shape, c_dtype, a_dtype = ctx.shapes
        g_ = _f32(g)
        dev = g_.device
        with torch.cuda.device(dev):
            lib = _lib.load()
            gc = torch.empty(B, T, K, dtype=torch.float32, device=dev)
            ga = torch.empty(B, T, dtype=torch.float32, device=dev)
            scratch = _bytes(lib.ddsp_osc_backward_scratch_bytes(B, T, K), dev)
            _lib.call("ddsp_osc_backward", g_.data_ptr(), f0_.data_ptr(), a_.data_ptr(), z.data_ptr(), state.data_ptr(),
                      scratch.data_ptr(), gc.data_ptr(), ga.data_ptr(), B, T, K, hop, sample_rate, _stream())
        # an accumulated-into `out` passes its gradient straight through to whoever produced it
        g_out = g if ctx.needs_input_grad[5] else None
        return None, gc.view(c_shape).to(c_dtype), ga.view(a_shape).to(a_dtype), None, None, g_out


def oscillator_bank(f0: torch.Tensor, c: torch.Tensor, a: torch.Tensor, sample_rate: int, hop: int,
                    out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """f0 [B,T,1] Hz, c [B,T,K], a [B,T,1] -> audio [B,T*hop]  (OscillatorBank.forward, harmonic_oscillator.py:57-62).
    With ``out`` the result is added into that buffer (fuses the ``harmonics + noise`` of decoder.py:132)."""
    return _Oscillator.apply(f0, c, a, int(sample_rate), int(hop), out)


class _FilteredNoise(torch.autograd.Function):
    @staticmethod
    def forward(ctx, H, hop: int, seed: int, offset: int, voice_offset: int, noise: Optional[torch.Tensor],
                out: Optional[torch.Tensor]):
        H_ = _f32(H)
        B, T, NB = H_.shape
        dev = H_.device
        noise_ = None
        if noise is not None:
            noise_ = _f32(noise)
            if noise_.numel() != B * T * hop:
                raise RuntimeError(f"filtered_noise: noise {tuple(noise.shape)} is not [B,T,hop]=[{B},{T},{hop}]")
        with torch.cuda.device(dev):
            accumulate = out is not None
            y = out if accumulate else torch.empty(B, T * hop, dtype=torch.float32, device=dev)
            _lib.call("ddsp_noise_forward", H_.data_ptr(), noise_.data_ptr() if noise_ is not None else None, y.data_ptr(),
                      int(accumulate), seed, offset, voice_offset, B, T, NB, hop, _stream())
        ctx.args = (B, T, NB, hop, seed, offset, voice_offset, H.shape, H.dtype)
        ctx.noise = noise_
        if accumulate:
            ctx.mark_dirty(out)
        return y

    @staticmethod
    def backward(ctx, g):
        B, T, NB, hop, seed, offset, voice_offset, H_shape, H_dtype = ctx.args
        g_ = _f32(g)
        dev = g_.device
        with torch.cuda.device(dev):
            gH = torch.empty(B, T, NB, dtype=torch.float32, device=dev)
            _lib.call("ddsp_noise_backward", g_.data_ptr(), ctx.noise.data_ptr() if ctx.noise is not None else None,
                      gH.data_ptr(), seed, offset, voice_offset, B, T, NB, hop, _stream())
        g_out = g if ctx.needs_input_grad[6] else None
        return gH.view(H_shape).to(H_dtype), None, None, None, None, None, g_out


def filtered_noise(H: torch.Tensor, hop: int, seed: int = 0, offset: int = 0, voice_offset: int = 0,
                   noise: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """H [B,T,NB] -> [B,T*hop]  (FilteredNoise.forward, filtered_noise.py:40-53).  ``noise`` [B,T,hop] in [-1,1)
    replaces the in-kernel Philox stream (parity tests inject the reference's draw this way)."""
    return _FilteredNoise.apply(H, int(hop), int(seed) & (2 ** 64 - 1), int(offset) & 0xFFFFFFFF,
                                int(voice_offset) & 0xFFFFFFFF, noise, out)


def noise_uniform(B: int, T: int, hop: int, seed: int, offset: int = 0, voice_offset: int = 0, device="cuda") -> torch.Tensor:
    """The [-1,1) Philox block the noise kernel draws for these arguments (layout: include/ddsp_b200.h)."""
    out = torch.empty(B, T, hop, dtype=torch.float32, device=device)
    with torch.cuda.device(out.device):
        _lib.call("ddsp_noise_uniform", out.data_ptr(), int(seed) & (2 ** 64 - 1), int(offset) & 0xFFFFFFFF,
                  int(voice_offset) & 0xFFFFFFFF, B, T, hop, _stream())
    return out


def amp_to_impulse_response(amp: torch.Tensor, target_size: int) -> torch.Tensor:
    """filtered_noise.py:7-22 with the same signature: [..., NB] -> [..., target_size] (forward only)."""
    a_ = _f32(amp)
    NB = a_.shape[-1]
    rows = a_.numel() // NB
    ir = torch.empty(*a_.shape[:-1], int(target_size), dtype=torch.float32, device=a_.device)
    with torch.cuda.device(a_.device):
        _lib.call("ddsp_amp_to_impulse_response", a_.data_ptr(), ir.data_ptr(), rows, NB, int(target_size), _stream())
    return ir


def probe_fp32_fma_tflops(iters: int = 4096, reps: int = 5) -> float:
    """Measured FP32 FFMA throughput (TFLOP/s, FMA = 2 flops) of this GPU: the oscillator's roofline denominator."""
    props = torch.cuda.get_device_properties(torch.cuda.current_device())
    blocks, threads = props.multi_processor_count * 4, 512
    out = torch.empty(blocks * threads, dtype=torch.float32, device="cuda")
    best = 0.0
    for _ in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.call("ddsp_probe_fp32_fma", out.data_ptr(), blocks, threads, iters, _stream())
        e1.record()
        e1.synchronize()
        flops = 2.0 * 64 * iters * blocks * threads
        best = max(best, flops / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best
