"""autograd.Functions over the C ABI.  PyTorch supplies device memory, streams and autograd plumbing only;
all arithmetic happens in libddsp_b200.so.  Inputs may arrive in fp16/bf16 under autocast (train/train.py:50
runs ``precision=16``) and are cast to fp32 at this boundary."""
from __future__ import annotations

import os
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


def _check_out(out: torch.Tensor, B: int, N: int, device, who: str) -> None:
    """The fused-accumulate buffer is handed to the kernel as a raw pointer (float4 read-modify-write of B*N floats):
    anything but a contiguous fp32 CUDA tensor of exactly that shape on the inputs' device would write out of bounds."""
    if not isinstance(out, torch.Tensor) or not out.is_cuda or out.dtype != torch.float32 or not out.is_contiguous() \
            or tuple(out.shape) != (B, N) or out.device != device:
        desc = f"{tuple(out.shape)} {out.dtype} on {out.device}" if isinstance(out, torch.Tensor) else type(out).__name__
        raise RuntimeError(f"{who}: out= must be a contiguous float32 CUDA tensor of shape ({B}, {N}) on {device}, got {desc}")
    if out.data_ptr() % 16:
        raise RuntimeError(f"{who}: out= must be 16-byte aligned (a view at an odd offset cannot take float4 stores)")


class _Oscillator(torch.autograd.Function):
    @staticmethod
    def forward(ctx, f0, c, a, sample_rate: int, hop: int, out: Optional[torch.Tensor]):
        f0_, c_, a_ = _f32(f0), _f32(c), _f32(a)
        B, T, K = c_.shape
        if f0_.numel() != B * T or a_.numel() != B * T:
            raise RuntimeError(f"oscillator: f0 {tuple(f0.shape)} / a {tuple(a.shape)} do not match c {tuple(c.shape)}")
        dev = c_.device
        need_grad = c.requires_grad or a.requires_grad or f0.requires_grad
        accumulate = out is not None
        if accumulate:
            _check_out(out, B, T * hop, dev, "oscillator_bank")
        ctx.saved_any = need_grad
        with torch.cuda.device(dev):
            lib = _lib.load()
            y = out if accumulate else torch.empty(B, T * hop, dtype=torch.float32, device=dev)
            z = torch.empty(B, T * hop, dtype=torch.float32, device=dev) if need_grad else None
            state = _bytes(lib.ddsp_osc_state_bytes(B, T, K), dev)
            _lib.call("ddsp_osc_forward", f0_.data_ptr(), c_.data_ptr(), a_.data_ptr(), y.data_ptr(),
                      z.data_ptr() if z is not None else None, state.data_ptr(), int(accumulate), B, T, K, hop,
                      sample_rate, _stream())
        if need_grad:
            ctx.save_for_backward(f0_, a_, z, state)
            ctx.dims = (B, T, K, hop, sample_rate)
            ctx.shapes = (c.shape, a.shape, c.dtype, a.dtype, f0.shape, f0.dtype)
        if accumulate:
            ctx.mark_dirty(out)
        return y

    @staticmethod
    def backward(ctx, g):
        g_out = g if ctx.needs_input_grad[5] else None
        if not ctx.saved_any:                   # c and a are frozen; only an accumulated-into `out` carries a gradient
            return None, None, None, None, None, g_out
        f0_, a_, z, state = ctx.saved_tensors
        B, T, K, hop, sample_rate = ctx.dims
        c_shape, a_shape, c_dtype, a_dtype, f0_shape, f0_dtype = ctx.shapes
        g_ = _f32(g)
        dev = g_.device
        gc = ga = gf0 = None
        with torch.cuda.device(dev):
            lib = _lib.load()
            if ctx.needs_input_grad[1] or ctx.needs_input_grad[2]:
                gc = torch.empty(B, T, K, dtype=torch.float32, device=dev)
                ga = torch.empty(B, T, dtype=torch.float32, device=dev)
                scratch = _bytes(lib.ddsp_osc_backward_scratch_bytes(B, T, K), dev)
                _lib.call("ddsp_osc_backward", g_.data_ptr(), f0_.data_ptr(), a_.data_ptr(), z.data_ptr(), state.data_ptr(),
                          scratch.data_ptr(), gc.data_ptr(), ga.data_ptr(), B, T, K, hop, sample_rate, _stream())
                gc = gc.view(c_shape).to(c_dtype) if ctx.needs_input_grad[1] else None
                ga = ga.view(a_shape).to(a_dtype) if ctx.needs_input_grad[2] else None
            if ctx.needs_input_grad[0]:
                # the reference differentiates through the pitch as well (harmonic_oscillator.py:26-43 is plain autograd)
                gf0 = torch.empty(B, T, dtype=torch.float32, device=dev)
                scratch0 = _bytes(lib.ddsp_osc_backward_f0_scratch_bytes(B, T, hop), dev)
                _lib.call("ddsp_osc_backward_f0", g_.data_ptr(), f0_.data_ptr(), a_.data_ptr(), state.data_ptr(),
                          scratch0.data_ptr(), gf0.data_ptr(), B, T, K, hop, sample_rate, _stream())
                gf0 = gf0.view(f0_shape).to(f0_dtype)
        # an accumulated-into `out` passes its gradient straight through to whoever produced it
        return gf0, gc, ga, None, None, g_out


def oscillator_bank(f0: torch.Tensor, c: torch.Tensor, a: torch.Tensor, sample_rate: int, hop: int,
                    out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """f0 [B,T,1] Hz, c [B,T,K], a [B,T,1] -> audio [B,T*hop]  (OscillatorBank.forward, harmonic_oscillator.py:57-62).
    With ``out`` the result is added into that buffer (fuses the ``harmonics + noise`` of decoder.py:132)."""
    return _Oscillator.apply(f0, c, a, int(sample_rate), int(hop), out)


class _FilteredNoise(torch.autograd.Function):
    @staticmethod
    def forward(ctx, H, hop: int, seed: int, offset: int, voice_offset: int, noise: Optional[torch.Tensor],
                out: Optional[torch.Tensor], offset_tensor: Optional[torch.Tensor]):
        H_ = _f32(H)
        B, T, NB = H_.shape
        dev = H_.device
        noise_ = None
        if noise is not None:
            noise_ = _f32(noise)
            if noise_.numel() != B * T * hop:
                raise RuntimeError(f"filtered_noise: noise {tuple(noise.shape)} is not [B,T,hop]=[{B},{T},{hop}]")
        accumulate = out is not None
        if accumulate:
            _check_out(out, B, T * hop, dev, "filtered_noise")
        with torch.cuda.device(dev):
            y = out if accumulate else torch.empty(B, T * hop, dtype=torch.float32, device=dev)
            _lib.call("ddsp_noise_forward", H_.data_ptr(), noise_.data_ptr() if noise_ is not None else None, y.data_ptr(),
                      int(accumulate), seed, offset, offset_tensor.data_ptr() if offset_tensor is not None else None,
                      voice_offset, B, T, NB, hop, _stream())
        ctx.args = (B, T, NB, hop, seed, offset, voice_offset, H.shape, H.dtype)
        ctx.noise = noise_
        # the device-side call counter may advance before backward runs (two forwards, one backward): keep the value this
        # forward drew its Philox block with (a one-element clone is capturable in a CUDA graph)
        ctx.offset_tensor = offset_tensor.clone() if (offset_tensor is not None and H.requires_grad) else None
        if accumulate:
            ctx.mark_dirty(out)
        return y

    @staticmethod
    def backward(ctx, g):
        B, T, NB, hop, seed, offset, voice_offset, H_shape, H_dtype = ctx.args
        g_out = g if ctx.needs_input_grad[6] else None
        if not ctx.needs_input_grad[0]:
            return None, None, None, None, None, None, g_out, None
        g_ = _f32(g)
        dev = g_.device
        with torch.cuda.device(dev):
            gH = torch.empty(B, T, NB, dtype=torch.float32, device=dev)
            _lib.call("ddsp_noise_backward", g_.data_ptr(), ctx.noise.data_ptr() if ctx.noise is not None else None,
                      gH.data_ptr(), seed, offset, ctx.offset_tensor.data_ptr() if ctx.offset_tensor is not None else None,
                      voice_offset, B, T, NB, hop, _stream())
        return gH.view(H_shape).to(H_dtype), None, None, None, None, None, g_out, None


def filtered_noise(H: torch.Tensor, hop: int, seed: int = 0, offset: int = 0, voice_offset: int = 0,
                   noise: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                   offset_tensor: Optional[torch.Tensor] = None) -> torch.Tensor:
    """H [B,T,NB] -> [B,T*hop]  (FilteredNoise.forward, filtered_noise.py:40-53).  ``noise`` [B,T,hop] in [-1,1)
    replaces the in-kernel Philox stream (parity tests inject the reference's draw this way).  ``offset_tensor``:
    a one-element int32 CUDA tensor added to ``offset`` on the device (a call counter a CUDA graph can advance)."""
    if offset_tensor is not None and not (offset_tensor.is_cuda and offset_tensor.dtype == torch.int32 and offset_tensor.numel() == 1):
        raise RuntimeError("filtered_noise: offset_tensor must be a one-element int32 CUDA tensor")
    return _FilteredNoise.apply(H, int(hop), int(seed) & (2 ** 64 - 1), int(offset) & 0xFFFFFFFF,
                                int(voice_offset) & 0xFFFFFFFF, noise, out, offset_tensor)


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


# ------------------------------------------------------------------------------------------------------
# Reverb
# ------------------------------------------------------------------------------------------------------
class _ReverbImpulse(torch.autograd.Function):
    @staticmethod
    def forward(ctx, noise, decay, wet, t):
        n_, d_, w_, t_ = _f32(noise).view(-1), _f32(decay).view(1), _f32(wet).view(1), _f32(t).view(-1)
        L = n_.numel()
        ir = torch.empty(1, L, dtype=torch.float32, device=n_.device)
        with torch.cuda.device(n_.device):
            _lib.call("ddsp_reverb_impulse", n_.data_ptr(), t_.data_ptr(), d_.data_ptr(), w_.data_ptr(), ir.data_ptr(), L, _stream())
        ctx.save_for_backward(n_, d_, w_, t_)
        ctx.meta = (noise.shape, decay.shape, wet.shape)
        return ir

    @staticmethod
    def backward(ctx, gir):
        n_, d_, w_, t_ = ctx.saved_tensors
        L = n_.numel()
        g_ = _f32(gir).view(-1)
        gn = torch.empty(L, dtype=torch.float32, device=n_.device)
        gd = torch.empty(1, dtype=torch.float32, device=n_.device)
        gw = torch.empty(1, dtype=torch.float32, device=n_.device)
        with torch.cuda.device(n_.device):
            _lib.call("ddsp_reverb_impulse_backward", g_.data_ptr(), n_.data_ptr(), t_.data_ptr(), d_.data_ptr(), w_.data_ptr(),
                      gn.data_ptr(), gd.data_ptr(), gw.data_ptr(), L, _stream())
        ns, ds, ws = ctx.meta
        return gn.view(ns), gd.view(ds), gw.view(ws), None


def reverb_impulse(noise, decay, wet, t) -> torch.Tensor:
    """Reverb.build_impulse (reverb.py:24-29): -> [1,L]."""
    return _ReverbImpulse.apply(noise, decay, wet, t)


class _CausalConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, ir):
        x_, ir_ = _f32(x), _f32(ir).view(-1)
        B, N = x_.shape
        Lk = ir_.numel()
        y = torch.empty(B, N, dtype=torch.float32, device=x_.device)
        with torch.cuda.device(x_.device):
            scratch = _bytes(_lib.load().ddsp_conv_scratch_bytes(B, N, Lk), x_.device)
            _lib.call("ddsp_conv_forward", x_.data_ptr(), ir_.data_ptr(), y.data_ptr(), scratch.data_ptr(), B, N, Lk, _stream())
        ctx.save_for_backward(x_, ir_)
        ctx.meta = (x.dtype, ir.shape, ir.dtype)
        # the block spectra of x and ir stay in `scratch`; the first backward reuses them (and then overwrites them)
        ctx.scratch = scratch if (x.requires_grad or ir.requires_grad) else None
        return y

    @staticmethod
    def backward(ctx, g):
        x_, ir_ = ctx.saved_tensors
        B, N = x_.shape
        Lk = ir_.numel()
        g_ = _f32(g)
        need_x, need_ir = ctx.needs_input_grad
        gx = torch.empty(B, N, dtype=torch.float32, device=x_.device) if need_x else None
        gir = torch.empty(Lk, dtype=torch.float32, device=x_.device) if need_ir else None
        with torch.cuda.device(x_.device):
            scratch, reuse = getattr(ctx, "scratch", None), 1
            ctx.scratch = None
            if scratch is None:
                scratch, reuse = _bytes(_lib.load().ddsp_conv_scratch_bytes(B, N, Lk), x_.device), 0
            _lib.call("ddsp_conv_backward", g_.data_ptr(), x_.data_ptr(), ir_.data_ptr(),
                      gx.data_ptr() if need_x else None, gir.data_ptr() if need_ir else None, scratch.data_ptr(), B, N, Lk,
                      reuse, _stream())
        xd, irs, ird = ctx.meta
        return (gx.to(xd) if need_x else None), (gir.view(irs).to(ird) if need_ir else None)


def causal_conv(x: torch.Tensor, ir: torch.Tensor) -> torch.Tensor:
    """x [B,N], ir [1,Lk] or [Lk] shared by the batch -> y[b,n] = sum_{q<=n} ir[q] x[b,n-q]
    (fft_convolve as Reverb.forward uses it, reverb.py:31-38; taps at lags >= N are irrelevant, so the
    reference's crop/pad of the IR to N is implicit)."""
    return _CausalConv.apply(x, ir)


class _RowConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, signal, kernel):
        s_, k_ = _f32(signal), _f32(kernel)
        n = s_.shape[-1]
        rows = s_.numel() // n
        y = torch.empty_like(s_)
        with torch.cuda.device(s_.device):
            _lib.call("ddsp_rowconv_forward", s_.data_ptr(), k_.data_ptr(), y.data_ptr(), rows, n, _stream())
        ctx.save_for_backward(s_, k_)
        ctx.meta = (signal.dtype, kernel.dtype)
        return y

    @staticmethod
    def backward(ctx, g):
        s_, k_ = ctx.saved_tensors
        n = s_.shape[-1]
        rows = s_.numel() // n
        g_ = _f32(g)
        need_s, need_k = ctx.needs_input_grad
        gs = torch.empty_like(s_) if need_s else None
        gk = torch.empty_like(k_) if need_k else None
        with torch.cuda.device(s_.device):
            _lib.call("ddsp_rowconv_backward", g_.data_ptr(), s_.data_ptr(), k_.data_ptr(), gs.data_ptr() if need_s else None,
                      gk.data_ptr() if need_k else None, rows, n, _stream())
        sd, kd = ctx.meta
        return (gs.to(sd) if need_s else None), (gk.to(kd) if need_k else None)


def row_conv(signal: torch.Tensor, kernel: torch.Tensor) -> torch.Tensor:
    """signal [...,n], kernel of the SAME shape -> y[..., m] = sum_{q<=m} kernel[..., q] * signal[..., m-q]
    (fft_convolve with one kernel per row, filtered_noise.py:25-32); n <= 4096."""
    if signal.shape != kernel.shape:
        raise RuntimeError(f"row_conv: signal {tuple(signal.shape)} and kernel {tuple(kernel.shape)} must have the same shape")
    return _RowConv.apply(signal, kernel)


def reverb_live(buffer: torch.Tensor, x: torch.Tensor, ir: torch.Tensor) -> torch.Tensor:
    """Reverb.live_forward (reverb.py:40-49): updates ``buffer`` [1,L] in place, returns the last len(x) outputs [1,n]."""
    x_, ir_ = _f32(x).view(-1), _f32(ir).view(-1)
    L, n = buffer.numel(), x_.numel()
    if buffer.dtype != torch.float32 or not buffer.is_cuda or not buffer.is_contiguous():
        raise RuntimeError("reverb_live: buffer must be a contiguous fp32 CUDA tensor")
    y = torch.empty(1, n, dtype=torch.float32, device=buffer.device)
    with torch.cuda.device(buffer.device):
        scratch = torch.empty(L, dtype=torch.float32, device=buffer.device)
        _lib.call("ddsp_reverb_live", buffer.data_ptr(), x_.data_ptr(), ir_.data_ptr(), y.data_ptr(), scratch.data_ptr(), L, n,
                  _stream())
    return y


# ------------------------------------------------------------------------------------------------------
# Multi-scale spectral loss
# ------------------------------------------------------------------------------------------------------
class _MSSLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x_pred, x_true, n_ffts, hops, alpha: float, eps: float, windows):
        import ctypes as C
        xp, xt = _f32(x_pred), _f32(x_true)
        if xp.dim() != 2 or xp.shape != xt.shape:
            raise RuntimeError(f"mss_loss: x_pred {tuple(x_pred.shape)} and x_true {tuple(x_true.shape)} must both be [B,N]")
        B, N = xp.shape
        ns = len(n_ffts)
        dev = xp.device
        need_grad = x_pred.requires_grad
        terms = torch.empty(1 + 2 * ns, dtype=torch.float64, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        nf = (C.c_int * ns)(*[int(v) for v in n_ffts])
        hp = (C.c_int * ns)(*[int(v) for v in hops])
        wins = None
        if windows is not None:
            keep = [None if w is None else w.detach().to(device=dev, dtype=torch.float32).contiguous() for w in windows]
            for w, n in zip(keep, n_ffts):
                if w is not None and w.numel() != n:
                    raise RuntimeError(f"mss_loss: window of {w.numel()} samples given for n_fft={n}")
            wins = (C.c_void_p * ns)(*[None if w is None else w.data_ptr() for w in keep])
        with torch.cuda.device(dev):
            nbytes = _lib.load().ddsp_mss_scratch_bytes(C.cast(nf, C.c_void_p), C.cast(hp, C.c_void_p), ns, B, N)
            if nbytes == 0:
                raise RuntimeError(f"mss_loss: unsupported arguments (n_ffts {list(n_ffts)} must be powers of two in [4,4096] with "
                                   f"n_fft/2 < N={N}; at most 16 scales)")
            scratch = _bytes(nbytes, dev)
            _lib.call("ddsp_mss_forward", xp.data_ptr(), xt.data_ptr(), terms.data_ptr(), loss.data_ptr(), scratch.data_ptr(),
                      int(need_grad), C.cast(nf, C.c_void_p), C.cast(hp, C.c_void_p),
                      C.cast(wins, C.c_void_p) if wins is not None else None, ns, float(alpha), float(eps), B, N, _stream())
        if need_grad:
            ctx.save_for_backward(scratch)
        ctx.meta = (x_pred.dtype, B, N, tuple(int(v) for v in n_ffts), tuple(int(v) for v in hops))
        ctx.terms = terms
        return loss

    @staticmethod
    def backward(ctx, gloss):
        import ctypes as C
        (scratch,) = ctx.saved_tensors
        dt, B, N, n_ffts, hops = ctx.meta
        ns = len(n_ffts)
        gl = _f32(gloss).view(1)
        out = torch.empty(B, N, dtype=torch.float32, device=scratch.device)
        nf, hp = (C.c_int * ns)(*n_ffts), (C.c_int * ns)(*hops)
        with torch.cuda.device(scratch.device):
            _lib.call("ddsp_mss_backward", gl.data_ptr(), scratch.data_ptr(), out.data_ptr(), C.cast(nf, C.c_void_p),
                      C.cast(hp, C.c_void_p), ns, B, N, _stream())
        return out.to(dt), None, None, None, None, None, None


_HANN_CACHE = {}


def _torch_hann(n: int, device) -> torch.Tensor:
    """torch.hann_window's own fp32 values (what torchaudio's Spectrogram holds, mss_loss.py:23).  The kernel's built-in
    Hann is rounded from fp64 and differs from these by an ulp in places -- enough to move the near-empty bins that
    dominate d log2(S + eps)/dx by 1e-4, so the default is the reference's table, not the kernel's."""
    key = (int(n), str(device))
    if key not in _HANN_CACHE:
        _HANN_CACHE[key] = torch.hann_window(int(n), dtype=torch.float32).to(device)
    return _HANN_CACHE[key]


def mss_loss(x_pred, x_true, n_ffts, alpha: float = 1.0, overlap: float = 0.75, eps: float = 1e-7, windows=None) -> torch.Tensor:
    """MSSLoss.forward (mss_loss.py:57-68) on tensors [B,N] -> 0-dim fp32 loss; gradient flows to x_pred only.
    ``windows``: optional per-scale fp32 analysis windows (the modules pass their ``stft.window`` buffers); default:
    torch.hann_window(n) as the reference's Spectrogram builds it."""
    hops = [int(n * (1 - overlap)) for n in n_ffts]          # mss_loss.py:21
    if windows is None:
        windows = [_torch_hann(n, x_pred.device) for n in n_ffts]
    return _MSSLoss.apply(x_pred, x_true, tuple(int(n) for n in n_ffts), tuple(hops), float(alpha), float(eps), tuple(windows))


def dftgemm_sss_terms(x_pred: torch.Tensor, x_true: torch.Tensor, n_fft: int, eps: float = 1e-7, splits: int = 3,
                      window: Optional[torch.Tensor] = None, return_spectra: bool = False, workspace: Optional[torch.Tensor] = None,
                      prepare: bool = True):
    """The tcgen05 DFT-as-GEMM experiment (csrc/ddsp_dftgemm.cu): raw sums (sum|Sp-St|, sum|log2 St+eps - log2 Sp+eps|) of one
    scale, n_fft in {64, 128}, hop = n_fft/4, as a float64 tensor [2]; with return_spectra also the power spectrograms
    [2, B, F, n_fft/2+1] (true, pred).  Forward only; not on the product path -- `mss_loss` uses the register-tiled FFT."""
    xp, xt = _f32(x_pred), _f32(x_true)
    B, N = xp.shape
    dev = xp.device
    with torch.cuda.device(dev):
        lib = _lib.load()
        nbytes = lib.ddsp_dftgemm_workspace_bytes(B, N, int(n_fft))
        if nbytes == 0:
            raise RuntimeError(f"dftgemm_sss_terms: n_fft={n_fft} must be 64 or 128 with n_fft/2 < N={N}")
        ws = workspace if workspace is not None else _bytes(nbytes, dev)
        if window is None:
            window = _torch_hann(int(n_fft), dev)
        w_ = window.detach().to(device=dev, dtype=torch.float32).contiguous()
        if prepare:
            _lib.call("ddsp_dftgemm_prepare", xp.data_ptr(), xt.data_ptr(), w_.data_ptr(), ws.data_ptr(), int(n_fft), B, N, _stream())
        terms = torch.empty(2, dtype=torch.float64, device=dev)
        F_ = 1 + N // (int(n_fft) // 4)
        S = torch.empty(2, B, F_, int(n_fft) // 2 + 1, dtype=torch.float32, device=dev) if return_spectra else None
        _lib.call("ddsp_dftgemm_sss_forward", ws.data_ptr(), terms.data_ptr(), S.data_ptr() if S is not None else None, int(n_fft),
                  int(splits), float(eps), B, N, _stream())
    return (terms, S) if return_spectra else terms


class _Rescale(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, hop: int):
        x_ = _f32(x)
        if x_.dim() != 3:
            raise RuntimeError(f"rescale: expected [B,T,C], got {tuple(x.shape)}")
        B, T, C = x_.shape
        y = torch.empty(B, T * hop, C, dtype=torch.float32, device=x_.device)
        with torch.cuda.device(x_.device):
            _lib.call("ddsp_rescale_forward", x_.data_ptr(), y.data_ptr(), B, T, C, hop, _stream())
        ctx.dims = (B, T, C, hop, x.dtype)
        return y

    @staticmethod
    def backward(ctx, g):
        B, T, C, hop, dtype = ctx.dims
        g_ = _f32(g)
        gx = torch.empty(B, T, C, dtype=torch.float32, device=g_.device)
        with torch.cuda.device(g_.device):
            _lib.call("ddsp_rescale_backward", g_.data_ptr(), gx.data_ptr(), B, T, C, hop, _stream())
        return gx.to(dtype), None


def rescale(x: torch.Tensor, hop: int) -> torch.Tensor:
    """[B,T,C] -> [B,T*hop,C]: ``OscillatorBank.rescale`` (harmonic_oscillator.py:52-55), linear upsampling with half-pixel
    centres.  Not used by the synthesis ops (they fold the interpolation into their kernels); it backs the reference's
    public helper methods."""
    return _Rescale.apply(x, int(hop))


def oscillator_live(f0, c, a, last_phases, sample_rate: int, hop: int):
    """OscillatorBank.live (harmonic_oscillator.py:64-75), no gradient: returns (audio [B,T*hop], new last_phases [K] fp32)."""
    f0_, c_, a_ = _f32(f0), _f32(c), _f32(a)
    B, T, K = c_.shape
    last_in = last_phases.detach().to(device=c_.device, dtype=torch.float32).contiguous()
    last_out = torch.empty(K, dtype=torch.float32, device=c_.device)
    y = torch.empty(B, T * hop, dtype=torch.float32, device=c_.device)
    with torch.cuda.device(c_.device):
        state = _bytes(_lib.load().ddsp_osc_state_bytes(B, T, K), c_.device)
        _lib.call("ddsp_osc_live", f0_.data_ptr(), c_.data_ptr(), a_.data_ptr(), last_in.data_ptr(), last_out.data_ptr(),
                  y.data_ptr(), state.data_ptr(), B, T, K, int(hop), int(sample_rate), _stream())
    return y, last_out


def loudness(signal: torch.Tensor, a_weight: torch.Tensor, n_fft: int, hop: int) -> torch.Tensor:
    """signal [B,N] -> loudness [B,F,1]  (LoudnessEncoder.forward, encoder.py:141-156).  No gradient: the encoder runs
    under no_grad in the reference (encoder.py:51)."""
    x, aw = _f32(signal), _f32(a_weight)
    B, N = x.shape
    if aw.numel() != n_fft // 2 + 1:
        raise RuntimeError(f"loudness: a_weight has {aw.numel()} entries, n_fft={n_fft} needs {n_fft // 2 + 1}")
    with torch.cuda.device(x.device):
        lib = _lib.load()
        F_ = lib.ddsp_loudness_frames(N, n_fft, hop)
        out = torch.empty(B, max(F_, 0), 1, dtype=torch.float32, device=x.device)
        _lib.call("ddsp_loudness_forward", x.data_ptr(), aw.data_ptr(), out.data_ptr(), B, N, n_fft, hop, _stream())
    return out


# Rows from which the Controller tail runs on the tensor cores (csrc/ddsp_ctrl.cu: tcgen05 path); below it the split
# pre-passes' launches cost more than the product.  0 disables the path (tests set it either way; the environment variable
# DDSP_B200_CTRL_TC_MIN_ROWS sets it for a process).
CTRL_TC_MIN_ROWS = int(os.environ.get("DDSP_B200_CTRL_TC_MIN_ROWS", "256"))


def _ctrl_use_tensor_cores(R: int, U: int) -> bool:
    return CTRL_TC_MIN_ROWS > 0 and R >= CTRL_TC_MIN_ROWS and U >= 64 and U % 8 == 0


class _ControllerHeads(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w_c, b_c, w_a, b_a, w_h, b_h):
        x_ = _f32(x)
        params = [_f32(p) for p in (w_c, b_c, w_a, b_a, w_h, b_h)]
        U = x_.shape[-1]
        K, NB = params[0].shape[0], params[4].shape[0]
        if params[0].shape[1] != U or params[2].shape != (1, U) or params[4].shape[1] != U:
            raise RuntimeError(f"controller_heads: weights {tuple(w_c.shape)}/{tuple(w_a.shape)}/{tuple(w_h.shape)} do not "
                               f"match the {U}-wide input")
        lead = x_.shape[:-1]
        R = x_.numel() // U
        dev = x_.device
        need_grad = any(t.requires_grad for t in (x, w_c, b_c, w_a, b_a, w_h, b_h))
        with torch.cuda.device(dev):
            lib = _lib.load()
            c = torch.empty(*lead, K, dtype=torch.float32, device=dev)
            a = torch.empty(*lead, 1, dtype=torch.float32, device=dev)
            H = torch.empty(*lead, NB, dtype=torch.float32, device=dev)
            z = torch.empty(R, lib.ddsp_ctrl_heads_ldz(K, NB), dtype=torch.float32, device=dev) if need_grad else None
            tc = _ctrl_use_tensor_cores(R, U)
            if tc:
                ws = _bytes(lib.ddsp_ctrl_heads_tc_workspace_bytes(R, U, K, NB, 0), dev)
                _lib.call("ddsp_ctrl_heads_forward_tc", x_.data_ptr(), *(p.data_ptr() for p in params), c.data_ptr(), a.data_ptr(),
                          H.data_ptr(), z.data_ptr() if z is not None else None, ws.data_ptr(), R, U, K, NB, _stream())
            else:
                _lib.call("ddsp_ctrl_heads_forward", x_.data_ptr(), *(p.data_ptr() for p in params), c.data_ptr(), a.data_ptr(),
                          H.data_ptr(), z.data_ptr() if z is not None else None, R, U, K, NB, _stream())
        if need_grad:
            ctx.save_for_backward(x_, params[0], params[2], params[4], z)
            ctx.dims = (R, U, K, NB)
            ctx.tc = tc
            ctx.meta = (x.shape, x.dtype, x.requires_grad, [p.dtype for p in (w_c, b_c, w_a, b_a, w_h, b_h)])
        return c, a, H

    @staticmethod
    def backward(ctx, g_c, g_a, g_h):
        x_, w_c, w_a, w_h, z = ctx.saved_tensors
        R, U, K, NB = ctx.dims
        x_shape, x_dtype, x_req, pdt = ctx.meta
        dev = x_.device
        gs = [None if g is None else _f32(g) for g in (g_c, g_a, g_h)]
        with torch.cuda.device(dev):
            lib = _lib.load()
            mk = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)  # noqa: E731
            gx = mk(R, U) if x_req else None
            outs = [mk(K, U), mk(K), mk(1, U), mk(1), mk(NB, U), mk(NB)]
            ws = _bytes(lib.ddsp_ctrl_heads_tc_workspace_bytes(R, U, K, NB, 1) if ctx.tc else
                        lib.ddsp_ctrl_heads_workspace_bytes(R, U, K, NB), dev)
            _lib.call("ddsp_ctrl_heads_backward_tc" if ctx.tc else "ddsp_ctrl_heads_backward", x_.data_ptr(), w_c.data_ptr(), w_a.data_ptr(), w_h.data_ptr(), z.data_ptr(),
                      *(None if g is None else g.data_ptr() for g in gs), None if gx is None else gx.data_ptr(),
                      *(o.data_ptr() for o in outs), ws.data_ptr(), R, U, K, NB, _stream())
        gx = None if gx is None else gx.view(x_shape).to(x_dtype)
        return (gx, *(o.to(d) for o, d in zip(outs, pdt)))


def controller_heads(x, w_c, b_c, w_a, b_a, w_h, b_h):
    """x [..,U] -> (c [..,K], a [..,1], H [..,NB]) = modified_sigmoid of the three Linear heads (decoder.py:96-100,110-116)
    in one launch, fp32 outputs in the layout the synthesis kernels read; nn.Linear weight layout [out,in]."""
    return _ControllerHeads.apply(x, w_c, b_c, w_a, b_a, w_h, b_h)
