"""Host-resident controls -> oscillator bank + filtered noise, forward and backward -> host-resident gradients.

The synthesis sweep (SURVEY.md §8.0 cfg 4) is bound by PCIe once the controls start on the host: 1.28 GB go in and
0.49 GB come out per step against 11 ms of kernels.  This driver cuts the voices into chunks and runs three CUDA
streams -- copy-in, compute, copy-out -- so the three overlap (PCIe is full duplex); the arithmetic is the same
`ops.oscillator_bank` / `ops.filtered_noise` calls as the module API (decoder.py:129-132), chunk by chunk, and the
Philox counters are keyed by global voice index, so the result does not depend on the chunking.
"""
from __future__ import annotations

import torch

from . import ops


class HostSynthPipeline:
    def __init__(self, conf, device, chunk_voices: int = 64, slots: int = 3, seed: int = 0, voice_offset: int = 0):
        self.sr, self.hop = conf.sample_rate, conf.hop_length
        self.K, self.NB = conf.n_harmonics, conf.n_noise_filters
        self.device = torch.device(device)
        self.chunk, self.slots = int(chunk_voices), int(slots)
        self.seed, self.voice_offset = int(seed), int(voice_offset)
        self.calls = 0
        self.s_in, self.s_run, self.s_out = (torch.cuda.Stream(self.device) for _ in range(3))
        self._bufs = None

    def _alloc(self, T: int):
        if self._bufs is not None and self._bufs[0]["f0"].shape[1] == T:
            return
        C, N, dev = self.chunk, T * self.hop, self.device
        mk = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)  # noqa: E731
        self._bufs = [{"f0": mk(C, T, 1), "c": mk(C, T, self.K), "a": mk(C, T, 1), "H": mk(C, T, self.NB), "g": mk(C, N)}
                      for _ in range(self.slots)]

    def step(self, f0, c, a, H, g, gc_out, ga_out, gH_out):
        """All arguments are PINNED host tensors: f0 [B,T,1], c [B,T,K], a [B,T,1], H [B,T,NB], g [B,N] in;
        gc_out, ga_out, gH_out receive d sum(g*y)/d c, a, H.  Returns when the host owns the gradients."""
        B, T = c.shape[0], c.shape[1]
        self._alloc(T)
        offset = self.calls
        self.calls += 1
        free = [None] * self.slots                      # event: the slot's device inputs may be overwritten
        keep = []                                       # tensors the copy-out stream still reads
        done = None
        for i, lo in enumerate(range(0, B, self.chunk)):
            hi = min(B, lo + self.chunk)
            n, slot = hi - lo, i % self.slots
            buf = self._bufs[slot]
            with torch.cuda.stream(self.s_in):
                if free[slot] is not None:
                    self.s_in.wait_event(free[slot])
                for name, src in (("f0", f0), ("c", c), ("a", a), ("H", H), ("g", g)):
                    buf[name][:n].copy_(src[lo:hi], non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(self.s_in)
            with torch.cuda.stream(self.s_run):
                self.s_run.wait_event(ready)
                cd = buf["c"][:n].detach().requires_grad_(True)
                ad = buf["a"][:n].detach().requires_grad_(True)
                Hd = buf["H"][:n].detach().requires_grad_(True)
                y = ops.filtered_noise(Hd, self.hop, seed=self.seed, offset=offset, voice_offset=self.voice_offset + lo)
                y = ops.oscillator_bank(buf["f0"][:n], cd, ad, self.sr, self.hop, out=y)
                y.backward(buf["g"][:n])
                computed = torch.cuda.Event()
                computed.record(self.s_run)
                free[slot] = computed
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(computed)
                gc_out[lo:hi].copy_(cd.grad, non_blocking=True)
                ga_out[lo:hi].copy_(ad.grad, non_blocking=True)
                gH_out[lo:hi].copy_(Hd.grad, non_blocking=True)
                keep.append((cd, ad, Hd, y))
                done = torch.cuda.Event()
                done.record(self.s_out)
        if done is not None:
            done.synchronize()                          # the host now owns the gradients
        del keep
