"""Headless real-time block pipeline: what rt/synth.py:21-55 does around
`Decoder.forward_live` (decoder.py:137-147), without JACK.  The CREPE/loudness encoder is out of scope, so a block
enters as the feature frames the encoder would emit (`f0`, `normalized_cents`, `loudness`, each [1, frames, 1]) and
leaves as `frames * hop` float32 samples in a pinned host buffer.  State carried across blocks: the GRU hidden
state, the oscillator's per-harmonic phases and the reverb's 1-second ring buffer -- all inside the modules, as in
the reference."""
from __future__ import annotations

import time
from typing import Dict, List

import numpy as np
import torch

from . import ops
from .controller import Decoder


class RealtimeSynth:
    def __init__(self, conf, device="cuda:0", state_dict=None, frames_per_block: int = 4, noise_seed: int = 0,
                 graph: bool = False):
        """graph=True: from the third block on, the whole block (controller, synthesis, copy to the pinned output buffer)
        is one CUDA-graph replay; the first two blocks run eagerly (they also warm up cuDNN/cuBLAS for the capture).
        All state -- oscillator phases, reverb buffer, Philox call counter -- is updated in place on the device, so the
        eager and the replayed blocks form one continuous stream."""
        self.conf, self.device = conf, torch.device(device)
        self.graph = bool(graph)
        self.decoder = Decoder(conf, noise_seed=noise_seed, device_counter=True).to(self.device).eval()
        if state_dict is not None:
            self.decoder.load_state_dict(state_dict, strict=True)          # rt/synth.py:17
        self.frames = int(frames_per_block)
        self.block = self.frames * conf.hop_length
        g = torch.Generator().manual_seed(0)
        self.hidden = torch.randn(conf.decoder_gru_layers, 1, conf.decoder_gru_units, generator=g).to(self.device)  # rt/synth.py:22
        self.out = torch.empty(self.block, dtype=torch.float32).pin_memory()
        self.deadline_s = self.block / conf.sample_rate                    # rt/synth.py:54
        self.latencies_ms: List[float] = []
        self.missed = 0
        self.blocks = 0
        self._stream = torch.cuda.Stream(self.device)
        self._z = {k: torch.zeros(1, self.frames, 1, device=self.device) for k in ("f0", "normalized_cents", "loudness")}
        self._g = None

    def _block(self) -> None:
        """static features -> `self.out`; every state update in place (graph-safe)."""
        dec = self.decoder
        ctrl, _ = dec.controller(self._z, self.hidden)
        # the reference's controller hands back the hidden state it was given (decoder.py:106-107), so rt/synth.py's
        # `hidden[...] = ...` is a no-op; the state that does advance lives in the three modules below
        bank = dec.harmonics
        if bank.last_phases.dtype == torch.float32:                        # after the first block: keep the state's address
            dry, last = ops.oscillator_live(ctrl["f0"], ctrl["c"], ctrl["a"], bank.last_phases.data, bank.sample_rate,
                                            bank.hop_size)
            bank.last_phases.data.copy_(last)
        else:
            dry = bank.live(ctrl)
        audio = dec.reverb.live_forward(dry + dec.noise(ctrl))
        self.out.copy_(audio.view(-1), non_blocking=True)

    @torch.no_grad()
    def process(self, z: Dict[str, torch.Tensor]) -> np.ndarray:
        """One block: features (host or device tensors [1, frames, 1]) in, `block` samples out (a view of the pinned
        output buffer, valid until the next call)."""
        t0 = time.perf_counter()
        with torch.cuda.stream(self._stream):
            for k, dst in self._z.items():
                dst.copy_(z[k].reshape(1, self.frames, 1), non_blocking=True)
            if self.graph and self._g is None and self.blocks >= 2:
                self._stream.synchronize()
                self._g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._g, stream=self._stream):
                    self._block()
            if self._g is not None:
                self._g.replay()
            else:
                self._block()
        self._stream.synchronize()                                         # the audio callback needs the samples now
        self.blocks += 1
        dt = time.perf_counter() - t0
        self.latencies_ms.append(dt * 1e3)
        self.missed += dt >= self.deadline_s
        return self.out.numpy()

    def stats(self) -> Dict[str, float]:
        v = sorted(self.latencies_ms)
        q = lambda p: v[min(len(v) - 1, int(p * len(v)))] if v else float("nan")  # noqa: E731
        return {"blocks": len(v), "deadline_ms": self.deadline_s * 1e3, "p50_ms": q(0.5), "p99_ms": q(0.99),
                "max_ms": v[-1] if v else float("nan"), "missed": int(self.missed)}
