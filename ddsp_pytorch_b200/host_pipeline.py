"""Host-resident controls -> oscillator bank + filtered noise, forward and backward -> host-resident audio and gradients.

The synthesis sweep (SURVEY.md §8.0 cfg 4) is bound by PCIe once the controls start on the host: per 1024 voices 1.28 GB
go in (f0, c, a, H and the upstream gradient g) and 1.28 GB come out (the synthesised audio y and the gradients of c, a,
H) against 10 ms of kernels.  This driver cuts the voices into chunks and runs three CUDA streams -- copy-in, compute,
copy-out -- so that the three overlap (PCIe is full duplex).  The arithmetic is what `ops.filtered_noise` +
`ops.oscillator_bank(out=...)` + backward do (decoder.py:129-132), issued straight through the C ABI into buffers owned
by the pipeline's slots: with chunks of 8-64 voices the step is made of 16+ chunks and the per-chunk host cost (autograd
bookkeeping, allocations) would otherwise bound a rank that only has 128 voices.  Philox counters are keyed by the global
voice index, so the result does not depend on the chunking (tests/test_gpu_pipeline.py).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import _lib


def _gpu_cpu_set(device_index: int):
    """(numa node or None, set of CPUs local to the GPU): sysfs first, NVML's ideal-CPU mask when the PCI device reports
    no node (virtual machines)."""
    props = torch.cuda.get_device_properties(device_index)
    try:
        path = f"/sys/bus/pci/devices/{props.pci_domain_id:04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node >= 0:
            cpus = set()
            for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
            return node, cpus
    except (OSError, ValueError, AttributeError):
        pass
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            uuid = str(props.uuid)
            handle = None
            for i in range(pynvml.nvmlDeviceGetCount()):
                h = pynvml.nvmlDeviceGetHandleByIndex(i)
                u = pynvml.nvmlDeviceGetUUID(h)
                u = u.decode() if isinstance(u, bytes) else u
                if uuid in u or u.replace("GPU-", "") == uuid:
                    handle = h
                    break
            if handle is None:
                handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            words = (os.cpu_count() + 63) // 64
            mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
            cpus = {w * 64 + bit for w, m in enumerate(mask) for bit in range(64) if (int(m) >> bit) & 1}
            node = None
            try:
                nm = pynvml.nvmlDeviceGetMemoryAffinity(handle, 4, pynvml.NVML_AFFINITY_SCOPE_NODE)
                nodes = [w * 64 + bit for w, m in enumerate(nm) for bit in range(64) if (int(m) >> bit) & 1]
                node = nodes[0] if nodes else None
            except Exception:
                pass
            return node, cpus
        finally:
            pynvml.nvmlShutdown()
    except Exception:
        return None, set()


def bind_to_gpu_numa_node(device_index: int) -> Optional[int]:
    """Pin the calling process to the CPUs next to the GPU, so that pinned host buffers allocated afterwards (first touch)
    are local to the GPU's PCIe root.  Returns the NUMA node (None when it is unknown or nothing was bound).  Eight ranks
    that all allocate from the node they happened to start on push every copy of half the GPUs across the socket
    interconnect."""
    try:
        node, cpus = _gpu_cpu_set(device_index)
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return node if node is not None else -1
    except (OSError, RuntimeError, AttributeError):
        return None


def default_chunk(voices: int) -> int:
    """About 16 chunks per step whatever the shard size (enough for the three streams to overlap), at least 4 and at
    most 64 voices per chunk (larger chunks only lengthen the un-overlapped first copy-in and last copy-out)."""
    return max(4, min(64, -(-int(voices) // 16)))


class HostSynthPipeline:
    def __init__(self, conf, device, chunk_voices: Optional[int] = None, slots: int = 3, seed: int = 0, voice_offset: int = 0):
        self.sr, self.hop = int(conf.sample_rate), int(conf.hop_length)
        self.K, self.NB = int(conf.n_harmonics), int(conf.n_noise_filters)
        self.device = torch.device(device)
        self.chunk = None if chunk_voices is None else int(chunk_voices)
        self.slots = int(slots)
        self.seed, self.voice_offset = int(seed) & (2 ** 64 - 1), int(voice_offset)
        self.calls = 0
        self.s_in, self.s_run, self.s_out = (torch.cuda.Stream(self.device) for _ in range(3))
        self._bufs = None
        self._key = None
        self.launches_per_step = 0

    def _alloc(self, B: int, T: int):
        C = self.chunk if self.chunk is not None else default_chunk(B)
        if self._key == (C, T):
            return C
        N, dev = T * self.hop, self.device
        lib = _lib.load()
        mk = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)  # noqa: E731
        raw = lambda n: torch.empty(max(int(n), 1), dtype=torch.uint8, device=dev)  # noqa: E731
        with torch.cuda.device(dev):
            self._bufs = [{"f0": mk(C, T, 1), "c": mk(C, T, self.K), "a": mk(C, T, 1), "H": mk(C, T, self.NB), "g": mk(C, N),
                           "y": mk(C, N), "z": mk(C, N), "gc": mk(C, T, self.K), "ga": mk(C, T, 1), "gH": mk(C, T, self.NB),
                           "state": raw(lib.ddsp_osc_state_bytes(C, T, self.K)),
                           "scratch": raw(lib.ddsp_osc_backward_scratch_bytes(C, T, self.K))}
                          for _ in range(self.slots)]
        self._key = (C, T)
        return C

    def step(self, f0, c, a, H, g, gc_out, ga_out, gH_out, y_out=None, copy_only: bool = False):
        """All arguments are PINNED host tensors: f0 [B,T,1], c [B,T,K], a [B,T,1], H [B,T,NB], g [B,N] in; y_out [B,N]
        (optional) receives the synthesised audio, gc_out / ga_out / gH_out receive d sum(g*y)/d c, a, H.  Returns when
        the host owns the results.  copy_only (measurement): the same copies in the same order with no kernel between
        them -- the host-link roofline of the step."""
        B, T = c.shape[0], c.shape[1]
        if B == 0:
            return
        C = self._alloc(B, T)
        K, NB, hop, sr = self.K, self.NB, self.hop, self.sr
        offset = self.calls & 0xFFFFFFFF
        self.calls += 1
        lib_call = _lib.call
        count0 = _lib.load().ddsp_launch_count()
        in_free = [None] * self.slots                   # event: the slot's device inputs may be overwritten
        out_free = [None] * self.slots                  # event: the slot's device results have left for the host
        done = None
        P = lambda t: t.data_ptr()  # noqa: E731
        with torch.cuda.device(self.device):
            for i, lo in enumerate(range(0, B, C)):
                hi = min(B, lo + C)
                n, slot = hi - lo, i % self.slots
                buf = self._bufs[slot]
                with torch.cuda.stream(self.s_in):
                    if in_free[slot] is not None:
                        self.s_in.wait_event(in_free[slot])
                    for name, src in (("f0", f0), ("c", c), ("a", a), ("H", H), ("g", g)):
                        buf[name][:n].copy_(src[lo:hi], non_blocking=True)
                    ready = torch.cuda.Event()
                    ready.record(self.s_in)
                with torch.cuda.stream(self.s_run):
                    self.s_run.wait_event(ready)
                    if out_free[slot] is not None:
                        self.s_run.wait_event(out_free[slot])
                    st = self.s_run.cuda_stream
                    voff = (self.voice_offset + lo) & 0xFFFFFFFF
                    if not copy_only:
                        # y = noise(H); y += osc(f0, c, a)  (decoder.py:132), then the two backward passes from g
                        lib_call("ddsp_noise_forward", P(buf["H"]), None, P(buf["y"]), 0, self.seed, offset, None, voff, n, T, NB, hop, st)
                        lib_call("ddsp_osc_forward", P(buf["f0"]), P(buf["c"]), P(buf["a"]), P(buf["y"]), P(buf["z"]), P(buf["state"]),
                                 1, n, T, K, hop, sr, st)
                        lib_call("ddsp_osc_backward", P(buf["g"]), P(buf["f0"]), P(buf["a"]), P(buf["z"]), P(buf["state"]),
                                 P(buf["scratch"]), P(buf["gc"]), P(buf["ga"]), n, T, K, hop, sr, st)
                        lib_call("ddsp_noise_backward", P(buf["g"]), None, P(buf["gH"]), self.seed, offset, None, voff, n, T, NB, hop, st)
                    computed = torch.cuda.Event()
                    computed.record(self.s_run)
                    in_free[slot] = computed
                with torch.cuda.stream(self.s_out):
                    self.s_out.wait_event(computed)
                    if y_out is not None:
                        y_out[lo:hi].copy_(buf["y"][:n], non_blocking=True)
                    gc_out[lo:hi].copy_(buf["gc"][:n], non_blocking=True)
                    ga_out[lo:hi].copy_(buf["ga"][:n], non_blocking=True)
                    gH_out[lo:hi].copy_(buf["gH"][:n], non_blocking=True)
                    done = torch.cuda.Event()
                    done.record(self.s_out)
                    out_free[slot] = done
        if not copy_only:                               # (the copy-only measurement launches nothing and must not reset the count)
            self.launches_per_step = int(_lib.load().ddsp_launch_count() - count0)
        if done is not None:
            done.synchronize()                          # the host now owns the audio and the gradients
