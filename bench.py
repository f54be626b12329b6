#!/usr/bin/env python
"""Benchmark of the DDSP synthesis hot path on B200 (contract: see the task statement / DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[3], SURVEY.md §8.0 cfg 4): 1024 voices x 4 s @ 48 kHz, hop 512 (T=375,
N=192000), K=256 harmonics, NB=65 noise bands; one step = oscillator bank + filtered noise forward and
backward (gradients to c, a, H from a fixed upstream gradient).  The 1024 voices are sharded over the
ranks (strong scaling, no data-path collective).  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "synth_audio_samples_per_sec_fwd_bwd"
UNIT = "samples/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--voices", type=int, default=1024, help="total voices over all ranks")
    p.add_argument("--harmonics", type=int, default=256)
    p.add_argument("--bands", type=int, default=65)
    p.add_argument("--sample-rate", type=int, default=48000)
    p.add_argument("--seconds", type=float, default=4.0)
    p.add_argument("--hop", type=int, default=512)
    p.add_argument("--cpu-voices", type=int, default=8, help="voices in the bounded CPU sample (SURVEY 8d: an 8-voice chunk)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying a CUDA graph")
    p.add_argument("--e2e-chunk", type=int, default=0, help="voices per chunk of the host pipeline (0: about 16 chunks per rank)")
    p.add_argument("--overlap-noise", action="store_true", help="issue the noise synthesiser on a second stream (its backward then overlaps the oscillator's)")
    p.add_argument("--equal-count-shards", action="store_true", help="shard by voice count instead of by work")
    p.add_argument("--shard-of", default="", help="diagnostic: R/W = run alone, on one GPU, the shard rank R of a W-rank job would get")
    return p.parse_args()


def shapes(args):
    T = int(args.seconds * args.sample_rate) // args.hop
    return dict(B=args.voices, T=T, N=T * args.hop, K=args.harmonics, NB=args.bands, sr=args.sample_rate, hop=args.hop)


def workload_config(args, sh, extra=None):
    cfg = {"workload": f"cfg4 large-batch synthesis sweep: {sh['B']} voices x {args.seconds:g} s @ {sh['sr']} Hz, "
                       f"K={sh['K']}, NB={sh['NB']}, hop={sh['hop']}; oscillator bank + filtered noise, fwd+bwd",
           "voices": sh["B"], "frames": sh["T"], "samples_per_voice": sh["N"], "harmonics": sh["K"],
           "noise_bands": sh["NB"], "sample_rate": sh["sr"], "hop": sh["hop"],
           "cache": "per-step working set (c 0.39 GB + g/y/z 0.79 GB each at 1 GPU) exceeds the 126 MB L2; no explicit flush",
           "parallelism": f"voices sharded over {args.gpus} rank(s), no collective",
           "launch": ("eager" if args.no_graph else "CUDA graph replay of the step") + (", noise synthesiser on a second stream" if args.overlap_noise else "")}
    if extra:
        cfg.update(extra)
    return cfg


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.samples = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-i", str(index), "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                                         text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        rows = [ln for (ts, ln) in self.samples if t0 <= ts <= t1 + 0.05] or [ln for (_, ln) in self.samples[-3:]]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in rows:
            parts = [x.strip() for x in ln.split(",")]
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (torch restatement of the reference's own CPU path) on host cores
# ------------------------------------------------------------------------------------------------
def cpu_step(O, f0, c, a, H, g, sh):
    """One fwd+bwd of oscillator + filtered noise exactly as the reference runs it on CPU, including its
    CPU noise draw (filtered_noise.py:44-48)."""
    c = c.detach().requires_grad_(True)
    a = a.detach().requires_grad_(True)
    H = H.detach().requires_grad_(True)
    noise = torch.rand(H.shape[0], H.shape[1], sh["hop"]) * 2 - 1
    y = O.oscillator_bank(f0, c, a, sh["K"], sh["sr"], sh["hop"]) + O.filtered_noise(H, sh["hop"], noise)
    y.backward(g)
    return y


def cpu_sample_voices(args, sh) -> int:
    """Voices in the bounded CPU sample: --cpu-voices, cut down so the fp32 autograd graph of the oracle (about a dozen
    live [voices, N, K] tensors) stays below a third of the host memory that is free right now."""
    want = max(1, min(args.cpu_voices, sh["B"]))
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 64 << 30
    per_voice = 14 * 4.0 * sh["N"] * sh["K"]
    return max(1, min(want, int(avail / 3 / per_voice)))


def cpu_measure(args, sh, steps: int, warmup: int):
    from oracle import ddsp_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    Bc = cpu_sample_voices(args, sh)
    f0, c, a, H = O.synthetic_controls(Bc, sh["T"], sh["K"], sh["NB"], sh["sr"], sh["hop"], seed=1234 + 4)
    g = torch.randn(Bc, sh["N"], generator=torch.Generator().manual_seed(99))
    for _ in range(warmup):
        cpu_step(O, f0, c, a, H, g, sh)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        cpu_step(O, f0, c, a, H, g, sh)
        times.append(time.perf_counter() - t0)
    mean = sum(times) / len(times)
    return {"value": Bc * sh["N"] / mean, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{Bc} of the {sh['B']} voices (same shapes), oracle/ddsp_oracle.py in fp32 torch eager incl. the reference's CPU "
                      f"noise draw, mean of {steps} step(s) after {warmup} warm-up ({mean:.2f} s/step, best {min(times):.2f} s)",
            "ms_per_step": mean * 1e3, "best_ms_per_step": min(times) * 1e3, "steps_run": steps, "warmup_run": warmup,
            "voices": Bc}


_JSON_FD = None


def claim_stdout():
    """Keep the real stdout for the ONE JSON line and send everything else written to fd 1 (NCCL's version banner on the first
    communicator, library chatter) to stderr."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    payload = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _JSON_FD is None:
        sys.stdout.buffer.write(payload)
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, payload)


def run_reference(args):
    """The reference's CPU implementation of the path (the oracle port: /root/reference does not travel to the GPU box) on
    all host threads; every step is one fwd+bwd of an 8-voice chunk of the workload and exactly --steps steps
    after --warmup warm-up steps are run and reported."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sh = shapes(args)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    res = cpu_measure(args, sh, steps, warmup)
    line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, sh),
            "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "step_definition": f"one step = fwd+bwd of a {res['voices']}-voice chunk; value = chunk samples / mean step time"}
    emit(line)


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def active_harmonic_samples(f0, K, sr, hop):
    """Sum over (voice, hop segment) of samples x harmonics that are not Nyquist-masked in BOTH frames of the
    segment -- the harmonic-samples the synthesis / gradient kernels actually evaluate."""
    B, T = f0.shape[0], f0.shape[1]
    k = torch.arange(1, K + 1, device=f0.device, dtype=torch.float32)
    kact = ((k.view(1, 1, -1) * f0.view(B, T, 1)) <= float(sr // 2)).sum(-1)              # [B,T], mask is monotone in k
    left = torch.cat([kact[:, :1], kact], dim=1)                                          # segment s: frames s-1 | s
    right = torch.cat([kact, kact[:, -1:]], dim=1)
    kmax = torch.maximum(left, right).double()                                            # [B,T+1]
    n = torch.full((T + 1,), float(hop), dtype=torch.float64, device=f0.device)
    n[0] = n[-1] = hop / 2
    return float((kmax * n).sum())


def time_call(fn, reps=10, warm=2):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


def run_b200(args):
    import torch.distributed as dist
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200 import _lib, ops
    from ddsp_pytorch_b200.host_pipeline import HostSynthPipeline, bind_to_gpu_numa_node, default_chunk
    from ddsp_pytorch_b200.sharding import balanced_voice_shards, voice_shard, voice_work
    from ddsp_pytorch_b200.synthetic import synthetic_controls, synthetic_f0

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local)                 # before any pinned allocation: host buffers local to the GPU's PCIe root
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    sh = shapes(args)
    T, N, K, NB, sr, hop = sh["T"], sh["N"], sh["K"], sh["NB"], sh["sr"], sh["hop"]
    conf = d.Config(sample_rate=sr, hop_length=hop, n_harmonics=K, n_noise_filters=NB)
    data_seed = 1234 + 4

    # ---- voice shards: contiguous ranges of equal WORK (active harmonics depend on pitch), no data-path collective ----
    work = voice_work(synthetic_f0(sh["B"], T, sr, hop, seed=data_seed), K, sr, hop)      # every rank: all voices' pitch tracks
    if args.equal_count_shards:
        ranges = [voice_shard(sh["B"], world, r) for r in range(world)]
    else:
        ranges = balanced_voice_shards(work.tolist(), world)
    if args.shard_of:
        er, ew = (int(v) for v in args.shard_of.split("/"))
        ranges = [(balanced_voice_shards(work.tolist(), ew) if not args.equal_count_shards else [voice_shard(sh["B"], ew, r) for r in range(ew)])[er]]
    loads = [float(work[lo:hi].sum()) for lo, hi in ranges]
    job_voices = sum(hi - lo for lo, hi in ranges)      # = sh["B"] except under --shard-of
    v_lo, v_hi = ranges[rank]
    Bl = v_hi - v_lo

    # ---- synthetic inputs, generated on the host, this rank's voices only ------------
    f0h, ch, ah, Hh = synthetic_controls(Bl, T, K, NB, sr, hop, seed=data_seed, voice_offset=v_lo)
    gh = torch.empty(Bl, N)
    for b in range(Bl):                                 # upstream gradient keyed by global voice: the same for every sharding
        gh[b] = torch.randn(N, generator=torch.Generator().manual_seed(99_000_000 + v_lo + b))
    host = [t.pin_memory() for t in (f0h, ch, ah, Hh, gh)]
    f0, c, a, H, g = (t.to(dev) for t in host)
    c.requires_grad_(True); a.requires_grad_(True); H.requires_grad_(True)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_floats(vals):
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        if world == 1:
            return [t.tolist()]
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [o.tolist() for o in out]

    seed, voff = 7, v_lo
    calls = torch.zeros(1, dtype=torch.int32, device=dev)      # device-side call counter: fresh Philox offset per step
    lib = _lib.load()

    side_noise = torch.cuda.Stream(dev) if args.overlap_noise else None

    def step_body():
        """resident inputs -> y = noise(H); y += osc(f0,c,a); backward from g."""
        calls.add_(1)
        if side_noise is not None:
            # the noise synthesiser on its own stream: autograd runs a node's backward on its forward's stream, so the noise
            # gradient kernel runs beside the oscillator's gradient kernels and fills their tails
            cur = torch.cuda.current_stream()
            side_noise.wait_stream(cur)
            with torch.cuda.stream(side_noise):
                y = ops.filtered_noise(H, hop, seed=seed, offset=0, voice_offset=voff, offset_tensor=calls)
            cur.wait_stream(side_noise)
            y.record_stream(cur)
        else:
            y = ops.filtered_noise(H, hop, seed=seed, offset=0, voice_offset=voff, offset_tensor=calls)
        y = ops.oscillator_bank(f0, c, a, sr, hop, out=y)
        y.backward(g)
        return y

    def eager_step():
        c.grad = a.grad = H.grad = None
        return step_body()

    n0 = lib.ddsp_launch_count()
    eager_step()
    launches_per_step = int(lib.ddsp_launch_count() - n0)       # kernels of libddsp_b200.so in one step (counted by the library)
    step = eager_step
    if not args.no_graph:
        # the step is launch-bound once the voices are spread over 8 GPUs (1.3 ms of kernels): capture it once and
        # replay it; the gradients land in the same buffers on every replay, the noise counter advances on the device
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                eager_step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        c.grad = a.grad = H.grad = None
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            step_body()
        step = graph.replay

    sampler = ClockSampler(local) if rank == 0 else None
    t_load0 = time.time()
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    my_ms = e0.elapsed_time(e1)
    ms = max_over_ranks(my_ms)
    clocks = sampler.stop(t_load0, time.time()) if sampler else None
    value = job_voices * N * args.steps / (ms * 1e-3)
    rank_ms = [v[0] / args.steps for v in gather_floats([my_ms])]

    # ---- end to end through the package's host pipeline: pinned host buffers in, pinned host buffers out ----------
    e2e = None
    if not args.no_e2e:
        chunk = args.e2e_chunk if args.e2e_chunk > 0 else default_chunk(Bl)
        pipe = HostSynthPipeline(conf, dev, chunk_voices=chunk, seed=seed, voice_offset=voff)
        outs = [torch.empty(s_, dtype=torch.float32).pin_memory() for s_ in ((Bl, T, K), (Bl, T, 1), (Bl, T, NB))]
        y_host = torch.empty(Bl, N, dtype=torch.float32).pin_memory()
        h2d = sum(t.numel() * 4 for t in host)
        d2h = sum(t.numel() * 4 for t in outs) + y_host.numel() * 4

        def e2e_step():
            # public API call: pinned host controls + upstream gradient in; synthesised audio and the three gradients
            # out to pinned host memory; returns once the host owns them
            pipe.step(*host, *outs, y_out=y_host)

        for _ in range(max(args.warmup, 3)):
            e2e_step()
        barrier()
        e0.record()
        for _ in range(args.steps):
            e2e_step()
        e1.record()
        barrier()
        my_ems = e0.elapsed_time(e1)
        ems = max_over_ranks(my_ems)
        # host-link roofline of the same step: the same chunked copies on the same streams with no kernel between them
        for _ in range(2):
            pipe.step(*host, *outs, y_out=y_host, copy_only=True)
        barrier()
        e0.record()
        for _ in range(args.steps):
            pipe.step(*host, *outs, y_out=y_host, copy_only=True)
        e1.record()
        barrier()
        copy_ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
        per_rank = gather_floats([my_ems / args.steps, float(h2d), float(d2h), float(Bl), -2.0 if numa is None else float(numa)])
        e2e = {"value": job_voices * N * args.steps / (ems * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(sum(r[1] for r in per_rank)), "d2h_bytes_per_step": int(sum(r[2] for r in per_rank)),
               "ms_per_step": ems / args.steps, "copy_only_ms_per_step": copy_ms,
               "frac_of_copy_roofline": copy_ms / (ems / args.steps),
               "bound": "host link: the step's H2D + D2H copies alone, chunked the same way on the same streams, take copy_only_ms_per_step",
               "chunk_voices": chunk, "chunks_per_rank": -(-Bl // chunk),
               "launches_per_step_rank0": pipe.launches_per_step,
               "per_rank": [{"voices": int(r[3]), "ms_per_step": r[0], "h2d_GBps": r[1] / (r[0] * 1e-3) / 1e9,
                             "d2h_GBps": r[2] / (r[0] * 1e-3) / 1e9, "cpu_binding": "none" if r[4] < -1.5 else ("GPU-local CPUs (NVML mask)" if r[4] < 0 else f"NUMA node {int(r[4])}")} for r in per_rank],
               "result": "synthesised audio y and the gradients of c, a, H (the forward's and the backward's outputs) in pinned host memory",
               "path": "HostSynthPipeline.step: pinned host f0,c,a,H,g -> voice chunks on copy-in / compute / copy-out streams "
                       "(ddsp_noise_forward + ddsp_osc_forward(accumulate) + ddsp_osc_backward + ddsp_noise_backward per chunk) "
                       "-> y, gc, ga, gH in pinned host memory"}

    # ---- per-kernel device times (this rank's shard) and the roofline of the dominant kernel -------
    st = torch.cuda.current_stream().cuda_stream
    state = torch.empty(lib.ddsp_osc_state_bytes(Bl, T, K), dtype=torch.uint8, device=dev)
    scratch = torch.empty(lib.ddsp_osc_backward_scratch_bytes(Bl, T, K), dtype=torch.uint8, device=dev)
    y = torch.empty(Bl, N, device=dev); z = torch.empty(Bl, N, device=dev)
    gc = torch.empty(Bl, T, K, device=dev); ga = torch.empty(Bl, T, device=dev); gH = torch.empty(Bl, T, NB, device=dev)
    P = lambda t: t.data_ptr()  # noqa: E731
    cd_, ad_, Hd_ = c.detach(), a.detach(), H.detach()
    kcalls = {
        "osc_prepare_kernel": lambda: _lib.call("ddsp_osc_prepare", P(f0), P(cd_), P(state), Bl, T, K, hop, sr, st),
        "osc_synth_kernel": lambda: _lib.call("ddsp_osc_synthesize", P(f0), P(ad_), P(state), P(y), P(z), 0, Bl, T, K, hop, sr, st),
        "osc_grad_kernel": lambda: _lib.call("ddsp_osc_grad_segments", P(g), P(f0), P(ad_), P(z), P(state), P(scratch), Bl, T, K, hop, sr, st),
        "osc_finalize_kernel": lambda: _lib.call("ddsp_osc_grad_finalize", P(f0), P(state), P(scratch), P(gc), P(ga), Bl, T, K, hop, sr, st),
        "noise_fwd_kernel": lambda: _lib.call("ddsp_noise_forward", P(Hd_), None, P(y), 0, seed, 0, None, voff, Bl, T, NB, hop, st),
        "noise_bwd_kernel": lambda: _lib.call("ddsp_noise_backward", P(g), None, P(gH), seed, 0, None, voff, Bl, T, NB, hop, st),
    }
    kernels_ms = {name: time_call(fn) for name, fn in kcalls.items()}
    fp32_peak = ops.probe_fp32_fma_tflops()
    props = torch.cuda.get_device_properties(dev)
    sm_max_mhz = (clocks or {}).get("sm_max_mhz") or 1965.0
    fp32_nominal = props.multi_processor_count * 128 * 2 * sm_max_mhz * 1e6 / 1e12       # 128 FMA lanes per SM per clock
    active = active_harmonic_samples(f0.view(Bl, T), K, sr, hop)
    dom = max(kernels_ms, key=kernels_ms.get)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    # DRAM bytes of one launch of the dominant kernel from the ncu --set full capture of THIS shard size (profiles/traffic_r02.json
    # is keyed by kernel and voices per launch); absent for shard sizes that were not captured
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic_r02.json"))).get(dom, {}).get(str(Bl))
    except (OSError, ValueError, AttributeError):
        pass
    roofline = None
    # SURVEY §8d counts 6 flop (forward) / 12 flop (backward) per harmonic-sample; both kernels EXECUTE 4 FMA-pipe lane operations
    # per harmonic-sample (recurrence FFMA2 + FADD2, two dot-product FFMA2), i.e. 8 flop of pipe capacity.  `frac` is the executed
    # figure (the one ncu's sm__pipe_fma_cycles_active must agree with); the SURVEY-count figure is printed beside it.
    survey_flop = {"osc_synth_kernel": 6.0, "osc_grad_kernel": 12.0}.get(dom)
    if survey_flop:
        sec = kernels_ms[dom] * 1e-3
        executed = 8.0 * active / sec / 1e12
        algo_bytes = 4.0 * (Bl * N * 2 + Bl * T * (K + 2))
        roofline = {"kernel": dom, "bound": "fp32", "achieved": executed, "peak": fp32_peak, "unit": "TFLOP/s",
                    "frac": executed / fp32_peak, "traffic": traffic, "launch_ms": kernels_ms[dom],
                    "peak_source": "FFMA-chain probe (ddsp_probe_fp32_fma) measured in this run; MEASURED_PEAKS.json has no FP32 figure",
                    "peak_nominal": fp32_nominal, "frac_of_nominal": executed / fp32_nominal,
                    "work": f"8 flop of FMA-pipe capacity (4 lane operations) x {active:.4g} active harmonic-samples on this rank "
                            f"(nominal K would be {Bl * N * K:.4g}; Nyquist-masked harmonics are skipped, as they contribute 0 in the reference)",
                    "frac_survey_count": survey_flop * active / sec / 1e12 / fp32_peak,
                    "survey_count": {"flop_per_harmonic_sample": survey_flop, "tflops": survey_flop * active / sec / 1e12},
                    "hbm": {"algorithmic_bytes": algo_bytes, "achieved_GBps": algo_bytes / sec / 1e9, "peak_GBps": peaks.get("hbm_gbs")}}

    # fused oscillator + noise, forward + backward (the north star's >= 50 % target): SURVEY §8d flop count / this rank's step time
    M = 2 * (NB - 1)
    fused_flop = 18.0 * active + 2.0 * (M + M * NB / hop) * Bl * N
    my_step_s = my_ms / args.steps * 1e-3
    fused = {"achieved": fused_flop / my_step_s / 1e12, "peak": fp32_peak, "unit": "TFLOP/s",
             "frac": fused_flop / my_step_s / 1e12 / fp32_peak, "frac_of_nominal": fused_flop / my_step_s / 1e12 / fp32_nominal,
             "work": "SURVEY 8d count: 18 flop x active harmonic-samples + 2(M + M*NB/hop) flop x samples on rank 0, over rank 0's step"}
    if world > 1:
        dist.barrier(device_ids=[local])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    rt = None
    if world == 1:      # BASELINE metric, third part: real-time block latency (cfg 3), a 1-second side measurement
        import types
        import bench_extra
        r = bench_extra.bench_rt(types.SimpleNamespace(rt_calls=500), dev)
        rt = {k: r[k] for k in ("workload", "deadline_ms", "host_p50_ms", "host_p99_ms", "device_p50_ms", "device_p99_ms")}
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        r = cpu_measure(args, sh, steps=3, warmup=1)
        cpu = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    mean_load = sum(loads) / len(loads)
    sharding = {"policy": "equal voice counts" if args.equal_count_shards else "contiguous ranges of equal work (sharding.balanced_voice_shards)",
                "voices_per_rank": [hi - lo for lo, hi in ranges], "work_max_over_mean": max(loads) / mean_load,
                "work_per_rank_over_mean": [x / mean_load for x in loads], "ms_per_step_per_rank": rank_ms}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args, sh), "clocks": clocks, "e2e": e2e,
            "gpu_launches": launches_per_step * args.steps, "gpu_launches_per_step": launches_per_step,
            # bits 0 / 1 / 2: synthesis / gradient / pitch-gradient kernel run from the hand-scheduled image (sass_sched.py)
            "tuned_sass": int(lib.ddsp_osc_tuned_sass(-1)),
            "roofline": roofline, "roofline_fused": fused, "cpu_baseline": cpu, "kernels_ms": kernels_ms,
            "fp32_fma_peak_tflops_measured": fp32_peak, "fp32_fma_peak_tflops_nominal": fp32_nominal, "sharding": sharding,
            "rt_block_latency": rt}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a_ = parse_args()
    claim_stdout()
    if a_.impl == "reference":
        run_reference(a_)
    else:
        run_b200(a_)
