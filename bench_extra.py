#!/usr/bin/env python
"""Side measurements for the SURVEY.md §8d rows that are not the headline bench line (bench.py keeps the driver
contract).  One JSON line per workload:

    python bench_extra.py [--workloads mss,reverb,path,rt] [--reps 20]

  mss     cfg 5: MSS loss alone, FFT sizes 2048..64 on 64 x 4 s @ 16 kHz clips, forward + backward
  reverb  cfg 1' shapes (B=16, N=88064, L=44100) and a large batch (B=256, N=192000, L=48000): fwd + bwd (gx, gir)
  path    cfg 1': oscillator + noise + reverb + 6-scale MSS, forward + backward through the module API
  ctrl    row X1: fused Controller tail vs the torch expression at cfg 2's shapes
  rtdec   cfg 3 with the Controller in front (RealtimeSynth), eager vs CUDA-graph replay
  rt      cfg 3: OscillatorBank.live + FilteredNoise + Reverb.live_forward per 2048-sample block, p50/p99 latency
  cfg1    cfg 1 as BASELINE names it (16 kHz, K=100, NB=65, T=125): whole path fwd+bwd
  rt3p    cfg 3' (64 / 128 / 256-sample blocks, one frame per block)
  sweep   cfg 4 at K = 100, 128, 180, 256
  dftgemm north_star item 3: tcgen05 DFT-as-GEMM forward of one scale vs the register-tiled FFT kernel
  cpu     the reference's CPU path (oracle port) for cfg 1, 1', 5 and cfg 3 p50/p99 -- the cpu_baseline leg of the side rows
No CPU fallback: needs the CUDA library and a GPU.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def ev_time(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        return {}


def bench_mss(args, dev):
    from ddsp_pytorch_b200 import ops
    B, N, n_ffts = 64, 64000, [2048, 1024, 512, 256, 128, 64]
    g = torch.Generator().manual_seed(1234 + 5)
    xt = (torch.randn(B, N, generator=g) * 0.1).clamp(-1, 1)
    xp = (0.5 * xt + torch.randn(B, N, generator=g) * 0.05)
    xt, xp = xt.to(dev), xp.to(dev).requires_grad_(True)

    def step():
        xp.grad = None
        ops.mss_loss(xp, xt, n_ffts).backward()

    ms = ev_time(step, args.reps)
    fwd_only = ev_time(lambda: ops.mss_loss(xp.detach(), xt, n_ffts), args.reps)
    samples = B * N
    flops = (1020 + 510) * samples            # FFT flops fwd (2 signals) + one inverse per pred frame
    byts = (48 + 52) * samples
    pk = peaks()
    return {"workload": "cfg5 MSS alone: 64 x 4 s @ 16 kHz, n_fft 2048..64, fwd+bwd", "ms": ms, "fwd_only_ms": fwd_only,
            "samples_per_s": samples / (ms * 1e-3), "alg_tflops": flops / (ms * 1e-3) / 1e12,
            "alg_GBps": byts / (ms * 1e-3) / 1e9, "hbm_peak_GBps": pk.get("hbm_gbs")}


def bench_reverb(args, dev):
    import ddsp_pytorch_b200 as d
    out = []
    shapes = ((16, 88064, 44100), (256, 192000, 48000))
    for (B, N, sr) in (shapes[1:] if args.reverb_big_only else shapes):
        conf = d.Config(sample_rate=sr, hop_length=512, n_harmonics=100, n_noise_filters=65)
        rv = d.Reverb(conf).to(dev)
        x = (torch.randn(B, N, generator=torch.Generator().manual_seed(3)) * 0.1).to(dev).requires_grad_(True)
        gy = torch.randn(B, N, generator=torch.Generator().manual_seed(4)).to(dev)

        def step():
            x.grad = None
            for p in rv.parameters():
                p.grad = None
            rv(x).backward(gy)

        ms = ev_time(step, args.reps)
        with torch.no_grad():
            fwd = ev_time(lambda: rv(x), args.reps)
        byts = 24.0 * B * N                   # 8 fwd + 8 gx + 8 gir bytes per sample
        out.append({"workload": f"reverb B={B} N={N} L={sr}: fwd+bwd (gx, g_noise, g_decay, g_wet)", "ms": ms, "fwd_only_ms": fwd,
                    "samples_per_s": B * N / (ms * 1e-3), "alg_GBps": byts / (ms * 1e-3) / 1e9,
                    "fwd_alg_GBps": 8.0 * B * N / (fwd * 1e-3) / 1e9, "hbm_peak_GBps": peaks().get("hbm_gbs")})
    return out


def bench_path(args, dev, conf=None, T=172, name="cfg1'"):
    """cfg 1' (shipped config/default.py) or cfg 1 (BASELINE's 16 kHz / K=100 / NB=65): decoder tail exactly as
    decoder.py:129-133 + MSSLoss, fwd+bwd."""
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200.synthetic import synthetic_controls
    conf = conf or d.SHIPPED
    B, sr, hop, K, NB = 16, conf.sample_rate, conf.hop_length, conf.n_harmonics, conf.n_noise_filters
    N = T * hop
    f0, c, a, H = (t.to(dev) for t in synthetic_controls(B, T, K, NB, sr, hop, seed=1234 + 1))
    target = (torch.randn(B, N, generator=torch.Generator().manual_seed(9)) * 0.1).clamp(-1, 1).to(dev)
    bank, fn, rv = d.OscillatorBank(conf).to(dev), d.FilteredNoise(conf, seed=7), d.Reverb(conf).to(dev)
    loss_fn = d.MSSLoss([2048, 1024, 512, 256, 128, 64]).to(dev)
    leaves = [c.requires_grad_(True), a.requires_grad_(True), H.requires_grad_(True)]

    def step():
        for t in leaves + list(rv.parameters()):
            t.grad = None
        ctrl = {"f0": f0, "c": c, "a": a, "H": H}
        dry = bank(ctrl) + fn(ctrl)
        loss_fn(rv(dry), target).backward()

    ms = ev_time(step, args.reps)
    # the eager step is ~35 launches in ~0.7 ms: host-launch bound on a slow host.  Replay it as a CUDA graph for the device time
    # (ev_time above already ran the step on this side stream's allocator pool, so capture allocates nothing new)
    graph_ms = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                step()
        torch.cuda.current_stream().wait_stream(side)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            step()
        graph_ms = ev_time(gr.replay, args.reps)
    except Exception as e:                              # report, do not hide: the eager number stands on its own
        graph_ms = f"capture failed: {type(e).__name__}: {e}"[:200]
    return {"workload": f"{name} path: B={B} T={T} sr={sr} K={K} NB={NB}: osc + noise + reverb + 6-scale MSS, fwd+bwd",
            "ms": ms, "graph_replay_ms": graph_ms, "samples_per_s": B * N / (ms * 1e-3)}


def bench_cfg1(args, dev):
    import ddsp_pytorch_b200 as d
    return bench_path(args, dev, conf=d.CFG1, T=125, name="cfg1")


def bench_rt3p(args, dev):
    """cfg 3' (BASELINE: 64-256-sample blocks): one frame per block, hop = block size, NB capped at hop/2+1."""
    import ddsp_pytorch_b200 as d
    out = []
    for hop in (64, 128, 256):
        conf = d.SHIPPED.with_(hop_length=hop, n_noise_filters=min(195, hop // 2 + 1))
        out.append(bench_rt(args, dev, conf=conf, T=1, name="cfg3'"))
    return out


def bench_rt(args, dev, conf=None, T=4, name="cfg3"):
    """cfg 3 (rt/synth.py:23, autoencoder.py:26-32): one 2048-sample block = 4 frames, batch 1, live path."""
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200.synthetic import synthetic_controls
    conf = conf or d.SHIPPED
    sr, hop, K, NB = conf.sample_rate, conf.hop_length, conf.n_harmonics, conf.n_noise_filters
    bank, fn, rv = d.OscillatorBank(conf).to(dev), d.FilteredNoise(conf, seed=7), d.Reverb(conf).to(dev)
    ctrls = []
    for i in range(16):
        f0, c, a, H = (t.to(dev) for t in synthetic_controls(1, T, K, NB, sr, hop, seed=50 + i))
        ctrls.append({"f0": f0, "c": c, "a": a, "H": H})
    n_calls = args.rt_calls
    host_ms, dev_ms = [], []
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_calls)]
    with torch.no_grad():
        for i in range(50 + n_calls):
            ctrl = ctrls[i % len(ctrls)]
            t0 = time.perf_counter()
            if i >= 50:
                evs[i - 50][0].record()
            y = rv.live_forward(bank.live(ctrl) + fn(ctrl))
            if i >= 50:
                evs[i - 50][1].record()
            out = y.cpu()                                  # decoder.py:147: the JACK callback needs host samples
            if i >= 50:
                host_ms.append((time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
    dev_ms = [a.elapsed_time(b) for a, b in evs]
    q = lambda v, p: sorted(v)[min(len(v) - 1, int(p * len(v)))]  # noqa: E731
    return {"workload": f"{name} rt: batch 1, {T * hop}-sample blocks @ {sr} Hz, live osc + noise + live reverb, {n_calls} calls",
            "deadline_ms": T * hop / sr * 1e3, "host_p50_ms": q(host_ms, 0.5), "host_p99_ms": q(host_ms, 0.99),
            "host_max_ms": max(host_ms), "device_p50_ms": q(dev_ms, 0.5), "device_p99_ms": q(dev_ms, 0.99),
            "out_shape": list(out.shape)}


def bench_rtdec(args, dev):
    """cfg 3 end to end (row X4): features -> Controller (GRU) -> live synthesis -> pinned host samples through
    RealtimeSynth, eager and as a CUDA-graph replay; host wall-clock per 2048-sample block."""
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200.realtime import RealtimeSynth
    conf = d.SHIPPED
    out = []
    for graph in (False, True):
        torch.manual_seed(0)
        rt = RealtimeSynth(conf, dev, frames_per_block=4, noise_seed=7, graph=graph)
        g = torch.Generator().manual_seed(1)
        feats = [{"f0": (torch.rand(1, 4, 1, generator=g) * 400 + 100).pin_memory(),
                  "normalized_cents": torch.rand(1, 4, 1, generator=g).pin_memory(),
                  "loudness": torch.rand(1, 4, 1, generator=g).pin_memory()} for _ in range(16)]
        for i in range(50):
            rt.process(feats[i % 16])
        rt.latencies_ms.clear()
        rt.missed = 0
        for i in range(args.rt_calls):
            rt.process(feats[i % 16])
        st = rt.stats()
        out.append({"workload": f"cfg3 rt decoder block ({rt.block} samples @ {conf.sample_rate} Hz): controller + live synth + D2H, "
                                f"graph={graph}", **st})
    return out


def bench_fft(args, dev):
    from ddsp_pytorch_b200 import _lib
    out = []
    for n, batch in ((4096, 12032), (2048, 65536), (512, 262144), (64, 2097152)):
        x = torch.randn(batch, n, dtype=torch.complex64, device=dev)
        y = torch.empty_like(x)
        st = torch.cuda.current_stream().cuda_stream
        ms = ev_time(lambda: _lib.call("ddsp_fft_c2c", x.data_ptr(), y.data_ptr(), n, batch, 0, 0, st), args.reps)
        ref = ev_time(lambda: torch.fft.fft(x, out=y), args.reps)
        byts = 16.0 * n * batch
        out.append({"workload": f"c2c fft n={n} batch={batch} fp32", "ms": ms, "GBps": byts / (ms * 1e-3) / 1e9,
                    "cufft_ms": ref, "hbm_peak_GBps": peaks().get("hbm_gbs")})
    return out


def bench_ctrl(args, dev):
    """Row X1: the Controller tail (three Linear heads + modified_sigmoid) as the fused kernel vs the torch expression
    (cuBLAS fp32 + elementwise kernels), forward and forward+backward, at cfg 2's shapes."""
    from ddsp_pytorch_b200 import ops
    from ddsp_pytorch_b200.controller import modified_sigmoid
    out = []
    U, K, NB = 512, 180, 195
    for B, T in ((32, 172), (4, 172)):
        x = torch.randn(B, T, U, device=dev, requires_grad=True)
        heads = [torch.nn.Linear(U, n).to(dev) for n in (K, 1, NB)]
        ps = [t for h in heads for t in (h.weight, h.bias)]
        gs = [torch.randn(B, T, n, device=dev) for n in (K, 1, NB)]

        def fused(bwd):
            c, a, H = ops.controller_heads(x, *ps)
            if bwd:
                torch.autograd.backward((c, a, H), gs)

        def stock(bwd):
            c, a, H = (modified_sigmoid(h(x)) for h in heads)
            if bwd:
                torch.autograd.backward((c, a, H), gs)

        R, J = B * T, K + 1 + NB
        res = {"workload": f"controller tail R={R} U={U} J={J}", "flop_fwd": 2.0 * R * U * J}
        # "fused" = what ops.controller_heads picks for this R (tcgen05 from ops.CTRL_TC_MIN_ROWS rows); "ffma" = the CUDA-core
        # product forced; each as eager calls (host launch time included) and as a CUDA-graph replay (device time)
        auto_rows = ops.CTRL_TC_MIN_ROWS
        res["fused_path"] = "tcgen05" if ops._ctrl_use_tensor_cores(R, U) else "ffma"
        for name, fn, rows in (("fused", fused, auto_rows), ("ffma", fused, 0), ("torch", stock, auto_rows)):
            ops.CTRL_TC_MIN_ROWS = rows
            with torch.no_grad():
                res[f"{name}_fwd_ms"] = ev_time(lambda: fn(False), args.reps)
            res[f"{name}_fwd_bwd_ms"] = ev_time(lambda: fn(True), args.reps)
            if name != "torch":
                for bwd, key in ((False, "fwd"), (True, "fwd_bwd")):
                    x.grad = None
                    for t in ps:
                        t.grad = None
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph):
                        if bwd:
                            fn(True)
                        else:
                            with torch.no_grad():
                                fn(False)
                    res[f"{name}_{key}_graph_ms"] = ev_time(graph.replay, args.reps)
        ops.CTRL_TC_MIN_ROWS = auto_rows
        res["fused_fwd_TFLOPs"] = res["flop_fwd"] / (res["fused_fwd_graph_ms"] * 1e-3) / 1e12
        out.append(res)
    return out


def bench_sweep(args, dev):
    """cfg 4 over the harmonic counts BASELINE names (K = 100, 128, 180, 256): 1024 voices x 4 s @ 48 kHz, NB = 65, oscillator +
    filtered noise forward + backward as a CUDA-graph replay (bench.py's step), with the fused-step FP32 roofline per K."""
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200 import ops
    from ddsp_pytorch_b200.sharding import voice_work
    from ddsp_pytorch_b200.synthetic import synthetic_controls
    B, T, NB, sr, hop = args.sweep_voices, 375, 65, 48000, 512
    N = T * hop
    peak = ops.probe_fp32_fma_tflops()
    out = []
    g = torch.randn(B, N, generator=torch.Generator().manual_seed(99)).to(dev)
    for K in (100, 128, 180, 256):
        f0, c, a, H = (t.to(dev) for t in synthetic_controls(B, T, K, NB, sr, hop, seed=1234 + 4))
        c.requires_grad_(True); a.requires_grad_(True); H.requires_grad_(True)
        calls = torch.zeros(1, dtype=torch.int32, device=dev)

        def body():
            calls.add_(1)
            y = ops.filtered_noise(H, hop, seed=7, offset=0, offset_tensor=calls)
            y = ops.oscillator_bank(f0, c, a, sr, hop, out=y)
            y.backward(g)

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                c.grad = a.grad = H.grad = None
                body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        c.grad = a.grad = H.grad = None
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            body()
        ms = ev_time(gr.replay, args.reps)
        active = float(voice_work(f0, K, sr, hop, per_sample_cost=0.0).sum())
        M = 2 * (NB - 1)
        flop = 18.0 * active + 2.0 * (M + M * NB / hop) * B * N
        out.append({"workload": f"cfg4 sweep K={K}: {B} voices x 4 s @ {sr} Hz, NB={NB}, osc + noise fwd+bwd (graph replay)",
                    "ms": ms, "samples_per_s": B * N / (ms * 1e-3), "active_harmonic_samples": active,
                    "nominal_harmonic_samples": float(B) * N * K, "fused_survey_tflops": flop / (ms * 1e-3) / 1e12,
                    "fp32_peak_tflops": peak, "fused_frac": flop / (ms * 1e-3) / 1e12 / peak,
                    "executed_pipe_frac_osc_only": 16.0 * active / (ms * 1e-3) / 1e12 / peak})
        del gr
    return out


def bench_dftgemm(args, dev):
    """north_star item 3: the tcgen05 DFT-as-GEMM forward of one scale (csrc/ddsp_dftgemm.cu) against the register-tiled FFT
    kernel of the product path on the same frames (cfg 5: 64 x 4 s @ 16 kHz), forward only (loss terms, no gradient)."""
    from ddsp_pytorch_b200 import _lib, ops
    B, N = 64, 64000
    g = torch.Generator().manual_seed(1234 + 5)
    xt = (torch.randn(B, N, generator=g) * 0.1).clamp(-1, 1)
    xp = (0.5 * xt + torch.randn(B, N, generator=g) * 0.05)
    xt, xp = xt.to(dev), xp.to(dev)
    out = []
    lib = _lib.load()
    for n in (64, 128):
        ws = torch.empty(lib.ddsp_dftgemm_workspace_bytes(B, N, n), dtype=torch.uint8, device=dev)
        fft_ms = ev_time(lambda: ops.mss_loss(xp, xt, [n]), args.reps)
        prep_ms = ev_time(lambda: ops.dftgemm_sss_terms(xp, xt, n, workspace=ws), args.reps)
        res = {"workload": f"SSS forward n_fft={n}, {B} x {N} samples ({B * (1 + N // (n // 4))} frames x 2 signals)",
               "fft_kernel_ms": fft_ms, "dftgemm_prepare_plus_forward_ms": prep_ms}
        for splits in (3, 1):
            res[f"dftgemm_forward_{splits}split_ms"] = ev_time(
                lambda: ops.dftgemm_sss_terms(xp, xt, n, splits=splits, workspace=ws, prepare=False), args.reps)
        frames = 2.0 * B * (1 + N // (n // 4))
        res["gemm_flops_3split"] = 6 * 2.0 * frames * n * n
        res["gemm_tflops_3split"] = res["gemm_flops_3split"] / (res["dftgemm_forward_3split_ms"] * 1e-3) / 1e12
        ref = ops.mss_loss(xp, xt, [n])
        t3 = ops.dftgemm_sss_terms(xp, xt, n, workspace=ws)
        cnt = B * (n // 2 + 1) * (1 + N // (n // 4))
        res["loss_fft_path"] = float(ref)
        res["loss_dftgemm_3split"] = float((t3[0] + t3[1]) / cnt)
        out.append(res)
    return out


def bench_cpu(args, dev):
    """cpu_baseline leg for the side rows: the oracle port of the reference's torch path in fp32 on all host threads
    -- cfg 1, cfg 1' and cfg 5 whole (forward + backward), cfg 3 call by call (p50 / p99).  The only place this file touches
    oracle/; nothing here runs on the GPU."""
    from oracle import ddsp_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    out = []

    def timed(fn, reps=3):
        fn()
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
        return statistics.median(ts) * 1e3

    n_ffts = [2048, 1024, 512, 256, 128, 64]
    for name, (B, T, K, NB, sr) in (("cfg1", (16, 125, 100, 65, 16000)), ("cfg1'", (16, 172, 180, 195, 44100))):
        hop = 512
        f0, c, a, H = O.synthetic_controls(B, T, K, NB, sr, hop, seed=1234 + 1)
        target = (torch.randn(B, T * hop, generator=torch.Generator().manual_seed(9)) * 0.1).clamp(-1, 1)
        rn = torch.rand(sr, generator=torch.Generator().manual_seed(3)) * 2 - 1

        def step():
            cc, aa, HH = (t.detach().requires_grad_(True) for t in (c, a, H))
            noise, decay, wet = rn.clone().requires_grad_(True), torch.tensor(5.0, requires_grad=True), torch.tensor(0.0, requires_grad=True)
            u = torch.rand(B, T, hop) * 2 - 1                               # the reference's CPU noise draw (filtered_noise.py:44)
            y = O.synth_path(f0, cc, aa, HH, u, noise, decay, wet, K, sr, hop)
            O.mss_loss(y, target, n_ffts).backward()

        ms = timed(step)
        out.append({"workload": f"CPU {name} path: B={B} T={T} sr={sr} K={K} NB={NB}: osc + noise + reverb + 6-scale MSS, fwd+bwd",
                    "kind": "port", "cores": torch.get_num_threads(), "ms": ms, "samples_per_s": B * T * hop / (ms * 1e-3)})
    B, N = 64, 64000
    g = torch.Generator().manual_seed(1234 + 5)
    xt = (torch.randn(B, N, generator=g) * 0.1).clamp(-1, 1)
    xp = (0.5 * xt + torch.randn(B, N, generator=g) * 0.05)

    def mss_step():
        x = xp.detach().requires_grad_(True)
        O.mss_loss(x, xt, n_ffts).backward()

    ms = timed(mss_step)
    out.append({"workload": "CPU cfg5 MSS alone: 64 x 4 s @ 16 kHz, n_fft 2048..64, fwd+bwd", "kind": "port",
                "cores": torch.get_num_threads(), "ms": ms, "samples_per_s": B * N / (ms * 1e-3)})
    # cfg 3: live oscillator + noise + live reverb per 2048-sample block, call by call
    sr, hop, K, NB, T = 44100, 512, 180, 195, 4
    last, buf = torch.zeros(K), torch.zeros(1, sr)
    ir = O.reverb_ir(torch.rand(sr) * 2 - 1, torch.tensor(5.0), torch.tensor(0.0), sr)
    ctrls = [O.synthetic_controls(1, T, K, NB, sr, hop, seed=50 + i) for i in range(16)]
    lat = []
    with torch.no_grad():
        for i in range(20 + args.cpu_rt_calls):
            f0, c, a, H = ctrls[i % 16]
            t0 = time.perf_counter()
            harm, last = O.oscillator_bank_live(f0, c, a, last, K, sr, hop)
            y, buf = O.reverb_live(buf, harm + O.filtered_noise(H, hop, torch.rand(1, T, hop) * 2 - 1), ir)
            if i >= 20:
                lat.append((time.perf_counter() - t0) * 1e3)
    lat.sort()
    out.append({"workload": f"CPU cfg3 rt: batch 1, {T * hop}-sample blocks @ {sr} Hz, live osc + noise + live reverb, {len(lat)} calls",
                "kind": "port", "cores": torch.get_num_threads(), "deadline_ms": T * hop / sr * 1e3, "host_p50_ms": lat[len(lat) // 2],
                "host_p99_ms": lat[int(0.99 * len(lat))], "host_max_ms": lat[-1]})
    return out


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--workloads", default="mss,reverb,path,rt")
    p.add_argument("--reps", type=int, default=20)
    p.add_argument("--rt-calls", type=int, default=2000)
    p.add_argument("--reverb-big-only", action="store_true", help="reverb: only the B=256 shape (profiling runs)")
    p.add_argument("--sweep-voices", type=int, default=1024)
    p.add_argument("--cpu-rt-calls", type=int, default=10000)
    args = p.parse_args()
    if not torch.cuda.is_available():
        raise RuntimeError("bench_extra.py needs a CUDA device; there is no CPU fallback")
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    table = {"fft": bench_fft, "mss": bench_mss, "reverb": bench_reverb, "path": bench_path, "rt": bench_rt, "ctrl": bench_ctrl, "rtdec": bench_rtdec,
             "cfg1": bench_cfg1, "rt3p": bench_rt3p, "sweep": bench_sweep, "dftgemm": bench_dftgemm, "cpu": bench_cpu}
    for w in args.workloads.split(","):
        res = table[w](args, dev)
        for r in (res if isinstance(res, list) else [res]):
            print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
