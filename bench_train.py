#!/usr/bin/env python
"""cfg 2 of BASELINE.json: autoencoder training step -- Controller -> harmonic + noise + reverb ->
6-scale MSS loss -> backward -> gradient all-reduce -> Adam -- global batch 32 of 2-s clips @ 44.1 kHz, sharded over
the ranks.  Report-only row: samples/s = 32 * 88064 / step time, with a per-phase breakdown.

    python bench_train.py [--steps 20] [--amp]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench_train.py
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--global-batch", type=int, default=32)
    p.add_argument("--amp", action="store_true", help="fp16 autocast of the Controller (reference: precision=16)")
    p.add_argument("--graph", action="store_true", help="capture the whole training step as a CUDA graph and replay it")
    p.add_argument("--gpus", type=int, default=None, help="accepted for symmetry with bench.py; the world size comes from torchrun")
    args = p.parse_args()
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200.sharding import max_over_ranks, voice_shard
    from ddsp_pytorch_b200.train import Trainer, allreduce_mean_, synthetic_batch
    world, rank, local = (int(os.environ.get(k, d_)) for k, d_ in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
    if not torch.cuda.is_available():
        raise RuntimeError("bench_train.py needs CUDA devices; there is no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    conf = d.SHIPPED
    lo, hi = voice_shard(args.global_batch, world, rank)
    torch.manual_seed(0)
    tr = Trainer(conf, dev, amp=args.amp, noise_seed=7, voice_offset=lo, graph=args.graph)
    batch = synthetic_batch(conf, hi - lo, dev, seed=2, voice_offset=lo)
    T = batch["f0"].shape[1]
    N = T * conf.hop_length

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        tr.training_step(batch)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = tr.training_step(batch)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1), dev) / args.steps
    # the same step with the gradient all-reduces switched off: the difference is the collective time the step EXPOSES
    # (what backward could not hide); measurement only -- the ranks' weights drift apart, so it comes last
    ms_nocomm = None
    if world > 1:
        tr_nc = tr
        if args.graph:      # a captured step keeps what it captured: time a second trainer whose capture has no collectives
            torch.manual_seed(0)
            tr_nc = Trainer(conf, dev, amp=args.amp, noise_seed=7, voice_offset=lo, graph=True)
        tr_nc.buckets.skip_collectives = True
        for _ in range(max(3, args.warmup)):
            tr_nc.training_step(batch)
        barrier()
        e0.record()
        for _ in range(args.steps):
            tr_nc.training_step(batch)
        e1.record()
        barrier()
        ms_nocomm = max_over_ranks(e0.elapsed_time(e1), dev) / args.steps
        tr_nc.buckets.skip_collectives = False

    # phase breakdown on rank 0's shard (separate pass, events around each phase)
    def timed(fn, reps=10):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        b.synchronize()
        return a.elapsed_time(b) / reps

    zak = tr.zak
    with torch.no_grad():
        ctrl = zak.model.controller(batch)
    ctrl = {k: (v.detach() if torch.is_tensor(v) else v) for k, v in ctrl.items()}
    dry = (zak.model.harmonics(ctrl) + zak.model.noise(ctrl)).detach()
    wet = zak.model.reverb(dry).detach()

    def synth_fb():
        c = {k: (v.clone().requires_grad_(True) if k in ("c", "a", "H") else v) for k, v in ctrl.items()}
        y = zak.model.harmonics(c) + zak.model.noise(c)
        y.backward(dry)

    def reverb_fb():
        x = dry.clone().requires_grad_(True)
        zak.model.reverb(x).backward(dry)
        zak.zero_grad(set_to_none=True)

    def loss_fb():
        x = wet.clone().requires_grad_(True)
        zak.loss(x, batch).backward()

    def controller_fb():
        out = zak.model.controller(batch)
        (out["c"].sum() + out["a"].sum() + out["H"].sum()).backward()
        zak.zero_grad(set_to_none=True)

    def comm():                 # all buckets reduced back to back with nothing to hide behind
        tr.buckets.zero(arm=False)
        tr.buckets.finish()

    phases = {"controller_fwd_bwd_ms": timed(controller_fb), "osc_noise_fwd_bwd_ms": timed(synth_fb),
              "reverb_fwd_bwd_ms": timed(reverb_fb), "mss_fwd_bwd_ms": timed(loss_fb), "grad_allreduce_ms": timed(comm)}
    if rank == 0:
        n_params = sum(p_.numel() for p_ in zak.parameters() if p_.requires_grad)
        print(json.dumps({"workload": f"cfg2 training step: global batch {args.global_batch} x {N} samples @ {conf.sample_rate} Hz, "
                                      f"K={conf.n_harmonics}, NB={conf.n_noise_filters}, 6-scale MSS, Adam, amp={args.amp}, graph={args.graph}",
                          "n_gpus": world, "ms_per_step": ms, "samples_per_s": args.global_batch * N / (ms * 1e-3),
                          "loss": float(loss), "trainable_params": n_params, "allreduce_bytes": 4 * n_params,
                          "ms_per_step_without_allreduce": ms_nocomm,
                          "allreduce_exposed_ms": None if ms_nocomm is None else ms - ms_nocomm,
                          "grad_buckets": len(tr.buckets.flats), "collectives_per_step": tr.last_collectives,
                          "graph_mode": tr.graph_mode,
                          "phases_rank0": phases}), flush=True)
    sys.stdout.flush()
    if world > 1:
        tr.release_graphs()
        if ms_nocomm is not None and args.graph:
            tr_nc.release_graphs()
        dist.barrier(device_ids=[local])
        torch.cuda.synchronize()
        if args.graph:
            os._exit(0)         # captured NCCL kernels: do not risk the communicator's teardown path (it hung once at 8 ranks)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
