"""Training driver equivalent to the reference's Lightning module `Zak` (train/train.py:15-51) without Lightning
(SURVEY.md §8f row X2): Decoder + MSSLoss([2048..64]), Adam(lr=1e-3), ReduceLROnPlateau(patience=5) on the training
loss, optional fp16 autocast of the Controller (the reference trains with precision=16), one process per GPU with
the gradients averaged by a bucketed NCCL all-reduce, and checkpoints in Lightning's key layout so that
`rt/utils.py:7-24` (which strips the ``model.`` prefix) still loads them.

The batch dimension is the only thing sharded: every rank runs the same step on its B/G voices; all loss terms
are means, so with equal shards the averaged gradient equals the single-GPU gradient.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from .controller import Decoder
from .modules import MSSLoss

N_FFTS = [2048, 1024, 512, 256, 128, 64]            # train/train.py:19


class Zak(nn.Module):
    """Same attribute names as the reference's LightningModule: `model` (Decoder) and `loss` (MSSLoss)."""

    def __init__(self, conf, noise_seed=None, voice_offset: int = 0, synth=None, loss=None, device_counter: bool = False):
        super().__init__()
        self.model = Decoder(conf, noise_seed=noise_seed, voice_offset=voice_offset, synth=synth, device_counter=device_counter)
        self.loss = loss if loss is not None else MSSLoss(N_FFTS)

    def forward(self, batch):
        return self.model(batch)

    def training_loss(self, batch):
        return self.loss(self.model(batch), batch)      # MSSLoss unwraps batch['audio'] (mss_loss.py:57-68)


def allreduce_mean_(params, bucket_bytes: int = 25 << 20) -> int:
    """Average .grad over ranks in flat fp32 buckets (one collective per bucket).  Returns the number of collectives."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return 0
    world = dist.get_world_size()
    grads = [p.grad for p in params if p.grad is not None]
    n_coll, i = 0, 0
    while i < len(grads):
        bucket, size = [], 0
        while i < len(grads) and (not bucket or size + grads[i].numel() * 4 <= bucket_bytes):
            bucket.append(grads[i])
            size += grads[i].numel() * 4
            i += 1
        flat = torch.cat([g.reshape(-1).float() for g in bucket])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat.div_(world)
        off = 0
        for g in bucket:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
        n_coll += 1
    return n_coll


class Trainer:
    def __init__(self, conf, device, amp: bool = False, lr: float = 1e-3, noise_seed: Optional[int] = None,
                 voice_offset: int = 0, synth=None, loss=None, graph: bool = False,
                 device_counter: Optional[bool] = None):
        """graph=True: the whole step (forward, backward, gradient all-reduce, Adam) is captured once as a CUDA graph on
        the first call and replayed afterwards (fp32 only; batches are copied into static buffers)."""
        self.device = torch.device(device)
        self.graph = bool(graph) and self.device.type == "cuda" and not amp
        self.zak = Zak(conf, noise_seed=noise_seed, voice_offset=voice_offset, synth=synth, loss=loss,
                       device_counter=self.graph if device_counter is None else device_counter).to(self.device)
        if self.graph:      # capturable Adam keeps step counts and the learning rate on the device
            self.opt = torch.optim.Adam(self.zak.parameters(), lr=torch.tensor(lr, device=self.device), capturable=True)
        else:
            self.opt = torch.optim.Adam(self.zak.parameters(), lr=lr)                    # train/train.py:22
        self._g = None
        self.sched = torch.optim.lr_scheduler.ReduceLROnPlateau(self.opt, patience=5)    # train/train.py:23
        self.amp = bool(amp) and self.device.type == "cuda"
        self.scaler = torch.amp.GradScaler("cuda", enabled=self.amp)
        self.steps = 0
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            for p in self.zak.parameters():             # same initial weights everywhere
                dist.broadcast(p.data, src=0)

    def _graphed_step(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        params = list(self.zak.parameters())
        if self._g is None:
            self._static = {k: v.clone() for k, v in batch.items()}

            def fwd_bwd():
                loss = self.zak.training_loss(self._static)
                loss.backward()
                return loss.detach()

            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):       # warm-up outside capture (allocator, cuDNN plans, twiddle tables, Adam state)
                for _ in range(3):
                    self.opt.zero_grad(set_to_none=True)
                    fwd_bwd()
                    allreduce_mean_(params)
                    self.opt.step()
            torch.cuda.current_stream(self.device).wait_stream(side)
            torch.cuda.synchronize(self.device)
            self.opt.zero_grad(set_to_none=True)            # gradients are (re)created inside the graph's memory pool
            # Two graphs with the NCCL all-reduce between them (eager, on the same stream) when there are several ranks;
            # with one rank the optimizer step is captured into the same graph.
            self._g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._g):
                self._loss = fwd_bwd()
                if not multi:
                    self.opt.step()
            self._g_opt = None
            if multi:
                allreduce_mean_(params)
                self._g_opt = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._g_opt, pool=self._g.pool()):
                    self.opt.step()
            self.steps += 3
        for k, v in batch.items():
            self._static[k].copy_(v)
        self._g.replay()
        if self._g_opt is not None:
            allreduce_mean_(params)
            self._g_opt.replay()
        self.steps += 1
        return self._loss

    def training_step(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        if self.graph:
            return self._graphed_step(batch)
        self.opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.float16, enabled=self.amp):
            loss = self.zak.training_loss(batch)
        self.scaler.scale(loss).backward()
        allreduce_mean_(list(self.zak.parameters()))    # before unscale_: every rank then sees the same inf/nan verdict
        if self.amp:
            self.scaler.unscale_(self.opt)
        self.scaler.step(self.opt)
        self.scaler.update()
        self.steps += 1
        return loss.detach()

    def end_epoch(self, mean_train_loss: float) -> None:
        self.sched.step(mean_train_loss)                # 'monitor': 'train_loss'

    # ---- checkpoints in Lightning's layout: {'state_dict': {'model.<...>': t, 'loss.<...>': t}, ...} ----------------
    def save_checkpoint(self, path: str, epoch: int = 0) -> None:
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        torch.save({"epoch": epoch, "global_step": self.steps, "state_dict": self.zak.state_dict(),
                    "optimizer_states": [self.opt.state_dict()], "lr_schedulers": [self.sched.state_dict()]}, path)

    def load_checkpoint(self, path: str) -> None:
        ck = torch.load(path, map_location=self.device)
        self.zak.load_state_dict(ck["state_dict"], strict=True)
        if ck.get("optimizer_states"):
            self.opt.load_state_dict(ck["optimizer_states"][0])
        if ck.get("lr_schedulers"):
            self.sched.load_state_dict(ck["lr_schedulers"][0])
        self.steps = int(ck.get("global_step", 0))


def synthetic_batch(conf, B: int, device, seed: int = 0, voice_offset: int = 0) -> Dict[str, torch.Tensor]:
    """A batch shaped like PLHDataset.__getitem__ (audio_dataset.py:109-110): CREPE/loudness features are synthetic."""
    T = (conf.example_duration * conf.sample_rate) // conf.hop_length
    N = T * conf.hop_length
    out = {k: [] for k in ("f0", "normalized_cents", "loudness", "audio")}
    for b in range(B):
        g = torch.Generator().manual_seed(seed * 1_000_003 + voice_offset + b)
        base = 80.0 * (1000.0 / 80.0) ** torch.rand(1, generator=g)
        t = torch.arange(T) * conf.hop_length / conf.sample_rate
        f0 = base * (1.0 + 0.01 * torch.sin(2 * torch.pi * 5.0 * t)) + torch.randn(T, generator=g)
        out["f0"].append(f0.view(T, 1))
        out["normalized_cents"].append((torch.log2(f0 / 440.0) * 12 / 60 + 0.5).view(T, 1))
        out["loudness"].append((torch.randn(T, generator=g) * 0.3).view(T, 1))
        out["audio"].append((torch.randn(N, generator=g) * 0.1).clamp(-1, 1))
    return {k: torch.stack(v).to(device) for k, v in out.items()}
