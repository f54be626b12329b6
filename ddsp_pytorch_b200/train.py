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

    def __init__(self, conf, noise_seed=None, voice_offset: int = 0, device_counter: bool = False):
        super().__init__()
        self.model = Decoder(conf, noise_seed=noise_seed, voice_offset=voice_offset, device_counter=device_counter)
        self.loss = MSSLoss(N_FFTS)

    def forward(self, batch):
        return self.model(batch)

    def training_loss(self, batch):
        return self.loss(self.model(batch), batch)      # MSSLoss unwraps batch['audio'] (mss_loss.py:57-68)


def allreduce_mean_(params, bucket_bytes: int = 25 << 20) -> int:
    """Average .grad over ranks in flat fp32 buckets (one collective per bucket) -- the plain, blocking form for gradients
    that live in their own tensors.  Returns the number of collectives.  The Trainer itself uses `GradBuckets`."""
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


class GradBuckets:
    """Gradient averaging the way stock DDP does it, without DDP: the trainable parameters are packed, in
    REVERSE registration order (the order backward produces their gradients), into persistent flat fp32 buckets and every
    `.grad` is a view into its bucket -- no concatenate / copy-back round trip.  In an eager multi-rank step a
    post-accumulate hook counts a bucket's gradients in and launches its all-reduce asynchronously the moment the last
    one has been written, so the collective runs on NCCL's stream under the rest of backward; `finish()` launches whatever
    was not triggered, waits, and divides by the world size.  The CUDA-graph step keeps the hooks disarmed and reduces the
    buckets between its two graphs."""

    def __init__(self, params, bucket_bytes: int = 25 << 20):
        self.params = [p for p in params if p.requires_grad]
        self.flats, self.members, self.views = [], [], {}
        cur, size = [], 0
        for p in reversed(self.params):
            if cur and size + p.numel() * 4 > bucket_bytes:
                self._close(cur)
                cur, size = [], 0
            cur.append(p)
            size += p.numel() * 4
        if cur:
            self._close(cur)
        self.bucket_of = {id(p): b for b, ps in enumerate(self.members) for p in ps}
        self.pending = [0] * len(self.flats)
        self.launched = [False] * len(self.flats)
        self.works = []
        self.armed = False
        self.n_collectives = 0
        self.skip_collectives = False        # measurement only: time a step without its all-reduces (gradients stay local)
        for p in self.params:
            p.register_post_accumulate_grad_hook(self._hook)

    def _close(self, ps):
        dev, n = ps[0].device, sum(p.numel() for p in ps)
        flat = torch.zeros(n, dtype=torch.float32, device=dev)
        off = 0
        for p in ps:
            self.views[id(p)] = flat[off:off + p.numel()].view_as(p)
            off += p.numel()
        self.flats.append(flat)
        self.members.append(ps)

    @staticmethod
    def _multi() -> bool:
        return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1

    def zero(self, arm: bool = True) -> None:
        """Start of a step: zero the buckets, make every .grad the bucket view (autograd accumulates into it in place)."""
        for flat in self.flats:
            flat.zero_()
        for p in self.params:
            v = self.views[id(p)]
            if p.grad is None or p.grad.data_ptr() != v.data_ptr():
                p.grad = v
        self.pending = [len(ps) for ps in self.members]
        self.launched = [False] * len(self.flats)
        self.works = []
        self.armed = bool(arm) and self._multi() and not self.skip_collectives

    def _launch(self, b: int) -> None:
        self.launched[b] = True
        self.works.append(dist.all_reduce(self.flats[b], op=dist.ReduceOp.SUM, async_op=True))
        self.n_collectives += 1

    def _hook(self, p) -> None:
        if not self.armed:
            return
        b = self.bucket_of[id(p)]
        self.pending[b] -= 1
        if self.pending[b] == 0 and not self.launched[b]:
            self._launch(b)

    def finish(self) -> int:
        """After backward: every bucket reduced (those no hook triggered are launched now, in bucket order on all ranks),
        waited for and divided by the world size.  Returns the number of collectives of this step."""
        if not self._multi() or self.skip_collectives:
            self.armed = False
            return 0
        for b in range(len(self.flats)):
            if not self.launched[b]:
                self._launch(b)
        for w in self.works:
            w.wait()
        world = dist.get_world_size()
        for flat in self.flats:
            flat.div_(world)
        n, self.works, self.armed = len(self.works), [], False
        return n


class Trainer:
    def __init__(self, conf, device, amp: bool = False, lr: float = 1e-3, noise_seed: Optional[int] = None,
                 voice_offset: int = 0, graph: bool = False, device_counter: Optional[bool] = None,
                 global_batch: Optional[int] = None, bucket_bytes: int = 25 << 20, zak: Optional[nn.Module] = None,
                 graph_collectives: bool = True):
        """graph=True: the whole step (forward, backward, gradient all-reduce, Adam) is captured once as a CUDA graph on
        the first call and replayed afterwards (fp32 only; batches are copied into static buffers).
        global_batch: total voices over all ranks; needed only when the ranks' shards differ in size -- the local loss is
        then weighted by B_local * world / global_batch so that the averaged gradient is the full-batch gradient.
        zak: the module to train (anything with `training_loss(batch)`); default: `Zak(conf, ...)`, the reference's model.
        bucket_bytes: gradient bucket size (stock DDP's 25 MB by default: the 19.9 MB of gradients travel as one all-reduce
        launched from the last gradient's hook; smaller buckets overlap more of backward but cost a host call each, which
        an eager step that is already launch-bound on the host cannot afford -- measured on 8 B200: six 4 MB buckets
        exposed 1.97 ms in a 9.5 ms eager step).
        graph_collectives: with graph=True on several ranks, capture the bucket all-reduces INSIDE the step's CUDA graph
        (NCCL's stream forks off the capture at each gradient hook and joins before Adam), so a replay is one launch and
        the collectives overlap backward; False, or a failed capture, falls back to two graphs with the all-reduces
        issued eagerly between them."""
        self.device = torch.device(device)
        self.graph = bool(graph) and self.device.type == "cuda" and not amp
        if zak is None:
            zak = Zak(conf, noise_seed=noise_seed, voice_offset=voice_offset,
                      device_counter=self.graph if device_counter is None else device_counter)
        self.zak = zak.to(self.device)
        if self.graph:      # capturable Adam keeps step counts and the learning rate on the device
            self.opt = torch.optim.Adam(self.zak.parameters(), lr=torch.tensor(lr, device=self.device), capturable=True)
        else:
            self.opt = torch.optim.Adam(self.zak.parameters(), lr=lr)                    # train/train.py:22
        self._g = None
        self.sched = torch.optim.lr_scheduler.ReduceLROnPlateau(self.opt, patience=5)    # train/train.py:23
        self.amp = bool(amp) and self.device.type == "cuda"
        self.scaler = torch.amp.GradScaler("cuda", enabled=self.amp)
        self.steps = 0
        self.global_batch = global_batch
        self.last_collectives = 0
        self.graph_collectives = bool(graph_collectives)
        self.graph_mode = None
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            for p in self.zak.parameters():             # same initial weights everywhere
                dist.broadcast(p.data, src=0)
        self.buckets = GradBuckets(self.zak.parameters(), bucket_bytes=bucket_bytes)

    def _loss_weight(self, batch) -> float:
        if self.global_batch is None or not GradBuckets._multi():
            return 1.0
        return float(batch["f0"].shape[0]) * dist.get_world_size() / float(self.global_batch)

    def _graphed_step(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        multi = GradBuckets._multi()
        if self._g is None:
            self._static = {k: v.clone() for k, v in batch.items()}
            weight = self._loss_weight(batch)

            def fwd_bwd():
                self.buckets.zero(arm=False)                # in-place: the gradients keep their addresses across replays
                loss = self.zak.training_loss(self._static)
                (loss * weight if weight != 1.0 else loss).backward()
                return loss.detach()

            # warm-up outside capture (allocator, cuDNN plans, twiddle tables, Adam state) on the real objects, then put
            # parameters, optimizer state and the noise counter back: the captured step is the FIRST step of the trajectory,
            # exactly as in the eager trainer
            keep_p = [p.detach().clone() for p in self.zak.parameters()]
            keep_b = [b_.detach().clone() for b_ in self.zak.buffers()]
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                for _ in range(3):
                    fwd_bwd()
                    self.buckets.finish()
                    self.opt.step()
            torch.cuda.current_stream(self.device).wait_stream(side)
            torch.cuda.synchronize(self.device)
            with torch.no_grad():
                for p, k in zip(self.zak.parameters(), keep_p):
                    p.copy_(k)
                for b_, k in zip(self.zak.buffers(), keep_b):
                    b_.copy_(k)
                for st in self.opt.state.values():          # Adam moments and the device-side step counter back to zero
                    for v in st.values():
                        if torch.is_tensor(v):
                            v.zero_()
            self._g_opt = None
            self._g = None
            if multi and self.graph_collectives:
                # one graph: forward, backward with the bucket all-reduces launched from the gradient hooks (NCCL's stream is
                # forked from / joined to the capturing stream by torch's work objects), division, Adam
                def whole_step():
                    self.buckets.zero(arm=True)
                    loss = self.zak.training_loss(self._static)
                    (loss * weight if weight != 1.0 else loss).backward()
                    self.buckets.finish()
                    self.opt.step()
                    return loss.detach()
                try:
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        self._loss = whole_step()
                    self._g, self.graph_mode = g, "one graph, collectives captured"
                    self.last_collectives = len(self.buckets.flats)
                except Exception as e:                  # capture of the collective refused: report it and use two graphs
                    self.graph_mode = f"two graphs (capture with collectives failed: {type(e).__name__})"
                    torch.cuda.synchronize(self.device)
                    self.buckets.works, self.buckets.armed = [], False
            if self._g is None:
                # Two graphs with the NCCL all-reduce between them (eager, on the same stream) when there are several ranks;
                # with one rank the optimizer step is captured into the same graph.
                self._g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._g):
                    self._loss = fwd_bwd()
                    if not multi:
                        self.opt.step()
                if multi:
                    self._g_opt = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(self._g_opt, pool=self._g.pool()):
                        self.opt.step()
                    if self.graph_mode is None:
                        self.graph_mode = "two graphs, eager collectives between"
                else:
                    self.graph_mode = "one graph"
        for k, v in batch.items():
            self._static[k].copy_(v)
        self._g.replay()
        if self._g_opt is not None:
            self.buckets.launched = [False] * len(self.buckets.flats)
            self.buckets.works = []
            self.last_collectives = self.buckets.finish()
            self._g_opt.replay()
        self.steps += 1
        return self._loss.clone()

    def training_step(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        if self.graph:
            return self._graphed_step(batch)
        self.buckets.zero(arm=True)
        weight = self._loss_weight(batch)
        with torch.autocast("cuda", dtype=torch.float16, enabled=self.amp):
            loss = self.zak.training_loss(batch)
        # the buckets' all-reduces start from the gradient hooks while backward is still running; they carry the SCALED
        # gradients, so every rank sees the same inf/nan verdict in unscale_
        self.scaler.scale(loss * weight if weight != 1.0 else loss).backward()
        self.last_collectives = self.buckets.finish()
        if self.amp:
            self.scaler.unscale_(self.opt)
        self.scaler.step(self.opt)
        self.scaler.update()
        self.steps += 1
        return loss.detach()

    def release_graphs(self) -> None:
        """Drop the captured CUDA graphs (and with them the NCCL kernels captured inside).  Call before
        `dist.destroy_process_group()`: tearing the communicator down while a graph still holds its collectives hung the
        8-rank run at exit."""
        if torch.cuda.is_available() and self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        self._g = None
        self._g_opt = None
        self._static = None
        self._loss = None

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
