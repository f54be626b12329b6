"""Training step on the GPU: the whole Decoder + MSSLoss step against the fp64 oracle with the same weights and noise,
and a few optimizer steps."""
import copy
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_training_step_matches_oracle_and_learns():
    import test_train_gloo as H
    from ddsp_pytorch_b200.train import Trainer, synthetic_batch
    conf = H.small_conf()
    conf.n_noise_filters, conf.hop_length, conf.sample_rate = 9, 64, 4000
    torch.manual_seed(0)
    tr = Trainer(conf, "cuda:0", noise_seed=3)
    tr.zak.loss = type(tr.zak.loss)([64, 128]).to("cuda:0")
    batch = synthetic_batch(conf, 3, "cuda:0", seed=4)

    # fp64 oracle twin with identical weights
    ref = H._make(conf, 0)
    sd = {k: v.detach().cpu() for k, v in tr.zak.state_dict().items() if not k.startswith("loss.")}
    sd = {k: v for k, v in sd.items() if k in ref.zak.state_dict()}
    ref.zak.load_state_dict(sd, strict=False)
    ref.zak.model.reverb.noise.data = tr.zak.model.reverb.noise.detach().cpu().clone()
    ref.zak.model.reverb.decay.data = tr.zak.model.reverb.decay.detach().cpu().clone()
    ref.zak.model.reverb.wet.data = tr.zak.model.reverb.wet.detach().cpu().clone()
    ref.zak.double()
    cpu_batch = {k: v.detach().cpu().double() for k, v in batch.items()}
    want = ref.zak.training_loss(cpu_batch)
    want.backward()

    tr.opt.zero_grad(set_to_none=True)
    got = tr.zak.training_loss(batch)
    got.backward()
    assert abs(float(got) - float(want)) <= 1e-4 * abs(float(want))
    for name in ("model.reverb.wet", "model.reverb.decay", "model.controller.dense_harmonic.weight",
                 "model.controller.mlp_f0.mlp_layer1.0.weight"):
        g = dict(tr.zak.named_parameters())[name].grad.double().cpu()
        w = dict(ref.zak.named_parameters())[name].grad
        err = float((g - w).abs().max() / w.abs().max())
        print(f"{name}: rel grad err {err:.2e}")
        assert err <= 2e-3, name                        # the fp32 Controller (cuBLAS) sits between the kernels and these

    losses = [float(tr.training_step(batch)) for _ in range(8)]
    assert all(l == l for l in losses) and losses[-1] < losses[0]


def test_graphed_step_matches_eager():
    """The CUDA-graph replay of the whole step (forward, backward, Adam) follows the eager step with the same noise."""
    import test_train_gloo as H
    from ddsp_pytorch_b200.train import Trainer, synthetic_batch
    conf = H.small_conf()
    conf.n_noise_filters, conf.hop_length, conf.sample_rate = 9, 64, 4000
    torch.manual_seed(0)
    tr = Trainer(conf, "cuda:0", noise_seed=3, graph=True)
    tr.zak.loss = type(tr.zak.loss)([64, 128]).to("cuda:0")
    graphed = [float(tr.training_step(synthetic_batch(conf, 3, "cuda:0", seed=10 + i))) for i in range(7)]
    # the graphed trainer warms up on a snapshot and restores parameters, Adam state and the noise counter before the
    # capture, so its trajectory is the eager trainer's from the first step on
    torch.manual_seed(0)
    tr_e = Trainer(conf, "cuda:0", noise_seed=3, graph=False, device_counter=True)
    tr_e.zak.loss = type(tr_e.zak.loss)([64, 128]).to("cuda:0")
    want = [float(tr_e.training_step(synthetic_batch(conf, 3, "cuda:0", seed=10 + i))) for i in range(7)]
    for g, w in zip(graphed, want):
        assert abs(g - w) <= 2e-3 * abs(w), (graphed, want)
    assert tr.steps == tr_e.steps == 7
    # every call hands back its own loss tensor (the static one is overwritten by the next replay)
    l0 = tr.training_step(synthetic_batch(conf, 3, "cuda:0", seed=30))
    v0 = float(l0)
    tr.training_step(synthetic_batch(conf, 3, "cuda:0", seed=31))
    assert float(l0) == v0


def test_amp_step_trains_and_the_loss_scale_settles():
    """The reference trains with precision=16 (train/train.py:50): fp16 autocast of the torch Controller, fp32 at the
    boundary of this package's kernels, GradScaler around Adam.  The random-weight loss is large, so the first steps overflow
    in fp16 and are skipped while the scale halves; it must settle (no skipped step in the last 15) and the loss must fall."""
    import test_train_gloo as H
    from ddsp_pytorch_b200.train import Trainer, synthetic_batch
    conf = H.small_conf()
    conf.n_noise_filters, conf.hop_length, conf.sample_rate = 9, 64, 4000
    torch.manual_seed(0)
    tr = Trainer(conf, "cuda:0", amp=True, noise_seed=3)
    tr.zak.loss = type(tr.zak.loss)([64, 128]).to("cuda:0")
    batch = synthetic_batch(conf, 3, "cuda:0", seed=4)
    losses, scales, moved = [], [], []
    for _ in range(60):
        before = [p.detach().clone() for p in list(tr.zak.parameters())[:4]]
        losses.append(float(tr.training_step(batch)))
        scales.append(tr.scaler.get_scale())
        moved.append(any(not torch.equal(b, p.detach()) for b, p in zip(before, list(tr.zak.parameters())[:4])))
    print("amp: loss", losses[0], "->", losses[-1], " scale", scales[0], "->", scales[-1], " applied steps", sum(moved))
    assert all(l == l and abs(l) != float("inf") for l in losses)
    assert all(moved[-15:]), "optimizer steps are still being skipped after 45 steps"
    assert len(set(scales[-15:])) == 1 and scales[-1] > 0
    assert min(losses[-5:]) < losses[0]
    # the Controller's GEMMs really run in fp16 under autocast (LayerNorm hands fp32 on; our ops cast at their boundary)
    with torch.autocast("cuda", dtype=torch.float16):
        out = tr.zak.model.controller.mlp_f0.mlp_layer1[0](batch["normalized_cents"])
    assert out.dtype == torch.float16
