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
    # the graphed trainer's first call runs three eager warm-up steps on its first batch before the captured one
    torch.manual_seed(0)
    tr_e = Trainer(conf, "cuda:0", noise_seed=3, graph=False, device_counter=True)
    tr_e.zak.loss = type(tr_e.zak.loss)([64, 128]).to("cuda:0")
    first = synthetic_batch(conf, 3, "cuda:0", seed=10)
    want = [float(tr_e.training_step(first)) for _ in range(4)][-1:]
    want += [float(tr_e.training_step(synthetic_batch(conf, 3, "cuda:0", seed=10 + i))) for i in range(1, 7)]
    for g, w in zip(graphed, want):
        assert abs(g - w) <= 2e-3 * abs(w), (graphed, want)
