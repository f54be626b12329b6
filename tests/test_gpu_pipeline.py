"""HostSynthPipeline (chunked, three streams, C-ABI calls into its own slots) must give exactly the audio and the gradients
of one un-chunked call through the autograd ops."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_pipeline_matches_single_call():
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200 import ops
    from ddsp_pytorch_b200.host_pipeline import HostSynthPipeline
    from oracle import ddsp_oracle as O
    B, T, K, NB, sr, hop = 11, 9, 100, 65, 16000, 512
    conf = d.Config(sample_rate=sr, hop_length=hop, n_harmonics=K, n_noise_filters=NB)
    f0, c, a, H = O.synthetic_controls(B, T, K, NB, sr, hop, seed=3)
    g = torch.randn(B, T * hop, generator=torch.Generator().manual_seed(1))
    host = [t.contiguous().pin_memory() for t in (f0, c, a, H, g)]
    outs = [torch.empty_like(t).pin_memory() for t in (c, a, H)]
    y_host = torch.empty(B, T * hop).pin_memory()
    pipe = HostSynthPipeline(conf, "cuda:0", chunk_voices=4, slots=2, seed=5, voice_offset=7)
    pipe.step(*host, *outs, y_out=y_host)
    assert pipe.launches_per_step == 6 * 3                 # 3 chunks x (noise fwd, prepare, synth, grad, finalize, noise bwd)

    dev = torch.device("cuda:0")
    cd, ad, Hd = (t.to(dev).requires_grad_(True) for t in (c, a, H))
    y = ops.filtered_noise(Hd, hop, seed=5, offset=0, voice_offset=7)
    y = ops.oscillator_bank(f0.to(dev), cd, ad, sr, hop, out=y)
    y.backward(g.to(dev))
    assert torch.equal(y_host, y.detach().cpu())
    for got, want in zip(outs, (cd.grad, ad.grad, Hd.grad)):
        assert torch.equal(got, want.cpu())
    # default chunking (about 16 chunks per step, 4..64 voices each) gives the same bits
    outs2 = [torch.empty_like(t).pin_memory() for t in (c, a, H)]
    HostSynthPipeline(conf, "cuda:0", seed=5, voice_offset=7).step(*host, *outs2)
    assert all(torch.equal(p, q) for p, q in zip(outs, outs2))
    # a second step draws fresh noise (call counter) and still matches
    pipe.step(*host, *outs)
    Hd.grad = None
    y = ops.filtered_noise(Hd, hop, seed=5, offset=1, voice_offset=7)
    y.backward(g.to(dev))
    assert torch.equal(outs[2], Hd.grad.cpu())


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_reduced_precision_inputs_are_cast_at_the_boundary(dtype):
    """Under the reference's ``precision=16`` (train/train.py:50) the Controller hands c, a, H over in half precision: the ops
    cast to fp32 at the autograd boundary, compute as for the same values in fp32 and return gradients in the
    caller's dtype."""
    from ddsp_pytorch_b200 import ops
    from ddsp_pytorch_b200.synthetic import synthetic_controls
    B, T, K, NB, sr, hop = 2, 6, 40, 65, 16000, 512
    f0, c, a, H = (t.cuda() for t in synthetic_controls(B, T, K, NB, sr, hop, seed=5))
    g = torch.randn(B, T * hop, device="cuda")
    lo = [t.to(dtype).requires_grad_(True) for t in (c, a, H)]
    hi = [t.detach().float().requires_grad_(True) for t in lo]
    outs = []
    for cc, aa, HH in (lo, hi):
        y = ops.filtered_noise(HH, hop, seed=3) + ops.oscillator_bank(f0, cc, aa, sr, hop)
        y.backward(g)
        outs.append(y.detach())
    assert outs[0].dtype == torch.float32 and torch.equal(outs[0], outs[1])
    for l, h in zip(lo, hi):
        assert l.grad.dtype == dtype and l.grad.shape == l.shape
        assert torch.equal(l.grad, h.grad.to(dtype))
    xp = (torch.randn(2, 3000, device="cuda") * 0.1).to(dtype).requires_grad_(True)
    xt = torch.randn(2, 3000, device="cuda") * 0.1
    ops.mss_loss(xp, xt, [256, 64]).backward()
    xp32 = xp.detach().float().requires_grad_(True)
    ops.mss_loss(xp32, xt, [256, 64]).backward()
    assert xp.grad.dtype == dtype and torch.equal(xp.grad, xp32.grad.to(dtype))
