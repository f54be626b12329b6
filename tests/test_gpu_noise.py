"""Parity of the CUDA filtered-noise path.  The reference draws its noise on the CPU generator
(filtered_noise.py:44-48), so parity injects the identical noise block (tolerance 1e-5);
the in-kernel Philox stream is checked bit for bit against oracle/philox.py."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(got, want):
    want = torch.as_tensor(want).double()
    return float((got.double().cpu() - want).abs().max() / want.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("name", ["ref_noise_small.npz", "ref_noise_full.npz", "ref_noise_default.npz"])
def test_golden_with_injected_noise(golden, name):
    from ddsp_pytorch_b200 import ops
    g = golden(name)
    hop = int(g["hop"])
    H = torch.from_numpy(g["H"]).cuda().requires_grad_(True)
    y = ops.filtered_noise(H, hop, noise=torch.from_numpy(g["noise"]).cuda())
    err = float((y.detach().double().cpu() - torch.from_numpy(g["y64"])).abs().max())
    print(f"{name}: |cuda-ref64|={err:.2e} |ref32-ref64|={float(np.abs(g['y32'] - g['y64']).max()):.2e}")
    assert err <= 1e-5
    (y * torch.from_numpy(g["gup"]).cuda()).sum().backward()
    assert rel(H.grad, g["gH"]) <= 1e-4
    ir = ops.amp_to_impulse_response(H.detach(), hop)
    assert float((ir.double().cpu() - torch.from_numpy(g["ir64"])).abs().max()) <= 1e-6


@pytest.mark.parametrize("B,T,NB,hop", [(2, 5, 65, 512), (1, 3, 195, 512), (2, 2, 33, 64), (1, 4, 65, 128), (1, 2, 129, 256),
                                        (1, 2, 5, 8), (2, 3, 2, 16),
                                        # hops that are not a multiple of 8 / 4 / 2 (masked tail group, unaligned rows)
                                        (2, 3, 17, 100), (1, 4, 101, 441), (3, 2, 3, 5), (1, 1, 2, 2), (2, 2, 4, 7), (1, 3, 9, 20)])
def test_against_oracle(B, T, NB, hop):
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    _, _, _, H = O.synthetic_controls(B, T, 4, NB, 16000, hop, seed=NB)
    gen = torch.Generator().manual_seed(hop)
    noise = torch.rand(B, T, hop, generator=gen) * 2 - 1
    gup = torch.randn(B, T * hop, generator=gen)
    Hd = H.double().requires_grad_(True)
    want = O.filtered_noise(Hd, hop, noise.double())
    (want * gup.double()).sum().backward()
    Hc = H.cuda().requires_grad_(True)
    y = ops.filtered_noise(Hc, hop, noise=noise.cuda())
    (y * gup.cuda()).sum().backward()
    assert float((y.detach().double().cpu() - want.detach()).abs().max()) <= 1e-5
    assert rel(Hc.grad, Hd.grad) <= 1e-4


def test_philox_block_is_bit_exact_and_shard_invariant():
    from ddsp_pytorch_b200 import ops
    from oracle import philox
    for (B, T, hop, seed, off, vo) in [(3, 4, 512, 1234, 0, 0), (2, 3, 64, 2 ** 40 + 7, 9, 5), (1, 2, 10, 1, 2 ** 32 - 1, 0)]:
        got = ops.noise_uniform(B, T, hop, seed, off, vo).cpu().numpy()
        assert np.array_equal(got, philox.noise_uniform(B, T, hop, seed, off, vo))
    full = ops.noise_uniform(8, 4, 512, 77, 3, 0)
    assert torch.equal(ops.noise_uniform(4, 4, 512, 77, 3, 4), full[4:])


def test_in_kernel_noise_equals_injected_and_is_reproducible():
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    B, T, NB, hop = 4, 6, 65, 512
    _, _, _, H = O.synthetic_controls(B, T, 4, NB, 16000, hop, seed=3)
    H = H.cuda().requires_grad_(True)
    gup = torch.randn(B, T * hop, device="cuda")
    y1 = ops.filtered_noise(H, hop, seed=99, offset=5)
    (y1 * gup).sum().backward()
    g1 = H.grad.clone()
    H.grad = None
    y2 = ops.filtered_noise(H, hop, noise=ops.noise_uniform(B, T, hop, 99, 5))
    (y2 * gup).sum().backward()
    assert torch.equal(y1, y2) and torch.equal(g1, H.grad)
    assert not torch.equal(ops.filtered_noise(H.detach(), hop, seed=99, offset=6), y1)
    # sharded halves reproduce the full batch bit for bit (voice_offset = global voice index)
    lo = ops.filtered_noise(H.detach()[:2], hop, seed=99, offset=5, voice_offset=0)
    hi = ops.filtered_noise(H.detach()[2:], hop, seed=99, offset=5, voice_offset=2)
    assert torch.equal(torch.cat([lo, hi]), y1.detach())
    u = ops.noise_uniform(64, 16, 512, 5).double()
    assert abs(float(u.mean())) < 5e-3 and abs(float(u.var()) - 1 / 3) < 5e-3


def test_module_draws_fresh_noise_each_call():
    import ddsp_pytorch_b200 as d
    mod = d.FilteredNoise(d.CFG1, seed=1)
    H = torch.rand(2, 3, 65).cuda()
    a, b = mod({"H": H}), mod({"H": H})
    assert a.shape == (2, 3 * 512) and not torch.equal(a, b)


@pytest.mark.parametrize("shape", [(2, 3, 64), (5, 100), (1, 4096), (3, 1), (2, 2, 7), (4, 5, 512)])
def test_fft_convolve_per_row_kernels(shape):
    """modules.fft_convolve in the call shape FilteredNoise uses in the reference ([B,T,n] x [B,T,n],
    filtered_noise.py:50): per-row causal convolution and both adjoints against the oracle in fp64."""
    from ddsp_pytorch_b200 import modules
    from oracle import ddsp_oracle as O
    gen = torch.Generator().manual_seed(sum(shape))
    s = torch.randn(*shape, generator=gen)
    k = torch.randn(*shape, generator=gen) * 0.3
    gup = torch.randn(*shape, generator=gen)
    sd, kd = s.double().requires_grad_(True), k.double().requires_grad_(True)
    want = O.causal_conv(sd, kd)
    (want * gup.double()).sum().backward()
    sc, kc = s.cuda().requires_grad_(True), k.cuda().requires_grad_(True)
    got = modules.fft_convolve(sc, kc)
    (got * gup.cuda()).sum().backward()
    scale = float(want.abs().max()) + 1e-12
    assert got.shape == want.shape
    assert float((got.double().cpu() - want.detach()).abs().max()) <= 1e-5 * scale
    assert float((sc.grad.double().cpu() - sd.grad).abs().max()) <= 1e-5 * float(sd.grad.abs().max() + 1e-12)
    assert float((kc.grad.double().cpu() - kd.grad).abs().max()) <= 1e-5 * float(kd.grad.abs().max() + 1e-12)


def test_fft_convolve_broadcast_kernel():
    """A [T,n] kernel shared over the batch broadcasts like the reference's elementwise product of spectra."""
    from ddsp_pytorch_b200 import modules
    from oracle import ddsp_oracle as O
    gen = torch.Generator().manual_seed(3)
    s, k = torch.randn(3, 4, 32, generator=gen), torch.randn(4, 32, generator=gen)
    want = O.causal_conv(s.double(), k.double())
    got = modules.fft_convolve(s.cuda(), k.cuda())
    assert float((got.double().cpu() - want).abs().max()) <= 1e-5 * float(want.abs().max())
