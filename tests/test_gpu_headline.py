"""Parity of waveform AND gradients at the exact shapes BASELINE.json quotes:

  cfg 4  1024 x 4 s @ 48 kHz, T=375, K=256, NB=65, hop 512  -- 2 voices of it (voices are independent; batching is
         checked bit for bit elsewhere), gc / ga / gH against autograd through the fp64 oracle
  cfg 1  16 x 4 s @ 16 kHz, T=125, K=100, NB=65              -- the whole batch, oracle evaluated voice by voice
  cfg 3' batch 1, one frame per block, hop in {64,128,256}   -- OscillatorBank.live + FilteredNoise + Reverb.live_forward
         over 10 consecutive blocks with carried phases / reverb buffer

The CUDA side goes through the C ABI (ops.* -> libddsp_b200.so).  Tolerances: SURVEY.md §8c."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

WAVE_TOL = 1e-4
GRAD_TOL = 1e-4


def _rel(got, want):
    want = want.double()
    return float((got.double().cpu() - want).abs().max() / want.abs().max().clamp_min(1e-30))


def _nrm(got, want):
    want = want.double()
    return float((got.double().cpu() - want).norm() / want.norm().clamp_min(1e-30))


def _oracle_voice_by_voice(f0, c, a, H, noise, gup, K, sr, hop):
    """fp64 oracle forward + autograd, one voice at a time (a [1,N,K] fp64 tensor is 0.4 GB at cfg 4)."""
    from oracle import ddsp_oracle as O
    ys, gcs, gas, gHs = [], [], [], []
    for b in range(f0.shape[0]):
        cd = c[b:b + 1].double().requires_grad_(True)
        ad = a[b:b + 1].double().requires_grad_(True)
        Hd = H[b:b + 1].double().requires_grad_(True)
        y = O.oscillator_bank(f0[b:b + 1].double(), cd, ad, K, sr, hop) + O.filtered_noise(Hd, hop, noise[b:b + 1].double())
        (y * gup[b:b + 1].double()).sum().backward()
        ys.append(y.detach()); gcs.append(cd.grad); gas.append(ad.grad); gHs.append(Hd.grad)
    return tuple(torch.cat(v) for v in (ys, gcs, gas, gHs))


def _cuda_step(f0, c, a, H, gup, sr, hop, seed):
    """The bench's own step: y = noise(H); y += osc(f0, c, a); backward from gup."""
    from ddsp_pytorch_b200 import ops
    f0c, gc_ = f0.cuda(), gup.cuda()
    cc, ac, Hc = (t.cuda().requires_grad_(True) for t in (c, a, H))
    y = ops.filtered_noise(Hc, hop, seed=seed, offset=2, voice_offset=5)
    y = ops.oscillator_bank(f0c, cc, ac, sr, hop, out=y)
    y.backward(gc_)
    return y.detach(), cc.grad, ac.grad, Hc.grad


@pytest.mark.parametrize("name,B,T,K,NB,sr,hop", [("cfg4", 2, 375, 256, 65, 48000, 512), ("cfg1", 16, 125, 100, 65, 16000, 512),
                                                  ("cfg1prime", 2, 172, 180, 195, 44100, 512)])
def test_waveform_and_all_gradients_at_the_quoted_shapes(name, B, T, K, NB, sr, hop):
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    f0, c, a, H = O.synthetic_controls(B, T, K, NB, sr, hop, seed=1234 + 4)
    gup = torch.randn(B, T * hop, generator=torch.Generator().manual_seed(99))
    seed = 7
    y, gc, ga, gH = _cuda_step(f0, c, a, H, gup, sr, hop, seed)
    noise = ops.noise_uniform(B, T, hop, seed, 2, 5).cpu()
    wy, wgc, wga, wgH = _oracle_voice_by_voice(f0, c, a, H, noise, gup, K, sr, hop)
    err = float((y.double().cpu() - wy).abs().max())
    print(f"{name}: |y-ref64|={err:.2e}  gc max-rel {_rel(gc, wgc):.2e} nrm {_nrm(gc, wgc):.2e}  ga {_rel(ga, wga):.2e}  "
          f"gH {_rel(gH, wgH):.2e} nrm {_nrm(gH, wgH):.2e}")
    assert err <= WAVE_TOL
    assert _rel(gc, wgc) <= GRAD_TOL and _nrm(gc, wgc) <= GRAD_TOL
    assert _rel(ga, wga) <= GRAD_TOL
    assert _rel(gH, wgH) <= GRAD_TOL and _nrm(gH, wgH) <= GRAD_TOL


@pytest.mark.parametrize("hop", [64, 128, 256])
def test_cfg3prime_live_blocks(hop):
    """rt block pipeline at BASELINE's block sizes: one frame per block, 10 blocks, state carried by the modules."""
    import ddsp_pytorch_b200 as d
    from oracle import ddsp_oracle as O
    from oracle.philox import noise_uniform
    sr, K = 44100, 180
    NB = min(195, hop // 2 + 1)                                           # pad hop - M must be >= 0
    conf = d.Config(sample_rate=sr, hop_length=hop, n_harmonics=K, n_noise_filters=NB)
    torch.manual_seed(11)
    bank, fn, rv = d.OscillatorBank(conf).cuda(), d.FilteredNoise(conf, seed=21), d.Reverb(conf, initial_wet=1.0).cuda()
    last = torch.zeros(K, dtype=torch.float64)
    buf = torch.zeros(1, sr, dtype=torch.float64)
    ir = O.reverb_ir(rv.noise.detach().cpu().double(), rv.decay.detach().cpu().double(), rv.wet.detach().cpu().double(), sr)
    worst = 0.0
    for blk in range(10):
        f0, c, a, H = O.synthetic_controls(1, 1, K, NB, sr, hop, seed=300 + blk)
        with torch.no_grad():
            got = rv.live_forward(bank.live({"f0": f0.cuda(), "c": c.cuda(), "a": a.cuda()}) + fn({"H": H.cuda()}))
        harm, last = O.oscillator_bank_live(f0.double(), c.double(), a.double(), last, K, sr, hop)
        u = torch.from_numpy(noise_uniform(1, 1, hop, 21, blk, 0)).double()
        want, buf = O.reverb_live(buf, harm + O.filtered_noise(H.double(), hop, u), ir)
        assert got.shape == (1, hop)
        scale = max(1.0, float(want.abs().max()))
        worst = max(worst, float((got.double().cpu() - want).abs().max()) / scale)
        dphi = (bank.last_phases.double().cpu() - last).abs()
        assert float(torch.minimum(dphi, 2 * np.pi - dphi).max()) <= 1e-4
    print(f"hop {hop}: worst |cuda-oracle64| over 10 blocks = {worst:.2e}")
    assert worst <= WAVE_TOL
    assert float((rv.buffer.double().cpu() - buf).abs().max()) <= 1e-5
