"""The oracle (oracle/ddsp_oracle.py) against fixtures produced by running the reference itself
(tests/golden/make_golden.py) and against the closed-form known answers of SURVEY.md §4."""
import numpy as np
import pytest
import torch

from oracle import ddsp_oracle as O
from oracle import philox

T64 = lambda x: torch.from_numpy(np.asarray(x)).double()   # noqa: E731


def maxabs(a, b):
    return float((torch.as_tensor(a).double() - torch.as_tensor(b).double()).abs().max())


@pytest.mark.parametrize("name", ["ref_osc_small.npz", "ref_osc_default.npz", "ref_osc_nyquist.npz", "ref_osc_long.npz"])
def test_oscillator_matches_reference_fp64_and_grads(golden, name):
    g = golden(name)
    K, sr, hop = int(g["K"]), int(g["sr"]), int(g["hop"])
    f0 = T64(g["f0"])
    c = T64(g["c"]).requires_grad_(True)
    a = T64(g["a"]).requires_grad_(True)
    y = O.oscillator_bank(f0, c, a, K, sr, hop)
    assert maxabs(y.detach(), g["y64"]) < 1e-12
    (y * T64(g["gup"])).sum().backward()
    assert maxabs(c.grad, g["gc"]) < 1e-9
    assert maxabs(a.grad, g["ga"]) < 1e-9
    # fp32 run of the oracle reproduces the reference's fp32 bits (same op sequence)
    y32 = O.oscillator_bank(f0.float(), c.detach().float(), a.detach().float(), K, sr, hop)
    assert maxabs(y32, g["y32"]) < 1e-6


def test_oscillator_live_state(golden):
    g = golden("ref_osc_live.npz")
    K, sr, hop = int(g["K"]), int(g["sr"]), int(g["hop"])
    last = torch.zeros(K, dtype=torch.float64)
    for i in range(int(g["calls"])):
        y, last = O.oscillator_bank_live(T64(g[f"f0_{i}"]), T64(g[f"c_{i}"]), T64(g[f"a_{i}"]), last, K, sr, hop)
        assert maxabs(y, g[f"y64_{i}"]) < 1e-12
        assert maxabs(last, g[f"phase64_{i}"]) < 1e-12


@pytest.mark.parametrize("name", ["ref_noise_small.npz", "ref_noise_full.npz", "ref_noise_default.npz"])
def test_filtered_noise_matches_reference(golden, name):
    g = golden(name)
    hop = int(g["hop"])
    H = T64(g["H"]).requires_grad_(True)
    assert maxabs(O.amp_to_ir(H.detach(), hop), g["ir64"]) < 1e-14
    assert maxabs(O.amp_to_ir_closed_form(H.detach(), hop), g["ir64"]) < 1e-14
    y = O.filtered_noise(H, hop, T64(g["noise"]))
    assert maxabs(y.detach(), g["y64"]) < 1e-12
    (y * T64(g["gup"])).sum().backward()
    assert maxabs(H.grad, g["gH"]) < 1e-10


@pytest.mark.parametrize("name", ["ref_reverb_long.npz", "ref_reverb_crop.npz"])
def test_reverb_matches_reference(golden, name):
    g = golden(name)
    noise = T64(g["noise"]).requires_grad_(True)
    decay = torch.tensor(float(g["decay"]), dtype=torch.float64, requires_grad=True)
    wet = torch.tensor(float(g["wet"]), dtype=torch.float64, requires_grad=True)
    x = T64(g["x"]).requires_grad_(True)
    ir = O.reverb_ir(noise, decay, wet, int(g["sr"]))
    assert maxabs(ir.detach(), g["ir64"]) < 1e-14
    y = O.reverb_apply(x, ir)
    assert maxabs(y.detach(), g["y64"]) < 1e-11
    (y * T64(g["gup"])).sum().backward()
    assert maxabs(x.grad, g["gx"]) < 1e-10
    assert maxabs(noise.grad, g["gnoise"]) < 1e-10
    assert abs(float(decay.grad) - float(g["gdecay"])) < 1e-9 * max(1.0, abs(float(g["gdecay"])))
    assert abs(float(wet.grad) - float(g["gwet"])) < 1e-9 * max(1.0, abs(float(g["gwet"])))


def test_reverb_live(golden):
    g = golden("ref_reverb_live.npz")
    sr = int(g["sr"])
    ir = O.reverb_ir(T64(g["noise"]), torch.tensor(float(g["decay"])).double(), torch.tensor(float(g["wet"])).double(), sr)
    buf = torch.zeros(1, sr, dtype=torch.float64)
    for i in range(int(g["calls"])):
        y, buf = O.reverb_live(buf, T64(g[f"x_{i}"]), ir)
        assert maxabs(y, g[f"y64_{i}"]) < 1e-12
        assert maxabs(buf, g[f"buf64_{i}"]) == 0.0


def test_mss_matches_reference(golden):
    g = golden("ref_mss.npz")
    xp = T64(g["x_pred"]).requires_grad_(True)
    xt = T64(g["x_true"])
    n_ffts = [int(n) for n in g["n_ffts"]]
    for n, want in zip(n_ffts, g["per_scale64"]):
        assert abs(float(O.sss_loss(xp.detach(), xt, n)) - float(want)) < 1e-10 * abs(float(want))
    loss = O.mss_loss(xp, xt, n_ffts)
    assert abs(float(loss) - float(g["loss64"])) < 1e-10 * float(g["loss64"])
    loss.backward()
    assert maxabs(xp.grad, g["gx"]) < 1e-9 * float(np.abs(g["gx"]).max())


def test_whole_path_matches_reference(golden):
    g = golden("ref_path.npz")
    K, NB, sr, hop = int(g["K"]), int(g["NB"]), int(g["sr"]), int(g["hop"])
    c, a, H = (T64(g[k]).requires_grad_(True) for k in ("c", "a", "H"))
    rn = T64(g["rev_noise"]).requires_grad_(True)
    rd = torch.tensor(float(g["rev_decay"]), dtype=torch.float64, requires_grad=True)
    rw = torch.tensor(float(g["rev_wet"]), dtype=torch.float64, requires_grad=True)
    y = O.synth_path(T64(g["f0"]), c, a, H, T64(g["noise"]), rn, rd, rw, K, sr, hop)
    assert maxabs(y.detach(), g["y64"]) < 1e-11
    loss = O.mss_loss(y, T64(g["target"]), [int(n) for n in g["n_ffts"]])
    assert abs(float(loss) - float(g["loss64"])) < 1e-10 * float(g["loss64"])
    loss.backward()
    for got, key in ((c.grad, "gc"), (a.grad, "ga"), (H.grad, "gH"), (rn.grad, "g_rev_noise")):
        assert maxabs(got, g[key]) < 1e-8 * max(1.0, float(np.abs(g[key]).max()))


# ---- closed-form known answers -------------------------------------------------
def test_kat_interpolate():
    x = torch.arange(4, dtype=torch.float64).view(1, 4, 1)
    want = torch.tensor([0, 0, .125, .375, .625, .875, 1.125, 1.375, 1.625, 1.875, 2.125, 2.375, 2.625, 2.875, 3, 3])
    assert maxabs(O.upsample_linear(x, 4).flatten(), want) < 1e-15
    assert maxabs(O.upsample_linear_explicit(x, 4).flatten(), want) < 1e-15
    y = torch.randn(2, 7, 3, dtype=torch.float64)
    assert maxabs(O.upsample_linear(y, 64), O.upsample_linear_explicit(y, 64)) < 1e-14


def test_kat_causal_conv_is_truncated_linear_convolution():
    rng = np.random.default_rng(0)
    s, k = rng.standard_normal(512), rng.standard_normal(512)
    got = O.causal_conv(torch.from_numpy(s), torch.from_numpy(k)).numpy()
    assert np.abs(got - np.convolve(s, k)[:512]).max() < 1e-11


def test_kat_constant_f0_phase():
    sr, hop, K, T = 16000, 64, 8, 5
    f0 = torch.full((1, T, 1), 440.0, dtype=torch.float64)
    om, _ = O.osc_prepare(f0, torch.ones(1, T, K, dtype=torch.float64), K, sr, hop)
    ph = O.osc_phases(om)
    n = torch.arange(1, T * hop + 1, dtype=torch.float64).view(-1, 1)
    k = torch.arange(1, K + 1, dtype=torch.float64).view(1, -1)
    want = torch.remainder(k * 2 * np.pi * 440.0 * n / sr, 2 * np.pi)
    d = (ph[0] - want).abs()
    assert float(torch.minimum(d, 2 * np.pi - d).max()) < 1e-9


def test_all_masked_frame_is_nan():
    f0 = torch.full((1, 3, 1), 9000.0, dtype=torch.float64)      # k*f0 > sr//2 for every k
    y = O.oscillator_bank(f0, torch.ones(1, 3, 4).double(), torch.ones(1, 3, 1).double(), 4, 16000, 8)
    assert torch.isnan(y).all()


# ---- Philox ---------------------------------------------------------------------------------
def test_philox_known_answers():
    # Random123 kat_vectors for philox4x32-10
    assert [int(v) for v in philox.philox4x32_10(0, 0, 0, 0, 0, 0)] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    m = 0xffffffff
    assert [int(v) for v in philox.philox4x32_10(m, m, m, m, m, m)] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert [int(v) for v in philox.philox4x32_10(0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344, 0xa4093822, 0x299f31d0)] \
        == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_philox_noise_properties():
    u = philox.noise_uniform(4, 8, 512, seed=1234, offset=3)
    assert u.dtype == np.float32 and u.min() >= -1.0 and u.max() < 1.0
    assert abs(float(u.mean())) < 0.02 and abs(float(u.var()) - 1 / 3) < 0.02
    # sharding invariance: voices 2..3 generated alone equal the slice of the full batch
    assert np.array_equal(philox.noise_uniform(2, 8, 512, 1234, 3, voice_offset=2), u[2:])
    assert not np.array_equal(philox.noise_uniform(4, 8, 512, 1234, 4), u)
