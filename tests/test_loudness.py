"""Loudness encoder (row X3): oracle pins on CPU, kernel parity on the GPU."""
import numpy as np
import pytest
import torch


def test_a_weighting_known_values_and_package_table():
    from ddsp_pytorch_b200.encoder import a_weighting_db
    from oracle import ddsp_oracle as O
    got = O.a_weighting([1000.0, 100.0, 2500.0, 10000.0, 20.0, 0.0])
    # IEC 61672 table: 0.0, -19.1, +1.3, -2.5, -50.5 dB; DC clips at min_db
    np.testing.assert_allclose(got, [0.0, -19.1, 1.3, -2.5, -50.5, -80.0], atol=0.15)
    f = np.linspace(0, 22050, 1025, dtype="float32")
    np.testing.assert_allclose(a_weighting_db(f), O.a_weighting(f), rtol=0, atol=1e-9)
    assert O.loudness_a_weight(2048, 44100).dtype == torch.float32


def test_oracle_loudness_closed_form():
    from oracle import ddsp_oracle as O
    # a unit impulse at the start of every frame has |X_k| = 1 for all bins: loudness = mean(a_weight)/90 + 1
    n, hop, F = 64, 16, 5
    x = torch.zeros(1, n + hop * (F - 1), dtype=torch.float64)
    x[0, ::hop] = 0.0
    x[0, 0] = 1.0
    aw = O.loudness_a_weight(n, 8000).double()
    out = O.loudness_encoder(x, aw, n, hop)
    assert out.shape == (1, F, 1)
    assert abs(float(out[0, 0, 0]) - float(aw.mean() / 90 + 1)) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("B,N,n_fft,hop,sr", [(3, 20000, 2048, 512, 44100), (2, 4096, 4096, 512, 48000), (1, 5000, 1024, 256, 16000),
                                              (4, 3000, 512, 128, 16000), (2, 999, 256, 100, 8000), (1, 700, 64, 16, 8000),
                                              (2, 100, 8, 3, 8000), (1, 2048, 2048, 512, 44100)])
def test_kernel_matches_oracle(B, N, n_fft, hop, sr):
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    g = torch.Generator().manual_seed(N)
    x = torch.randn(B, N, generator=g) * 0.1 * torch.rand(B, 1, generator=g)
    aw = O.loudness_a_weight(n_fft, sr)
    want = O.loudness_encoder(x.double(), aw.double(), n_fft, hop)
    ref32 = O.loudness_encoder(x, aw, n_fft, hop)
    got = ops.loudness(x.cuda(), aw.cuda(), n_fft, hop).cpu().double()
    assert got.shape == want.shape
    err, drift = float((got - want).abs().max()), float((ref32.double() - want).abs().max())
    print(f"n_fft={n_fft}: |cuda-ref64|={err:.2e}  reference fp32 |ref32-ref64|={drift:.2e}")
    assert err <= 1e-5


@pytest.mark.gpu
def test_module_state_dict_and_shape():
    import ddsp_pytorch_b200 as d
    from ddsp_pytorch_b200.encoder import LoudnessEncoder
    enc = LoudnessEncoder(d.SHIPPED).cuda()
    assert set(enc.state_dict()) == {"a_weight"} and enc.a_weight.shape == (1025,) and not enc.a_weight.requires_grad
    out = enc(torch.randn(2, 2048 + 3 * 512, device="cuda"))      # rt: 4 frames from a 3584-sample buffer (autoencoder.py:29)
    assert out.shape == (2, 4, 1) and bool(torch.isfinite(out).all())
