"""The register-tiled FFT (csrc/ddsp_fft8.cuh) through the C ABI against numpy's fp64 FFT."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _fft(x: torch.Tensor, inverse: bool) -> torch.Tensor:
    from ddsp_pytorch_b200 import _lib
    x = x.contiguous()
    out = torch.empty_like(x)
    n, batch = x.shape[-1], x.numel() // x.shape[-1]
    _lib.call("ddsp_fft_c2c", x.data_ptr(), out.data_ptr(), n, batch, int(inverse), int(x.dtype == torch.complex128),
              torch.cuda.current_stream().cuda_stream)
    return out


@pytest.mark.parametrize("n", [8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096])
@pytest.mark.parametrize("dtype,tol", [(torch.complex64, 4e-7), (torch.complex128, 1e-15)])
def test_fft_matches_numpy(n, dtype, tol):
    rng = np.random.default_rng(n)
    batch = 37                                            # not a multiple of the groups per CTA
    x = (rng.standard_normal((batch, n)) + 1j * rng.standard_normal((batch, n)))
    xt = torch.from_numpy(x).to(dtype)
    ref_in = xt.to(torch.complex128).numpy()
    for inverse in (False, True):
        want = np.fft.ifft(ref_in, axis=-1) * n if inverse else np.fft.fft(ref_in, axis=-1)
        got = _fft(xt.cuda(), inverse).cpu().to(torch.complex128).numpy()
        err = np.linalg.norm(got - want) / np.linalg.norm(want)
        print(f"n={n} {dtype} inverse={inverse}: rel l2 err {err:.2e}")
        assert err <= tol


def test_fft_impulse_and_roundtrip():
    n = 2048
    x = torch.zeros(3, n, dtype=torch.complex64)
    x[0, 0] = 1
    x[1, 1] = 1
    x[2] = torch.randn(n) + 1j * torch.randn(n)
    X = _fft(x.cuda(), False)
    assert torch.allclose(X[0].cpu(), torch.ones(n, dtype=torch.complex64), atol=1e-6)
    k = torch.arange(n, dtype=torch.float64)
    want = torch.exp(-2j * np.pi * k / n).to(torch.complex64)
    assert torch.allclose(X[1].cpu(), want, atol=1e-6)
    back = _fft(X, True).cpu() / n
    assert torch.allclose(back[2], x[2], atol=2e-6)
