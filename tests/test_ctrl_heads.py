"""Controller tail: oracle against the fixture produced by the reference's own Controller
(tests/golden/make_golden.py ctrl_heads_case), and the fused kernel against both."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from oracle import ddsp_oracle as O  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_ctrl_heads.npz")
PARAMS = ("w_c", "b_c", "w_a", "b_a", "w_h", "b_h")


def _gold(dtype, device="cpu"):
    g = np.load(GOLD)
    t = lambda k: torch.tensor(g[k], dtype=dtype, device=device)  # noqa: E731
    return g, t


def test_oracle_matches_reference_fixture():
    g, t = _gold(torch.float64)
    x = t("x").requires_grad_(True)
    ps = [t(k).requires_grad_(True) for k in PARAMS]
    c, a, H = O.controller_heads(x, *ps)
    for got, key in ((c, "c"), (a, "a"), (H, "H")):
        assert np.abs(got.detach().numpy() - g[key]).max() <= 1e-14
    ((c * t("g_c")).sum() + (a * t("g_a")).sum() + (H * t("g_h")).sum()).backward()
    assert np.abs(x.grad.numpy() - g["gx"]).max() <= 1e-12 * np.abs(g["gx"]).max()
    for p, k in zip(ps, PARAMS):
        assert np.abs(p.grad.numpy() - g["g" + k]).max() <= 1e-12 * np.abs(g["g" + k]).max(), k


def _run_kernel(x, ps, gs):
    from ddsp_pytorch_b200 import ops
    x = x.clone().requires_grad_(True)
    ps = [p.clone().requires_grad_(True) for p in ps]
    c, a, H = ops.controller_heads(x, *ps)
    ((c * gs[0]).sum() + (a * gs[1]).sum() + (H * gs[2]).sum()).backward()
    return (c.detach(), a.detach(), H.detach()), x.grad, [p.grad for p in ps]


def _rel(got, want):
    want = want.double()
    return float((got.double() - want).abs().max() / want.abs().max().clamp_min(1e-30))


@pytest.mark.gpu
@pytest.mark.parametrize("path", ["ffma", "tcgen05"])
def test_kernel_matches_reference_fixture(path, monkeypatch):
    """Both product paths against the fixture the reference's own Controller produced (150 rows, U = 64, J = 35), same bounds."""
    from ddsp_pytorch_b200 import ops
    monkeypatch.setattr(ops, "CTRL_TC_MIN_ROWS", 0 if path == "ffma" else 1)
    g, t = _gold(torch.float32, "cuda")
    assert ops._ctrl_use_tensor_cores(150, 64) == (path == "tcgen05")
    (c, a, H), gx, gps = _run_kernel(t("x"), [t(k) for k in PARAMS], [t("g_c"), t("g_a"), t("g_h")])
    _, t64 = _gold(torch.float64, "cuda")
    for got, key in ((c, "c"), (a, "a"), (H, "H")):
        want = t64(key)
        err = float(((got.double() - want).abs() / (want.abs() + 1e-6)).max())
        print(key, "max rel err", err)
        assert got.shape == want.shape and err <= 2e-5, key          # fp32 product of 64 terms + fp32 exp/log
    assert _rel(gx, t64("gx")) <= 1e-5
    for gp, k in zip(gps, PARAMS):
        e = _rel(gp, t64("g" + k))
        print("g" + k, e)
        assert e <= 1e-5, k


@pytest.mark.gpu
@pytest.mark.parametrize("B,T,U,K,NB", [(32, 172, 512, 180, 195),    # cfg 2 (shipped config): J = 376
                                       (4, 172, 512, 180, 195),     # one rank of eight
                                       (16, 250, 512, 100, 65),     # config/default.py: J = 166, ld 168
                                       (1, 1, 512, 180, 195),       # rt/synth.py: one frame
                                       (2, 37, 24, 3, 2),           # ragged everything
                                       (1, 300, 128, 256, 65),
                                       (1, 4, 512, 180, 195),       # rt/synth.py: a block of four frames (small-R forward kernel)
                                       (2, 8, 64, 21, 13),          # 16 rows: the small kernel's last size
                                       (1, 17, 128, 5, 3),          # 17 rows: the tiled kernel's first
                                       (4, 100, 64, 3, 2),          # J = 6: far fewer output columns than a tile
                                       (3, 130, 72, 50, 20)])       # U not a multiple of the 64-wide K panel
@pytest.mark.parametrize("path", ["ffma", "tcgen05", "auto"])
def test_kernel_matches_fp64_torch(B, T, U, K, NB, path, monkeypatch):
    from ddsp_pytorch_b200 import ops
    # "tcgen05": every shape the tensor-core path accepts (U >= 64, U % 8 == 0) goes through it, down to one row
    monkeypatch.setattr(ops, "CTRL_TC_MIN_ROWS", {"ffma": 0, "tcgen05": 1, "auto": ops.CTRL_TC_MIN_ROWS}[path])
    if path == "tcgen05":
        assert ops._ctrl_use_tensor_cores(B * T, U) == (U >= 64 and U % 8 == 0)
    torch.manual_seed(B * 1000 + T)
    dev = "cuda"
    x = torch.randn(B, T, U, device=dev)
    ps = [torch.randn(K, U, device=dev) * 0.2, torch.randn(K, device=dev), torch.randn(1, U, device=dev) * 0.2,
          torch.randn(1, device=dev), torch.randn(NB, U, device=dev) * 0.2, torch.randn(NB, device=dev)]
    gs = [torch.randn(B, T, K, device=dev), torch.randn(B, T, 1, device=dev), torch.randn(B, T, NB, device=dev)]
    outs, gx, gps = _run_kernel(x, ps, gs)
    xd = x.double().requires_grad_(True)
    pd = [p.double().requires_grad_(True) for p in ps]
    want = O.controller_heads(xd, *pd)
    (sum((w * g.double()).sum() for w, g in zip(want, gs))).backward()
    for got, w in zip(outs, want):
        assert got.shape == w.shape and got.dtype == torch.float32
        assert float(((got.double() - w).abs() / (w.abs() + 1e-6)).max()) <= 1e-4
    assert _rel(gx, xd.grad) <= 2e-5
    for gp, p in zip(gps, pd):
        assert gp.shape == p.shape and _rel(gp, p.grad) <= 2e-5
    # deterministic: a second run gives the same bits
    outs2, gx2, gps2 = _run_kernel(x, ps, gs)
    assert all(torch.equal(a, b) for a, b in zip(outs + (gx,) + tuple(gps), outs2 + (gx2,) + tuple(gps2)))


@pytest.mark.gpu
def test_controller_uses_fused_tail_and_matches_torch_tail():
    from ddsp_pytorch_b200.config import Config
    from ddsp_pytorch_b200.controller import Controller
    conf = Config()
    torch.manual_seed(1)
    ctl = Controller(conf).cuda()
    batch = {k: torch.rand(2, 20, 1, device="cuda") for k in ("normalized_cents", "loudness", "f0")}
    res = {}
    for fused in (True, False):
        ctl.fused_tail = fused
        ctl.zero_grad(set_to_none=True)
        out = ctl(batch)
        (out["c"].sum() + 3 * out["a"].sum() + out["H"].square().sum()).backward()
        res[fused] = (out, {n: p.grad.clone() for n, p in ctl.named_parameters()})
    for k in ("c", "a", "H"):
        assert torch.allclose(res[True][0][k], res[False][0][k], rtol=2e-5, atol=1e-7)
    for n in res[True][1]:
        a, b = res[True][1][n], res[False][1][n]
        assert float((a - b).abs().max()) <= 1e-4 * float(b.abs().max()) + 1e-9, n
    # gradient flows to part of the outputs only (g_a = g_H = None path)
    ctl.fused_tail = True
    ctl.zero_grad(set_to_none=True)
    ctl(batch)["c"].sum().backward()
    assert ctl.dense_filter.weight.grad.abs().max() == 0 and ctl.dense_harmonic.weight.grad.abs().max() > 0

@pytest.mark.gpu
def test_tensor_core_tail_random_shapes(monkeypatch):
    """Twenty-four random (R, U, K, NB) through the tcgen05 kernels against fp64 torch: ragged tiles in every dimension (R, J and U
    not multiples of 128 / 128 / 64), K-split counts from 1 up, single rows, J below one tile and above three."""
    from ddsp_pytorch_b200 import ops
    monkeypatch.setattr(ops, "CTRL_TC_MIN_ROWS", 1)
    rng = np.random.default_rng(2025)
    dev = "cuda"
    for case in range(24):
        R = int(rng.choice([1, 7, 127, 128, 129, 300, 777, 1500, 2049]))
        U = int(rng.choice([64, 72, 128, 200, 512, 520]))
        K = int(rng.integers(1, 300))
        NB = int(rng.integers(1, 260))
        torch.manual_seed(case)
        x = torch.randn(R, U, device=dev)
        ps = [torch.randn(K, U, device=dev) * 0.1, torch.randn(K, device=dev), torch.randn(1, U, device=dev) * 0.1,
              torch.randn(1, device=dev), torch.randn(NB, U, device=dev) * 0.1, torch.randn(NB, device=dev)]
        gs = [torch.randn(R, K, device=dev), torch.randn(R, 1, device=dev), torch.randn(R, NB, device=dev)]
        assert ops._ctrl_use_tensor_cores(R, U)
        outs, gx, gps = _run_kernel(x, ps, gs)
        xd = x.double().requires_grad_(True)
        pd = [p.double().requires_grad_(True) for p in ps]
        want = O.controller_heads(xd, *pd)
        (sum((w * g.double()).sum() for w, g in zip(want, gs))).backward()
        tag = f"case {case}: R={R} U={U} K={K} NB={NB}"
        for got, w in zip(outs, want):
            assert float(((got.double() - w.detach()).abs() / (w.detach().abs() + 1e-6)).max()) <= 1e-4, tag
        assert _rel(gx, xd.grad) <= 2e-5, tag
        for gp, p in zip(gps, pd):
            assert _rel(gp, p.grad) <= 2e-5, tag
