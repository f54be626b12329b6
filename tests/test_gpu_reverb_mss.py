"""Parity of the CUDA reverb, spectral loss and streaming paths with the reference's fp64 outputs (golden
fixtures) and with the oracle.  Tolerances per SURVEY.md §8c: reverb <= 1e-4 of max|y|; MSS loss and all
gradients <= 1e-4 relative (norm-wise)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-4


def rel(got, want):
    want = torch.as_tensor(want).double()
    return float((got.detach().double().cpu() - want).abs().max() / want.abs().max().clamp_min(1e-30))


def l2rel(got, want):
    want = torch.as_tensor(want).double()
    return float((got.detach().double().cpu() - want).norm() / want.norm().clamp_min(1e-30))


def cu(x, grad=False):
    return torch.as_tensor(np.asarray(x), dtype=torch.float32).cuda().requires_grad_(grad)


@pytest.mark.parametrize("name", ["ref_reverb_long.npz", "ref_reverb_crop.npz"])
def test_reverb_golden(golden, name):
    from ddsp_pytorch_b200 import ops
    g = golden(name)
    noise, x = cu(g["noise"], True), cu(g["x"], True)
    decay, wet = cu(g["decay"], True), cu(g["wet"], True)
    ir = ops.reverb_impulse(noise, decay, wet, cu(g["t"]))
    assert rel(ir, g["ir64"]) <= 1e-6
    y = ops.causal_conv(x, ir)
    assert rel(y, g["y64"]) <= TOL
    (y * cu(g["gup"])).sum().backward()
    assert rel(x.grad, g["gx"]) <= TOL
    assert rel(noise.grad, g["gnoise"]) <= TOL
    assert abs(float(decay.grad) - float(g["gdecay"])) <= TOL * abs(float(g["gdecay"]))
    assert abs(float(wet.grad) - float(g["gwet"])) <= TOL * abs(float(g["gwet"]))


@pytest.mark.parametrize("B,N,L", [(3, 9000, 5000), (2, 4096, 4096), (1, 2048, 44100), (5, 20000, 16000), (2, 6000, 2049),
                                   (1, 100, 7),
                                   # 13 partitions (the second lane of a pair holds one), 25 and 28 partitions (a second MAC /
                                   # kernel-gradient pass that accumulates onto the first)
                                   (4, 30000, 26000), (3, 52000, 50000), (2, 60000, 56000)])
def test_conv_against_oracle(B, N, L):
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    gen = torch.Generator().manual_seed(N + L)
    x = torch.randn(B, N, generator=gen) * 0.3
    ir = torch.randn(1, L, generator=gen) * torch.exp(-torch.arange(L) / (0.2 * L))
    gup = torch.randn(B, N, generator=gen)
    xd, ird = x.double().requires_grad_(True), ir.double().requires_grad_(True)
    want = O.reverb_apply(xd, ird)
    (want * gup.double()).sum().backward()
    xc, irc = x.cuda().requires_grad_(True), ir.cuda().requires_grad_(True)
    y = ops.causal_conv(xc, irc)
    (y * gup.cuda()).sum().backward()
    assert rel(y, want.detach()) <= TOL
    assert rel(xc.grad, xd.grad) <= TOL
    assert rel(irc.grad, ird.grad) <= TOL


def test_conv_backward_twice_and_partial_grads():
    """The first backward reuses the spectra the forward left in its scratch buffer and overwrites them; a second
    backward (retain_graph) recomputes them -- both give the same bits.  Also gradient w.r.t. only one of the inputs."""
    from ddsp_pytorch_b200 import ops
    gen = torch.Generator().manual_seed(5)
    x = (torch.randn(3, 9000, generator=gen) * 0.3).cuda().requires_grad_(True)
    ir = (torch.randn(1, 5000, generator=gen) * torch.exp(-torch.arange(5000) / 1000.0)).cuda().requires_grad_(True)
    gup = torch.randn(3, 9000, generator=gen).cuda()
    y = ops.causal_conv(x, ir)
    g1 = torch.autograd.grad((y * gup).sum(), (x, ir), retain_graph=True)
    g2 = torch.autograd.grad((y * gup).sum(), (x, ir))
    assert torch.equal(g1[0], g2[0]) and torch.equal(g1[1], g2[1])
    y = ops.causal_conv(x.detach(), ir)                 # kernel gradient only
    assert torch.equal(torch.autograd.grad((y * gup).sum(), ir)[0], g1[1])
    y = ops.causal_conv(x, ir.detach())                 # signal gradient only
    assert torch.equal(torch.autograd.grad((y * gup).sum(), x)[0], g1[0])


def test_reverb_module_forward_and_live(golden):
    import ddsp_pytorch_b200 as d
    g = golden("ref_reverb_live.npz")
    sr = int(g["sr"])
    rv = d.Reverb(d.Config(sample_rate=sr), initial_wet=float(g["wet"]), initial_decay=float(g["decay"])).cuda()
    rv.noise.data = cu(g["noise"])
    assert set(rv.state_dict()) == {"noise", "decay", "wet", "t", "buffer"}
    for i in range(int(g["calls"])):
        y = rv.live_forward(cu(g[f"x_{i}"]))
        assert rel(y, g[f"y64_{i}"]) <= TOL
        assert float((rv.buffer.double().cpu() - torch.from_numpy(g[f"buf64_{i}"])).abs().max()) < 1e-7


def test_mss_golden(golden):
    from ddsp_pytorch_b200 import ops
    import ddsp_pytorch_b200 as d
    g = golden("ref_mss.npz")
    n_ffts = [int(n) for n in g["n_ffts"]]
    xp, xt = cu(g["x_pred"], True), cu(g["x_true"])
    loss = d.MSSLoss(n_ffts).cuda()(xp, xt)
    assert loss.dim() == 0 and loss.dtype == torch.float32
    assert abs(float(loss) - float(g["loss64"])) <= TOL * float(g["loss64"])
    loss.backward()
    # norm-wise; the fp32 reference itself sits at 5e-5..2.5e-4 here (d log2(S+eps) is ill-conditioned)
    assert l2rel(xp.grad, g["gx"]) <= TOL
    for n, want in zip(n_ffts, g["per_scale64"]):
        assert abs(float(ops.mss_loss(xp.detach(), xt, [n])) - float(want)) <= TOL * float(want)


@pytest.mark.parametrize("B,N,n_ffts", [(3, 5000, [64, 128, 256]), (1, 2050, [4096, 2048]), (2, 777, [16, 32, 8]),
                                        (1, 1001, [4, 8, 512, 1024]), (5, 333, [128]),
                                        (4, 64000, [2048, 1024, 512, 256, 128, 64])])
def test_mss_against_oracle(B, N, n_ffts):
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    gen = torch.Generator().manual_seed(N)
    xt = (torch.randn(B, N, generator=gen) * 0.1).clamp(-1, 1)
    xp = 0.5 * xt + torch.randn(B, N, generator=gen) * 0.05
    xpd = xp.double().requires_grad_(True)
    want = O.mss_loss(xpd, xt.double(), n_ffts)
    (want * 0.7).backward()
    xp32 = xp.clone().requires_grad_(True)
    (O.mss_loss(xp32, xt, n_ffts) * 0.7).backward()          # the reference's own fp32 arithmetic, for scale
    ref32_err = l2rel(xp32.grad, xpd.grad)
    xpc = xp.cuda().requires_grad_(True)
    got = ops.mss_loss(xpc, xt.cuda(), n_ffts, windows=[torch.hann_window(n) for n in n_ffts])
    (got * 0.7).backward()
    err = l2rel(xpc.grad, xpd.grad)
    print(f"MSS grad {n_ffts}: |cuda-ref64|/|ref64| = {err:.2e}, reference fp32 vs fp64 = {ref32_err:.2e}")
    assert abs(float(got) - float(want)) <= TOL * float(want)
    assert err <= TOL


@pytest.mark.parametrize("N,n", [(4096, 8), (4096, 16), (4096, 64), (1000, 32), (5000, 128)])
def test_mss_gradient_on_noise_like_input(N, n):
    """White noise against white noise: some bins of S_pred are nearly empty, d log2(S+eps)/dx is dominated by them
    and any ulp of difference in the analysis window shows up at 1e-4.  The functional API defaults to
    torch.hann_window's own fp32 table (found by a randomised shape sweep: the in-kernel Hann gave 2e-5..7e-4)."""
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    gen = torch.Generator().manual_seed(N + n)
    xp, xt = torch.randn(2, N, generator=gen) * 0.5, torch.randn(2, N, generator=gen) * 0.5
    xpd = xp.double().requires_grad_(True)
    want = O.mss_loss(xpd, xt.double(), [n])
    want.backward()
    xpc = xp.cuda().requires_grad_(True)
    got = ops.mss_loss(xpc, xt.cuda(), [n])
    got.backward()
    assert abs(float(got) - float(want)) <= 1e-6 * float(want)
    err = float((xpc.grad.double().cpu() - xpd.grad).norm() / xpd.grad.norm())
    print(f"N={N} n={n}: grad rel l2 err {err:.2e}")
    assert err <= 2e-6


@pytest.mark.parametrize("eps", [1e-7, 1e-3, 1e-39])
def test_mss_eps_values(eps):
    """The register-tiled kernels evaluate log2(S+eps) and 1/(S+eps) with MUFU for a normal eps; a denormal eps
    (1e-39) takes the generic kernel with the library functions.  Both against the fp64 oracle at the same eps."""
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    gen = torch.Generator().manual_seed(77)
    xp, xt = torch.randn(3, 3000, generator=gen) * 0.3, torch.randn(3, 3000, generator=gen) * 0.3
    xpd = xp.double().requires_grad_(True)
    want = O.mss_loss(xpd, xt.double(), [256, 64], eps=eps)
    want.backward()
    xpc = xp.cuda().requires_grad_(True)
    got = ops.mss_loss(xpc, xt.cuda(), [256, 64], eps=eps)
    got.backward()
    assert abs(float(got) - float(want)) <= 1e-5 * float(want)
    err = float((xpc.grad.double().cpu() - xpd.grad).norm() / xpd.grad.norm())
    print(f"eps={eps:g}: grad rel l2 err {err:.2e}")
    assert err <= 1e-4


def test_mss_dict_inputs_and_identity():
    import ddsp_pytorch_b200 as d
    x = torch.randn(2, 4096).cuda() * 0.1
    m = d.MSSLoss([256, 64])
    assert float(m({"audio_reverb": x, "hidden": None}, {"audio": x})) < 1e-6     # pred goes through fp64, true through fp32


def test_osc_live_golden(golden):
    import ddsp_pytorch_b200 as d
    g = golden("ref_osc_live.npz")
    K, sr, hop = int(g["K"]), int(g["sr"]), int(g["hop"])
    bank = d.OscillatorBank(d.Config(n_harmonics=K, sample_rate=sr, hop_length=hop)).cuda()
    assert bank.last_phases.dtype == torch.int64
    for i in range(int(g["calls"])):
        y = bank.live({"f0": cu(g[f"f0_{i}"]), "c": cu(g[f"c_{i}"]), "a": cu(g[f"a_{i}"])})
        assert float((y.double().cpu() - torch.from_numpy(g[f"y64_{i}"])).abs().max()) <= TOL
        dphi = (bank.last_phases.double().cpu() - torch.from_numpy(g[f"phase64_{i}"])).abs()
        assert float(torch.minimum(dphi, 2 * np.pi - dphi).max()) <= 1e-4
    assert bank.last_phases.dtype == torch.float32
    # rt-length clip: also within 1e-3 of the fp32 reference
    assert float((y.cpu() - torch.from_numpy(g[f"y32_{int(g['calls']) - 1}"])).abs().max()) <= 1e-3


def test_whole_path_golden(golden):
    """harmonics + noise -> reverb -> MSS with gradients to every trainable input (decoder.py:129-133, train.py:33-34)."""
    import ddsp_pytorch_b200 as d
    g = golden("ref_path.npz")
    K, NB, sr, hop = int(g["K"]), int(g["NB"]), int(g["sr"]), int(g["hop"])
    conf = d.Config(n_harmonics=K, n_noise_filters=NB, sample_rate=sr, hop_length=hop)
    bank, fn, rv = d.OscillatorBank(conf).cuda(), d.FilteredNoise(conf), d.Reverb(conf, float(g["rev_wet"]), float(g["rev_decay"])).cuda()
    rv.noise.data = cu(g["rev_noise"])
    c, a, H = cu(g["c"], True), cu(g["a"], True), cu(g["H"], True)
    ctrl = {"f0": cu(g["f0"]), "c": c, "a": a, "H": H}
    dry = bank(ctrl) + fn(ctrl, noise=cu(g["noise"]))
    assert rel(dry, g["dry64"]) <= TOL
    y = rv(dry)
    assert rel(y, g["y64"]) <= TOL
    loss = d.MSSLoss([int(n) for n in g["n_ffts"]])(y, cu(g["target"]))
    assert abs(float(loss) - float(g["loss64"])) <= TOL * float(g["loss64"])
    loss.backward()
    for got, key in ((c.grad, "gc"), (a.grad, "ga"), (H.grad, "gH"), (rv.noise.grad, "g_rev_noise")):
        assert l2rel(got, g[key]) <= 2e-4, key
    assert abs(float(rv.decay.grad) - float(g["g_rev_decay"])) <= 2e-4 * abs(float(g["g_rev_decay"]))
    assert abs(float(rv.wet.grad) - float(g["g_rev_wet"])) <= 2e-4 * abs(float(g["g_rev_wet"]))


def _tied_bin_gradients(xp64, xt64, ffts, thresh=1e-5, alpha=1.0, eps=1e-7):
    """For every (scale, bin, frame) whose predicted and true power agree to `thresh` relative: the fp64 gradient of that
    one bin's two L1 terms (mss_loss.py:29-30) w.r.t. x_pred -- flipping the bin's sign changes the total gradient by
    -2 times this."""
    from oracle import ddsp_oracle as O
    out = []
    for n in ffts:
        x = xp64.detach().clone().requires_grad_(True)
        sp, st = O.power_spectrogram(x, n, n // 4), O.power_spectrogram(xt64, n, n // 4)
        tie = ((sp.detach() - st).abs() / st) < thresh
        for idx in tie.nonzero().tolist():
            b, k, f = idx
            term = ((sp[b, k, f] - st[b, k, f]).abs() + alpha * ((st[b, k, f] + eps).log2() - (sp[b, k, f] + eps).log2()).abs()) / sp.numel()
            (g,) = torch.autograd.grad(term, x, retain_graph=True)
            out.append(g)
    return out


def test_mss_gradient_exceedances_are_sign_ties():
    """The two L1 terms put a sign() into d loss / d S_pred.  A bin whose predicted and true power agree to ~1e-6 relative makes
    that sign -- and one windowed sinusoid of the gradient -- arbitrary in fp32, for this kernel and for the reference's own fp32
    run alike (about one random input in 600 at this shape; found by profiles/probes/fuzz.py, settled by profiles/probes/mss_ties.py).
    Strict form: for every case over the 1e-4 bound, the fp64 oracle gradient with the sign of (a subset of) its tied bins
    flipped must match the kernel within the same 1e-4 -- a tie explains exactly that much and nothing else."""
    import itertools
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    gen = torch.Generator().manual_seed(12345)
    ffts, over = [512, 1024], 0
    for it in range(250):
        xp, xt = torch.randn(1, 4096, generator=gen) * 0.5, torch.randn(1, 4096, generator=gen) * 0.5
        xpd = xp.double().requires_grad_(True)
        O.mss_loss(xpd, xt.double(), ffts).backward()
        xc = xp.cuda().requires_grad_(True)
        ops.mss_loss(xc, xt.cuda(), ffts).backward()
        got, want = xc.grad.double().cpu(), xpd.grad
        err = float((got - want).abs().max() / want.abs().max())
        if err > 1e-4:
            over += 1
            ties = _tied_bin_gradients(xp.double(), xt.double(), ffts)
            assert 1 <= len(ties) <= 6, f"case {it}: gradient error {err:.2e} with {len(ties)} tied bins"
            best = min(float((got - (want - 2 * sum(sub))).abs().max() / want.abs().max())
                       for r in range(1, len(ties) + 1) for sub in itertools.combinations(ties, r))
            print(f"case {it}: err {err:.2e}, {len(ties)} tied bin(s), flip-corrected err {best:.2e}")
            assert best <= 1e-4, f"case {it}: gradient error {err:.2e}; {best:.2e} after flipping the tied bins"
    assert over <= 3


def test_mss_is_bit_reproducible_and_batch_invariant():
    """No floating-point atomics in the loss (csrc/ddsp_mss.cu): 20 calls give identical loss and gradient bits, and a voice's
    gradient inside a batch of 4 is bit for bit half its gradient inside a batch of 2 (the means differ by exactly 2)."""
    from ddsp_pytorch_b200 import ops
    gen = torch.Generator().manual_seed(5)
    ffts = [2048, 1024, 512, 256, 128, 64]
    xp, xt = (torch.randn(4, 16000, generator=gen) * 0.3).cuda(), (torch.randn(4, 16000, generator=gen) * 0.3).cuda()

    def run(p, t):
        p = p.clone().requires_grad_(True)
        loss = ops.mss_loss(p, t, ffts)
        loss.backward()
        return loss.detach().clone(), p.grad.clone()

    l0, g0 = run(xp, xt)
    for _ in range(20):
        l, g = run(xp, xt)
        assert torch.equal(l, l0) and torch.equal(g, g0)
    for lo, hi in ((0, 2), (2, 4)):
        _, gs = run(xp[lo:hi], xt[lo:hi])
        assert torch.equal(gs, 2 * g0[lo:hi])
    assert float(g0.abs().max()) > 0


@pytest.mark.parametrize("n,overlap,N", [(64, 0.5, 3000), (256, 0.875, 5000), (128, 0.0, 2000), (4, 0.75, 500), (32, 0.6, 1111)])
def test_mss_other_overlaps_take_the_generic_kernel(n, overlap, N):
    """hop != n_fft/4 (SSSLoss's `overlap` argument, mss_loss.py:16-21) and n_fft = 4 run the shared-memory kernel whose
    frames are overlap-added by the gather; same parity bar, same reproducibility."""
    from ddsp_pytorch_b200 import ops
    from oracle import ddsp_oracle as O
    gen = torch.Generator().manual_seed(n + N)
    xt = torch.randn(2, N, generator=gen) * 0.3
    xp = 0.5 * xt + torch.randn(2, N, generator=gen) * 0.05
    xpd = xp.double().requires_grad_(True)
    want = O.sss_loss(xpd, xt.double(), n, overlap=overlap)
    want.backward()
    outs = []
    for _ in range(2):
        xc = xp.cuda().requires_grad_(True)
        got = ops.mss_loss(xc, xt.cuda(), [n], overlap=overlap)
        got.backward()
        outs.append((got.detach().clone(), xc.grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert abs(float(got) - float(want)) <= TOL * float(want)
    assert l2rel(xc.grad, xpd.grad) <= TOL
