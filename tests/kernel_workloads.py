"""Every kernel family of the library once, forward and backward, on small shapes.  Used by tests/test_gpu_guards.py
(guard bands + poisoned scratch: the stand-in for compute-sanitizer, which this pool does not run) and by
profiles/probes/sanitize_all.py (the same workload under `compute-sanitizer --tool ...` on a box that allows it).

The shapes are small and still take the paths the large configurations take: hop 512 with K = 100 launches the
hand-scheduled oscillator loops (synthesis 8 samples per lane, gradient 16, phase gradient), hop 64 / 256 the compiled-in
ones; the loss runs every register-tiled scale, the generic shared-memory kernel (overlap 0.5) and both gather kernels; the
reverb runs the partitioned 4096-point path (impulse longer than one block) and the per-row and live variants.

Each workload takes a seeded CPU generator and returns the list of tensors a caller would read (outputs and gradients)."""
import torch

from ddsp_pytorch_b200 import _lib, ops


def _dev():
    return torch.device("cuda")


def _rnd(g, *shape, scale=1.0):
    return (torch.randn(*shape, generator=g) * scale).to(_dev())


def _uni(g, *shape):
    return torch.rand(*shape, generator=g).to(_dev())


def osc(g):
    out = []
    for hop, K, T, B, sr in ((512, 100, 6, 2, 16000), (64, 37, 9, 3, 16000), (256, 256, 4, 1, 48000), (100, 5, 3, 2, 8000)):
        f0 = (_uni(g, B, T, 1) * 400 + 80).requires_grad_(True)
        c, a = _uni(g, B, T, K).requires_grad_(True), _uni(g, B, T, 1).requires_grad_(True)
        y = ops.oscillator_bank(f0, c, a, sr, hop)
        (y * _rnd(g, B, T * hop)).sum().backward()
        yl, last = ops.oscillator_live(f0.detach()[:1], c.detach()[:1], a.detach()[:1], torch.zeros(K, device=_dev()), sr, hop)
        cr = c.detach().clone().requires_grad_(True)
        up = ops.rescale(cr, hop)
        (up * _rnd(g, *up.shape)).sum().backward()
        out += [y.detach(), f0.grad, c.grad, a.grad, yl, last, up.detach(), cr.grad]
    return out


def noise(g):
    out = []
    for hop, NB, T, B in ((512, 65, 6, 2), (64, 33, 9, 3), (256, 129, 5, 1), (100, 17, 3, 2)):
        H = _uni(g, B, T, NB).requires_grad_(True)
        y = ops.filtered_noise(H, hop, seed=3)
        (y * _rnd(g, B, T * hop)).sum().backward()
        out += [y.detach(), H.grad, ops.noise_uniform(B, T, hop, 3), ops.amp_to_impulse_response(H.detach(), hop)]
    return out


def reverb(g):
    B, N, L = 3, 3 * 4096 + 100, 5000
    x = _rnd(g, B, N).requires_grad_(True)
    nz = _rnd(g, L).requires_grad_(True)
    decay = torch.tensor(2.0, device=_dev(), requires_grad=True)
    wet = torch.tensor(0.1, device=_dev(), requires_grad=True)
    t = torch.linspace(0, 1, L, device=_dev()).view(1, L)
    ir = ops.reverb_impulse(nz, decay, wet, t)
    y = ops.causal_conv(x, ir)
    (y * _rnd(g, B, N)).sum().backward()
    xs, ks = _rnd(g, 4, 6, 1024).requires_grad_(True), _rnd(g, 4, 6, 1024).requires_grad_(True)
    yr = ops.row_conv(xs, ks)
    (yr * _rnd(g, 4, 6, 1024)).sum().backward()
    buf = torch.zeros(1, 4096, device=_dev())
    yl = ops.reverb_live(buf, _rnd(g, 512), _rnd(g, 4096, scale=0.01))
    return [ir.detach(), y.detach(), x.grad, nz.grad, decay.grad, wet.grad, yr.detach(), xs.grad, ks.grad, yl, buf]


def mss(g):
    out = []
    for B, N, scales, overlap in ((3, 9000, [2048, 1024, 512, 256, 128, 64], 0.75), (3, 9000, [512, 64], 0.5),
                                  (2, 8192, [1024, 256], 0.75)):          # N % 4 == 0: the four-samples-per-thread gather
        xp, xt = _rnd(g, B, N).requires_grad_(True), _rnd(g, B, N)
        loss = ops.mss_loss(xp, xt, scales, overlap=overlap)
        loss.backward()
        out += [loss.detach(), xp.grad]
    return out


def features(g):
    n_fft, hop = 2048, 512
    ld = ops.loudness(_rnd(g, 2, 16000), _rnd(g, n_fft // 2 + 1), n_fft, hop)
    out = [ld]
    for U, K, NB, R in ((512, 100, 65, 70),          # FP32-pipe Controller tail
                        (264, 100, 65, 300)):        # tcgen05 tail (R >= ops.CTRL_TC_MIN_ROWS): ragged in R, J and U (TMA zero fill)
        x = _rnd(g, R, U).requires_grad_(True)
        ws = [_rnd(g, K, U, scale=0.05), _rnd(g, K), _rnd(g, 1, U, scale=0.05), _rnd(g, 1), _rnd(g, NB, U, scale=0.05), _rnd(g, NB)]
        ws = [w.requires_grad_(True) for w in ws]
        c, a, H = ops.controller_heads(x, *ws)
        ((c * _rnd(g, R, K)).sum() + (a * _rnd(g, R, 1)).sum() + (H * _rnd(g, R, NB)).sum()).backward()
        out += [c.detach(), a.detach(), H.detach(), x.grad] + [w.grad for w in ws]
    return out


def fft(g):
    out = []
    for n in (8, 64, 512, 4096):
        for dt in (torch.complex64, torch.complex128):
            z = torch.view_as_complex(torch.randn(5, n, 2, generator=g, dtype=torch.float64)).to(_dev()).to(dt)
            for inv in (0, 1):
                res = torch.empty_like(z)
                _lib.call("ddsp_fft_c2c", z.data_ptr(), res.data_ptr(), n, 5, inv, int(dt == torch.complex128),
                          torch.cuda.current_stream().cuda_stream)
                out.append(torch.view_as_real(res))
    return out


def dftgemm(g):
    return [ops.dftgemm_sss_terms(_rnd(g, 2, 4096), _rnd(g, 2, 4096), n) for n in (64, 128)]


WORKLOADS = {"osc": osc, "noise": noise, "reverb": reverb, "mss": mss, "features": features, "fft": fft, "dftgemm": dftgemm}
