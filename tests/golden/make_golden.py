"""Generate the golden fixtures in this directory by RUNNING THE REFERENCE ITSELF.

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    python tests/golden/make_golden.py

It imports kureta/ddsp-pytorch's own modules from ``/root/reference`` (read-only), drives them
with small seeded inputs (all inputs are fp32-representable so the fp32 CUDA path sees the
same values) in fp64 (truth) and fp32 (the reference's own drift), and stores inputs and
outputs as ``ref_*.npz``.  Gradients come from autograd through the fp64 reference.
The reference ships no golden vectors of its own; these files are the pin.
"""
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

REF = os.environ.get("DDSP_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
from model.ddsp.harmonic_oscillator import OscillatorBank      # noqa: E402
from model.ddsp import filtered_noise as ref_fn                # noqa: E402
from model.ddsp.reverb import Reverb                           # noqa: E402
from loss.mss_loss import MSSLoss                              # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle.ddsp_oracle import synthetic_controls              # noqa: E402


def conf(K, sr, hop, NB=65):
    return SimpleNamespace(n_harmonics=K, sample_rate=sr, hop_length=hop, n_noise_filters=NB)


def f32(x):
    return x.detach().to(torch.float32).numpy().copy()


def f64(x):
    return x.detach().to(torch.float64).numpy().copy()


class supplied_rand:
    """Make ``torch.rand`` inside the reference return a chosen tensor (filtered_noise.py:44)."""

    def __init__(self, u01):
        self.u01 = u01

    def __enter__(self):
        self._orig = torch.rand
        torch.rand = lambda *shape, **kw: self.u01.clone()
        return self

    def __exit__(self, *exc):
        torch.rand = self._orig


def osc_case(name, B, T, K, sr, hop, seed, f0_override=None):
    f0, c, a, _ = synthetic_controls(B, T, K, 5, sr, hop, seed)
    if f0_override is not None:
        f0 = f0_override(f0)
    g = torch.Generator().manual_seed(seed + 17)
    gup = torch.randn(B, T * hop, generator=g)
    out = {"B": B, "T": T, "K": K, "sr": sr, "hop": hop, "f0": f32(f0), "c": f32(c), "a": f32(a), "gup": f32(gup)}
    with torch.no_grad():
        y32 = OscillatorBank(conf(K, sr, hop))({"f0": f0, "c": c, "a": a})
    bank = OscillatorBank(conf(K, sr, hop)).double()
    f0d = f0.double().requires_grad_(True)
    cd = c.double().requires_grad_(True)
    ad = a.double().requires_grad_(True)
    y64 = bank({"f0": f0d, "c": cd, "a": ad})
    (y64 * gup.double()).sum().backward()
    out.update(y64=f64(y64), y32=f32(y32), gc=f64(cd.grad), ga=f64(ad.grad), gf0=f64(f0d.grad))
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "max|y32-y64| =", float((y32.double() - y64).abs().max()))


def osc_live_case(name, T, K, sr, hop, seed, calls=3):
    out = {"T": T, "K": K, "sr": sr, "hop": hop, "calls": calls}
    bank32 = OscillatorBank(conf(K, sr, hop))
    bank64 = OscillatorBank(conf(K, sr, hop)).double()
    for i in range(calls):
        f0, c, a, _ = synthetic_controls(1, T, K, 5, sr, hop, seed + i)
        f0 = f0 * 0 + f0[:, :1] * (1.0 + 0.02 * i)           # steady pitch per block, changes between blocks
        with torch.no_grad():
            y32 = bank32.live({"f0": f0, "c": c, "a": a})
            y64 = bank64.live({"f0": f0.double(), "c": c.double(), "a": a.double()})
        out.update({f"f0_{i}": f32(f0), f"c_{i}": f32(c), f"a_{i}": f32(a), f"y64_{i}": f64(y64),
                    f"y32_{i}": f32(y32), f"phase64_{i}": f64(bank64.last_phases.data),
                    f"phase32_{i}": f32(bank32.last_phases.data)})
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "ok")


def noise_case(name, B, T, NB, hop, seed):
    _, _, _, H = synthetic_controls(B, T, 4, NB, 16000, hop, seed)
    g = torch.Generator().manual_seed(seed + 5)
    u01 = torch.rand(B, T, hop, generator=g)
    gup = torch.randn(B, T * hop, generator=g)
    Hd = H.double().requires_grad_(True)
    mod = ref_fn.FilteredNoise(conf(4, 16000, hop, NB))
    with supplied_rand(u01):
        y64 = mod({"H": Hd})
    (y64 * gup.double()).sum().backward()
    with supplied_rand(u01), torch.no_grad():
        y32 = mod({"H": H})
    ir64 = ref_fn.amp_to_impulse_response(H.double(), hop)
    np.savez_compressed(os.path.join(HERE, name), B=B, T=T, NB=NB, hop=hop, H=f32(H), noise=f32(u01 * 2 - 1),
                        gup=f32(gup), y64=f64(y64), y32=f32(y32), gH=f64(Hd.grad), ir64=f64(ir64))
    print(name, "max|y32-y64| =", float((y32.double() - y64).abs().max()))


def reverb_case(name, B, N, sr, seed, wet=0.3, decay=4.0):
    torch.manual_seed(seed)
    rv = Reverb(conf(4, sr, 64), initial_wet=wet, initial_decay=decay)
    g = torch.Generator().manual_seed(seed + 3)
    x = torch.randn(B, N, generator=g) * 0.3
    gup = torch.randn(B, N, generator=g)
    with torch.no_grad():
        y32 = rv(x)
    rv64 = Reverb(conf(4, sr, 64), initial_wet=wet, initial_decay=decay).double()
    rv64.noise.data = rv.noise.data.double()
    xd = x.double().requires_grad_(True)
    y64 = rv64(xd)
    (y64 * gup.double()).sum().backward()
    np.savez_compressed(os.path.join(HERE, name), B=B, N=N, sr=sr, x=f32(x), gup=f32(gup), noise=f32(rv.noise),
                        decay=float(rv64.decay), wet=float(rv64.wet), t=f32(rv.t), ir64=f64(rv64.build_impulse()), y64=f64(y64), y32=f32(y32),
                        gx=f64(xd.grad), gnoise=f64(rv64.noise.grad), gdecay=f64(rv64.decay.grad),
                        gwet=f64(rv64.wet.grad))
    print(name, "max|y32-y64| =", float((y32.double() - y64).abs().max()))


def reverb_live_case(name, sr, block, calls, seed, wet=0.3, decay=4.0):
    torch.manual_seed(seed)
    rv = Reverb(conf(4, sr, 64), initial_wet=wet, initial_decay=decay)
    rv64 = Reverb(conf(4, sr, 64), initial_wet=wet, initial_decay=decay).double()
    rv64.noise.data = rv.noise.data.double()
    out = {"sr": sr, "block": block, "calls": calls, "noise": f32(rv.noise), "decay": float(rv64.decay), "wet": float(rv64.wet),
           "t": f32(rv.t)}
    g = torch.Generator().manual_seed(seed + 9)
    with torch.no_grad():
        for i in range(calls):
            x = torch.randn(1, block, generator=g) * 0.3
            y64 = rv64.live_forward(x.double())
            out.update({f"x_{i}": f32(x), f"y64_{i}": f64(y64), f"buf64_{i}": f64(rv64.buffer)})
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "ok")


def mss_case(name, B, N, n_ffts, seed):
    g = torch.Generator().manual_seed(seed)
    xt = (torch.randn(B, N, generator=g) * 0.1).clamp(-1, 1)
    xp = 0.5 * xt + torch.randn(B, N, generator=g) * 0.05
    loss_mod = MSSLoss(list(n_ffts)).double()
    xpd = xp.double().requires_grad_(True)
    l64 = loss_mod(xpd, xt.double())
    l64.backward()
    with torch.no_grad():
        l32 = MSSLoss(list(n_ffts))(xp, xt)
        per_scale = [float(m(xpd, xt.double())) for m in loss_mod.losses]
    np.savez_compressed(os.path.join(HERE, name), B=B, N=N, n_ffts=np.array(n_ffts), x_pred=f32(xp), x_true=f32(xt),
                        loss64=float(l64), loss32=float(l32), per_scale64=np.array(per_scale), gx=f64(xpd.grad))
    print(name, "loss64 =", float(l64), "loss32 =", float(l32))


def path_case(name, B, T, K, NB, sr, hop, seed):
    """harmonics + noise -> reverb -> MSS, as decoder.py:129-133 and train.py:33-34 wire it."""
    f0, c, a, H = synthetic_controls(B, T, K, NB, sr, hop, seed)
    g = torch.Generator().manual_seed(seed + 1)
    u01 = torch.rand(B, T, hop, generator=g)
    target = (torch.randn(B, T * hop, generator=g) * 0.1).clamp(-1, 1)
    torch.manual_seed(seed)
    rv = Reverb(conf(K, sr, hop), initial_wet=0.5, initial_decay=4.0).double()
    bank = OscillatorBank(conf(K, sr, hop)).double()
    fn = ref_fn.FilteredNoise(conf(K, sr, hop, NB))
    n_ffts = [256, 128, 64]
    loss_mod = MSSLoss(n_ffts).double()
    cd, ad, Hd = (t.double().requires_grad_(True) for t in (c, a, H))
    with supplied_rand(u01):
        dry = bank({"f0": f0.double(), "c": cd, "a": ad}) + fn({"H": Hd})
    y = rv(dry)
    loss = loss_mod(y, target.double())
    loss.backward()
    np.savez_compressed(os.path.join(HERE, name), B=B, T=T, K=K, NB=NB, sr=sr, hop=hop, n_ffts=np.array(n_ffts),
                        f0=f32(f0), c=f32(c), a=f32(a), H=f32(H), noise=f32(u01 * 2 - 1), target=f32(target),
                        rev_noise=f32(rv.noise), rev_decay=float(rv.decay), rev_wet=float(rv.wet), dry64=f64(dry), y64=f64(y),
                        loss64=float(loss), gc=f64(cd.grad), ga=f64(ad.grad), gH=f64(Hd.grad),
                        g_rev_noise=f64(rv.noise.grad), g_rev_decay=f64(rv.decay.grad), g_rev_wet=f64(rv.wet.grad))
    print(name, "loss64 =", float(loss))


def ctrl_heads_case(name, B, T, U, K, NB, seed):
    """The reference's own Controller (decoder.py:41-116), fp64: the input of its three heads (output of mlp_gru, taken
    with a forward hook), the head parameters, c / a / H and the gradients of a random linear functional of them."""
    from model.autoencoder.decoder import Controller
    torch.manual_seed(seed)
    cf = SimpleNamespace(n_harmonics=K, n_noise_filters=NB, decoder_mlp_units=U, decoder_mlp_layers=2,
                         decoder_gru_units=U, decoder_gru_layers=1)
    ctl = Controller(cf)
    with torch.no_grad():       # spread the pre-activations so both tails of the sigmoid are exercised
        for head in (ctl.dense_harmonic, ctl.dense_loudness, ctl.dense_filter):
            head.weight.mul_(6.0)
            head.bias.add_(torch.randn_like(head.bias))
            head.weight.copy_(head.weight.float())      # fp32-representable parameters
            head.bias.copy_(head.bias.float())
    ctl = ctl.double()
    seen = {}

    def hook(_m, _inp, out):
        out.retain_grad()
        seen["x"] = out
    ctl.mlp_gru.register_forward_hook(hook)
    batch = dict(normalized_cents=torch.rand(B, T, 1, dtype=torch.float64), loudness=torch.rand(B, T, 1, dtype=torch.float64),
                 f0=torch.rand(B, T, 1, dtype=torch.float64) * 300 + 80)
    out = ctl(batch)
    # the fp32 kernel sees the fp32 rounding of the latent: re-run the heads on that so the pin is exact
    x = seen["x"].detach().float().double().requires_grad_(True)
    c = ctl.modified_sigmoid(ctl.dense_harmonic(x))
    a = ctl.modified_sigmoid(ctl.dense_loudness(x))
    H = ctl.modified_sigmoid(ctl.dense_filter(x))
    assert torch.allclose(c, out["c"], atol=1e-6) and torch.allclose(H, out["H"], atol=1e-6)
    g = [torch.randn_like(t).float().double() for t in (c, a, H)]
    for p_ in ctl.parameters():
        p_.grad = None
    ((c * g[0]).sum() + (a * g[1]).sum() + (H * g[2]).sum()).backward()
    np.savez_compressed(os.path.join(HERE, name), B=B, T=T, U=U, K=K, NB=NB, x=f32(x),
                        w_c=f32(ctl.dense_harmonic.weight), b_c=f32(ctl.dense_harmonic.bias),
                        w_a=f32(ctl.dense_loudness.weight), b_a=f32(ctl.dense_loudness.bias),
                        w_h=f32(ctl.dense_filter.weight), b_h=f32(ctl.dense_filter.bias),
                        c=f64(c), a=f64(a), H=f64(H), g_c=f32(g[0]), g_a=f32(g[1]), g_h=f32(g[2]), gx=f64(x.grad),
                        gw_c=f64(ctl.dense_harmonic.weight.grad), gb_c=f64(ctl.dense_harmonic.bias.grad),
                        gw_a=f64(ctl.dense_loudness.weight.grad), gb_a=f64(ctl.dense_loudness.bias.grad),
                        gw_h=f64(ctl.dense_filter.weight.grad), gb_h=f64(ctl.dense_filter.bias.grad))
    print(name, "c range", float(c.min()), float(c.max()))


if __name__ == "__main__":
    torch.set_num_threads(4)
    if len(sys.argv) > 1 and sys.argv[1] == "ctrl":     # only the controller-tail fixtures (added with row X1)
        ctrl_heads_case("ref_ctrl_heads.npz", B=3, T=50, U=64, K=21, NB=13, seed=14)   # R=150: ragged tiles, J=35 (ld 36)
        sys.exit(0)
    osc_case("ref_osc_small.npz", B=2, T=6, K=24, sr=16000, hop=64, seed=1)
    osc_case("ref_osc_default.npz", B=1, T=8, K=180, sr=44100, hop=512, seed=2)
    # high pitch: most harmonics masked, some frames cross Nyquist between neighbours
    osc_case("ref_osc_nyquist.npz", B=2, T=10, K=100, sr=16000, hop=128, seed=3,
             f0_override=lambda f: (f * 0 + torch.linspace(300.0, 3900.0, f.shape[1]).view(1, -1, 1)))
    # long clip: the fp32 reference's phase drift becomes visible
    osc_case("ref_osc_long.npz", B=1, T=125, K=32, sr=16000, hop=512, seed=4)
    osc_live_case("ref_osc_live.npz", T=4, K=60, sr=44100, hop=256, seed=5)
    noise_case("ref_noise_small.npz", B=2, T=3, NB=17, hop=64, seed=6)
    noise_case("ref_noise_full.npz", B=2, T=2, NB=33, hop=64, seed=7)          # M == hop: zero pad length
    noise_case("ref_noise_default.npz", B=1, T=2, NB=195, hop=512, seed=8)     # shipped config, M=388
    reverb_case("ref_reverb_long.npz", B=2, N=1536, sr=1000, seed=9)           # N > L
    reverb_case("ref_reverb_crop.npz", B=2, N=512, sr=1000, seed=10)           # N < L: IR cropped (reverb.py:34)
    reverb_live_case("ref_reverb_live.npz", sr=1000, block=256, calls=3, seed=11)
    mss_case("ref_mss.npz", B=2, N=4096, n_ffts=[2048, 1024, 512, 256, 128, 64], seed=12)
    path_case("ref_path.npz", B=2, T=12, K=20, NB=17, sr=4000, hop=64, seed=13)
    ctrl_heads_case("ref_ctrl_heads.npz", B=3, T=50, U=64, K=21, NB=13, seed=14)
