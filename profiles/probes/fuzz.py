"""Randomised shape sweep of every op against the fp64 oracle (scratch tool; regressions go to tests/)."""
import sys, random, traceback
sys.path.insert(0, ".")
import torch
from ddsp_pytorch_b200 import ops
from oracle import ddsp_oracle as O

rnd = random.Random(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n_cases = int(sys.argv[2]) if len(sys.argv) > 2 else 40
bad = 0

def rel(a, b):
    b = b.double().cpu(); a = a.double().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-9))

def check(name, cfg, pairs, tol):
    global bad
    worst = max(rel(a, b) for a, b in pairs)
    finite = all(torch.isfinite(a).all() == torch.isfinite(b).all() for a, b in pairs)
    if not (worst <= tol) and finite:
        bad += 1
        print("FAIL", name, cfg, [f"{rel(a, b):.2e}" for a, b in pairs], "max|ref|", [f"{float(b.abs().max()):.2e}" for a, b in pairs],
              "max|diff|", [f"{float((a.double().cpu() - b.double().cpu()).abs().max()):.2e}" for a, b in pairs])

for it in range(n_cases):
    try:
        # oscillator
        hop = rnd.choice([1, 2, 3, 4, 5, 6, 7, 8, 10, 15, 16, 24, 32, 48, 63, 64, 100, 128, 255, 256, 384, 441, 512, 1024])
        B, T, K = rnd.randint(1, 4), rnd.randint(1, 12), rnd.choice([1, 2, 3, 7, 16, 33, 100, 180, 256])
        sr = rnd.choice([4000, 16000, 44100, 48000])
        f0, c, a, H = O.synthetic_controls(B, T, K, 9, sr, hop, seed=it)
        if rnd.random() < 0.3:
            f0 = f0 * rnd.choice([0.1, 3.0, 8.0])
            f0 = f0.clamp(1.0, sr / 2 - 1)
        g = torch.randn(B, T * hop)
        cd, ad = c.double().requires_grad_(True), a.double().requires_grad_(True)
        want = O.oscillator_bank(f0.double(), cd, ad, K, sr, hop)
        (want * g.double()).sum().backward()
        cc, ac = c.cuda().requires_grad_(True), a.cuda().requires_grad_(True)
        got = ops.oscillator_bank(f0.cuda(), cc, ac, sr, hop)
        (got * g.cuda()).sum().backward()
        ok = torch.isfinite(want).all()
        if ok and K > 1:        # K = 1: c/sum(c) == 1, the gradient is exactly zero
            check("osc", (B, T, K, hop, sr), [(got, want.detach()), (cc.grad, cd.grad), (ac.grad, ad.grad)], 2e-4)
        # noise
        hopn = rnd.choice([16, 32, 64, 80, 128, 256, 512, 40, 100, 441, 7, 30, 250])
        NB = rnd.choice([2, 3, 5, 9, 17, 33, 65, 129, 195])
        if 2 * (NB - 1) > hopn:
            NB = hopn // 2 + 1
        Bn, Tn = rnd.randint(1, 3), rnd.randint(1, 6)
        Hn = torch.rand(Bn, Tn, NB) + 0.01
        u = torch.rand(Bn, Tn, hopn) * 2 - 1
        gn = torch.randn(Bn, Tn * hopn)
        Hd = Hn.double().requires_grad_(True)
        wantn = O.filtered_noise(Hd, hopn, u.double())
        (wantn * gn.double()).sum().backward()
        Hc = Hn.cuda().requires_grad_(True)
        gotn = ops.filtered_noise(Hc, hopn, noise=u.cuda())
        (gotn * gn.cuda()).sum().backward()
        check("noise", (Bn, Tn, NB, hopn), [(gotn, wantn.detach()), (Hc.grad, Hd.grad)], 5e-5)
        # conv
        Bc, N, L = rnd.randint(1, 5), rnd.choice([1, 5, 100, 2047, 2048, 2049, 4096, 5000, 12000]), rnd.choice([1, 2, 50, 2048, 2049, 4097, 9000])
        x = torch.randn(Bc, N) * 0.3; ir = torch.randn(1, L) * torch.exp(-torch.arange(L) / (0.3 * L + 1)); gc = torch.randn(Bc, N)
        xd, ird = x.double().requires_grad_(True), ir.double().requires_grad_(True)
        wantc = O.reverb_apply(xd, ird); (wantc * gc.double()).sum().backward()
        xc, irc = x.cuda().requires_grad_(True), ir.cuda().requires_grad_(True)
        gotc = ops.causal_conv(xc, irc); (gotc * gc.cuda()).sum().backward()
        check("conv", (Bc, N, L), [(gotc, wantc.detach()), (xc.grad, xd.grad), (irc.grad, ird.grad)], 1e-4)
        # mss
        Bm, Nm = rnd.randint(1, 3), rnd.choice([64, 100, 257, 1000, 4096, 5000])
        ffts = rnd.sample([8, 16, 32, 64, 128, 256, 512, 1024, 2048], rnd.randint(1, 3))
        ffts = [n for n in ffts if n // 2 < Nm] or [8]
        xp, xt = torch.randn(Bm, Nm) * 0.5, torch.randn(Bm, Nm) * 0.5
        xpd = xp.double().requires_grad_(True)
        wantm = O.mss_loss(xpd, xt.double(), ffts); wantm.backward()
        xpc = xp.cuda().requires_grad_(True)
        gotm = ops.mss_loss(xpc, xt.cuda(), ffts); gotm.backward()
        before = bad
        check("mss", (Bm, Nm, ffts), [(gotm.detach().view(1), wantm.detach().view(1)), (xpc.grad, xpd.grad)], 1e-4)
        if bad > before:                               # keep the case: a sign(0) bin of the L1 terms or a real defect?
            import os
            os.makedirs("gpurun_out", exist_ok=True)
            torch.save({"xp": xp, "xt": xt, "ffts": ffts, "got_grad": xpc.grad.cpu(), "got_loss": gotm.detach().cpu()},
                       f"gpurun_out/fuzz_mss_fail_{it}.pt")
    except Exception as e:
        bad += 1
        print("EXC", it, repr(e)[:300])
        traceback.print_exc(limit=2)
torch.cuda.synchronize()
print("cases", n_cases, "failures", bad)
