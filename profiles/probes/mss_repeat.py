"""Run-to-run stability of the spectral-loss gradient: the only legitimate source of variation is the order of the fp32
atomic overlap-adds (~1e-7 relative); anything larger is a race."""
import sys
sys.path.insert(0, ".")
import torch
from ddsp_pytorch_b200 import ops

torch.manual_seed(0)
worst = {}
for n in (8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096):
    for (B, N) in ((1, 4096), (3, 5000), (2, 9000)):
        if n // 2 >= N:
            continue
        xp, xt = (torch.randn(B, N) * 0.5).cuda(), (torch.randn(B, N) * 0.5).cuda()
        ref = None
        dev = 0.0
        ldev = 0.0
        for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 100):
            x = xp.clone().requires_grad_(True)
            loss = ops.mss_loss(x, xt, [n])
            loss.backward()
            g = x.grad
            if ref is None:
                ref, lref = g.clone(), float(loss)
            else:
                dev = max(dev, float((g - ref).abs().max() / ref.abs().max()))
                ldev = max(ldev, abs(float(loss) - lref) / abs(lref))
        worst[(n, B, N)] = (dev, ldev)
        flag = "  <-- RACE?" if dev > 1e-5 or ldev > 1e-6 else ""
        print(f"n_fft={n:5d} B={B} N={N}: max grad deviation {dev:.2e}, loss deviation {ldev:.2e}{flag}", flush=True)
