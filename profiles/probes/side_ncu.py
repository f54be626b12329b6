"""One reverb forward + backward (B = 256, N = 192 000) and one spectral-loss forward + backward (cfg 5) for ncu."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
import ddsp_pytorch_b200 as d
from ddsp_pytorch_b200 import ops
dev = torch.device("cuda")
what = sys.argv[1] if len(sys.argv) > 1 else "both"
if what in ("reverb", "both"):
    B, N, L = 256, 192000, 48000
    x = torch.randn(B, N, device=dev, requires_grad=True)
    ir = (torch.randn(L, device=dev) * 0.01).requires_grad_(True)
    for _ in range(2):
        x.grad = None; ir.grad = None
        y = ops.causal_conv(x, ir)
        y.backward(torch.ones_like(y))
    torch.cuda.synchronize()
if what in ("mss", "both"):
    B, N = 64, 64000
    xp = torch.randn(B, N, device=dev, requires_grad=True)
    xt = torch.randn(B, N, device=dev)
    for _ in range(2):
        xp.grad = None
        l = ops.mss_loss(xp, xt, [2048, 1024, 512, 256, 128, 64])
        l.backward()
    torch.cuda.synchronize()
