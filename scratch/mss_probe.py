import sys
sys.path.insert(0, ".")
import torch
from ddsp_pytorch_b200 import ops
from oracle import ddsp_oracle as O
torch.manual_seed(0)
for N in (4096, 1000):
    for n in (4, 8, 16, 32, 64, 128, 256):
        xp, xt = torch.randn(2, N) * 0.5, torch.randn(2, N) * 0.5
        xpd = xp.double().requires_grad_(True)
        want = O.mss_loss(xpd, xt.double(), [n]); want.backward()
        xpc = xp.cuda().requires_grad_(True)
        got = ops.mss_loss(xpc, xt.cuda(), [n], windows=[torch.hann_window(n).cuda()] if len(sys.argv) > 1 else None); got.backward()
        d = (xpc.grad.double().cpu() - xpd.grad).abs()
        i = int(d.view(-1).argmax())
        # the reference's own fp32 run
        xpf = xp.clone().requires_grad_(True)
        wf = O.mss_loss(xpf, xt, [n]); wf.backward()
        df = (xpf.grad.double() - xpd.grad).abs()
        print(f"N={N} n={n}: loss rel {abs(float(got)-float(want))/float(want):.1e} grad max err {float(d.max()):.2e} / max|g| {float(xpd.grad.abs().max()):.2e} at idx {i % N} ; l2 rel {float(d.norm()/xpd.grad.norm()):.2e}; fp32-torch max err {float(df.max()):.2e}")
