"""One oscillator forward + backward with the hand-scheduled loops, then one with ptxas's schedule (for ncu)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from ddsp_pytorch_b200 import _lib, ops
from ddsp_pytorch_b200.synthetic import synthetic_f0
B, T, K, sr, hop = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, 375, 256, 48000, 512
lib = _lib.load()
f0 = synthetic_f0(B, T, sr, hop, 0)[..., 0].cuda()
c = torch.rand(B, T, K, device="cuda").requires_grad_(True)
a = torch.rand(B, T, device="cuda").requires_grad_(True)
g = torch.randn(B, T * hop, device="cuda")
for mode in (1, 0):
    lib.ddsp_osc_tuned_sass(mode)
    y = ops.oscillator_bank(f0, c, a, sr, hop)
    y.backward(g)
    torch.cuda.synchronize()
