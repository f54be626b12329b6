"""One filtered-noise forward + backward at the cfg 4 shape (for ncu)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from ddsp_pytorch_b200 import ops
B, T, NB, hop = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, 375, 65, 512
H = torch.rand(B, T, NB, device="cuda").requires_grad_(True)
g = torch.randn(B, T * hop, device="cuda")
for _ in range(2):
    H.grad = None
    y = ops.filtered_noise(H, hop, seed=7, offset=0)
    y.backward(g)
torch.cuda.synchronize()
