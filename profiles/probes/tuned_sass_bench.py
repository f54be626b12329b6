"""Times the oscillator kernels with the hand-scheduled loops on and off (cfg 4 shape, 1024 voices by default).
usage: python profiles/probes/tuned_sass_bench.py [voices]"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from ddsp_pytorch_b200 import _lib, ops
from ddsp_pytorch_b200.synthetic import synthetic_f0

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
T, K, sr, hop = 375, 256, 48000, 512
lib = _lib.load()
f0 = synthetic_f0(B, T, sr, hop, 0)[..., 0].cuda()
c = torch.rand(B, T, K, device="cuda").requires_grad_(True)
a = torch.rand(B, T, device="cuda").requires_grad_(True)
g = torch.randn(B, T * hop, device="cuda")
for mode in ((1,) if len(sys.argv) > 2 else (1, 0, 1, 0)):
    print("tuned mask:", lib.ddsp_osc_tuned_sass(mode))
    ts = {"fwd": [], "bwd": []}
    for it in range(8):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        c.grad = None; a.grad = None
        e[0].record()
        y = ops.oscillator_bank(f0, c, a, sr, hop)
        e[1].record()
        y.backward(g)
        e[2].record()
        torch.cuda.synchronize()
        if it >= 3:
            ts["fwd"].append(e[0].elapsed_time(e[1])); ts["bwd"].append(e[1].elapsed_time(e[2]))
    print("mode", mode, {k: round(sorted(v)[len(v) // 2], 4) for k, v in ts.items()}, "ms (prepare+synth | grad+finalize)")
