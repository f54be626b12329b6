"""Controller tail: CUDA-core FFMA product against the tcgen05 path on the same inputs -- event times of
forward and forward + backward, and the largest difference between the two paths' results."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import json
import torch
from ddsp_pytorch_b200 import ops

dev = torch.device("cuda")
out = []
SHAPES = ((5504, 512, 180, 195), (688, 512, 180, 195), (4000, 512, 100, 65), (256, 512, 180, 195))
if len(sys.argv) > 1:
    SHAPES = SHAPES[:int(sys.argv[1])]
for R, U, K, NB in SHAPES:
    torch.manual_seed(R)
    x = torch.randn(R, U, device=dev, requires_grad=True)
    ps = [torch.randn(K, U, device=dev) * 0.05, torch.randn(K, device=dev), torch.randn(1, U, device=dev) * 0.05,
          torch.randn(1, device=dev), torch.randn(NB, U, device=dev) * 0.05, torch.randn(NB, device=dev)]
    ps = [p.requires_grad_(True) for p in ps]
    gs = [torch.randn(R, K, device=dev), torch.randn(R, 1, device=dev), torch.randn(R, NB, device=dev)]
    res, row = {}, {"R": R, "U": U, "K": K, "NB": NB}
    for name, minrows in (("ffma", 0), ("tcgen05", 1)):
        ops.CTRL_TC_MIN_ROWS = minrows

        def fwd():
            with torch.no_grad():
                return ops.controller_heads(x, *ps)

        def fwdbwd():
            for t in [x] + ps:
                t.grad = None
            c, a, H = ops.controller_heads(x, *ps)
            torch.autograd.backward([c, a, H], gs)
            return (c, a, H, x.grad, *[p.grad for p in ps])

        if len(sys.argv) > 2 and sys.argv[2] == "once":          # for an ncu launch list: one forward + backward per path
            fwdbwd()
            torch.cuda.synchronize()
            continue
        for fn, key in ((fwd, "fwd_ms"), (fwdbwd, "fwdbwd_ms")):
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    r = fn()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()            # graph replay: device time, not the host's launch time
            with torch.cuda.graph(graph):
                r = fn()
            for _ in range(3):
                graph.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 50
            e0.record()
            for _ in range(n):
                graph.replay()
            e1.record()
            torch.cuda.synchronize()
            row[f"{name}_{key}"] = round(e0.elapsed_time(e1) / n, 4)
        res[name] = [t.detach().clone() for t in r]
    if len(sys.argv) > 2 and sys.argv[2] == "once":
        continue
    row["max_rel_diff"] = max(float((a - b).abs().max() / b.abs().max().clamp_min(1e-30)) for a, b in zip(res["tcgen05"], res["ffma"]))
    # torch (cuBLAS fp32 + elementwise) on the same shapes
    def torch_fb():
        for t in [x] + ps:
            t.grad = None
        z = [torch.nn.functional.linear(x, ps[0], ps[1]), torch.nn.functional.linear(x, ps[2], ps[3]), torch.nn.functional.linear(x, ps[4], ps[5])]
        o = [2 * torch.sigmoid(t) ** 2.3026 + 1e-7 for t in z]
        torch.autograd.backward(o, gs)
    for _ in range(5):
        torch_fb()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        torch_fb()
    e1.record()
    torch.cuda.synchronize()
    row["torch_fwdbwd_ms"] = round(e0.elapsed_time(e1) / 50, 4)
    print(json.dumps(row), flush=True)
