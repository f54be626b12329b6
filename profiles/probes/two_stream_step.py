"""Does running the two independent branches of the cfg-4 step (oscillator / noise) on two streams inside the step's CUDA graph
hide the kernels' tails and the launch gaps?  Three graphs over the same buffers: the autograd step bench.py replays, the same
C-ABI calls on one stream, and on two streams (prepare beside the noise forward, the noise gradient beside the oscillator's)."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from ddsp_pytorch_b200 import _lib, ops
from ddsp_pytorch_b200.synthetic import synthetic_controls

dev = torch.device("cuda")
lib = _lib.load()
T, K, NB, sr, hop = 375, 256, 65, 48000, 512
N = T * hop
P = lambda t: t.data_ptr()  # noqa: E731


def timeit(fn, reps=30):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for B in [int(v) for v in (sys.argv[1:] or ["128", "1024"])]:
    f0, c, a, H = (t.to(dev) for t in synthetic_controls(B, T, K, NB, sr, hop, seed=1234 + 4))
    g = torch.randn(B, N, generator=torch.Generator().manual_seed(99)).to(dev)
    c.requires_grad_(True); a.requires_grad_(True); H.requires_grad_(True)
    seed = 7

    def auto_step():
        y = ops.filtered_noise(H, hop, seed=seed, offset=0, voice_offset=0)
        y = ops.oscillator_bank(f0, c, a, sr, hop, out=y)
        y.backward(g)
        return y

    state = torch.empty(lib.ddsp_osc_state_bytes(B, T, K), dtype=torch.uint8, device=dev)
    scratch = torch.empty(lib.ddsp_osc_backward_scratch_bytes(B, T, K), dtype=torch.uint8, device=dev)
    y2, z2 = torch.empty(B, N, device=dev), torch.empty(B, N, device=dev)
    gc, ga, gH = torch.empty(B, T, K, device=dev), torch.empty(B, T, device=dev), torch.empty(B, T, NB, device=dev)
    cd_, ad_, Hd_ = c.detach(), a.detach(), H.detach()
    s2 = torch.cuda.Stream(dev)

    def raw_step(two):
        cur = torch.cuda.current_stream()
        st = cur.cuda_stream
        if two:
            s2.wait_stream(cur)
            with torch.cuda.stream(s2):
                st2 = s2.cuda_stream
                _lib.call("ddsp_osc_prepare", P(f0), P(cd_), P(state), B, T, K, hop, sr, st2)
                _lib.call("ddsp_noise_backward", P(g), None, P(gH), seed, 0, None, 0, B, T, NB, hop, st2)
            _lib.call("ddsp_noise_forward", P(Hd_), None, P(y2), 0, seed, 0, None, 0, B, T, NB, hop, st)
            # prepare is the first thing on s2: wait for it alone through an event
            ev = torch.cuda.Event()
        else:
            _lib.call("ddsp_osc_prepare", P(f0), P(cd_), P(state), B, T, K, hop, sr, st)
            _lib.call("ddsp_noise_forward", P(Hd_), None, P(y2), 0, seed, 0, None, 0, B, T, NB, hop, st)
        if two:
            cur.wait_stream(s2)      # (joins after the noise gradient too; the variant below splits the wait)
        _lib.call("ddsp_osc_synthesize", P(f0), P(ad_), P(state), P(y2), P(z2), 1, B, T, K, hop, sr, st)
        _lib.call("ddsp_osc_grad_segments", P(g), P(f0), P(ad_), P(z2), P(state), P(scratch), B, T, K, hop, sr, st)
        _lib.call("ddsp_osc_grad_finalize", P(f0), P(state), P(scratch), P(gc), P(ga), B, T, K, hop, sr, st)
        if not two:
            _lib.call("ddsp_noise_backward", P(g), None, P(gH), seed, 0, None, 0, B, T, NB, hop, st)

    def raw_step_split():
        """two streams, the synthesis kernel waits for prepare only; the noise gradient joins at the end"""
        cur = torch.cuda.current_stream()
        st, st2 = cur.cuda_stream, s2.cuda_stream
        s2.wait_stream(cur)
        ev = torch.cuda.Event()
        with torch.cuda.stream(s2):
            _lib.call("ddsp_osc_prepare", P(f0), P(cd_), P(state), B, T, K, hop, sr, st2)
            ev.record(s2)
            _lib.call("ddsp_noise_backward", P(g), None, P(gH), seed, 0, None, 0, B, T, NB, hop, st2)
        _lib.call("ddsp_noise_forward", P(Hd_), None, P(y2), 0, seed, 0, None, 0, B, T, NB, hop, st)
        cur.wait_event(ev)
        _lib.call("ddsp_osc_synthesize", P(f0), P(ad_), P(state), P(y2), P(z2), 1, B, T, K, hop, sr, st)
        _lib.call("ddsp_osc_grad_segments", P(g), P(f0), P(ad_), P(z2), P(state), P(scratch), B, T, K, hop, sr, st)
        _lib.call("ddsp_osc_grad_finalize", P(f0), P(state), P(scratch), P(gc), P(ga), B, T, K, hop, sr, st)
        cur.wait_stream(s2)

    res = {"voices": B}
    side = torch.cuda.Stream(dev)
    variants = (("autograd", auto_step), ("raw_one_stream", lambda: raw_step(False)), ("raw_two_streams_join_early", lambda: raw_step(True)),
                ("raw_two_streams", raw_step_split))
    outs = {}
    for name, fn in variants:
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                c.grad = a.grad = H.grad = None
                r = fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        c.grad = a.grad = H.grad = None
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            r = fn()
        graph.replay()
        torch.cuda.synchronize()
        outs[name] = [t.clone() for t in ((r.detach(), c.grad, a.grad.view(B, T), H.grad) if name == "autograd" else (y2, gc, ga, gH))]
        res[name + "_ms"] = round(timeit(graph.replay), 4)
    for name in outs:
        res[name + "_equal"] = all(torch.equal(x.reshape(-1), w.reshape(-1)) for x, w in zip(outs[name], outs["autograd"]))
    print(json.dumps(res), flush=True)
