"""Does running the filtered-noise kernels beside the oscillator kernels (two streams) beat running them back to back?
The oscillator kernels are FMA-pipe bound (71-76 % of its cycles), the noise kernels issue-bound (50 % of the pipe):
scratch measurement at cfg 4 through the per-kernel C ABI entry points."""
import sys
sys.path.insert(0, ".")
import torch
import ddsp_pytorch_b200 as d  # noqa: F401
from ddsp_pytorch_b200 import _lib
from ddsp_pytorch_b200.synthetic import synthetic_controls

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
T, K, NB, sr, hop = 375, 256, 65, 48000, 512
N = T * hop
dev = torch.device("cuda", 0)
f0, c, a, H = (t.to(dev) for t in synthetic_controls(B, T, K, NB, sr, hop, seed=1238))
g = torch.randn(B, N, device=dev)
lib = _lib.load()
state = torch.empty(lib.ddsp_osc_state_bytes(B, T, K), dtype=torch.uint8, device=dev)
scratch = torch.empty(lib.ddsp_osc_backward_scratch_bytes(B, T, K), dtype=torch.uint8, device=dev)
y = torch.empty(B, N, device=dev); z = torch.empty(B, N, device=dev); yn = torch.empty(B, N, device=dev)
gc = torch.empty(B, T, K, device=dev); ga = torch.empty(B, T, device=dev); gH = torch.empty(B, T, NB, device=dev)
P = lambda t: t.data_ptr()  # noqa: E731


def osc_fwd(st):
    _lib.call("ddsp_osc_prepare", P(f0), P(c), P(state), B, T, K, hop, sr, st)
    _lib.call("ddsp_osc_synthesize", P(f0), P(a), P(state), P(y), P(z), 0, B, T, K, hop, sr, st)


def osc_bwd(st):
    _lib.call("ddsp_osc_grad_segments", P(g), P(f0), P(a), P(z), P(state), P(scratch), B, T, K, hop, sr, st)
    _lib.call("ddsp_osc_grad_finalize", P(f0), P(state), P(scratch), P(gc), P(ga), B, T, K, hop, sr, st)


def noise_fwd(st):
    _lib.call("ddsp_noise_forward", P(H), None, P(yn), 0, 7, 0, None, 0, B, T, NB, hop, st)


def noise_bwd(st):
    _lib.call("ddsp_noise_backward", P(g), None, P(gH), 7, 0, None, 0, B, T, NB, hop, st)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


main = torch.cuda.current_stream()


def serial():
    st = main.cuda_stream
    noise_fwd(st); osc_fwd(st); noise_bwd(st); osc_bwd(st)


def make_overlap(prio, noise_first):
    side = torch.cuda.Stream(dev, priority=prio)

    def run():
        ev0 = torch.cuda.Event(); ev0.record(main)
        side.wait_event(ev0)
        if noise_first:
            noise_fwd(side.cuda_stream); noise_bwd(side.cuda_stream)
            osc_fwd(main.cuda_stream); osc_bwd(main.cuda_stream)
        else:
            osc_fwd(main.cuda_stream)
            noise_fwd(side.cuda_stream); noise_bwd(side.cuda_stream)
            osc_bwd(main.cuda_stream)
        ev1 = torch.cuda.Event(); ev1.record(side)
        main.wait_event(ev1)
    return run


print("serial               ms", round(timed(serial), 3))
for prio in (0, -1):
    for nf in (True, False):
        print(f"overlap prio={prio} noise_first={nf} ms", round(timed(make_overlap(prio, nf)), 3))
