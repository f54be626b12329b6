"""ddsp_osc_forward_noalloc (include/ddsp_b200.h: no counterpart in the reference):
the oscillator forward with its frame-rate state in a static device array inside the library.  Called through the C ABI,
twice in a row with different shapes (the second call finds the first call's state in the array), it must reproduce
ddsp_osc_forward with a caller-owned state buffer bit for bit -- the kernels are the same -- stay within the suite's
waveform tolerance of the fp64 oracle, allocate nothing, and refuse shapes beyond its built-in scratch.

Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

WAVE_TOL = 1e-4          # tests/test_gpu_osc.py


def _inputs(B, T, K, sr, hop, seed):
    from ddsp_pytorch_b200.synthetic import synthetic_controls
    f0, c, a, _ = synthetic_controls(B, T, K, 5, sr, hop, seed=seed)
    return f0, c, a


def test_noalloc_forward_matches_the_allocating_entry_point_twice_in_a_row():
    from ddsp_pytorch_b200 import _lib
    from oracle import ddsp_oracle as O
    lib = _lib.load()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    for call, (B, T, K, sr, hop) in enumerate([(16, 125, 100, 16000, 512), (2, 9, 180, 44100, 512), (1, 4, 180, 44100, 512)]):
        f0h, ch, ah = _inputs(B, T, K, sr, hop, seed=70 + call)
        f0, c, a = (t.to(dev).contiguous() for t in (f0h, ch, ah))
        N = T * hop
        # everything the calls need exists before them: the entry point under test must not allocate
        y_ref = torch.empty(B, N, device=dev)
        y_new = torch.full((B, N), float("nan"), device=dev)
        state = torch.empty(lib.ddsp_osc_state_bytes(B, T, K), dtype=torch.uint8, device=dev)
        _lib.call("ddsp_osc_forward", f0.data_ptr(), c.data_ptr(), a.data_ptr(), y_ref.data_ptr(), None, state.data_ptr(), 0, B, T, K,
                  hop, sr, st)
        torch.cuda.synchronize()
        before, reserved = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        _lib.call("ddsp_osc_forward_noalloc", f0.data_ptr(), c.data_ptr(), a.data_ptr(), y_new.data_ptr(), 0, B, T, K, hop, sr, st)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before and torch.cuda.memory_reserved(dev) == reserved
        assert torch.cuda.mem_get_info(dev)[0] >= free_before - (1 << 20)          # no device allocation behind torch's back
        assert torch.equal(y_new, y_ref), f"call {call}: static-scratch forward differs from ddsp_osc_forward"
        want = O.oscillator_bank(f0h.double(), ch.double(), ah.double(), K, sr, hop)
        assert float((y_new.double().cpu() - want).abs().max()) <= WAVE_TOL
        # accumulate != 0 adds into y, as in the allocating entry point
        _lib.call("ddsp_osc_forward_noalloc", f0.data_ptr(), c.data_ptr(), a.data_ptr(), y_new.data_ptr(), 1, B, T, K, hop, sr, st)
        assert torch.equal(y_new, y_ref + y_ref)


def test_noalloc_forward_rejects_shapes_beyond_its_scratch():
    from ddsp_pytorch_b200 import _lib
    lib = _lib.load()
    B, T, K, sr, hop = 64, 375, 256, 48000, 512            # 25 MB of frame-rate state > the 8 MiB built in
    assert lib.ddsp_osc_state_bytes(B, T, K) > (8 << 20)
    dev = torch.device("cuda:0")
    f0, c, a = torch.full((B, T), 200.0, device=dev), torch.ones(B, T, K, device=dev), torch.ones(B, T, device=dev)
    y = torch.zeros(B, T * hop, device=dev)
    with pytest.raises(RuntimeError, match="DDSP_OSC_NOALLOC_STATE_BYTES"):
        _lib.call("ddsp_osc_forward_noalloc", f0.data_ptr(), c.data_ptr(), a.data_ptr(), y.data_ptr(), 0, B, T, K, hop, sr,
                  torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert float(y.abs().max()) == 0.0                     # nothing was launched
