"""ctypes binding of libddsp_b200.so -- the only way this package reaches the GPU.

There is deliberately no fallback: if the library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

from .build import LIB_PATH

_p, _i, _u32, _u64, _sz = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_size_t

# name -> (restype, argtypes); mirrors include/ddsp_b200.h one to one (tests/test_abi.py checks it)
SIGNATURES = {
    "ddsp_abi_version": (_i, []),
    "ddsp_last_error": (C.c_char_p, []),
    "ddsp_launch_count": (C.c_ulonglong, []),
    "ddsp_osc_state_bytes": (_sz, [_i, _i, _i]),
    "ddsp_osc_forward": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "ddsp_osc_prepare": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ddsp_osc_synthesize": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "ddsp_osc_grad_segments": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ddsp_osc_grad_finalize": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ddsp_osc_backward_scratch_bytes": (_sz, [_i, _i, _i]),
    "ddsp_osc_backward": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ddsp_osc_forward_noalloc": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "ddsp_osc_tuned_sass": (_i, [_i]),
    "ddsp_osc_backward_f0_scratch_bytes": (_sz, [_i, _i, _i]),
    "ddsp_osc_backward_f0": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ddsp_osc_live": (_i, [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ddsp_rescale_forward": (_i, [_p, _p, _i, _i, _i, _i, _p]),
    "ddsp_rescale_backward": (_i, [_p, _p, _i, _i, _i, _i, _p]),
    "ddsp_noise_forward": (_i, [_p, _p, _p, _i, _u64, _u32, _p, _u32, _i, _i, _i, _i, _p]),
    "ddsp_noise_backward": (_i, [_p, _p, _p, _u64, _u32, _p, _u32, _i, _i, _i, _i, _p]),
    "ddsp_noise_uniform": (_i, [_p, _u64, _u32, _u32, _i, _i, _i, _p]),
    "ddsp_amp_to_impulse_response": (_i, [_p, _p, _i, _i, _i, _p]),
    "ddsp_reverb_impulse": (_i, [_p, _p, _p, _p, _p, _i, _p]),
    "ddsp_reverb_impulse_backward": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _i, _p]),
    "ddsp_conv_scratch_bytes": (_sz, [_i, _i, _i]),
    "ddsp_conv_forward": (_i, [_p, _p, _p, _p, _i, _i, _i, _p]),
    "ddsp_conv_backward": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p]),
    "ddsp_reverb_live": (_i, [_p, _p, _p, _p, _p, _i, _i, _p]),
    "ddsp_mss_count": (C.c_int64, [_i, _i, _i, _i]),
    "ddsp_mss_scratch_bytes": (_sz, [_p, _p, _i, _i, _i]),
    "ddsp_mss_forward": (_i, [_p, _p, _p, _p, _p, _i, _p, _p, _p, _i, C.c_float, C.c_float, _i, _i, _p]),
    "ddsp_mss_backward": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _p]),
    "ddsp_dftgemm_workspace_bytes": (_sz, [_i, _i, _i]),
    "ddsp_dftgemm_prepare": (_i, [_p, _p, _p, _p, _i, _i, _i, _p]),
    "ddsp_dftgemm_sss_forward": (_i, [_p, _p, _p, _i, _i, C.c_float, _i, _i, _p]),
    "ddsp_rowconv_forward": (_i, [_p, _p, _p, C.c_longlong, _i, _p]),
    "ddsp_rowconv_backward": (_i, [_p, _p, _p, _p, _p, C.c_longlong, _i, _p]),
    "ddsp_ctrl_heads_ldz": (_i, [_i, _i]),
    "ddsp_ctrl_heads_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "ddsp_ctrl_heads_forward": (_i, [_p] * 11 + [_i, _i, _i, _i, _p]),
    "ddsp_ctrl_heads_backward": (_i, [_p] * 16 + [_i, _i, _i, _i, _p]),
    "ddsp_ctrl_heads_tc_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "ddsp_ctrl_heads_forward_tc": (_i, [_p] * 12 + [_i, _i, _i, _i, _p]),
    "ddsp_ctrl_heads_backward_tc": (_i, [_p] * 16 + [_i, _i, _i, _i, _p]),
    "ddsp_loudness_frames": (_i, [_i, _i, _i]),
    "ddsp_loudness_forward": (_i, [_p, _p, _p, _i, _i, _i, _i, _p]),
    "ddsp_fft_c2c": (_i, [_p, _p, _i, C.c_longlong, _i, _i, _p]),
    "ddsp_probe_fp32_fma": (_i, [_p, _i, _i, _i, _p]),
}

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        path = os.environ.get("DDSP_B200_LIB", LIB_PATH)     # experiment builds (profiles/probes/variant_bench.py)
        if not os.path.exists(path):
            raise RuntimeError(
                f"ddsp_pytorch_b200: {path} is missing. Build it with `python -m ddsp_pytorch_b200.build` "
                "(needs nvcc). There is no CPU or eager fallback for this path.")
        lib = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)          # AttributeError if the library is stale
            fn.restype, fn.argtypes = res, args
        if lib.ddsp_abi_version() != 1:
            raise RuntimeError("ddsp_pytorch_b200: libddsp_b200.so ABI version mismatch; rebuild it")
        _lib = lib
    return _lib


def call(name: str, *args) -> None:
    """Invoke an int-returning entry point and raise on a non-zero status."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.ddsp_last_error().decode()}")
