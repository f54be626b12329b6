"""The C-ABI library builds, loads on a GPU-less host and exports exactly what include/ddsp_b200.h declares;
the ctypes table mirrors the header's argument counts.  No compute calls are made here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ddsp_b200.h")


def declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(ddsp_\w+)\s*\(([^)]*)\)\s*;", src):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out


@pytest.fixture(scope="module")
def lib_path():
    from ddsp_pytorch_b200.build import build_library
    return build_library()


def test_header_declares_something():
    d = declared()
    assert "ddsp_osc_forward" in d and "ddsp_noise_backward" in d and len(d) >= 10


def test_library_exports_every_declared_symbol(lib_path):
    lib = ctypes.CDLL(lib_path)
    for name in declared():
        assert hasattr(lib, name), f"{name} declared in include/ddsp_b200.h but not exported"
    lib.ddsp_abi_version.restype = ctypes.c_int
    assert lib.ddsp_abi_version() == 1


def test_ctypes_table_matches_header(lib_path):
    from ddsp_pytorch_b200 import _lib
    d = declared()
    assert set(_lib.SIGNATURES) == set(d)
    for name, (_, argtypes) in _lib.SIGNATURES.items():
        assert len(argtypes) == d[name], name
    _lib.load()


def test_size_queries_and_argument_validation_need_no_gpu(lib_path):
    from ddsp_pytorch_b200 import _lib
    lib = _lib.load()
    assert lib.ddsp_osc_state_bytes(2, 3, 4) >= 8 * 2 * 4 + 4 * 2 * 3 * 4 + 2 * 4 * 2 * 3
    assert lib.ddsp_osc_backward_scratch_bytes(2, 3, 4) >= 8 * 2 * 4 * 4 + 8 * 2 * 4
    assert lib.ddsp_osc_state_bytes(0, 3, 4) == 0
    # validation happens before any CUDA call, so these fail cleanly on a GPU-less host
    with pytest.raises(RuntimeError, match="hop"):
        _lib.call("ddsp_osc_forward", 1, 1, 1, 1, None, 1, 0, 1, 3, 4, 0, 16000, None)
    with pytest.raises(RuntimeError, match="null"):
        _lib.call("ddsp_osc_forward", None, None, None, None, None, None, 0, 1, 3, 4, 8, 16000, None)
    with pytest.raises(RuntimeError, match="FIR length"):
        _lib.call("ddsp_noise_forward", 1, None, 1, 0, 0, 0, None, 0, 1, 2, 65, 64, None)


def test_ops_refuse_cpu_tensors():
    import torch
    from ddsp_pytorch_b200 import ops
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.oscillator_bank(torch.ones(1, 2, 1), torch.ones(1, 2, 3), torch.ones(1, 2, 1), 16000, 8)
