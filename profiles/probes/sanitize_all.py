"""tests/kernel_workloads.py (every kernel family once, forward and backward, small shapes) as a plain script: the workload for
`compute-sanitizer --tool memcheck | racecheck | synccheck | initcheck` on a box that allows it (profiles/probes/sanitize.sh).

    compute-sanitizer --tool racecheck --error-exitcode 7 python profiles/probes/sanitize_all.py [osc noise reverb mss features fft dftgemm]

On this pool compute-sanitizer is closed (r02: every tool returned 86 before starting python); tests/test_gpu_guards.py checks the
same workloads with guard bands and poisoned buffers instead."""
import os
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

from ddsp_pytorch_b200 import _lib
from kernel_workloads import WORKLOADS

lib = _lib.load()
print("[sanitize] hand-scheduled kernels available:", lib.ddsp_osc_tuned_sass(-1), flush=True)
for name in (sys.argv[1:] or list(WORKLOADS)):
    print("[sanitize]", name, flush=True)
    outs = WORKLOADS[name](torch.Generator().manual_seed(0))
    torch.cuda.synchronize()
    assert all(bool(torch.isfinite(o).all()) for o in outs), name
print("[sanitize] done, launches", lib.ddsp_launch_count(), flush=True)
