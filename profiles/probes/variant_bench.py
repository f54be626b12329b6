"""Builds libddsp_b200 variants with -D experiment switches and times the six kernels of the headline step for each
(one process per variant).  python profiles/probes/variant_bench.py build|run"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
VARIANTS = {        # the NOISE_* switches exist in profiles/probes/noise_experiments_r02.cu.txt only (not in the shipped kernels)
    "base": [],
    "osc_sweep_packed": ["-DOSC_SWEEP_PACKED"],
}
OUT = os.path.join(ROOT, "profiles", "probes", "variants")

def build():
    from ddsp_pytorch_b200.build import NVCC_FLAGS, _sources, nvcc_path
    os.makedirs(OUT, exist_ok=True)
    procs = []
    for name, flags in VARIANTS.items():
        cmd = [nvcc_path(), *NVCC_FLAGS, *flags, "-o", os.path.join(OUT, f"lib_{name}.so"), *_sources()]
        procs.append((name, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for name, p in procs:
        out, _ = p.communicate()
        print(name, "rc", p.returncode, out[-300:] if p.returncode else "")

def child():
    import torch
    from ddsp_pytorch_b200 import _lib
    from ddsp_pytorch_b200.synthetic import synthetic_controls
    B, T, K, NB, sr, hop = 256, 375, 256, 65, 48000, 512
    N = T * hop
    dev = torch.device("cuda:0")
    f0, c, a, H = (t.to(dev) for t in synthetic_controls(B, T, K, NB, sr, hop, seed=1238))
    g = torch.randn(B, N, device=dev)
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    state = torch.empty(lib.ddsp_osc_state_bytes(B, T, K), dtype=torch.uint8, device=dev)
    scratch = torch.empty(lib.ddsp_osc_backward_scratch_bytes(B, T, K), dtype=torch.uint8, device=dev)
    y = torch.empty(B, N, device=dev); z = torch.empty(B, N, device=dev)
    gc = torch.empty(B, T, K, device=dev); ga = torch.empty(B, T, device=dev); gH = torch.empty(B, T, NB, device=dev)
    P = lambda t: t.data_ptr()
    calls = {
        "osc_synth": lambda: _lib.call("ddsp_osc_synthesize", P(f0), P(a), P(state), P(y), P(z), 0, B, T, K, hop, sr, st),
        "osc_grad": lambda: _lib.call("ddsp_osc_grad_segments", P(g), P(f0), P(a), P(z), P(state), P(scratch), B, T, K, hop, sr, st),
        "noise_fwd": lambda: _lib.call("ddsp_noise_forward", P(H), None, P(y), 0, 7, 0, None, 0, B, T, NB, hop, st),
        "noise_bwd": lambda: _lib.call("ddsp_noise_backward", P(g), None, P(gH), 7, 0, None, 0, B, T, NB, hop, st),
    }
    _lib.call("ddsp_osc_prepare", P(f0), P(c), P(state), B, T, K, hop, sr, st)
    res = {}
    for name, fn in calls.items():
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record(); e1.synchronize()
        res[name] = round(e0.elapsed_time(e1) / 20, 4)
    print(json.dumps({"variant": os.environ["DDSP_VARIANT"], **res}), flush=True)

def run():
    for name in VARIANTS:
        env = dict(os.environ, DDSP_B200_LIB=os.path.join(OUT, f"lib_{name}.so"), DDSP_VARIANT=name)
        subprocess.run([sys.executable, os.path.abspath(__file__), "child"], env=env)

if __name__ == "__main__":
    {"build": build, "run": run, "child": child}[sys.argv[1]]()
