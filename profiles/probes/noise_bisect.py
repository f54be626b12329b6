"""Scheduler debugging: which part of a reordered noise loop breaks the results?  Needs the experiment build of the library
(-DDDSP_NOISE_TUNED_EXPERIMENT) at profiles/probes/variants/libddsp_exp.so and images profiles/probes/variants/nb_*.cubin."""
import glob, os, subprocess, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, ROOT)
    import torch
    from ddsp_pytorch_b200 import _lib, ops
    lib = _lib.load()
    B, T, NB, hop = 64, 375, 65, 512
    H = torch.rand(B, T, NB, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    res = []
    for mode in (1, 0):
        lib.ddsp_osc_tuned_sass(mode)
        y = ops.filtered_noise(H, hop, seed=7, offset=1)
        torch.cuda.synchronize()
        res.append(y.clone())
    lib.ddsp_osc_tuned_sass(1)
    d = (res[0] - res[1]).abs()
    print("equal", torch.equal(res[0], res[1]), "n_diff", int((d > 0).sum()), "max", float(d.max()))
else:
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "probes", "variants", "nb_*.cubin"))):
        env = dict(os.environ, DDSP_B200_TUNED_NOISE_CUBIN=f, DDSP_B200_LIB=os.path.join(ROOT, "profiles", "probes", "variants", "libddsp_exp.so"))
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "child"], env=env, capture_output=True, text=True)
        print(os.path.basename(f), r.stdout.strip()[-90:] or r.stderr.strip()[-300:])
