"""Scheduler experiments: writes one tuned cubin per variant of sass_sched's knobs into profiles/probes/variants/ (run here,
after build.py), and -- with `run` -- times each on the GPU through DDSP_B200_TUNED_CUBIN (one process per image)."""
import os, subprocess, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "profiles", "probes", "variants")
VARIANTS = {"lat6": {"lat": 6}, "lat6_ygroup": {"lat": 6, "yield": "group"}, "lat6_order": {"lat": 6, "key": "order"},
            "lat6_order_ygroup": {"lat": 6, "key": "order", "yield": "group"}, "lat8": {"lat": 8}, "lat8_ygroup": {"lat": 8, "yield": "group"},
            "lat5": {"lat": 5}, "lat7": {"lat": 7}, "lat10": {"lat": 10}}
if len(sys.argv) > 1 and sys.argv[1] == "run":
    import glob
    for name in sorted(os.path.basename(f)[:-6] for f in glob.glob(os.path.join(OUT, "*.cubin"))):
        env = dict(os.environ, DDSP_B200_TUNED_CUBIN=os.path.join(OUT, name + ".cubin"))
        r = subprocess.run([sys.executable, os.path.join(ROOT, "profiles", "probes", "tuned_sass_bench.py"), "1024", "once"], env=env, capture_output=True, text=True)
        print(name, [l for l in r.stdout.splitlines() if l.startswith("mode 1")][:1], r.stderr[-200:] if r.returncode else "")
else:
    from ddsp_pytorch_b200 import sass_sched
    cubin = os.path.join(ROOT, "ddsp_pytorch_b200", "lib", "ddsp_osc.cubin")
    for name, o in VARIANTS.items():
        sass_sched.OPTS.update({"yield": "group", "key": "default"})
        sass_sched.OPTS.update({k: v for k, v in o.items() if k != "lat"})
        sass_sched.FP_LATENCY = o.get("lat", 6)
        lg = []
        sass_sched.tune(cubin, os.path.join(OUT, name + ".cubin"), ["osc_synth_kernelILi4E", "osc_grad_kernelILi8E"], lg)
        print(name, [l.split(": ")[1][:120] for l in lg])
