"""Turn an ncu launch list (--metrics gpu__time_duration.sum --csv) and an ncu-rep raw page into the
markdown summaries committed under profiles/.  Usage:
    python profiles/summarize.py launches.csv [prof.ncu-rep] > profiles/summary_rNN.md
"""
import collections
import csv
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
           "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.sum",
           "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__occupancy_limit_registers",
           "launch__occupancy_limit_shared_mem", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    H = rows[h]
    ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[h + 1:]:
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
        a = agg.setdefault(r[ki].split("(")[0][:70], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("## Launch list (cold-cache, serialised; compare shares)\n")
    print("| kernel | launches | total ms | avg ms | share |\n|---|---|---|---|---|")
    for n, a in sorted(agg.items(), key=lambda x: -x[1][1])[:16]:
        print(f"| `{n}` | {a[0]} | {a[1]:.3f} | {a[1] / a[0]:.4f} | {100 * a[1] / tot:.1f}% |")


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    H = rows[0]
    print("\n## `ncu --set full` per kernel (one launch each)\n")
    for r in rows[2:]:
        print(f"### `{r[H.index('Kernel Name')].split('(')[0]}`\n\n| metric | value | unit |\n|---|---|---|")
        for m in METRICS:
            if m in H:
                print(f"| {m} | {r[H.index(m)]} | {rows[1][H.index(m)]} |")
        print()


if __name__ == "__main__":
    launches(sys.argv[1])
    if len(sys.argv) > 2:
        raw(sys.argv[2])
