"""Turn what the GPU runs brought back under gpurun_out/ into tracked summaries under profiles/."""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
GO = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)

KEYS = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__cluster_size",
        "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"]


def launches():
    path = os.path.join(GO, "launches.csv")
    if not os.path.exists(path):
        return
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[hi]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg, tot = collections.OrderedDict(), 0.0
    for r in rows[hi + 1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v *= {"us": 1e3, "ms": 1e6}.get(r[ui], 1.0)
        name = re.sub(r"\(.*", "", r[ki])
        name = re.sub(r"void |<unnamed>::", "", name)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
    with open(os.path.join(OUT, "step_launches.md"), "w") as f:
        f.write("# One training step, every kernel launch (ncu gpu__time_duration, serialised, cold L2)\n\n"
                "GPT-2-small-shaped (12L, d=768, 12 heads, T=1024), batch 16, 1x B200, eager launch order "
                "(`tools/profile_step.py`). Compare SHARES, not absolutes.\n\n| ms | share | launches | kernel |\n|---|---|---|---|\n")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| {t / 1e6:.3f} | {100 * t / tot:.1f}% | {n} | `{k[:90]}` |\n")
        f.write(f"\nTotal {tot / 1e6:.3f} ms over {sum(n for n, _ in agg.values())} launches.\n")


def ncu_reports():
    for rep in sorted(f for f in os.listdir(GO) if f.endswith(".ncu-rep")):
        try:
            raw = subprocess.run(["ncu", "-i", os.path.join(GO, rep), "--page", "raw", "--csv"], capture_output=True,
                                 text=True, timeout=300).stdout
        except Exception as e:  # pragma: no cover
            print("ncu failed", e)
            continue
        rows = list(csv.reader(raw.splitlines()))
        if len(rows) < 3:
            continue
        h = rows[0]
        with open(os.path.join(OUT, rep.replace(".ncu-rep", "_ncu.md")), "w") as f:
            f.write(f"# ncu --set full --clock-control none: {rep}\n\n")
            for r in rows[2:]:
                d = dict(zip(h, r))
                f.write(f"## `{d.get('Kernel Name', '?')[:110]}`\n\n| metric | value |\n|---|---|\n")
                for k in KEYS:
                    if k in d:
                        f.write(f"| {k} | {d[k]} |\n")
                f.write("\n")


def kernel_bench():
    path = os.path.join(GO, "kernel_bench.json")
    if not os.path.exists(path):
        return
    d = json.load(open(path))
    with open(os.path.join(OUT, "kernel_bench.md"), "w") as f:
        p = d["peaks"]
        f.write("# Per-kernel micro-benchmarks (CUDA events, L2 flushed between iterations, median of 10)\n\n"
                f"Measured peaks (MEASURED_PEAKS.json): HBM copy {p['hbm_gbs']} GB/s, cuBLAS bf16 {p['bf16_tflops']} TFLOP/s burst / "
                f"{p.get('bf16_tflops_sustained')} sustained. Fractions are 'of measured'.\n\n")
        keys = []
        for r in d["rows"]:
            for k in r:
                if k not in keys:
                    keys.append(k)
        f.write("| " + " | ".join(keys) + " |\n|" + "---|" * len(keys) + "\n")
        for r in d["rows"]:
            f.write("| " + " | ".join(str(r.get(k, "")) for k in keys) + " |\n")


if __name__ == "__main__":
    launches()
    kernel_bench()
    ncu_reports()
    print(os.listdir(OUT))
