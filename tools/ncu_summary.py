#!/usr/bin/env python
"""Key metrics per captured kernel of an .ncu-rep (ncu --set full), or the
per-kernel-name aggregate of a launch-list CSV.
usage: ncu_summary.py report.ncu-rep | ncu_summary.py launches.csv"""
import collections
import csv
import io
import subprocess
import sys

WANT = [
    ("time_ms", "gpu__time_duration.sum"),
    ("dram_rd_GB", "dram__bytes_read.sum"),
    ("dram_wr_GB", "dram__bytes_write.sum"),
    ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("sm%", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("issue%", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("regs", "launch__registers_per_thread"),
    ("fp64%", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
    ("fma%", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    ("alu%", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    ("xu%", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
    ("lsu%", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
    ("warp_inst", "smsp__inst_executed.sum"),
    ("l2_GB", "lts__t_bytes.sum"),
    ("grid", "launch__grid_size"),
]

UNIT = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3,
        "byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0}


def report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    H, U = rows[0], rows[1]
    ki = H.index("Kernel Name")
    for r in rows[2:]:
        name = r[ki].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
        parts = [f"{name[:34]:34s}"]
        for label, m in WANT:
            if m not in H:
                continue
            i = H.index(m)
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                continue
            v *= UNIT.get(U[i], 1.0)
            parts.append(f"{label}={v:.4g}")
        print(" ".join(parts))


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hi = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    H = rows[hi]
    ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[hi + 1:]:
        n = r[ki].split("(")[0].replace("void ", "").replace("<unnamed>::", "")[:60]
        v = float(r[vi].replace(",", "")) * UNIT.get(r[ui], 1.0)
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"{'kernel':62s} {'n':>4s} {'total ms':>10s} {'avg ms':>9s} share")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n:62s} {c:4d} {t:10.3f} {t / c:9.4f} {100 * t / tot:5.1f}%")


if __name__ == "__main__":
    p = sys.argv[1]
    (launches if p.endswith(".csv") else report)(p)
