#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source cuda,sass` per CUDA source
line: share of executed instructions and of stall samples.
usage: ncu_lines.py report.ncu-rep kernel_regex [top_n]"""
import csv, subprocess, sys, io, collections
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source",
                      "cuda,sass", "--kernel-name", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur_file = None
agg = collections.OrderedDict()
H = None
first_fn = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if len(r) == 2 and r[0] == "Function Name":
        if first_fn is None: first_fn = r[1]
        cur_fn = r[1]; continue
    if len(r) > 5 and r[0] == "Line No":
        H = r; continue
    if H is None or len(r) < len(H) or cur_fn != first_fn: continue
    if r[2] != "-": continue          # SASS rows
    try:
        n = int(r[H.index("Instructions Executed")]); s = int(r[H.index("# Samples")])
    except ValueError:
        continue
    stalls = {h: int(v) for h, v in zip(H, r) if h.startswith("stall_") and "Not Issued" not in h and v.isdigit() and int(v)}
    k = (cur_file, int(r[0]))
    a = agg.setdefault(k, [0, 0, r[1], collections.Counter()])
    a[0] += n; a[1] += s; a[3].update(stalls)
ti = sum(a[0] for a in agg.values()) or 1
ts = sum(a[1] for a in agg.values()) or 1
print(f"kernel: {first_fn}\ntotal warp-instructions {ti}, samples {ts}")
bys = len(sys.argv) > 4 and sys.argv[4] == "samples"
for (f, l), (n, s, src, st) in sorted(agg.items(), key=lambda kv: -kv[1][1 if bys else 0])[:top]:
    top_st = ",".join(f"{k[6:]}:{v}" for k, v in st.most_common(2))
    print(f"{100*n/ti:5.1f}%i {100*s/ts:5.1f}%s {f}:{l:<4d} {src.strip()[:70]:70s} {top_st}")
