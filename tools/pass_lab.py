#!/usr/bin/env python
"""Tuning bench of the resident full pass (ocsb_hfun_pass on device buffers):
the bench workload (20000^2 DEM, contours at -5/0/5, delimiter) built once,
then one subprocess per library variant (``name=path[,ENV=VAL...]``; the
product library when none is given).  Prints ms per step, per-class kernel
times and a CRC of the result (every variant must produce the same raster).

    python tools/pass_lab.py [--size N] [--steps K] [name=/path/lib.so,ENV=1 ...]
"""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r'''
import sys, json, zlib, os
sys.path.insert(0, %r)
import numpy as np, torch
from ocsmesh_b200 import _lib, synthetic as S
n = int(sys.argv[1]); steps = int(sys.argv[2])
ctx = _lib.default_context(0)
z = torch.from_numpy(np.load("/dev/shm/pass_lab_dem.npy", mmap_mode="r")[:]).cuda()
pts = torch.from_numpy(np.load("/dev/shm/pass_lab_pts.npy")).cuda()
tr = S.dem_transform(n)
out = torch.empty((n, n), dtype=torch.float32, device="cuda")
win = _lib.make_win((tr[2], tr[5] + tr[4] * n, tr[2] + tr[0] * n, tr[5]), n, n, 0, 0)
ops = [("feature", pts, len(pts), 0.001, 100.0, True), ("limiter", 50.0, 1500.0, None, None, True)]
sel = os.environ.get("LAB_OPS", "both")
if sel == "feat": ops = ops[:1]
if sel == "lim": ops = ops[1:]
band = os.environ.get("LAB_BAND")
if band:
    # a rank's share of the raster: rows [rb, re) of the window, asynchronous call
    rb, re = (int(v) for v in band.split(":"))
    win = _lib.make_win((tr[2], tr[5] + tr[4] * n, tr[2] + tr[0] * n, tr[5]), n, n, rb, re)
    wins = (_lib.ocsb_win * 1)(win)
    off = np.zeros(1, np.int64)
    status = torch.zeros(2, dtype=torch.int32, device="cuda")
    def step():
        ctx.hfun_pass_windows(wins, off, off, z, _lib.OCSB_F32, n, n, n, None, n, out, n,
                              10.0, 10.0, 100.0, 1500.0, ops, status=status)
else:
    def step():
        ctx.hfun_pass(win, z, _lib.OCSB_F32, n, None, n, out, n, 10.0, 10.0, 100.0, 1500.0, ops, 0)
for _ in range(3): step()
torch.cuda.synchronize()
ctx.reset_timing(); ctx.set_timing(True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps): step()
e1.record(); torch.cuda.synchronize()
ctx.set_timing(False)
kt = {k: round(v[0] / steps, 4) for k, v in ctx.get_timing().items()}
crc = zlib.crc32(out[::7].cpu().numpy().tobytes())
print(json.dumps({"ms_step": round(e0.elapsed_time(e1) / steps, 4), "k": kt, "crc": crc, "nv": len(pts)}))
''' % ROOT


def main():
    args = sys.argv[1:]
    n, steps = 20000, 20
    variants = []
    while args:
        a = args.pop(0)
        if a == "--size":
            n = int(args.pop(0))
        elif a == "--steps":
            steps = int(args.pop(0))
        else:
            variants.append(a)
    if not variants:
        variants = ["product="]
    import numpy as np
    from ocsmesh_b200 import features as F, synthetic as S
    t0 = time.time()
    z = np.empty((n, n), np.float32)
    for s in range(0, n, 2000):
        z[s:s + 2000] = S.synth_dem(n, n, seed=2, rows=(s, min(n, s + 2000)))
    tr = S.dem_transform(n)
    x = np.linspace(tr[2], tr[2] + tr[0] * n, n)
    y = np.linspace(tr[5], tr[5] + tr[4] * n, n)
    lines = []
    for lv in (-5.0, 0.0, 5.0):
        lines.extend(F.marching_squares(x, y, z, lv))
    pts = F.feature_points(lines, 100.0)
    np.save("/dev/shm/pass_lab_dem.npy", z)
    np.save("/dev/shm/pass_lab_pts.npy", pts)
    print(f"inputs: {n}x{n}, {len(pts)} vertices, {time.time() - t0:.1f} s", flush=True)
    for v in variants:
        name, _, spec = v.partition("=")
        parts = [p for p in spec.split(",") if p]
        env = dict(os.environ)
        for p in parts:
            if "=" in p:
                k, _, val = p.partition("=")
                env[k] = val
            else:
                env["OCSB_LIB"] = p if os.path.isabs(p) else os.path.join(ROOT, p)
        r = subprocess.run([sys.executable, "-c", CHILD, str(n), str(steps)], env=env,
                           capture_output=True, text=True)
        line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ""
        print(f"{name:24s} {line}", flush=True)
        err = [l for l in r.stderr.splitlines() if l.startswith("PASS_STATS") or r.returncode]
        for l in err[-6:]:
            print(f"{'':24s} {l}", flush=True)
    os.remove("/dev/shm/pass_lab_dem.npy")
    os.remove("/dev/shm/pass_lab_pts.npy")


if __name__ == "__main__":
    main()
