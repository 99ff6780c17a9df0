#!/usr/bin/env python
"""Determinism check of the resident pass on a row band with a padded vertex
set (what a rank of a multi-GPU run executes): the same call repeated must give
the same raster every time, and the persistent kernel the same as the plain
tile kernel (OCSB_PASS_V2=0, run as a second process by the caller)."""
import sys, os, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from ocsmesh_b200 import _lib, features as F, synthetic as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
dup = int(sys.argv[3]) if len(sys.argv) > 3 else 12
rb, re = 0, n // 2
ctx = _lib.default_context(0)
z = np.empty((n, n), np.float32)
for s in range(0, n, 2000):
    z[s:s + 2000] = S.synth_dem(n, n, seed=2, rows=(s, min(n, s + 2000)))
tr = S.dem_transform(n)
x = np.linspace(tr[2], tr[2] + tr[0] * n, n); y = np.linspace(tr[5], tr[5] + tr[4] * n, n)
lines = []
for lv in (-5.0, 0.0, 5.0):
    lines.extend(F.marching_squares(x, y, z, lv))
p = F.feature_points(lines, 100.0)
p = np.concatenate([p, np.repeat(p[-1:], dup, axis=0)])
pts = torch.from_numpy(p).cuda()
zd = torch.from_numpy(z[rb:re + 1]).cuda()          # band + the halo row below
out = torch.empty((re - rb, n), dtype=torch.float32, device="cuda")
win = _lib.make_win((tr[2], tr[5] + tr[4] * n, tr[2] + tr[0] * n, tr[5]), n, n, rb, re)
wins = (_lib.ocsb_win * 1)(win)
ops = [("feature", pts, len(pts), 0.001, 100.0, True), ("limiter", 50.0, 1500.0, None, None, True)]
off = np.zeros(1, np.int64)
status = torch.zeros(2, dtype=torch.int32, device="cuda")
ref = None
bad = 0
for i in range(reps):
    out.fill_(-1.0)
    ctx.hfun_pass_windows(wins, off, off, zd, _lib.OCSB_F32, n, re - rb + 1, n, None, n, out, n,
                          10.0, 10.0, 100.0, 1500.0, ops, status=status)
    torch.cuda.synchronize()
    if ref is None:
        ref = out.clone()
    elif not torch.equal(ref, out):
        d = (ref != out).nonzero()
        bad += 1
        print("run", i, "differs at", len(d), "pixels, first", d[0].tolist(), flush=True)
print("status", status.tolist(), "nv", len(p), "bad runs", bad, "crc", zlib.crc32(ref.cpu().numpy().tobytes()))
