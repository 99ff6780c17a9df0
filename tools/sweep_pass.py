#!/usr/bin/env python
"""Streamed-pass e2e experiments: PCIe floor (H2D, D2H, both at once) and
ocsb_hfun_pass with pinned host buffers for several band heights."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from ocsmesh_b200 import _lib, features as F, synthetic as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
ctx = _lib.default_context(0)
zh = torch.empty((n, n), dtype=torch.float32).pin_memory()
for s in range(0, n, 2000):
    zh[s:s + 2000] = torch.from_numpy(S.synth_dem(n, n, seed=2, rows=(s, min(n, s + 2000))))
oh = torch.empty((n, n), dtype=torch.float32).pin_memory()
zd = torch.empty((n, n), dtype=torch.float32, device="cuda")
od = torch.empty((n, n), dtype=torch.float32, device="cuda")
def t(f, reps=3):
    f(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def h2d():
    with torch.cuda.stream(s1): zd.copy_(zh, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): oh.copy_(od, non_blocking=True)
def both(): h2d(); d2h()
gb = n * n * 4 / 1e9
for name, f in (("H2D", h2d), ("D2H", d2h), ("H2D+D2H concurrently", both)):
    ms = t(f); print(f"{name}: {ms:.2f} ms  {gb / ms * 1e3:.1f} GB/s per direction", flush=True)
tr = S.dem_transform(n)
x = np.linspace(tr[2], tr[2] + tr[0] * n, n); y = np.linspace(tr[5], tr[5] + tr[4] * n, n)
lines = []
for lv in (-5.0, 0.0, 5.0):
    lines.extend(F.marching_squares(x, y, zh.numpy(), lv))
pts = torch.from_numpy(F.feature_points(lines, 100.0)).pin_memory()
win = _lib.make_win((tr[2], tr[5] + tr[4] * n, tr[2] + tr[0] * n, tr[5]), n, n, 0, 0)
ops = [("feature", pts, len(pts), 0.001, 100.0, True), ("limiter", 50.0, 1500.0, None, None, True)]
for rows in (256, 512, 768, 1024, 2048, 4096):
    f = lambda: ctx.hfun_pass(win, zh, _lib.OCSB_F32, n, None, n, oh, n, 10.0, 10.0, 100.0, 1500.0, ops, rows)
    ms = t(f); print(f"pass band_rows={rows}: {ms:.2f} ms  {n * n / ms / 1e3:.0f} Mpx/s", flush=True)
