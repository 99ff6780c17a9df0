#!/usr/bin/env python
"""Where the HOST time of a band step goes (at 8 GPUs a step's kernels take
~0.25 ms: the enqueue path must stay below that).  One GPU, the band a rank of
an 8-rank run would own, steps through the public deferred API under cProfile.

    python tools/host_profile.py [--steps N] [--timing]
"""
import cProfile, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, distributed as D, features as F, synthetic as S

steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 300
H = W = 20000
part = D.RowBands(H, 8)
rank = 3
b0, b1 = part.band(rank)
s0, s1 = part.stored(rank)
tr = S.dem_transform(H)
z = S.synth_dem(H, W, seed=2, rows=(s0, s1))
x = np.linspace(tr[2], tr[2] + tr[0] * W, W)
y = np.linspace(tr[5], tr[5] + tr[4] * H, H)[s0:s1]
lines = []
for lv in (-5.0, 0.0, 5.0):
    lines.extend(F.marching_squares(x, y, z, lv))
pts = torch.from_numpy(F.feature_points(lines, 100.0)).cuda()
raster = D.band_raster(z[b0 - s0:b1 - s0], tr, H, part, rank, crs=32618, device=0)
hf = ob.HfunRaster(raster, hmin=100.0, hmax=1500.0, device=0, row_band=(b0, b1), deferred=True)
ctx = _lib.default_context(0)


def step():
    hf.reset()
    hf.add_feature_points(pts, 0.001, 100.0)
    hf.add_gradient_delimiter(hmin=50.0, hmax=1500.0)
    hf.flush()


for _ in range(5):
    step()
torch.cuda.synchronize()
if "--timing" in sys.argv:
    ctx.set_timing(True)
for label in ("plain", "cProfile"):
    pr = cProfile.Profile() if label == "cProfile" else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    t0 = time.perf_counter()
    if pr:
        pr.enable()
    for _ in range(steps):
        step()
    if pr:
        pr.disable()
    host = (time.perf_counter() - t0) * 1e3 / steps
    e1.record()
    torch.cuda.synchronize()
    print(f"{label}: host enqueue {host:.4f} ms per step, device {e0.elapsed_time(e1) / steps:.4f} "
          f"ms per step, {len(pts)} vertices", flush=True)
    ctx.reset_timing()
    if pr:
        pstats.Stats(pr).sort_stats("tottime").print_stats(22)
