#!/usr/bin/env python
"""Single process emulating rank `r` of `n` (no collectives): is a band slow
because of its data / geometry or because of the GPU it ran on?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, synthetic as S, distributed as D, features as F
size, world = int(sys.argv[1]), int(sys.argv[2])
H, W = size * world, size
tr = S.dem_transform(H)
part = D.RowBands(H, world)
allpts = []
zs = {}
for rank in range(world):
    r0, r1 = part.band(rank)
    z = np.concatenate([S.synth_dem(H, W, seed=2, rows=(s, min(r1, s + 2000))) for s in range(r0, r1, 2000)])
    zs[rank] = z
    x = np.linspace(tr[2], tr[2] + tr[0] * W, W); y = np.linspace(tr[5] + tr[4] * r0, tr[5] + tr[4] * r1, r1 - r0)
    lines = []
    for lv in (-5.0, 0.0, 5.0): lines.extend(F.marching_squares(x, y, z, lv))
    allpts.append(F.feature_points(lines, 100.0))
    print(f"rank {rank}: z min {z.min():.1f} max {z.max():.1f} verts {len(allpts[-1])}", flush=True)
verts = torch.from_numpy(np.concatenate(allpts)).cuda()
ctx = _lib.default_context(0)
for rank in range(world):
    raster = D.band_raster(zs[rank], tr, H, part, rank, crs=32618, device=0)
    hfun = ob.HfunRaster(raster, hmin=100, hmax=1500, device=0, row_band=part.band(rank))
    def step():
        hfun.reset(); hfun.add_feature_points(verts, 0.001, 100.0)
        hfun.add_gradient_delimiter(hmin=50, hmax=1500)
    for _ in range(3): step()
    ctx.reset_timing(); ctx.set_timing(True)
    for _ in range(5): step()
    t = ctx.get_timing(); ctx.set_timing(False)
    print(f"rank {rank}: " + ", ".join(f"{k} {v[0] / 5:.3f} ms" for k, v in t.items()), flush=True)
