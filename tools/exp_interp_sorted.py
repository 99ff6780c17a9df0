"""Experiment: Mesh.interpolate gather time vs node ordering (random vs
spatially bucketed).  python tools/exp_interp_sorted.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, synthetic as S
from ocsmesh_b200.mesh import EuclideanMesh2D

dev = 0
ctx = _lib.default_context(dev)
T, ovl, grid, N = 5000, 100, 4, 10_000_000
rasters = []
for k in range(grid * grid):
    ty, tx = divmod(k, grid)
    z = S.synth_dem(T, T, seed=20 + k)
    tr = list(S.dem_transform(T))
    tr[2] += (T - ovl) * S.PIXEL * tx
    tr[5] += (T - ovl) * S.PIXEL * ty
    r = ob.Raster.from_numpy(z, tr, crs=32618)
    r.device_band(1, dev)
    rasters.append(r)
ext = ((T - ovl) * (grid - 1) + T) * S.PIXEL
bbox = (S.X0, S.Y0, S.X0 + ext, S.Y0 + ext)
xy_h = S.synth_nodes(N, bbox, seed=400, outside_frac=0.02)


def run(xy_np, label, method="linear"):
    xy = torch.from_numpy(np.ascontiguousarray(xy_np)).cuda()
    vals = torch.full((len(xy_np),), float("nan"), dtype=torch.float64, device="cuda")
    touched = torch.zeros(len(xy_np), dtype=torch.uint8, device="cuda")
    for _ in range(3):
        EuclideanMesh2D._interp_device(ctx, rasters, xy, vals, touched, method, 1, dev)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        EuclideanMesh2D._interp_device(ctx, rasters, xy, vals, touched, method, 1, dev)
    e1.record()
    torch.cuda.synchronize()
    print(f"{label:28s} {method:8s} {e0.elapsed_time(e1) / 10:.3f} ms")
    return vals.cpu().numpy()


v0 = run(xy_h, "random order")
# bucket by 64x64-px cells of a global 10 m grid, row-major over buckets
cx = ((xy_h[:, 0] - S.X0) / (64 * S.PIXEL)).astype(np.int64)
cy = ((xy_h[:, 1] - S.Y0) / (64 * S.PIXEL)).astype(np.int64)
order = np.argsort(cy * 100000 + cx, kind="stable")
v1 = run(xy_h[order], "bucketed 64px cells")
assert np.array_equal(np.nan_to_num(v0[order], nan=-1), np.nan_to_num(v1, nan=-1))
order2 = np.lexsort((xy_h[:, 0], -xy_h[:, 1]))
run(xy_h[order2], "sorted by y then x")
run(xy_h[order], "bucketed 64px cells", "nearest")
# device-side sort cost
xy = torch.from_numpy(xy_h).cuda()
torch.cuda.synchronize()
t = time.perf_counter()
for _ in range(5):
    key = ((xy[:, 1] - S.Y0) / 640.0).to(torch.int64) * 100000 + ((xy[:, 0] - S.X0) / 640.0).to(torch.int64)
    o = torch.argsort(key)
    xs = xy[o]
torch.cuda.synchronize()
print(f"torch key+argsort+gather: {(time.perf_counter() - t) / 5 * 1e3:.3f} ms")
