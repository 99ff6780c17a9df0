#!/usr/bin/env python
"""Per-piece timing of the multi-GPU step (run under torchrun)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, synthetic as S, distributed as D, features as F
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{lr}"))
size = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
H, W = size * world, size
tr = S.dem_transform(H)
part = D.RowBands(H, world)
r0, r1 = part.band(rank)
z = np.concatenate([S.synth_dem(H, W, seed=2, rows=(s, min(r1, s + 2000))) for s in range(r0, r1, 2000)])
x = np.linspace(tr[2], tr[2] + tr[0] * W, W); y = np.linspace(tr[5] + tr[4] * r0, tr[5] + tr[4] * r1, r1 - r0)
lines = []
for lv in (-5.0, 0.0, 5.0): lines.extend(F.marching_squares(x, y, z, lv))
pts = torch.from_numpy(F.feature_points(lines, 100.0)).cuda()
raster = D.band_raster(z, tr, H, part, rank, crs=32618, device=lr)
hfun = ob.HfunRaster(raster, hmin=100, hmax=1500, device=lr, row_band=part.band(rank))
def timeit(name, f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n * 1e3
    print(f"[rank {rank}] {name}: {dt:.3f} ms", flush=True)
verts = D.all_gather_vertices(pts)
timeit("all_gather_vertices", lambda: D.all_gather_vertices(pts))
timeit("exchange_dem_halo", lambda: D.exchange_dem_halo(raster, part, rank, lr))
def feat():
    hfun.reset(); hfun.add_feature_points(verts, 0.001, 100.0)
timeit("reset+add_feature_points", feat)
timeit("add_gradient_delimiter", lambda: hfun.add_gradient_delimiter(hmin=50, hmax=1500))
def step():
    hfun.reset()
    v = D.all_gather_vertices(pts)
    D.exchange_dem_halo(raster, part, rank, lr)
    hfun.add_feature_points(v, 0.001, 100.0)
    hfun.add_gradient_delimiter(hmin=50, hmax=1500)
timeit("full step", step)
ctx = _lib.default_context(lr)
ctx.reset_timing(); ctx.set_timing(True)
for _ in range(5): step()
t = ctx.get_timing(); ctx.set_timing(False)
print(f"[rank {rank}] device-side: " + ", ".join(f"{k} {v[0] / 5:.3f} ms" for k, v in t.items()), flush=True)
print(f"[rank {rank}] hfun dev {hfun._hfun.device} dem dev {raster.device_band(1, lr).device} verts {verts.device} cur {torch.cuda.current_device()} affinity {len(os.sched_getaffinity(0))}", flush=True)
dist.destroy_process_group()
