"""BASELINE.json configs[2]: HfunCollector-style multi-raster run -- 8
overlapping synthetic DEMs (4 x 2, 10000 x 10000 px, 1000-px overlap), every
contour of every raster applied to every raster (the reference's exact
collector order, ocsmesh/hfun/collector.py:1536-1655) + gradient delimiter, then
the size functions min-combined over the overlaps.

    python tools/run_c3.py [--tile 10000] [--steps 5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
        --master-addr 127.0.0.1 --master-port 29544 tools/run_c3.py

Raster r lives on rank r % N; per step the ranks all-gather the contour
vertices and swap the overlap strips pairwise over NCCL.  Prints one JSON line
(rank 0).  Size-independent checks: after the combine both sides of every
overlap hold identical pixels, and a second combine changes nothing.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, distributed as D, features as F, synthetic as S

RATE, TARGET, HMIN, HMAX = 0.001, 100.0, 100.0, 1500.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tile", type=int, default=10000)
    ap.add_argument("--overlap", type=int, default=1000)
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{dev}"))
    ctx = _lib.default_context(dev)
    T, ovl, nx, ny = args.tile, args.overlap, 4, 2
    owners = [D.raster_owner(k, world) for k in range(nx * ny)]
    geoms, hfuns = [], {}
    for k in range(nx * ny):
        ty, tx = divmod(k, nx)
        x0 = S.X0 + (T - ovl) * S.PIXEL * tx
        y0 = S.Y0 - (T - ovl) * S.PIXEL * ty
        tr = S.dem_transform(T, x0=x0, y0=y0)
        geoms.append((tuple(tr), T, T))
        if owners[k] != rank:
            continue
        z = S.synth_dem(T, T, seed=10 + k, x0=x0, y0=y0)
        r = ob.Raster.from_numpy(z, tr, crs=32618)
        r.device_band(1, dev)
        hfuns[k] = ob.HfunRaster(r, hmin=HMIN, hmax=HMAX, device=dev)

    # contour extraction (8f next-1) on the owned rasters, outside the step
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    lines = []
    for k in sorted(hfuns):
        lines.extend(hfuns[k].raster.get_contour(0.0))
    pts = F.feature_points(lines, TARGET)
    contour_s = time.perf_counter() - t0
    verts_local = torch.from_numpy(np.ascontiguousarray(pts)).to(f"cuda:{dev}")

    def get_window(i, w):
        return hfuns[i].window_values(w)

    def combine(i, w, t):
        hfuns[i].min_combine_values(w, t)

    vsizes = D.gather_sizes(len(pts), f"cuda:{dev}") if world > 1 else None

    def step():
        verts = (D.all_gather_vertices(verts_local, sizes=vsizes) if world > 1
                 else verts_local)
        for k in sorted(hfuns):
            h = hfuns[k]
            h.reset()
            h.add_feature_points(verts, RATE, TARGET)
            h.add_gradient_delimiter(hmin=50.0, hmax=1500.0)
        n = D.min_combine_overlaps(geoms, owners, rank, get_window, combine)
        return n, int(verts.shape[0])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(2):
        pairs, nv = step()
    barrier()
    ctx.reset_timing()
    ctx.set_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ctx.set_timing(False)
    ms = e0.elapsed_time(e1) / args.steps
    kt = {k: round(v[0] / args.steps, 4) for k, v in ctx.get_timing().items()}
    if world > 1:
        tt = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{dev}")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())

    # checks: both sides of every overlap identical; combine is idempotent
    bad = torch.zeros(1, dtype=torch.int64, device=f"cuda:{dev}")

    def check_equal(i, w, t):
        if not torch.equal(hfuns[i].window_values(w), t):
            bad.add_(1)

    before = {k: h.device_values.clone() for k, h in hfuns.items()}
    D.min_combine_overlaps(geoms, owners, rank, get_window, check_equal)
    D.min_combine_overlaps(geoms, owners, rank, get_window, combine)
    for k, h in hfuns.items():
        if not torch.equal(before[k], h.device_values):
            bad.add_(1)
        lo, hi = float(h.device_values.min()), float(h.device_values.max())
        if lo < 50.0 or hi > HMAX:
            bad.add_(1)
    if world > 1:
        dist.all_reduce(bad)
    px = float(nx * ny) * T * T
    if rank == 0:
        print(json.dumps({
            "config": f"BASELINE.json configs[2]: {nx * ny} rasters {T}x{T} "
                      f"({nx}x{ny}, {ovl}-px overlap), raster r on rank r % {world}; "
                      "per step: vertex all-gather, add_feature (every contour "
                      "to every raster) + add_gradient_delimiter per raster, "
                      "pairwise overlap min-combine",
            "n_gpus": world, "Mpx_per_s": px / (ms * 1e-3) / 1e6, "ms_per_step": ms,
            "steps": args.steps, "overlapping_pairs": pairs, "n_vertices": nv,
            "kernel_ms_per_step_rank0": kt,
            "contour_extract_s_rank0": round(contour_s, 3),
            "checks_failed": int(bad.item())}), flush=True)
    if world > 1:
        dist.destroy_process_group()
    sys.exit(1 if int(bad.item()) else 0)


if __name__ == "__main__":
    main()
