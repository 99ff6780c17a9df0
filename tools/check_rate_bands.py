"""A rate constraint on a raster split into row bands over the ranks (NCCL):
every rank applies add_topo_bound_constraint(rate=...) to its band object --
mask step, all-gather of the bands' mask-boundary pixels, _apply_rate step --
and compares its rows with the whole raster done on its own GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
        --master-addr 127.0.0.1 --master-port 29555 tools/check_rate_bands.py
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ocsmesh_b200 as ob
from ocsmesh_b200 import distributed as D, synthetic as S

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dev = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{dev}"))
H, W = 3000, 2600
bad = 0
for chunk in (None, 1100):
    z = S.synth_dem(H, W, seed=6)
    tr = S.dem_transform(H)
    lines = S.synth_shoreline(H, W, seed=5, n_lines=2)
    part = D.RowBands(H, world)

    def build(row_band):
        r = ob.Raster.from_numpy(z, tr, crs=32618, chunk_size=chunk, overlap=0)
        h = ob.HfunRaster(r, hmin=50.0, hmax=2000.0, device=dev, row_band=row_band)
        h.add_feature(lines, 0.01, 60.0)
        return h

    # whole raster on this GPU (no process group involved: holds whole windows)
    full = build(None)
    full.add_topo_bound_constraint(400.0, upper_bound=3.0, lower_bound=-8.0,
                                   value_type="min", rate=0.02)
    want = full.get_values()
    band = build(part.band(rank))
    band.add_topo_bound_constraint(400.0, upper_bound=3.0, lower_bound=-8.0,
                                   value_type="min", rate=0.02)
    a, b = part.band(rank)
    got = band.get_values()
    if not np.array_equal(got, want[a:b]):
        bad += 1
        print(f"rank {rank} chunk {chunk}: {(got != want[a:b]).sum()} pixels differ",
              file=sys.stderr)
t = torch.tensor([bad], device=f"cuda:{dev}")
if world > 1:
    dist.all_reduce(t)
if rank == 0:
    print(json.dumps({"check": "rate constraint on row bands == whole raster",
                      "n_gpus": world, "raster": [H, W], "chunk_sizes": [None, 1100],
                      "mismatching_ranks": int(t.item())}), flush=True)
if world > 1:
    dist.destroy_process_group()
sys.exit(1 if int(t.item()) else 0)
