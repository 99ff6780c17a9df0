"""configs[1] read literally: the 20000 x 20000 DEM "tiled by raster windows"
(Raster chunk_size / overlap, ocsmesh/raster.py:1891-1939).  Every window is
its own semantic unit (own coordinates, own kd-tree radius, one-sided
differences at its edges), so the pass is one set of launches per
window.  Times the device-resident pass for several window sizes.

    python tools/time_windows.py [size]
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, features as F, synthetic as S

N = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
LEVELS = [-5.0, 0.0, 5.0]
ctx = _lib.default_context(0)
z = np.empty((N, N), np.float32)
for s in range(0, N, 2000):
    z[s:s + 2000] = S.synth_dem(N, N, seed=2, rows=(s, min(N, s + 2000)))
tr = S.dem_transform(N)
r0 = ob.Raster.from_numpy(z, tr, crs=32618)
lines = []
for lv in LEVELS:
    lines.extend(r0.get_contour(lv))
pts = torch.from_numpy(F.feature_points(lines, 100.0)).cuda()
dem = r0.device_band(1, 0)
for chunk, ovl in ((None, None), (10000, 0), (5000, 0), (5000, 2), (2000, 0), (1000, 0)):
    r = ob.Raster.from_numpy(z, tr, crs=32618, chunk_size=chunk, overlap=ovl)
    r.set_device_band(dem, 0, band=1, device=0)
    h = ob.HfunRaster(r, hmin=100.0, hmax=1500.0, device=0)
    nwin = len(list(r.iter_windows()))

    def step():
        h.reset()
        h.add_feature_points(pts, 0.001, 100.0)
        h.add_gradient_delimiter(hmin=50.0, hmax=1500.0)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    ctx.reset_timing()
    ctx.set_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        step()
    e1.record()
    torch.cuda.synchronize()
    ctx.set_timing(False)
    ms = e0.elapsed_time(e1) / 5
    print(json.dumps({"chunk_size": chunk, "overlap": ovl, "windows": nwin,
                      "ms_per_pass": round(ms, 3),
                      "Mpx_per_s": round(N * N / ms / 1e3, 1),
                      "kernel_ms": {k: round(v[0] / 5, 3) for k, v in ctx.get_timing().items()},
                      "launches": {k: v[1] // 5 for k, v in ctx.get_timing().items()}}),
          flush=True)
