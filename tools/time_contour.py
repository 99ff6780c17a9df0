#!/usr/bin/env python
"""Where the time of the device contour pipeline goes: Raster.get_contour_points
at 20000^2, wall / CUDA-event time per call and the per-class kernel time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, synthetic as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
z = np.empty((n, n), np.float32)
for s in range(0, n, 2000):
    z[s:s + 2000] = S.synth_dem(n, n, seed=2, rows=(s, min(n, s + 2000)))
r = ob.Raster.from_numpy(z, S.dem_transform(n), crs=32618)
r.device_band(1, 0)
ctx = _lib.default_context(0)
for _ in range(2):
    p = r.get_contour_points([-5.0, 0.0, 5.0], 100.0)
torch.cuda.synchronize()
ctx.reset_timing(); ctx.set_timing(True)
t0 = time.perf_counter()
for _ in range(reps):
    p = r.get_contour_points([-5.0, 0.0, 5.0], 100.0)
torch.cuda.synchronize()
t1 = time.perf_counter()
ctx.set_timing(False)
print("vertices", tuple(p.shape), "wall ms per call", round((t1 - t0) / reps * 1e3, 3),
      "timed kernel classes ms per call", {k: round(v[0] / reps, 3) for k, v in ctx.get_timing().items()})
