#!/usr/bin/env python
"""BASELINE.json configs[4]: add_feature distance sizing against a 5 M-vertex
synthetic shoreline set (2000 random-walk polylines x 2500 points, 40 m steps,
seed 5) on a 40000 x 40000 raster -- the vertex-set stress case of K1.  The
feature op needs no DEM (only the window geometry), so only the 6.4 GB hfun
raster lives on the device.  Prints one JSON line: device ms per call (median
of 3), per-kernel ms, and a bit-exact check of 20000 sampled pixels against an
exact scipy kd-tree query + the size law of hfun/raster.py:1240-1257."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from scipy.spatial import cKDTree
from ocsmesh_b200 import _lib, synthetic as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 40000
HMIN, HMAX, RATE, TARGET = 100.0, 1500.0, 0.001, 100.0
ctx = _lib.default_context(0)
tr = S.dem_transform(n)
x0, y1 = tr[2], tr[5]
x1, y0 = x0 + tr[0] * n, y1 + tr[4] * n
pts = S.synth_random_walk_features(2000, 2500, 40.0, (x0, y0, x1, y1), seed=5)
vd = torch.from_numpy(pts).cuda()
out = torch.empty((n, n), dtype=torch.float32, device="cuda")
win = _lib.make_win((x0, y0, x1, y1), n, n)


def step():
    ctx.hfun_feature(win, vd, len(pts), None, n, out, n, RATE, TARGET, HMIN, HMAX, True)


step(); torch.cuda.synchronize()
ts = []
for _ in range(3):
    ctx.reset_timing(); ctx.set_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); step(); e1.record(); torch.cuda.synchronize()
    ctx.set_timing(False)
    ts.append(e0.elapsed_time(e1))
    kt = {k: round(v[0], 3) for k, v in ctx.get_timing().items()}
ms = float(np.median(ts))
# ---- sampled exactness ------------------------------------------------------------
rng = np.random.default_rng(0)
rows, cols = rng.integers(0, n, 20000), rng.integers(0, n, 20000)
x = np.linspace(x0, x1, n); y = np.linspace(y1, y0, n)
xy = np.stack([x[cols], y[rows]], axis=1)
t0 = time.perf_counter()
tree = cKDTree(pts)
r = (HMAX - TARGET) / (RATE * TARGET)
d, _ = tree.query(xy, distance_upper_bound=r, workers=-1)
t_cpu = time.perf_counter() - t0
d = np.where(np.isinf(d), r, d)
v = (RATE * TARGET * d + TARGET).astype(np.float32)
v[v < HMIN] = HMIN; v[v > HMAX] = HMAX
ref = np.minimum(np.float32(np.finfo(np.float32).max), v)
got = out[rows, cols].cpu().numpy()
print(json.dumps({"config": f"BASELINE.json configs[4]: {n}x{n}, {len(pts)} vertices",
                  "ms_call": round(ms, 3), "Mpx_per_s": round(n * n / ms / 1e3, 1),
                  "kernel_ms": kt, "sampled_pixels": len(rows),
                  "bit_exact": bool(np.array_equal(got, ref)),
                  "mismatches": int((got != ref).sum()),
                  "frac_within_radius": round(float((d < r).mean()), 3),
                  "cpu_kdtree_build+20k_queries_s": round(t_cpu, 2)}), flush=True)
