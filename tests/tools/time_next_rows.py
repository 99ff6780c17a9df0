#!/usr/bin/env python
"""Device timings of the SURVEY.md 8(f) rows built so far, on the bench DEM
(20000 x 20000 f32, resident): contour segment extraction (next-1),
add_constant_value / topo constraint with rate / Courant constraint (next-2).
Prints one JSON line per op: CUDA-event ms per call (median of 5 after 2
warm-ups), per-kernel-class ms, algorithmic bytes/px and the resulting
fraction of the measured HBM peak; the host marching-squares sweep and the
numpy oracle are timed beside them on a crop for scale.

Test infrastructure (uses the oracle as the checker): lives under tests/."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, features as F, synthetic as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
peak = 6460.2
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
ctx = _lib.default_context(0)
z = np.empty((n, n), np.float32)
for s in range(0, n, 2000):
    z[s:s + 2000] = S.synth_dem(n, n, seed=2, rows=(s, min(n, s + 2000)))
tr = S.dem_transform(n)
r = ob.Raster.from_numpy(z, tr, crs=32618)
r.device_band(1, 0)
px = float(n) * n


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    out = []
    kt = None
    for _ in range(reps):
        ctx.reset_timing(); ctx.set_timing(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ctx.set_timing(False)
        out.append(e0.elapsed_time(e1))
        kt = {k: round(v[0], 4) for k, v in ctx.get_timing().items()}
    return float(np.median(out)), kt


def report(name, ms, kt, bytes_px, kclass, extra=None):
    kms = kt.get(kclass, ms)
    rec = {"op": name, "px": int(px), "ms_call": round(ms, 4), "kernel_ms": kt,
           "algorithmic_bytes_per_px": bytes_px,
           "roofline_frac_kernels": round(bytes_px * px / (kms * 1e-3) / 1e9 / peak, 4),
           "Mpx_per_s_call": round(px / (ms * 1e-3) / 1e6, 1)}
    if extra:
        rec.update(extra)
    print(json.dumps(rec), flush=True)


# ---- next-1: contour segments (count pass + emit pass = 2 x 4 B/px) --------------
win = _lib.make_win(r.get_window_bounds(ob.Window(0, 0, n, n)), n, n)
dem = r.device_band(1, 0)
segs = {}
def contour():
    e, p = ctx.contour_segments(win, dem, _lib.OCSB_F32, n, 0.0)
    segs["n"] = len(e)
ms, kt = timed(contour)
t0 = time.perf_counter(); lines = r.get_contour(0.0); t_full = time.perf_counter() - t0
c = min(n, 4000)
t0 = time.perf_counter()
F.marching_squares(np.linspace(0, 1, c), np.linspace(1, 0, c), z[:c, :c], 0.0)
t_host = time.perf_counter() - t0
report("contour_segments(level 0)", ms, kt, 8.0, "contour",
       {"segments": segs["n"], "polylines": len(lines),
        "get_contour_total_s(device sweep + host stitch)": round(t_full, 3),
        "host_numpy_sweep_Mpx_per_s(4000^2 crop)": round(c * c / t_host / 1e6, 1)})

# ---- next-2 ------------------------------------------------------------------------
hf = ob.Hfun(r, hmin=100.0, hmax=1500.0)
lines3 = []
for lv in (-5.0, 0.0, 5.0):
    lines3.extend(r.get_contour(lv))
pts = torch.from_numpy(F.feature_points(lines3, 100.0)).cuda()
hf.add_feature_points(pts, 0.001, 100.0)
base = hf.device_values.clone()

def restore():
    hf._hfun.copy_(base)

def const():
    hf.add_constant_value(250.0, -3.0, 2.5)
ms, kt = timed(const)
report("add_constant_value(250, -3, 2.5) [clamp fused]", ms, kt, 12.0, "constraint")

from ocsmesh_b200.constraint import TopoConstConstraint, CourantNumConstraint
cour = [CourantNumConstraint(0.9, 150, 2.0, "max")]
ms, kt = timed(lambda: hf.apply_constraints(cour))
report("apply_constraints([Courant max 0.9]) + clamp", ms, kt, 12.0, "constraint")

restore()
topo = [TopoConstConstraint(400.0, 0.0, -10.0, "min", None)]
ms, kt = timed(lambda: (restore(), hf.apply_constraints(topo)))
report("apply_constraints([TopoConst min 400 in (-10, 0), no rate]) + clamp (+ restore copy)",
       ms, kt, 12.0, "constraint")

topo_r = [TopoConstConstraint(400.0, 0.0, -10.0, "min", 0.01)]
ms, kt = timed(lambda: (restore(), hf.apply_constraints(topo_r)), reps=3, warm=1)
report("apply_constraints([TopoConst min 400 in (-10, 0), rate 0.01]) + clamp (+ restore copy)",
       ms, kt, 12.0 + 1.0 + 8.0, "feature_eval",
       {"note": "mask (13 B/px incl. u8 mask) + boundary compaction + K1 refine/eval with the "
                "rate epilogue (8 B/px hfun + 1 B/px mask)"})

# ---- next-3: msh_t sampling side ------------------------------------------------------
K = 10_000_000
x0, y0, x1, y1 = r.get_window_bounds(ob.Window(0, 0, n, n))
rng = np.random.default_rng(4)
nodes = torch.from_numpy(np.stack([x0 + rng.random(K) * (x1 - x0),
                                   y0 + rng.random(K) * (y1 - y0)], axis=1)).cuda()
restore()
ms, kt = timed(lambda: hf.sample_nearest(nodes, None, None))
rec = {"op": "sample_nearest(10 M random nodes, stride 1)", "nodes": K, "ms_call": round(ms, 4),
       "kernel_ms": kt, "algorithmic_bytes_per_node": 28.0,
       "roofline_frac_kernels": round(28.0 * K / (kt.get("sample", ms) * 1e-3) / 1e9 / peak, 4),
       "Mnodes_per_s_call": round(K / (ms * 1e-3) / 1e6, 1),
       "note": "random (unbucketed) nodes: one 32 B sector per 4 B texel"}
print(json.dumps(rec), flush=True)
ms, kt = timed(lambda: hf.sizing_source(None, 4))
rec = {"op": "sizing_source(stride 4): 25 M source points, coords + values", "ms_call": round(ms, 4),
       "kernel_ms": kt, "bytes_out": 25_000_000 * 24,
       "GBps_out": round(25_000_000 * 24 / (kt.get("sample", ms) * 1e-3) / 1e9, 1)}
print(json.dumps(rec), flush=True)
