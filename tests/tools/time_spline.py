"""Mesh.interpolate(method='spline') -- the reference's default method -- at
BASELINE.json configs[3] scale: 16 tiles 5000 x 5000, 10 M nodes.  Kernel
times (prefilter per tile, evaluation) from the library's own events, and the
agreement with scipy's RectBivariateSpline on one tile's sample.

    python tests/tools/time_spline.py [lib.so ...]     (OCSB_LIB variants, as sweep_interp)

Test infrastructure (uses the oracle as the checker): lives under tests/.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def child():
    import numpy as np
    import torch
    import ocsmesh_b200 as ob
    from ocsmesh_b200 import _lib, synthetic as S
    from ocsmesh_b200.mesh import EuclideanMesh2D
    from oracle import ocsmesh_oracle as O

    dev = 0
    ctx = _lib.default_context(dev)
    T, ovl, grid, N = 5000, 100, 4, 10_000_000
    rasters, host = [], []
    for k in range(grid * grid):
        ty, tx = divmod(k, grid)
        z = S.synth_dem(T, T, seed=20 + k)
        tr = list(S.dem_transform(T))
        tr[2] += (T - ovl) * S.PIXEL * tx
        tr[5] += (T - ovl) * S.PIXEL * ty
        r = ob.Raster.from_numpy(z, tr, crs=32618)
        r.device_band(1, dev)
        rasters.append(r)
        host.append((z, tr))
    ext = ((T - ovl) * (grid - 1) + T) * S.PIXEL
    bbox = (S.X0, S.Y0, S.X0 + ext, S.Y0 + ext)
    xy_h = S.synth_nodes(N, bbox, seed=400, outside_frac=0.02)
    xy = torch.from_numpy(xy_h).cuda()
    vals = torch.full((N,), float("nan"), dtype=torch.float64, device="cuda")
    touched = torch.zeros(N, dtype=torch.uint8, device="cuda")
    res = {"lib": _lib.LIB_PATH}
    if os.environ.get("OCSB_SWEEP_NCU"):
        # under ncu: two tiles, one call
        EuclideanMesh2D._interp_device(ctx, rasters[:2], xy, vals, touched, "spline", 1, dev)
        torch.cuda.synchronize()
        return
    # agreement with scipy on the last tile (it wins where tiles overlap)
    EuclideanMesh2D._interp_device(ctx, rasters, xy, vals, touched, "spline", 1, dev)
    torch.cuda.synchronize()
    got = vals.cpu().numpy()
    z, tr = host[-1]
    orr = O.OracleRaster(z, tr)
    b = O.array_bounds(T, T, tr)
    inside = np.nonzero((xy_h[:, 0] > b[0]) & (xy_h[:, 0] < b[2]) &
                        (xy_h[:, 1] > b[1]) & (xy_h[:, 1] < b[3]))[0][:100_000]
    want = O.mesh_interpolate(xy_h[inside], np.full(len(inside), np.nan), orr,
                              method="spline").reshape(-1)
    res["max_rel_err_vs_scipy_100k_nodes"] = float(
        np.max(np.abs(got[inside] - want) / np.maximum(np.abs(want), 1.0)))
    ctx.set_timing(True)
    ctx.reset_timing()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        EuclideanMesh2D._interp_device(ctx, rasters, xy, vals, touched, "spline", 1, dev)
    e1.record()
    torch.cuda.synchronize()
    tm = ctx.get_timing()
    ctx.set_timing(False)
    res["call_ms_16_tiles_10M_nodes"] = round(e0.elapsed_time(e1) / 3, 3)
    res["kernel_ms_per_call"] = {k: round(v[0] / 3, 3) for k, v in tm.items()}
    res["launches_per_call"] = {k: v[1] // 3 for k, v in tm.items()}
    pre = tm.get("spline_prefilter", (0, 1))[0] / 3 / 16
    res["prefilter_ms_per_tile"] = round(pre, 4)
    # prefilter: read DEM 4 B/px, write coefficients 8 B/px, twice (two axes)
    res["prefilter_GBs_at_16B_per_px"] = round(16.0 * T * T / (pre * 1e-3) / 1e9, 1)
    print(json.dumps(res), flush=True)


def main():
    for lib in (sys.argv[1:] or [""]):
        env = dict(os.environ, OCSB_CHILD="1")
        if lib:
            env["OCSB_LIB"] = os.path.abspath(lib)
        subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, check=False)


if __name__ == "__main__":
    child() if os.environ.get("OCSB_CHILD") else main()
