"""Interpolation gather (K3) timing + parity sweep over library builds.

    python tests/tools/sweep_interp.py [lib.so ...]

Each library (default: the product one) is loaded in a fresh process through
OCSB_LIB.  Workload = BASELINE.json configs[3]: 10 M nodes, 16 tiles 5000x5000
(4x4, 100-px overlap), nodes resident in bucket order.  A 200 k-node sample is
checked bit-for-bit against the scipy oracle before anything is timed.

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
    xy_raw = torch.from_numpy(xy_h).cuda()
    xy, perm = EuclideanMesh2D.bucket_nodes(ctx, rasters, xy_raw, dev)
    res = {"lib": _lib.LIB_PATH}
    if os.environ.get("OCSB_SWEEP_NCU"):
        # under ncu: two launches per method in bucket order, nothing else
        vals = torch.full((N,), float("nan"), dtype=torch.float64, device="cuda")
        touched = torch.zeros(N, dtype=torch.uint8, device="cuda")
        for method in ("linear", "nearest"):
            for _ in range(2):
                EuclideanMesh2D._interp_device(ctx, rasters, xy, vals, touched, method, 1, dev)
        torch.cuda.synchronize()
        return
    # parity on a sample (mesh order), oracle = scipy RGI per raster, last wins
    ns = 200_000
    sub = np.ascontiguousarray(xy_h[:ns])
    for method in ("linear", "nearest"):
        orasters = [O.OracleRaster(z, tr) for z, tr in host]
        want = O.mesh_interpolate(sub, np.full(ns, np.nan), orasters, method=method)[:, 0]
        xs = torch.from_numpy(sub).cuda()
        v = torch.full((ns,), float("nan"), dtype=torch.float64, device="cuda")
        t = torch.zeros(ns, dtype=torch.uint8, device="cuda")
        EuclideanMesh2D._interp_device(ctx, rasters, xs, v, t, method, 1, dev)
        got = v.cpu().numpy()
        same = (got == want) | (np.isnan(got) & np.isnan(want))
        res[f"parity_{method}"] = bool(same.all())
        res[f"mismatch_{method}"] = int((~same).sum())
    vals = torch.full((N,), float("nan"), dtype=torch.float64, device="cuda")
    touched = torch.zeros(N, dtype=torch.uint8, device="cuda")
    for method, bpn in (("linear", 40.0), ("nearest", 28.0)):
        for label, q, p in (("bucket", xy, None), ("bucket_scatter", xy, perm),
                            ("mesh_order", xy_raw, None)):
            for _ in range(3):
                EuclideanMesh2D._interp_device(ctx, rasters, q, vals, touched, method, 1, dev, perm=p)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                EuclideanMesh2D._interp_device(ctx, rasters, q, vals, touched, method, 1, dev, perm=p)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
            res[f"{method}_{label}_ms"] = round(ms, 4)
            # the kernel alone (events on the library's stream around the launch)
            ctx.set_timing(True)
            ctx.reset_timing()
            for _ in range(10):
                EuclideanMesh2D._interp_device(ctx, rasters, q, vals, touched, method, 1, dev, perm=p)
            torch.cuda.synchronize()
            tm = ctx.get_timing()
            ctx.set_timing(False)
            res[f"{method}_{label}_kernel_ms"] = {k: round(v[0] / v[1], 4) for k, v in tm.items()}
            if label == "bucket":
                res[f"{method}_GBs_algorithmic"] = round(bpn * N / ms / 1e6, 1)
    print(json.dumps(res), flush=True)


def main():
    libs = sys.argv[1:] or [""]
    for lib in libs:
        env = dict(os.environ, OCSB_CHILD="1")
        if lib:
            env["OCSB_LIB"] = os.path.abspath(lib)
        subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, check=False)


if __name__ == "__main__":
    if os.environ.get("OCSB_CHILD"):
        child()
    else:
        main()
