"""ocsb_interp_multi (one launch for a raster list, last valid hit wins) ==
ocsb_interp called once per raster in list order == the oracle's
EuclideanMesh2D.interpolate restatement (ocsmesh/mesh/mesh.py:365-376)."""
import numpy as np
import pytest

from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, synthetic as S

pytestmark = pytest.mark.gpu
METHOD = {"linear": _lib.INTERP_LINEAR, "nearest": _lib.INTERP_NEAREST,
          "spline": _lib.INTERP_SPLINE}


def make_list(n, h, w, mask_frac, dtype=np.float32):
    rasters = []
    for k in range(n):
        z = S.synth_dem(h, w, seed=60 + k).astype(dtype)
        tr = list(S.dem_transform(h))
        tr[2] += 0.6 * w * S.PIXEL * (k % 3)
        tr[5] += 0.5 * h * S.PIXEL * (k // 3)
        mask = None
        if mask_frac and k % 2 == 1:
            mask = np.random.default_rng(k).random((h, w)) < mask_frac
            z = z.copy()
            z[mask] = -99999.0
        rasters.append(O.OracleRaster(z, tr, mask=mask))
    return rasters


def entries_of(rasters, device):
    import torch
    keep, entries = [], []
    for r in rasters:
        z = np.ascontiguousarray(r.bands[0])
        h, w = z.shape
        m = None if r.mask is None else np.ascontiguousarray(r.mask[0]).astype(np.uint8)
        if device:
            z = torch.from_numpy(z).cuda()
            m = None if m is None else torch.from_numpy(m).cuda()
        keep.append((z, m))
        code = _lib.OCSB_F32 if r.bands.dtype == np.float32 else _lib.OCSB_F64
        win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h)
        entries.append((win, z, code, w, m, w))
    return entries, keep


@pytest.mark.parametrize("method", ["linear", "nearest"])
@pytest.mark.parametrize("device", [True, False])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_multi_equals_sequential_and_oracle(ctx, method, device, dtype):
    import torch
    rasters = make_list(5, 90, 110, 0.15, dtype)
    x0, y0, x1, y1 = rasters[0].get_window_bounds(rasters[0].full_window())
    xy = S.synth_nodes(40000, (x0 - 100, y0 - 100, x1 + 1500, y1 + 600), seed=9)
    init = np.full(len(xy), np.nan)
    ref = O.mesh_interpolate(xy, init, rasters, method=method)[:, 0]
    entries, keep = entries_of(rasters, device)
    K = len(xy)
    if device:
        xyd = torch.from_numpy(xy).cuda()
        vals = torch.from_numpy(init.copy()).cuda()
        touched = torch.zeros(K, dtype=torch.uint8, device="cuda")
        src = torch.full((K,), -1, dtype=torch.int32, device="cuda")
        ctx.interp_multi(entries, xyd, K, METHOD[method], touched, vals, src)
        ctx.sync()
        vals, touched, src = vals.cpu().numpy(), touched.cpu().numpy(), src.cpu().numpy()
    else:
        vals, touched = init.copy(), np.zeros(K, np.uint8)
        src = np.full(K, -1, np.int32)
        ctx.interp_multi(entries, xy, K, METHOD[method], touched, vals, src)
    nan = np.isnan(ref)
    np.testing.assert_array_equal(np.isnan(vals), nan)
    np.testing.assert_array_equal(vals[~nan], ref[~nan])
    np.testing.assert_array_equal(touched.astype(bool), ~nan)
    # source = index of the LAST raster in list order that covers the node
    exp_src = np.full(K, -1, np.int32)
    for k, r in enumerate(rasters):
        mask, _ = O.mesh_interpolate_worker(xy, r, None, method)
        exp_src[mask] = k
    np.testing.assert_array_equal(src, exp_src)


def test_multi_spline_and_errors(ctx):
    rasters = make_list(3, 70, 80, 0.0)
    x0, y0, x1, y1 = rasters[0].get_window_bounds(rasters[0].full_window())
    xy = S.synth_nodes(5000, (x0, y0, x1 + 900, y1), seed=2)
    init = np.full(len(xy), np.nan)
    ref = O.mesh_interpolate(xy, init, rasters, method="spline")[:, 0]
    entries, keep = entries_of(rasters, False)
    vals, touched = init.copy(), np.zeros(len(xy), np.uint8)
    ctx.interp_multi(entries, xy, len(xy), METHOD["spline"], touched, vals)
    nan = np.isnan(ref)
    np.testing.assert_array_equal(np.isnan(vals), nan)
    assert np.abs(vals[~nan] - ref[~nan]).max() < 1e-8
    with pytest.raises(_lib.OcsbError, match="Invalid value method"):
        ctx.interp_multi(entries, xy, len(xy), 9, touched, vals)
    # empty list / no nodes are no-ops
    ctx.interp_multi([], xy, len(xy), 0, touched, vals)
    ctx.interp_multi(entries, xy, 0, 0, touched, vals)


def test_bucketed_nodes_give_identical_results(ctx):
    """ocsb_bucket_nodes + perm: results are written at the original node
    index, bit-identical to the unbucketed call; perm is a permutation and the
    sorted coordinates are the permuted input."""
    import torch
    rasters = make_list(4, 120, 140, 0.1)
    boxes = np.array([r.get_window_bounds(r.full_window()) for r in rasters])
    bbox = (boxes[:, 0].min(), boxes[:, 1].min(), boxes[:, 2].max(), boxes[:, 3].max())
    xy = S.synth_nodes(300000, (bbox[0] - 300, bbox[1] - 300, bbox[2] + 300, bbox[3] + 300),
                       seed=5)
    xy[7] = np.nan                      # NaN coordinates stay untouched
    K = len(xy)
    entries, keep = entries_of(rasters, True)
    xyd = torch.from_numpy(xy).cuda()
    perm = torch.empty(K, dtype=torch.int32, device="cuda")
    xys = torch.empty_like(xyd)
    ctx.bucket_nodes(xyd, K, bbox, 64 * 10.0, perm, xys)
    ctx.sync()
    p = perm.cpu().numpy()
    np.testing.assert_array_equal(np.sort(p), np.arange(K))
    np.testing.assert_array_equal(xys.cpu().numpy(), xy[p])
    for method in ("linear", "nearest"):
        out = []
        for use_perm in (False, True):
            vals = torch.full((K,), float("nan"), dtype=torch.float64, device="cuda")
            touched = torch.zeros(K, dtype=torch.uint8, device="cuda")
            src = torch.full((K,), -1, dtype=torch.int32, device="cuda")
            if use_perm:
                ctx.interp_multi(entries, xys, K, METHOD[method], touched, vals, src, perm)
            else:
                ctx.interp_multi(entries, xyd, K, METHOD[method], touched, vals, src)
            ctx.sync()
            out.append((vals.cpu().numpy(), touched.cpu().numpy(), src.cpu().numpy()))
        np.testing.assert_array_equal(out[0][0], out[1][0])
        np.testing.assert_array_equal(out[0][1], out[1][1])
        np.testing.assert_array_equal(out[0][2], out[1][2])
    # host buffers through the C ABI
    ph = np.zeros(1000, np.int32)
    xh = np.zeros((1000, 2))
    ctx.bucket_nodes(xy[:1000].copy(), 1000, bbox, 640.0, ph, xh)
    np.testing.assert_array_equal(np.sort(ph), np.arange(1000))


def test_spline_coefficients_are_kept_with_the_raster(ctx):
    """Mesh.interpolate(method='spline') twice from the same rasters: the second
    call reuses the window coefficients computed by the first (ocsb_interp_raster
    .spline_coef) and gives the same values; dropping the device cache and
    interpolating again reproduces them too."""
    import ocsmesh_b200 as ob
    from ocsmesh_b200 import synthetic as S
    rasters = []
    for k in range(2):
        z = S.synth_dem(90, 110, seed=40 + k)
        tr = list(S.dem_transform(90))
        tr[2] += 600.0 * k
        rasters.append(ob.Raster.from_numpy(z, tr, crs=32618))
    bbox = (S.X0, S.Y0, S.X0 + 1700.0, S.Y0 + 900.0)
    xy = S.synth_nodes(5000, bbox, seed=9, outside_frac=0.05)
    out = []
    for rep in range(3):
        if rep == 2:
            for r in rasters:
                r.drop_device_cache()
        mesh = ob.Mesh(ob.MeshData(xy, values=np.full(len(xy), np.nan)))
        mesh.interpolate(rasters, method="spline")
        out.append(mesh.value.copy())
        keys = [k for k in rasters[0]._dev_cache if isinstance(k, tuple) and k[0] == "spline_coef"]
        assert len(keys) == 1
    np.testing.assert_array_equal(out[0], out[1])
    np.testing.assert_array_equal(out[0], out[2])
    assert np.isfinite(out[0]).sum() > 2000
