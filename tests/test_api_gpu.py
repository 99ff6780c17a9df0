"""Drop-in API parity: ocsmesh_b200.{Raster,Hfun,Mesh} vs the oracle's
restatement of HfunRaster / EuclideanMesh2D.interpolate on identical inputs
(BASELINE.json config 1 scaled down, chunked windows, overlap, multi-raster)."""
import numpy as np
import pytest

import ocsmesh_b200 as ob
from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def both(h, w, seed, chunk_size=None, overlap=None):
    z = S.synth_dem(h, w, seed=seed)
    tr = S.dem_transform(h)
    gr = ob.Raster.from_numpy(z, tr, crs=32618, chunk_size=chunk_size, overlap=overlap)
    orr = O.OracleRaster(z, tr, chunk_size=chunk_size, overlap=overlap)
    return gr, orr


@pytest.mark.parametrize("chunk,overlap", [(None, None), (200, 0), (256, 2), (150, 7)])
def test_config1_contour_limiter(ctx, chunk, overlap):
    gr, orr = both(500, 620, 1, chunk, overlap)
    hf = ob.Hfun(gr, hmin=100, hmax=1500)
    oh = O.OracleHfun(orr, hmin=100, hmax=1500)
    assert np.all(hf.get_values() == np.finfo(np.float32).max)
    hf.add_contour(level=0, expansion_rate=0.001, target_size=100)
    # parity boundary = the resampled vertex array (hfun/raster.py:1230): the
    # oracle consumes the contours the product traced
    oh.add_feature(gr.get_contour(0), 0.001, 100)
    oh._clamp_pass()
    np.testing.assert_array_equal(hf.get_values(), oh.values)
    hf.add_subtidal_flow_limiter(hmin=50, hmax=1500, lower_bound=-30, upper_bound=0)
    oh.add_subtidal_flow_limiter(50, 1500, -30, 0)
    np.testing.assert_array_equal(hf.get_values(), oh.values)
    hf.add_gradient_delimiter(hmin=50, hmax=1500)
    oh.add_subtidal_flow_limiter(50, 1500, None, None)
    np.testing.assert_array_equal(hf.get_values(), oh.values)
    hf.add_contour(level=[-5, 5], expansion_rate=0.002)     # target = hmin
    oh.add_feature(gr.get_contour(-5) + gr.get_contour(5), 0.002)
    oh._clamp_pass()
    np.testing.assert_array_equal(hf.get_values(), oh.values)


def test_no_contour_returns_silently(ctx):
    gr, _ = both(64, 64, 2)
    hf = ob.Hfun(gr, hmin=100, hmax=1500)
    hf.add_contour(level=1e6, expansion_rate=0.001, target_size=100)
    assert np.all(hf.get_values() == np.finfo(np.float32).max)


def test_hold_constraints(ctx):
    gr, orr = both(200, 200, 3)
    hf = ob.Hfun(gr, hmin=100, hmax=1500)
    lines = S.synth_shoreline(200, 200, seed=3, n_lines=1)
    with hf.hold_applying_added_constratins():
        hf.add_feature(lines, 0.001, 100)
        # clamp not applied yet: untouched pixels... add_feature itself clamps
        # its own values to [hmin, hmax]; the global re-pass is deferred
    oh = O.OracleHfun(orr, hmin=100, hmax=1500)
    oh.add_feature(lines, 0.001, 100)
    np.testing.assert_array_equal(hf.get_values(), oh.values)


def test_errors_match_reference(ctx):
    gr, _ = both(64, 64, 2)
    with pytest.raises(TypeError):
        ob.Hfun("not a raster")
    hf = ob.Hfun(gr)
    line = S.synth_shoreline(64, 64, seed=1, n_lines=1)
    with pytest.raises(ValueError, match="target_size must be specified"):
        hf.add_feature(line, 0.001)
    with pytest.raises(ValueError, match="greater than zero"):
        hf.add_feature(line, 0.001, target_size=0)
    with pytest.raises(TypeError):
        hf.add_feature(42, 0.001, 100)
    with pytest.raises(ValueError):
        ob.Raster.from_numpy(np.zeros((4, 4)), S.dem_transform(4), chunk_size=-1)
    mesh = ob.Mesh(ob.MeshData(np.zeros((3, 2))))
    with pytest.raises(ValueError, match="Invalid value method"):
        mesh.interpolate(gr, method="cubic")


@pytest.mark.parametrize("method", ["linear", "nearest", "spline"])
def test_mesh_interpolate_multi_raster(ctx, method):
    rasters, orasters = [], []
    for k in range(3):
        z = S.synth_dem(120, 160, seed=20 + k)
        tr = list(S.dem_transform(120))
        tr[2] += 1000.0 * k           # 100 px east each: 60 px overlap
        rasters.append(ob.Raster.from_numpy(z, tr, crs=32618))
        orasters.append(O.OracleRaster(z, tr))
    x0, y0, _, y1 = rasters[0].get_bbox()
    x1 = rasters[-1].get_bbox()[2]
    xy = S.synth_nodes(30000, (x0, y0, x1, y1), seed=400, outside_frac=0.02)
    init = np.full(len(xy), np.nan)
    mesh = ob.Mesh(ob.MeshData(xy, values=init, crs=32618))
    mesh.interpolate(rasters, method=method, nprocs=4)
    ref = O.mesh_interpolate(xy, init, orasters, method=method)
    assert mesh.value.shape == (len(xy), 1)
    nan = np.isnan(ref)
    np.testing.assert_array_equal(np.isnan(mesh.value), nan)
    if method == "spline":
        err = np.abs(mesh.value[~nan] - ref[~nan]) / np.maximum(np.abs(ref[~nan]), 1.0)
        assert err.max() < 1e-9
    else:
        np.testing.assert_array_equal(mesh.value[~nan], ref[~nan])


def test_collector_min_combine(ctx):
    # config 3 (scaled down): overlapping rasters, every contour to every
    # raster, overlap strips min-combined
    hs, ohs, zs = [], [], []
    for k in range(2):
        z = S.synth_dem(150, 200, seed=10 + k, x0=S.X0 + 1500.0 * k)
        tr = list(S.dem_transform(150))
        tr[2] += 1500.0 * k           # 150 px east: 50 px overlap
        hs.append(ob.Raster.from_numpy(z, tr, crs=32618))
        ohs.append(O.OracleHfun(O.OracleRaster(z, tr), 100, 1500))
    coll = ob.Hfun(hs, hmin=100, hmax=1500)
    coll.add_contour(0, 0.001, 100)
    contours = []
    for h in coll.hfuns:
        contours.extend(h.raster.get_contour(0))
    for oh in ohs:
        oh.add_feature(contours, 0.001, 100)
    for h, oh in zip(coll.hfuns, ohs):
        np.testing.assert_array_equal(h.get_values(), oh.values)
    a, b = coll.hfuns
    wa, wb = ob.Window(150, 0, 50, 150), ob.Window(0, 0, 50, 150)
    ref = O.min_combine(ohs[0].values[:, 150:], ohs[1].values[:, :50])
    a.min_combine(b, wa, wb)
    np.testing.assert_array_equal(a.get_values()[:, 150:], ref)


def test_config3_eight_rasters_min_combined(ctx):
    """BASELINE.json configs[2] scaled down: 8 overlapping rasters (4 x 2,
    corners covered by four), every contour to every raster, limiter, then the
    overlaps min-combined: each raster must equal its view of the mosaic min of
    the oracle's per-raster results."""
    th, tw, ovl, nx, ny = 130, 160, 30, 4, 2
    rs, ohs, offs = [], [], []
    for k in range(nx * ny):
        ty, tx = divmod(k, nx)
        x0 = S.X0 + (tw - ovl) * S.PIXEL * tx
        y0 = S.Y0 - (th - ovl) * S.PIXEL * ty
        z = S.synth_dem(th, tw, seed=10 + k, x0=x0, y0=y0)
        tr = S.dem_transform(th, x0=x0, y0=y0)
        rs.append(ob.Raster.from_numpy(z, tr, crs=32618))
        ohs.append(O.OracleHfun(O.OracleRaster(z, tr), 100, 1500))
        offs.append(((th - ovl) * ty, (tw - ovl) * tx))
    coll = ob.Hfun(rs, hmin=100, hmax=1500)
    coll.add_contour(0, 0.001, 100)
    coll.add_subtidal_flow_limiter(hmin=50, hmax=1500, lower_bound=-30, upper_bound=0)
    contours = []
    for h in coll.hfuns:
        contours.extend(h.raster.get_contour(0))
    assert contours
    for oh in ohs:
        oh.add_feature(contours, 0.001, 100)
        oh._clamp_pass()
        oh.add_subtidal_flow_limiter(50, 1500, -30, 0)
    assert coll.min_combine_overlaps() == 16     # 6 E-W + 4 N-S + 6 diagonal pairs
    H, W = (th - ovl) * (ny - 1) + th, (tw - ovl) * (nx - 1) + tw
    G = np.full((H, W), np.inf, dtype=np.float32)
    for (r, c), oh in zip(offs, ohs):
        G[r:r + th, c:c + tw] = O.min_combine(G[r:r + th, c:c + tw], oh.values)
    for (r, c), h in zip(offs, coll.hfuns):
        np.testing.assert_array_equal(h.get_values(), G[r:r + th, c:c + tw])
    # idempotent
    coll.min_combine_overlaps()
    for (r, c), h in zip(offs, coll.hfuns):
        np.testing.assert_array_equal(h.get_values(), G[r:r + th, c:c + tw])


@pytest.mark.parametrize("chunk,overlap", [(128, 0), (300, 5)])
def test_many_windows_one_library_call(ctx, chunk, overlap):
    """The window loop of add_feature / add_subtidal_flow_limiter as one
    ocsb_hfun_op_windows call (windows spread over the library's worker
    streams; in list order when they overlap): 99 windows of a 1100 x 1300
    raster against the oracle's sequential window loop, every pixel."""
    gr, orr = both(1100, 1300, 3, chunk, overlap)
    nwin = len(list(gr.iter_windows()))
    assert nwin >= 16
    hf = ob.Hfun(gr, hmin=100, hmax=1500)
    oh = O.OracleHfun(orr, hmin=100, hmax=1500)
    lines = gr.get_contour(0)
    for _ in range(2):          # pristine first op, then a read-modify-write op
        hf.add_feature(lines, 0.001, 100)
        oh.add_feature(lines, 0.001, 100)
        np.testing.assert_array_equal(hf.get_values(), oh.values)
        lines = gr.get_contour(5)
    hf.add_subtidal_flow_limiter(hmin=50, hmax=1500, lower_bound=-30, upper_bound=0)
    oh.add_subtidal_flow_limiter(50, 1500, -30, 0)
    np.testing.assert_array_equal(hf.get_values(), oh.values)


def test_op_windows_argument_checks(ctx):
    """ocsb_hfun_op_windows: resident buffers only, empty list is a no-op,
    bad ops are refused with the reference's messages."""
    import torch
    from ocsmesh_b200 import _lib
    gr, _ = both(300, 320, 4, 100, 0)
    W, H = gr.width, gr.height
    wins_l = list(gr.iter_windows())
    wins = (_lib.ocsb_win * len(wins_l))(*[
        _lib.make_win(gr.get_window_bounds(w), w.width, w.height) for w in wins_l])
    off = np.array([w.row_off * W + w.col_off for w in wins_l], dtype=np.int64)
    dev = torch.full((H, W), 3.0e38, dtype=torch.float32, device="cuda")
    host = np.zeros((H, W), np.float32)
    pts = torch.tensor([[S.X0 + 500.0, S.Y0 + 700.0]], dtype=torch.float64, device="cuda")
    op = ("feature", pts, 1, 0.001, 100.0, True)
    with pytest.raises(_lib.OcsbError, match="device memory"):
        ctx.hfun_op_windows(wins, None, off, None, _lib.OCSB_F32, W, None, W, host, W,
                            10.0, 10.0, 100.0, 1500.0, op)
    with pytest.raises(_lib.OcsbError, match="greater than zero"):
        ctx.hfun_op_windows(wins, None, off, None, _lib.OCSB_F32, W, None, W, dev, W,
                            10.0, 10.0, 100.0, 1500.0,
                            ("feature", pts, 1, 0.001, 0.0, True))
    with pytest.raises(_lib.OcsbError, match="device DEM"):
        ctx.hfun_op_windows(wins, None, off, None, _lib.OCSB_F32, W, None, W, dev, W,
                            10.0, 10.0, 100.0, 1500.0,
                            ("limiter", 50.0, 1500.0, None, None, True))
    before = dev.clone()
    empty = (_lib.ocsb_win * 0)()
    ctx.hfun_op_windows(empty, None, off[:0], None, _lib.OCSB_F32, W, None, W, dev, W,
                        10.0, 10.0, 100.0, 1500.0, op)
    assert torch.equal(dev, before)
    # and the real thing equals the per-window calls
    ctx.hfun_op_windows(wins, None, off, None, _lib.OCSB_F32, W, None, W, dev, W,
                        10.0, 10.0, 100.0, 1500.0, op)
    ref = torch.empty_like(dev)
    for w, cw in zip(wins_l, wins):
        ctx.hfun_feature(cw, pts, 1, None, W,
                         ref.data_ptr() + (w.row_off * W + w.col_off) * 4, W,
                         0.001, 100.0, 100.0, 1500.0, fuse_clamp=True)
    ctx.sync()
    assert torch.equal(dev, ref)
