"""ocsb_resample_nearest (the nearest warp of
HfunCollector._create_big_raster, ocsmesh/hfun/collector.py:2419-2438) and
HfunCollector(method='fast') (collector.py:2296-2571) against the oracle.
Bit-exact: the warp copies float32 values, everything after it is the path
already pinned elsewhere.  The GDAL rule itself is a restatement (parity
unpinned, see k_warp.cu)."""
import numpy as np
import pytest

import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, synthetic as S
from oracle import ocsmesh_oracle as O

pytestmark = pytest.mark.gpu


def _call(ctx, src, strf, dst, dtr, mask=None):
    code = _lib.OCSB_F32 if src.dtype == np.float32 else _lib.OCSB_F64
    ctx.resample_nearest(src, code, src.shape[1], src.shape[1], src.shape[0], strf,
                         mask, src.shape[1], dst, dst.shape[1], dst.shape[1],
                         dst.shape[0], dtr)
    ctx.sync()


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("case", [
    # (src shape, src transform, dst shape, dst transform)
    ((300, 420), (10.0, 0, 500000.0, 0, -10.0, 4103000.0),
     (131, 190), (25.0, 0, 499987.5, 0, -25.0, 4103012.5)),       # coarser, hmin/2 grid
    ((257, 300), (7.0, 0, 100.0, 0, -7.0, 4000.0),
     (400, 333), (3.3, 0, -50.0, 0, -3.3, 4100.0)),               # finer, partial cover
    ((64, 64), (10.0, 0, 0.0, 0, -10.0, 640.0),
     (64, 64), (10.0, 0, 0.0, 0, -10.0, 640.0)),                  # identity
    ((50, 70), (10.0, 0, 5000.0, 0, -10.0, 640.0),
     (64, 64), (10.0, 0, 0.0, 0, -10.0, 640.0)),                  # disjoint
])
def test_resample_nearest_host_buffers(ctx, dtype, case):
    (sh, strf, dh, dtr) = case
    rng = np.random.default_rng(11)
    src = rng.normal(size=sh).astype(dtype)
    mask = (rng.random(sh) < 0.1).astype(np.uint8)
    for m in (None, mask):
        dst = np.full(dh, -99999, np.float32)
        want = dst.copy()
        O.resample_nearest(want, dtr, src, strf, None if m is None else m.astype(bool))
        _call(ctx, src, strf, dst, dtr, m)
        np.testing.assert_array_equal(dst, want)


def test_resample_nearest_device_buffers_last_wins(ctx):
    import torch
    rng = np.random.default_rng(12)
    a = rng.normal(size=(500, 700)).astype(np.float32)
    b = rng.normal(size=(450, 640)).astype(np.float32)
    ta = (10.0, 0, 1000.0, 0, -10.0, 9000.0)
    tb = (10.0, 0, 6000.0, 0, -10.0, 7500.0)
    dtr = (22.0, 0, 900.0, 0, -22.0, 9100.0)
    want = np.full((330, 560), -99999, np.float32)
    O.resample_nearest(want, dtr, a, ta)
    O.resample_nearest(want, dtr, b, tb)
    dst = torch.full((330, 560), -99999.0, dtype=torch.float32, device="cuda")
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    n0 = ctx.launch_count
    for s, t in ((da, ta), (db, tb)):
        ctx.resample_nearest(s, _lib.OCSB_F32, s.shape[1], s.shape[1], s.shape[0], t,
                             None, 0, dst, 560, 560, 330, dtr)
    ctx.sync()
    assert ctx.launch_count - n0 == 2
    np.testing.assert_array_equal(dst.cpu().numpy(), want)
    # strided destination window
    big = torch.full((400, 600), 7.0, dtype=torch.float32, device="cuda")
    sub = big[30:360, 20:580]
    ctx.resample_nearest(da, _lib.OCSB_F32, 700, 700, 500, ta, None, 0,
                         sub.data_ptr(), 600, 560, 330, dtr)
    ctx.sync()
    w2 = np.full((330, 560), 7.0, np.float32)
    O.resample_nearest(w2, dtr, a, ta)
    got = big.cpu().numpy()
    np.testing.assert_array_equal(got[30:360, 20:580], w2)
    assert (got[:30] == 7.0).all() and (got[:, :20] == 7.0).all() \
        and (got[360:] == 7.0).all() and (got[:, 580:] == 7.0).all()


def test_resample_nearest_argument_checks(ctx):
    src = np.zeros((4, 4), np.float32)
    dst = np.zeros((4, 4), np.float32)
    t = (1.0, 0, 0.0, 0, -1.0, 4.0)
    with pytest.raises(_lib.OcsbError, match="rotated"):
        _call(ctx, src, (1.0, 0.1, 0.0, 0, -1.0, 4.0), dst, t)
    with pytest.raises(_lib.OcsbError, match="degenerate"):
        _call(ctx, src, t, dst, (0.0, 0, 0.0, 0, -1.0, 4.0))
    with pytest.raises(_lib.OcsbError, match="dtype"):
        ctx.resample_nearest(src, 99, 4, 4, 4, t, None, 0, dst, 4, 4, 4, t)


def _inputs():
    """Two overlapping 10 m DEMs side by side, one float64 with nodata."""
    h, w = 260, 330
    za = S.synth_dem(h, w, seed=21)
    zb = S.synth_dem(h, w, seed=22).astype(np.float64)
    ta = S.dem_transform(h)
    tb = (ta[0], 0.0, ta[2] + 10.0 * (w - 40), 0.0, ta[4], ta[5] - 300.0)
    mb = np.zeros((h, w), bool)
    mb[100:130, 50:90] = True
    return (za, ta, None), (zb, tb, mb)


@pytest.mark.parametrize("nprocs", [64, 1])
def test_fast_collector_matches_oracle(ctx, nprocs):
    (za, ta, _), (zb, tb, mb) = _inputs()
    ga = ob.Raster.from_numpy(za, ta, crs=32618)
    gb = ob.Raster.from_numpy(zb, tb, crs=32618, nodata=-9999.0, mask=mb)
    coll = ob.Hfun([ga, gb], hmin=30.0, hmax=900.0, method='fast', nprocs=nprocs)
    assert coll.method == 'fast'
    coll.add_contour([0.0, -4.0], 0.002, 40.0)
    coll.add_subtidal_flow_limiter(hmin=35.0, hmax=800.0, lower_bound=-40.0,
                                   upper_bound=-1.0)
    coll.add_constant_value(120.0, lower_bound=2.0, upper_bound=6.0)
    coll.add_topo_bound_constraint(60.0, upper_bound=-20.0, value_type='min')
    big = coll.apply()
    assert coll.apply() is big                   # kept until the next add_*

    # oracle: the same pipeline on the restated big raster
    oa = O.OracleRaster(za, ta)
    orb = O.OracleRaster(zb, tb, mask=mb)
    obig = O.create_big_raster([oa, orb], 30.0, nprocs)
    assert tuple(big.raster.transform)[:6] == obig.transform
    assert (big.raster.height, big.raster.width) == (obig.height, obig.width)
    assert big.raster.chunk_size == obig.chunk_size
    if nprocs == 1:
        assert obig.chunk_size is not None       # the chunked branch is exercised
    np.testing.assert_array_equal(big.raster.get_values(band=1), obig.bands[0])
    oh = O.OracleHfun(obig, hmin=30.0, hmax=900.0)
    for lv in (0.0, -4.0):
        for src in (ga, gb):
            lines = src.get_contour(lv)
            if lines:
                oh.add_feature(lines, 0.002, 40.0)
                oh._clamp_pass()
    oh.add_subtidal_flow_limiter(35.0, 800.0, -40.0, -1.0)
    oh._clamp_pass()
    oh.add_constant_value(120.0, 2.0, 6.0)
    oh.add_topo_bound_constraint(60.0, upper_bound=-20.0, value_type='min')
    np.testing.assert_array_equal(big.get_values(), oh.values)
    # a new refinement invalidates the applied state
    coll.add_constant_value(33.0, upper_bound=-30.0)
    big2 = coll.apply()
    assert big2 is not big
    oh.add_constant_value(33.0, None, -30.0)
    np.testing.assert_array_equal(big2.get_values(), oh.values)


def test_fast_collector_contours_only_keeps_deep_ocean(ctx):
    (za, ta, _), (zb, tb, mb) = _inputs()
    ga = ob.Raster.from_numpy(za, ta, crs=32618)
    gb = ob.Raster.from_numpy(zb, tb, crs=32618, mask=mb)
    coll = ob.Hfun([ga, gb], hmin=30.0, hmax=900.0, method='fast')
    coll.add_contour(0.0, 0.002, 40.0)
    big = coll.apply()
    # nothing refers to the DEM values: no input is warped (collector.py:2410-2426)
    assert (big.raster.get_values(band=1) == -99999).all()
    obig = O.create_big_raster([O.OracleRaster(za, ta), O.OracleRaster(zb, tb, mask=mb)],
                               30.0, 64, warp=[False, False])
    oh = O.OracleHfun(obig, hmin=30.0, hmax=900.0)
    for src in (ga, gb):
        oh.add_feature(src.get_contour(0.0), 0.002, 40.0)
        oh._clamp_pass()
    np.testing.assert_array_equal(big.get_values(), oh.values)


def test_fast_collector_errors(ctx):
    (za, ta, _), _ = _inputs()
    ga = ob.Raster.from_numpy(za, ta, crs=32618)
    with pytest.raises(ValueError, match="Invalid method"):
        ob.Hfun([ga], hmin=30.0, method='quick')
    with pytest.raises(ValueError, match="fast"):
        ob.Hfun([ga], hmin=30.0).apply()
    gc = ob.Raster.from_numpy(za, ta, crs=32619)
    coll = ob.Hfun([ga, gc], hmin=30.0, method='fast')
    coll.add_constant_value(50.0)
    with pytest.raises(NotImplementedError, match="one projected CRS"):
        coll.apply()


def test_fast_collector_dispatches_every_kind_in_reference_order(ctx):
    """Line features, patches and the constraint kinds go to the ONE big hfun in
    the order of collector.py:2444-2495 (contours, limiters, constant values,
    line features, patches, constraints): same raster as applying the calls by
    hand to an HfunRaster on the same big raster (each op is pinned to the
    oracle in its own test file)."""
    (za, ta, _), (zb, tb, mb) = _inputs()
    ga = ob.Raster.from_numpy(za, ta, crs=32618)
    gb = ob.Raster.from_numpy(zb, tb, crs=32618, mask=mb)
    x0, ytop = ta[2], ta[5]
    line = np.array([[x0 + 300.0, ytop - 400.0], [x0 + 2500.0, ytop - 900.0],
                     [x0 + 4200.0, ytop - 2100.0]])
    ring = np.array([[x0 + 800.0, ytop - 600.0], [x0 + 2200.0, ytop - 650.0],
                     [x0 + 2000.0, ytop - 1900.0], [x0 + 900.0, ytop - 1700.0]])
    coll = ob.Hfun([ga, gb], hmin=30.0, hmax=900.0, method='fast', nprocs=64)
    # recorded in an order that differs from the order of application
    coll.add_region_constraint(70.0, [ring], crs=32618, value_type='max', rate=0.02)
    coll.add_patch([ring], target_size=45.0)
    coll.add_feature([line], 0.004, 50.0)
    coll.add_topo_bound_constraint(300.0, upper_bound=-10.0, value_type='max', rate=0.01)
    coll.add_constant_value(200.0, lower_bound=0.0)
    coll.add_gradient_delimiter(hmin=40.0, hmax=700.0, source_index=1)
    coll.add_contour(0.0, 0.003, 40.0)
    big = coll.apply()

    ref = ob.HfunRaster(big.raster, hmin=30.0, hmax=900.0)
    for src in (ga, gb):
        lines = src.get_contour(0.0)
        if lines:
            ref.add_feature(lines, 0.003, 40.0)
    ref.add_subtidal_flow_limiter(40.0, 700.0, -99990, None)
    ref.add_constant_value(200.0, 0.0, None)
    ref.add_feature([line], 0.004, 50.0)
    ref.add_patch([ring], None, 45.0)
    ref.add_region_constraint(70.0, [ring], crs=32618, value_type='max', rate=0.02)
    ref.add_topo_bound_constraint(300.0, upper_bound=-10.0, value_type='max', rate=0.01)
    np.testing.assert_array_equal(big.get_values(), ref.get_values())
    vals = big.get_values()
    assert (vals < 900.0).any() and len(np.unique(vals)) > 50
    # only raster 1 is referred to by index, but the constant value and the
    # constraints name no source: both inputs were warped
    dem = big.raster.get_values(band=1)
    X = big.raster.transform[2] + (np.arange(big.raster.width) + 0.5) * big.raster.transform[0]
    assert (dem[:, X < tb[2]] != -99999).any()
