"""Polygon regions on the raster (SURVEY.md 8(f) next-2 remainder):
RegionConstraint (ocsmesh/features/constraint.py:318-395) and
HfunRaster.add_patch (ocsmesh/hfun/raster.py:803-933) against the oracle's
restatement of the two predicates (even-odd `within` of the pixel coordinate,
all_touched rasterisation).  shapely / GDAL are not importable here: what is
checked is the device path against the SAME stated rule plus the reference's
surrounding arithmetic (value assignment, _apply_rate, clamps, min) -- parity of
the predicates themselves with shapely / GDAL is unpinned."""
import numpy as np
import pytest

import ocsmesh_b200 as ob
from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, features as F, synthetic as S

pytestmark = pytest.mark.gpu

H, W = 180, 240


def shapes(tr):
    x0, ytop = tr[2], tr[5]
    def ring(cx, cy, rx, ry, n, phase=0.0):
        t = np.linspace(0, 2 * np.pi, n, endpoint=False) + phase
        return np.stack([x0 + cx + rx * np.cos(t) * (1 + 0.25 * np.sin(3 * t)),
                         ytop - cy + ry * np.sin(t)], axis=1)
    outer = ring(1200.0, 900.0, 700.0, 500.0, 41)
    hole = ring(1250.0, 880.0, 200.0, 150.0, 17, 0.3)
    island = ring(300.0, 300.0, 130.0, 90.0, 9)           # a second polygon
    sliver = np.array([[x0 + 1900.0, ytop - 100.0], [x0 + 2350.0, ytop - 104.0],
                       [x0 + 2350.0, ytop - 101.0]])       # thinner than a pixel
    return [(outer, [hole]), island, sliver]


class Poly:                     # a shapely-style object
    class _R:
        def __init__(self, c): self.coords = c
    def __init__(self, ext, holes=()):
        self.exterior = self._R(ext)
        self.interiors = [self._R(h) for h in holes]


@pytest.mark.parametrize("touched", [False, True])
def test_region_mask_equals_oracle_rule(ctx, touched):
    import torch
    tr = S.dem_transform(H)
    r = O.OracleRaster(np.zeros((H, W), np.float32), tr)
    rings = F.polygon_rings(shapes(tr))
    assert len(rings) == 4 and all(np.array_equal(g[0], g[-1]) for g in rings)
    edges = F.ring_edges(rings)
    win = r.full_window()
    b = r.get_window_bounds(win)
    cw = _lib.make_win(b, W, H)
    out = torch.empty((H, W), dtype=torch.float32, device="cuda")
    ctx.region_mask(cw, edges, len(edges), touched, out, W)
    got = out.cpu().numpy() > 0.5
    if touched:
        want = O.touched_mask(b, H, W, rings)
        assert want[9:11].any()                  # the sliver is caught by its edges
    else:
        want = O.points_in_rings(r.get_xy(win), rings).reshape(H, W)
    np.testing.assert_array_equal(got, want)
    assert 0.05 < got.mean() < 0.6
    # a row band of the window gives the same rows
    band = torch.full((60, W), -1.0, dtype=torch.float32, device="cuda")
    cb = _lib.make_win(b, W, H, 50, 110)
    ctx.region_mask(cb, edges, len(edges), touched, band.data_ptr() - 50 * W * 4, W)
    np.testing.assert_array_equal(band.cpu().numpy() > 0.5, want[50:110])


@pytest.mark.parametrize("chunk,overlap", [(None, None), (100, 0)])
def test_add_patch(ctx, chunk, overlap):
    z = S.synth_dem(H, W, seed=6)
    tr = S.dem_transform(H)
    gr = ob.Raster.from_numpy(z, tr, crs=32618, chunk_size=chunk, overlap=overlap)
    orr = O.OracleRaster(z, tr, chunk_size=chunk, overlap=overlap)
    lines = S.synth_shoreline(H, W, seed=6, n_lines=1)
    shp = shapes(tr)
    rings = F.polygon_rings(shp)
    for rate in (None, 0.05):
        hf = ob.Hfun(gr, hmin=30.0, hmax=800.0)
        oh = O.OracleHfun(orr, hmin=30.0, hmax=800.0)
        hf.add_feature(lines, 0.01, 100.0)
        oh.add_feature(lines, 0.01, 100.0); oh._clamp_pass()
        hf.add_patch(Poly(shp[0][0], shp[0][1]) if rate is None else shp,
                     expansion_rate=rate, target_size=45.0)
        use = F.polygon_rings(Poly(shp[0][0], shp[0][1])) if rate is None else rings
        if rate is not None:
            oh.add_feature(use, rate, 45.0); oh._clamp_pass()
        b = orr.get_window_bounds(orr.full_window())
        mask = O.touched_mask(b, H, W, use)
        oh.values = O.add_patch_window(oh.values, mask, 45.0, 30.0, 800.0)
        oh._clamp_pass()
        np.testing.assert_array_equal(hf.get_values(), oh.values)
        assert (hf.get_values() == 45.0).sum() > 500
    with pytest.raises(TypeError):
        ob.Hfun(gr, hmin=30.0, hmax=800.0).add_patch("nope", target_size=45.0)
    with pytest.raises(ValueError, match="greater than zero"):
        ob.Hfun(gr, hmin=30.0, hmax=800.0).add_patch(shp, target_size=-1.0)


@pytest.mark.parametrize("value_type,value,rate", [("min", 300.0, 0.01), ("max", 60.0, 0.02),
                                                    ("min", 250.0, None)])
def test_region_constraint(ctx, value_type, value, rate):
    z = S.synth_dem(H, W, seed=9)
    tr = S.dem_transform(H)
    gr = ob.Raster.from_numpy(z, tr, crs=32618)
    orr = O.OracleRaster(z, tr)
    lines = S.synth_shoreline(H, W, seed=9, n_lines=2)
    shp = shapes(tr)[:2]
    rings = F.polygon_rings(shp)
    hf = ob.Hfun(gr, hmin=30.0, hmax=800.0)
    oh = O.OracleHfun(orr, hmin=30.0, hmax=800.0)
    hf.add_feature(lines, 0.01, 100.0)
    oh.add_feature(lines, 0.01, 100.0); oh._clamp_pass()
    with pytest.warns(UserWarning, match="No CRS"):
        hf.add_region_constraint(value, shp, value_type=value_type, rate=rate)
    oh.apply_constraints([O.RegionConstraint(value, rings, value_type, rate)])
    np.testing.assert_array_equal(hf.get_values(), oh.values)
    inside = O.points_in_rings(orr.get_xy(orr.full_window()), rings).reshape(H, W)
    v = hf.get_values()
    assert (v[inside] >= value).all() if value_type == "min" else (v[inside] <= value).all()
    # the constraint is re-applied after the next op, like every constraint
    hf.add_subtidal_flow_limiter(40.0, 700.0, None, None)
    oh.add_subtidal_flow_limiter(40.0, 700.0, None, None)
    oh.apply_constraints([O.RegionConstraint(value, rings, value_type, rate)])
    np.testing.assert_array_equal(hf.get_values(), oh.values)
    with pytest.raises(ValueError, match="Invalid input shape type"):
        hf.add_region_constraint(1.0, 12345, crs=32618)
