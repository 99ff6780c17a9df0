"""the sampling side of HfunRaster.msh_t
(ocsmesh/hfun/raster.py:345-377) and utils.interpolate(method='nearest')
(ocsmesh/utils.py:1122-1140, scipy griddata over ALL source pixels) vs the
oracle."""
import numpy as np
import pytest

import ocsmesh_b200 as ob
from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def make(h, w, seed, hmin=80.0, hmax=900.0):
    z = S.synth_dem(h, w, seed=seed)
    tr = S.dem_transform(h)
    gr = ob.Raster.from_numpy(z, tr, crs=32618)
    orr = O.OracleRaster(z, tr)
    hf = ob.Hfun(gr, hmin=hmin, hmax=hmax)
    oh = O.OracleHfun(orr, hmin=hmin, hmax=hmax)
    lines = S.synth_shoreline(h, w, seed=seed, n_lines=2)
    hf.add_feature(lines, 0.02, 60.0)      # target < hmin: clamp is active
    oh.add_feature(lines, 0.02, 60.0)
    return gr, orr, hf, oh


@pytest.mark.parametrize("stride", [None, 1, 2, 3, 7])
@pytest.mark.parametrize("window", [None, (13, 9, 120, 77)])
def test_source_and_nearest(ctx, stride, window):
    gr, orr, hf, oh = make(150, 210, 4)
    win = None if window is None else ob.Window(*window)
    owin = O.Window(0, 0, 210, 150) if window is None else O.Window(*window)
    vals = oh.values[owin.row_off:owin.row_off + owin.height,
                     owin.col_off:owin.col_off + owin.width]
    xy = orr.get_xy(owin)
    # raw (unclamped) raster values differ from the clamp bounds somewhere
    src_c, src_v = O.msh_t_source(vals, xy, owin.height, owin.width, 80.0, 900.0, stride)
    c, v = hf.sizing_source(win, stride)
    np.testing.assert_array_equal(c.cpu().numpy(), src_c)
    np.testing.assert_array_equal(v.cpu().numpy(), src_v)
    assert v.dtype.itemsize == 8 and tuple(v.shape) == src_v.shape
    # nodes: random, outside the window, exactly on source points
    x0, y0, x1, y1 = orr.get_window_bounds(owin)
    nodes = S.synth_nodes(5000, (x0 - 40, y0 - 40, x1 + 40, y1 + 40), seed=9)
    nodes = np.concatenate([nodes, src_c[::97]])
    ref = O.interpolate_nearest(src_c, src_v, nodes)
    got = hf.sample_nearest(nodes, win, stride)
    assert got.dtype == np.float64
    bad = got != ref
    if bad.any():
        # only exact distance ties may differ (kd-tree traversal order)
        d = ((nodes[bad][:, None, :] - src_c[None, :, :]) ** 2).sum(-1)
        dmin = d.min(axis=1)
        assert np.all((d == dmin[:, None]).sum(axis=1) > 1)
    assert bad.mean() < 1e-3


def test_device_nodes_and_unclamped(ctx):
    import torch
    z = S.synth_dem(90, 130, seed=6)
    gr = ob.Raster.from_numpy(z, S.dem_transform(90), crs=32618)
    hf = ob.Hfun(gr)                       # no hmin / hmax
    hf.add_constant_value(55.5, -1.0, 3.0)
    vals = hf.get_values()
    orr = O.OracleRaster(z, S.dem_transform(90))
    xy = orr.get_xy(O.Window(0, 0, 130, 90))
    src_c, src_v = O.msh_t_source(vals, xy, 90, 130, None, None, 2)
    nodes = S.synth_nodes(3000, orr.get_window_bounds(O.Window(0, 0, 130, 90)), seed=3)
    got = hf.sample_nearest(torch.from_numpy(nodes).cuda(), None, 2)
    assert got.is_cuda
    ref = O.interpolate_nearest(src_c, src_v, nodes)
    np.testing.assert_array_equal(got.cpu().numpy(), ref)


def test_golden_reference_source(ctx):
    """Against utils.interpolate lifted from the reference
    (tests/golden/make_golden_constraints.py)."""
    from test_oracle_constraints import GOLD
    gold = np.load(GOLD)
    col, row, w, h = (int(v) for v in gold["msh_win"])
    r = ob.Raster.from_numpy(gold["dem"], tuple(gold["dem_transform"]), crs=32618)
    hf = ob.Hfun(r, hmin=80.0, hmax=900.0, initial_value=gold["msh_hfun"])
    for step in (1, 3):
        got = hf.sample_nearest(gold["msh_nodes"], ob.Window(col, row, w, h), step)
        ref = gold[f"msh_nearest_stride{step}"]
        assert (got != ref).mean() < 1e-3     # exact ties only
        np.testing.assert_allclose(got, ref, rtol=0, atol=900.0)
        assert np.count_nonzero(got != ref) <= 2
