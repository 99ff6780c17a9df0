"""Split form of a (feature, limiter) pass (k_pass.cu: limiter sweep on a
second stream beside the candidate pyramid, then the feature op in place on
the tiles that ended with candidates): bit-identical to the op-by-op calls and
to the oracle (reference ocsmesh/hfun/raster.py:1230-1257, 1396-1435), and
really taken where its condition holds."""
import numpy as np
import pytest

import ocsmesh_b200 as ob
from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, features as F, synthetic as S

pytestmark = pytest.mark.gpu


def _run(z, tr, pts, chunk, deferred, pristine, limiter_kw, hmin=100.0, hmax=1500.0,
         rate=0.001, target=100.0):
    r = ob.Raster.from_numpy(z, tr, crs=32618, chunk_size=chunk)
    r.device_band(1, 0)
    hf = ob.HfunRaster(r, hmin=hmin, hmax=hmax, deferred=False)
    if not pristine:
        hf.add_constant_value(900.0, -2.0, 2.0)
        hf.device_values
    hf._deferred = deferred
    hf.add_feature_points(pts, rate, target)
    hf.add_subtidal_flow_limiter(**limiter_kw)
    ctx = _lib.default_context(0)
    ctx.reset_timing()
    ctx.set_timing(True)
    try:
        hf.flush() if deferred else None
        vals = hf.device_values.cpu().numpy()
        classes = set(ctx.get_timing())
    finally:
        ctx.set_timing(False)
        ctx.reset_timing()
    return vals, classes


@pytest.mark.parametrize("chunk", [None, 200])
@pytest.mark.parametrize("pristine", [True, False])
@pytest.mark.parametrize("limiter_kw", [
    dict(hmin=50.0, hmax=1500.0, lower_bound=-30.0, upper_bound=0.0),
    dict(hmin=80.0, hmax=1200.0),
])
def test_split_pass_equals_eager(chunk, pristine, limiter_kw):
    h, w = 700, 900
    z = S.synth_dem(h, w, seed=31)
    tr = S.dem_transform(h)
    # a short feature: most tiles end without candidates (radius 1400 px is the
    # whole raster here, so shrink it through hmax)
    pts = F.feature_points(S.synth_shoreline(h, w, seed=31, n_lines=1), 100.0)[:40]
    kw = dict(hmin=100.0, hmax=160.0, rate=0.01, target=100.0)   # r = 60 m = 6 px
    got, classes = _run(z, tr, pts, chunk, True, pristine, limiter_kw, **kw)
    ref, _ = _run(z, tr, pts, chunk, False, pristine, limiter_kw, **kw)
    np.testing.assert_array_equal(got, ref)
    assert {"pass_limiter", "pass_feature"} <= classes, classes
    assert "pass_tile" not in classes


def test_split_pass_equals_oracle_and_skips_nothing_it_should_not():
    h, w = 640, 832
    z = S.synth_dem(h, w, seed=33)
    tr = S.dem_transform(h)
    lines = S.synth_shoreline(h, w, seed=33, n_lines=2)
    pts = F.feature_points(lines, 100.0)
    lim = dict(hmin=50.0, hmax=1500.0, lower_bound=-30.0, upper_bound=0.0)
    got, classes = _run(z, tr, pts, None, True, True, lim, hmax=300.0, rate=0.002)
    assert "pass_feature" in classes
    oh = O.OracleHfun(O.OracleRaster(z, tr), hmin=100.0, hmax=300.0)
    oh.add_feature_points(pts, 0.002, 100.0)
    oh._clamp_pass()
    oh.add_subtidal_flow_limiter(50.0, 1500.0, -30.0, 0.0)
    np.testing.assert_array_equal(got, oh.values)


def test_split_with_unbounded_radius():
    """No global hmax: the search radius is unbounded, every tile keeps
    candidates (the list holds all tiles) and the miss value is +inf."""
    h, w = 300, 400
    z = S.synth_dem(h, w, seed=35)
    tr = S.dem_transform(h)
    pts = F.feature_points(S.synth_shoreline(h, w, seed=35, n_lines=1), 100.0)
    lim = dict(hmin=50.0, hmax=1500.0)
    got, classes = _run(z, tr, pts, None, True, True, lim, hmax=None)
    ref, _ = _run(z, tr, pts, None, False, True, lim, hmax=None)
    np.testing.assert_array_equal(got, ref)
    assert "pass_feature" in classes
