"""K3/K4 parity: ocsb_interp vs scipy RegularGridInterpolator /
RectBivariateSpline through the oracle (reference ocsmesh/mesh/mesh.py:2285-2347)."""
import numpy as np
import pytest

from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, synthetic as S

pytestmark = pytest.mark.gpu

METHOD = {"linear": _lib.INTERP_LINEAR, "nearest": _lib.INTERP_NEAREST,
          "spline": _lib.INTERP_SPLINE}


def make_raster(h, w, seed, dtype=np.float32, mask_frac=0.0):
    z = S.synth_dem(h, w, seed=seed).astype(dtype)
    mask = None
    if mask_frac:
        rng = np.random.default_rng(seed + 50)
        mask = rng.random((h, w)) < mask_frac
        z = z.copy()
        z[mask] = -99999.0
    return O.OracleRaster(z, S.dem_transform(h), mask=mask)


def special_nodes(r, n, seed):
    """random nodes + nodes on grid lines, bbox corners/edges, cell midpoints,
    and outside the raster"""
    x0, y0, x1, y1 = r.get_window_bounds(r.full_window())
    xs, ys = r.get_x(), r.get_y()
    rng = np.random.default_rng(seed)
    xy = S.synth_nodes(n, (x0, y0, x1, y1), seed=seed, outside_frac=0.05)
    extra = [(x0, y0), (x0, y1), (x1, y0), (x1, y1), (xs[3], ys[5]),
             (0.5 * (xs[3] + xs[4]), 0.5 * (ys[5] + ys[6])),
             (xs[0], ys[2]), (xs[-1], ys[-2]), (xs[7], y0), (xs[7], y1),
             (np.nextafter(x1, np.inf), ys[3]), (np.nextafter(x0, -np.inf), ys[3])]
    gi = rng.integers(0, len(xs), 200)
    gj = rng.integers(0, len(ys), 200)
    on_grid = np.stack([xs[gi], ys[gj]], axis=1)
    mid = np.stack([0.5 * (xs[gi[:100] % (len(xs) - 1)] + xs[gi[:100] % (len(xs) - 1) + 1]),
                    ys[gj[:100]]], axis=1)
    return np.concatenate([xy, np.array(extra), on_grid, mid])


def run_gpu(ctx, r, xy, method, values0, device=True, band=1):
    import torch
    z = np.ascontiguousarray(r.bands[band - 1])
    h, w = z.shape
    win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h)
    code = _lib.OCSB_F32 if z.dtype == np.float32 else _lib.OCSB_F64
    K = len(xy)
    mask = None if r.mask is None else np.ascontiguousarray(r.mask[band - 1]).astype(np.uint8)
    if device:
        zd = torch.from_numpy(z).cuda()
        md = None if mask is None else torch.from_numpy(mask).cuda()
        xyd = torch.from_numpy(np.ascontiguousarray(xy)).cuda()
        vd = torch.from_numpy(values0.copy()).cuda()
        td = torch.zeros(K, dtype=torch.uint8, device="cuda")
        ctx.interp(win, zd, code, w, md, w, xyd, K, METHOD[method], td, vd)
        ctx.sync()
        return td.cpu().numpy().astype(bool), vd.cpu().numpy()
    v = values0.copy()
    t = np.zeros(K, np.uint8)
    ctx.interp(win, z, code, w, mask, w, np.ascontiguousarray(xy), K,
               METHOD[method], t, v)
    return t.astype(bool), v


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("method", ["linear", "nearest"])
@pytest.mark.parametrize("mask_frac", [0.0, 0.1, 0.3])
def test_linear_nearest_bit_exact(ctx, dtype, method, mask_frac):
    r = make_raster(120, 150, 3, dtype, mask_frac)
    xy = special_nodes(r, 20000, 100)
    ref_mask, ref_vals = O.mesh_interpolate_worker(xy, r, None, method)
    v0 = np.full(len(xy), np.nan)
    t, v = run_gpu(ctx, r, xy, method, v0)
    np.testing.assert_array_equal(t, ref_mask)
    # bit-exact: nearest picks exactly the reference's texel; linear follows
    # scipy's operation order (tolerance in north_star: 1e-5 relative)
    np.testing.assert_array_equal(v[t], ref_vals.astype(np.float64))
    assert np.all(np.isnan(v[~t]))


def test_spline(ctx):
    r = make_raster(90, 140, 6)
    xy = special_nodes(r, 20000, 101)
    ref_mask, ref_vals = O.mesh_interpolate_worker(xy, r, None, "spline")
    v0 = np.full(len(xy), np.nan)
    t, v = run_gpu(ctx, r, xy, "spline", v0)
    np.testing.assert_array_equal(t, ref_mask)
    # elevations within 1e-5 relative (north_star); FITPACK solves the same
    # collocation system by QR, we by LU -> agreement ~1e-12
    err = np.abs(v[t] - ref_vals) / np.maximum(np.abs(ref_vals), 1.0)
    assert err.max() < 1e-9, err.max()


def test_spline_coefficients(ctx):
    from scipy.interpolate import RectBivariateSpline
    r = make_raster(37, 53, 8, np.float64)
    z = np.ascontiguousarray(r.bands[0])
    win = _lib.make_win(r.get_window_bounds(r.full_window()), 53, 37)
    coef = np.zeros(37 * 53)
    ctx.spline_prefilter(win, z, _lib.OCSB_F64, 53, coef)
    xi, yi = r.get_x(), r.get_y()
    f = RectBivariateSpline(xi, np.ascontiguousarray(np.flip(yi)),
                            np.flipud(z).T, kx=3, ky=3, s=0)
    np.testing.assert_allclose(coef, f.get_coeffs(), rtol=0, atol=1e-9)


def test_constant_raster_known_answer(ctx):
    # reference tests/api/test_mesh.py:361-379: constant rasters 1.0 / 2.0
    bands = np.stack([np.ones((11, 8)), np.full((11, 8), 2.0)])
    r = O.OracleRaster(bands, (1.0, 0, -3.0, 0, -1.0, 8.0))
    x0, y0, x1, y1 = r.get_window_bounds(r.full_window())
    xy = S.synth_nodes(500, (x0, y0, x1, y1), seed=1)
    for method in ("spline", "linear", "nearest"):
        for band, val in ((1, 1.0), (2, 2.0)):
            t, v = run_gpu(ctx, r, xy, method, np.zeros(len(xy)), band=band)
            assert t.all()
            assert np.allclose(v, val, rtol=0, atol=1e-12)


def test_host_buffers_and_last_wins(ctx):
    r1 = make_raster(100, 100, 1)
    z2 = S.synth_dem(100, 100, seed=2)
    tr2 = list(S.dem_transform(100))
    tr2[2] += 400.0   # shifted east by 40 px: overlaps r1
    r2 = O.OracleRaster(z2, tr2)
    x0, y0, x1, y1 = r1.get_window_bounds(r1.full_window())
    xy = S.synth_nodes(5000, (x0, y0, x1 + 400.0, y1), seed=3, outside_frac=0.02)
    init = np.full(len(xy), -7.0)
    ref = O.mesh_interpolate(xy, init, [r1, r2], method="linear")
    v = init.copy()
    for r in (r1, r2):
        t, v = run_gpu(ctx, r, xy, "linear", v, device=False)
    np.testing.assert_array_equal(v.reshape(-1, 1), ref)


def test_errors(ctx):
    r = make_raster(16, 16, 1)
    xy = np.zeros((4, 2))
    z = np.ascontiguousarray(r.bands[0])
    win = _lib.make_win(r.get_window_bounds(r.full_window()), 16, 16)
    with pytest.raises(_lib.OcsbError, match="Invalid value method"):
        ctx.interp(win, z, _lib.OCSB_F32, 16, None, 16, xy, 4, 7,
                   np.zeros(4, np.uint8), np.zeros(4))
