"""K2 parity: ocsb_hfun_flow_limiter vs the numpy oracle
(reference ocsmesh/hfun/raster.py:1396-1435)."""
import itertools

import numpy as np
import pytest

from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, synthetic as S

pytestmark = pytest.mark.gpu

FLT_MAX = np.finfo(np.float32).max


def run_gpu(ctx, z, old, dx, dy, hmin, hmax, lb, ub, ghmin, ghmax, band=(0, 0),
            device=True, fuse=False):
    import torch
    h, w = z.shape
    win = _lib.make_win((0.0, 0.0, float(w), float(h)), w, h, band[0], band[1])
    code = _lib.OCSB_F32 if z.dtype == np.float32 else _lib.OCSB_F64
    if device:
        zd = torch.from_numpy(z).cuda()
        od = torch.from_numpy(old).cuda()
        nd = od.clone()
        ctx.hfun_flow_limiter(win, zd, code, w, od, w, nd, w, dx, dy, hmin,
                              hmax, lb, ub, ghmin, ghmax, fuse)
        ctx.sync()
        return nd.cpu().numpy()
    new = old.copy()
    ctx.hfun_flow_limiter(win, z, code, w, old, w, new, w, dx, dy, hmin, hmax,
                          lb, ub, ghmin, ghmax, fuse)
    return new


def check(out, ref):
    # sizes within 1e-5 relative in fp32 (BASELINE.json north_star); the
    # kernel follows numpy's operation order so it is in fact bit-exact
    ref = np.ma.getdata(ref)
    assert out.dtype == np.float32
    np.testing.assert_array_equal(out, ref)


@pytest.mark.parametrize("shape", [(2, 2), (3, 5), (64, 128), (257, 515), (33, 1000)])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_shapes_dtypes(ctx, shape, dtype):
    rng = np.random.default_rng(3)
    z = (rng.normal(size=shape) * 20).astype(dtype)
    old = np.full(shape, FLT_MAX, np.float32)
    old[rng.random(shape) < 0.3] = 700.0
    ref = O.flow_limiter_window(z, old, 10.0, 12.5, 50.0, 1500.0, -30.0, 0.0,
                                100.0, 1500.0)
    out = run_gpu(ctx, z, old, 10.0, 12.5, 50.0, 1500.0, -30.0, 0.0, 100.0, 1500.0)
    check(out, ref)


def test_all_bound_combinations(ctx):
    rng = np.random.default_rng(11)
    z = (rng.normal(size=(96, 200)) * 15).astype(np.float32)
    z[10:20, 30:60] = 4.0                       # plateau: zero gradient
    old = (rng.random(z.shape) * 2000 + 1).astype(np.float32)
    opts = [(None, 50.0), (None, 1500.0), (None, -30.0), (None, 0.0),
            (None, 100.0), (None, 1200.0)]
    for hmin, hmax, lb, ub, ghmin, ghmax in itertools.product(*opts):
        if (hmin is None or hmin <= 1 / 3) and (ghmin is None or ghmin <= 1 / 3):
            # zero-gradient pixels stay masked in the reference: what rasterio
            # writes there is outside the oracle -> compare
            # unmasked pixels only
            ref = O.flow_limiter_window(z, old, 10.0, 10.0, hmin, hmax, lb, ub,
                                        ghmin, ghmax)
            out = run_gpu(ctx, z, old, 10.0, 10.0, hmin, hmax, lb, ub, ghmin, ghmax)
            m = np.ma.getmaskarray(ref)
            np.testing.assert_array_equal(out[~m], np.ma.getdata(ref)[~m])
            continue
        ref = O.flow_limiter_window(z, old, 10.0, 10.0, hmin, hmax, lb, ub,
                                    ghmin, ghmax)
        out = run_gpu(ctx, z, old, 10.0, 10.0, hmin, hmax, lb, ub, ghmin, ghmax)
        check(out, ref)


def test_plateau_gives_hmin(ctx):
    z = np.full((40, 70), -5.0, np.float32)
    old = np.full(z.shape, FLT_MAX, np.float32)
    out = run_gpu(ctx, z, old, 10.0, 10.0, 50.0, 1500.0, None, None, 100.0, 1500.0)
    assert np.all(out == 100.0)   # raw 1/3 -> hmin 50 -> global hmin 100


def test_host_buffers_and_bands(ctx):
    z = S.synth_dem(300, 421, seed=4)
    old = np.full(z.shape, FLT_MAX, np.float32)
    ref = np.ma.getdata(O.flow_limiter_window(z, old, 10.0, 10.0, 50.0, 1500.0,
                                              -30.0, 0.0, 100.0, 1500.0))
    out = run_gpu(ctx, z, old, 10.0, 10.0, 50.0, 1500.0, -30.0, 0.0, 100.0,
                  1500.0, device=False)
    np.testing.assert_array_equal(out, ref)
    # row bands (device tiles inside one semantic window) == one call
    for device in (True, False):
        acc = old.copy()
        for rb, re in [(0, 1), (1, 97), (97, 299), (299, 300)]:
            part = run_gpu(ctx, z, old, 10.0, 10.0, 50.0, 1500.0, -30.0, 0.0,
                           100.0, 1500.0, band=(rb, re), device=device)
            assert np.array_equal(part[:rb], old[:rb]) and np.array_equal(part[re:], old[re:])
            acc[rb:re] = part[rb:re]
        np.testing.assert_array_equal(acc, ref)


def test_fused_clamp_equals_clamp_pass(ctx):
    z = S.synth_dem(128, 256, seed=9)
    old = np.full(z.shape, FLT_MAX, np.float32)
    ref = np.ma.getdata(O.flow_limiter_window(z, old, 10.0, 10.0, None, None,
                                              -30.0, 0.0, 100.0, 1500.0))
    ref = O.apply_constraints_clamp(ref, 100.0, 1500.0)
    out = run_gpu(ctx, z, old, 10.0, 10.0, None, None, -30.0, 0.0, 100.0, 1500.0,
                  fuse=True)
    np.testing.assert_array_equal(out, ref)


def test_errors(ctx):
    z = np.zeros((1, 8), np.float32)
    with pytest.raises(_lib.OcsbError):
        run_gpu(ctx, z, z.copy(), 1.0, 1.0, None, None, None, None, None, None,
                device=False)


@pytest.mark.parametrize("dxdy", [(7.3, 30.0), (1.0 / 3.0, 1e-3), (10.0, 10.0),
                                  (123456.789, 0.7)])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_constant_division_is_ieee(ctx, dxdy, dtype):
    """The gradient divides by loop-invariant constants with Markstein's
    fma sequence; it must stay bit-identical to numpy's true division, also
    for zero / tiny / huge differences (slow-path operands)."""
    rng = np.random.default_rng(17)
    z = (rng.normal(size=(300, 700)) * 50).astype(dtype)
    z[:40] = 3.0                                   # zero differences
    z[40:80] *= dtype(1e-36 if dtype == np.float32 else 1e-300)   # tiny
    z[80:120] *= dtype(1e30 if dtype == np.float32 else 1e290)    # huge
    old = np.full(z.shape, FLT_MAX, np.float32)
    with np.errstate(all="ignore"):
        ref = O.flow_limiter_window(z, old, dxdy[0], dxdy[1], 50.0, 1500.0,
                                    None, None, 100.0, 1500.0)
    out = run_gpu(ctx, z, old, dxdy[0], dxdy[1], 50.0, 1500.0, None, None,
                  100.0, 1500.0)
    check(out, ref)
    # unclamped: the raw |z| / (3 |grad z|) values themselves
    with np.errstate(all="ignore"):
        ref = O.flow_limiter_window(z, old, dxdy[0], dxdy[1], None, None,
                                    None, None, None, None)
    out = run_gpu(ctx, z, old, dxdy[0], dxdy[1], None, None, None, None, None, None)
    m = np.ma.getmaskarray(ref)
    a, b = out[~m], np.ma.getdata(ref)[~m]
    np.testing.assert_array_equal(np.isnan(a), np.isnan(b))
    np.testing.assert_array_equal(a[~np.isnan(a)], b[~np.isnan(b)])


@pytest.mark.parametrize("fuse", [False, True])
@pytest.mark.parametrize("band", [(None, None), (-30.0, 0.0)])
def test_fast_path_fallbacks(ctx, fuse, band):
    """The float32 fast path (unguarded sqrt/div/Markstein sequences + one
    range test per 4-pixel group) must hand every group it cannot prove to the
    IEEE path: plateaus that do not line up with the 4-pixel groups, exact
    zeros on a slope, subnormal / tiny / huge elevations, nodata-like
    -3.4e38 blocks, all inside an otherwise ordinary DEM."""
    rng = np.random.default_rng(23)
    z = S.synth_dem(333, 1031, seed=6)
    z[20:41, 33:71] = 2.5                      # plateau, ragged w.r.t. groups
    z[60:64, 101:107] = 0.0                    # zeros next to a slope
    z[90, 200:260] = 0.0                       # zero row on a slope
    z[100:130, 300:350] *= np.float32(1e-36)   # subnormal differences
    z[140:170, 400:460] *= np.float32(1e-22)   # tiny but normal
    z[180:200, 500:540] *= np.float32(1e28)    # huge
    z[210:240, 600:650] = np.float32(-3.4028235e38)   # nodata-like block
    z[250:260, 700:710] = rng.normal(size=(10, 10)).astype(np.float32) * np.float32(1e19)
    old = (rng.random(z.shape) * 3000 + 1).astype(np.float32)
    old[rng.random(z.shape) < 0.5] = FLT_MAX
    lb, ub = band
    with np.errstate(all="ignore"):
        ref = O.flow_limiter_window(z, old, 10.0, 10.0, 50.0, 1500.0, lb, ub,
                                    100.0, 1500.0)
        ref = np.ma.getdata(ref)
        if fuse:
            ref = O.apply_constraints_clamp(ref, 100.0, 1500.0)
    out = run_gpu(ctx, z, old, 10.0, 10.0, 50.0, 1500.0, lb, ub, 100.0, 1500.0,
                  fuse=fuse)
    np.testing.assert_array_equal(out, ref)
    # raw values, no bounds at all (nothing hides a wrong bit)
    with np.errstate(all="ignore"):
        ref = O.flow_limiter_window(z, old, 10.0, 10.0, None, None, None, None,
                                    None, None)
    out = run_gpu(ctx, z, old, 10.0, 10.0, None, None, None, None, None, None)
    m = np.ma.getmaskarray(ref)
    a, b = out[~m], np.ma.getdata(ref)[~m]
    np.testing.assert_array_equal(np.isnan(a), np.isnan(b))
    np.testing.assert_array_equal(a[~np.isnan(a)], b[~np.isnan(b)])


@pytest.mark.parametrize("fuse", [False, True])
def test_null_old_is_initial_state(ctx, fuse):
    import torch
    z = S.synth_dem(301, 517, seed=8)
    old = np.full(z.shape, FLT_MAX, np.float32)
    h, w = z.shape
    for lb, ub in [(None, None), (-30.0, 0.0)]:
        ref = run_gpu(ctx, z, old, 10.0, 10.0, 50.0, 1500.0, lb, ub, 100.0,
                      1500.0, fuse=fuse)
        win = _lib.make_win((0.0, 0.0, float(w), float(h)), w, h, 0, 0)
        nd = torch.full((h, w), -7.0, dtype=torch.float32, device="cuda")
        ctx.hfun_flow_limiter(win, torch.from_numpy(z).cuda(), _lib.OCSB_F32, w,
                              None, w, nd, w, 10.0, 10.0, 50.0, 1500.0, lb, ub,
                              100.0, 1500.0, fuse)
        ctx.sync()
        np.testing.assert_array_equal(nd.cpu().numpy(), ref)
        new = np.full((h, w), -7.0, np.float32)
        ctx.hfun_flow_limiter(win, z, _lib.OCSB_F32, w, None, w, new, w, 10.0,
                              10.0, 50.0, 1500.0, lb, ub, 100.0, 1500.0, fuse)
        np.testing.assert_array_equal(new, ref)
