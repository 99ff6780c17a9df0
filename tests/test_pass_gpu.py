"""ocsb_hfun_pass (one streamed pass over the op list, band by band with
H2D / kernels / D2H overlapped) must be bit-identical to the op-by-op calls,
which tests/test_feature_gpu.py and tests/test_limiter_gpu.py pin to the
oracle (reference ocsmesh/hfun/raster.py:1230-1257, 1396-1435)."""
import numpy as np
import pytest

import ocsmesh_b200 as ob
from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, features as F, synthetic as S

pytestmark = pytest.mark.gpu

FLT_MAX = np.finfo(np.float32).max


def make_inputs(h, w, seed, dtype=np.float32):
    z = S.synth_dem(h, w, seed=seed).astype(dtype)
    tr = S.dem_transform(h)
    r = O.OracleRaster(z, tr)
    lines = S.synth_shoreline(h, w, seed=seed, n_lines=2)
    pts = np.ascontiguousarray(F.feature_points(lines, 100.0))
    pts2 = np.ascontiguousarray(F.feature_points(
        S.synth_shoreline(h, w, seed=seed + 50, n_lines=1), 60.0))
    return z, r, pts, pts2


def op_by_op(ctx, z, r, ops, old, ghmin, ghmax, band=(0, 0)):
    """Reference sequence through the single-op entry points (host buffers)."""
    h, w = z.shape
    win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h, *band)
    code = _lib.OCSB_F32 if z.dtype == np.float32 else _lib.OCSB_F64
    cur = None if old is None else old.copy()
    new = np.full((h, w), -1.0, np.float32) if old is None else old.copy()
    for op in ops:
        if op[0] == "feature":
            _, pts, nv, rate, target, fuse = op
            ctx.hfun_feature(win, pts, nv, cur, w, new, w, rate, target, ghmin,
                             ghmax, fuse)
        else:
            _, hmin, hmax, lb, ub, fuse = op
            ctx.hfun_flow_limiter(win, z, code, w, cur, w, new, w, 10.0, 10.0,
                                  hmin, hmax, lb, ub, ghmin, ghmax, fuse)
        cur = new
    return new


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("band_rows", [0, 37, 64, 1000])
def test_pass_equals_op_by_op_host(ctx, dtype, band_rows):
    h, w = 301, 517
    z, r, pts, pts2 = make_inputs(h, w, 3, dtype)
    ops = [("feature", pts, len(pts), 0.001, 100.0, True),
           ("limiter", 50.0, 1500.0, -30.0, 0.0, True),
           ("feature", pts2, len(pts2), 0.01, 60.0, False),
           ("limiter", None, None, None, None, False)]
    ref = op_by_op(ctx, z, r, ops, None, 100.0, 1500.0)
    win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h, 0, 0)
    code = _lib.OCSB_F32 if dtype == np.float32 else _lib.OCSB_F64
    out = np.full((h, w), -1.0, np.float32)
    ctx.hfun_pass(win, z, code, w, None, w, out, w, 10.0, 10.0, 100.0, 1500.0,
                  ops, band_rows)
    np.testing.assert_array_equal(out, ref)


def test_pass_with_input_aliasing_and_strides(ctx):
    import torch
    h, w = 257, 300
    z, r, pts, _ = make_inputs(h, w, 5)
    rng = np.random.default_rng(0)
    old = (rng.random((h, w)) * 2000 + 1).astype(np.float32)
    ops = [("limiter", 50.0, 1500.0, None, None, False),
           ("feature", pts, len(pts), 0.001, 100.0, False)]
    ref = op_by_op(ctx, z, r, ops, old, None, 1500.0)
    win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h, 0, 0)
    # separate host in / out
    out = np.zeros((h, w), np.float32)
    ctx.hfun_pass(win, z, _lib.OCSB_F32, w, old, w, out, w, 10.0, 10.0, None,
                  1500.0, ops, 50)
    np.testing.assert_array_equal(out, ref)
    # aliased (in place), pinned
    buf = torch.from_numpy(old.copy()).pin_memory()
    ctx.hfun_pass(win, z, _lib.OCSB_F32, w, buf, w, buf, w, 10.0, 10.0, None,
                  1500.0, ops, 50)
    np.testing.assert_array_equal(buf.numpy(), ref)
    # padded rows (window inside a wider buffer), DEM on the device, out host
    W = w + 13
    zpad = np.zeros((h, W), np.float32); zpad[:, :w] = z
    opad = np.zeros((h, W), np.float32); opad[:, :w] = old
    res = np.full((h, W), -3.0, np.float32)
    zd = torch.from_numpy(zpad).cuda()
    ctx.hfun_pass(win, zd, _lib.OCSB_F32, W, opad, W, res, W, 10.0, 10.0, None,
                  1500.0, ops, 64)
    np.testing.assert_array_equal(res[:, :w], ref)
    assert np.all(res[:, w:] == -3.0)
    # everything on the device
    od = torch.from_numpy(old).cuda()
    nd = torch.empty_like(od)
    ctx.hfun_pass(win, zd, _lib.OCSB_F32, W, od, w, nd, w, 10.0, 10.0, None,
                  1500.0, ops, 0)
    ctx.sync()
    np.testing.assert_array_equal(nd.cpu().numpy(), ref)


def test_pass_row_band(ctx):
    h, w = 300, 260
    z, r, pts, _ = make_inputs(h, w, 8)
    ops = [("feature", pts, len(pts), 0.001, 100.0, True),
           ("limiter", 50.0, 1500.0, -30.0, 0.0, True)]
    ref = op_by_op(ctx, z, r, ops, None, 100.0, 1500.0)
    acc = np.full((h, w), -1.0, np.float32)
    for rb, re in [(0, 7), (7, 130), (130, 300)]:
        win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h, rb, re)
        ctx.hfun_pass(win, z, _lib.OCSB_F32, w, None, w, acc, w, 10.0, 10.0,
                      100.0, 1500.0, ops, 40)
    np.testing.assert_array_equal(acc, ref)


def test_pass_errors(ctx):
    h, w = 40, 40
    z, r, pts, _ = make_inputs(h, w, 1)
    win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h, 0, 0)
    out = np.zeros((h, w), np.float32)
    with pytest.raises(_lib.OcsbError):      # limiter without a DEM
        ctx.hfun_pass(win, None, _lib.OCSB_F32, w, None, w, out, w, 10.0, 10.0,
                      None, None, [("limiter", None, None, None, None, False)])
    with pytest.raises(_lib.OcsbError):      # target_size <= 0
        ctx.hfun_pass(win, z, _lib.OCSB_F32, w, None, w, out, w, 10.0, 10.0,
                      None, None, [("feature", pts, len(pts), 0.1, 0.0, False)])


@pytest.mark.parametrize("chunk", [None, 128])
def test_deferred_hfun_equals_eager(chunk):
    """Python API: HfunRaster(deferred=True) queues add_* and streams the pass
    into the caller's (pinned) buffer on get_values."""
    import torch
    h, w = 384, 512
    z = S.synth_dem(h, w, seed=4)
    tr = S.dem_transform(h)
    pts = F.feature_points(S.synth_shoreline(h, w, seed=4, n_lines=2), 100.0)

    def build(deferred):
        r = ob.Raster.from_numpy(z, tr, crs=32618, chunk_size=chunk)
        hf = ob.HfunRaster(r, hmin=100, hmax=1500, deferred=deferred)
        hf.add_feature_points(pts, 0.001, 100.0)
        hf.add_subtidal_flow_limiter(hmin=50, hmax=1500, lower_bound=-30,
                                     upper_bound=0)
        return hf

    ref = build(False).get_values()
    hd = build(True)
    out = torch.empty((h, w), dtype=torch.float32).pin_memory()
    got = hd.get_values(out=out)
    np.testing.assert_array_equal(got, ref)
    # the result can be looked at again and extended eagerly afterwards
    np.testing.assert_array_equal(hd.get_values(), ref)
    hd.add_gradient_delimiter(hmin=80, hmax=900)
    he = build(False)
    he.add_gradient_delimiter(hmin=80, hmax=900)
    np.testing.assert_array_equal(hd.get_values(), he.get_values())
    np.testing.assert_array_equal(hd.device_values.cpu().numpy(), he.get_values())


@pytest.mark.parametrize("ops_order", ["FL", "FLL", "FFL", "LF", "FLF"])
@pytest.mark.parametrize("pristine", [True, False])
def test_resident_pass_overlap_equals_eager(ops_order, pristine, monkeypatch):
    """Resident op lists that start with a feature op run the limiters first
    (on the context's stream) while the candidate pyramid is refined on a
    second stream: bit-identical to the eager, in-order calls."""
    monkeypatch.setenv("OCSB_PASS_OVERLAP", "1")      # opt-in knob, read per call
    h, w = 700, 900
    z = S.synth_dem(h, w, seed=14)
    tr = S.dem_transform(h)
    pa = F.feature_points(S.synth_shoreline(h, w, seed=14, n_lines=2), 100.0)
    pb = F.feature_points(S.synth_shoreline(h, w, seed=15, n_lines=1), 60.0)

    def build(deferred):
        r = ob.Raster.from_numpy(z, tr, crs=32618)
        r.device_band(1, 0)                        # resident DEM
        hf = ob.HfunRaster(r, hmin=60, hmax=1500, deferred=False)
        if not pristine:
            hf.add_constant_value(900.0, -2.0, 2.0)
            hf.device_values                       # materialised
        hf._deferred = deferred
        nf = nl = 0
        for o in ops_order:
            if o == "F":
                hf.add_feature_points(pa if nf == 0 else pb, 0.002, 100.0 if nf == 0 else 60.0)
                nf += 1
            else:
                if nl == 0:
                    hf.add_gradient_delimiter(hmin=80, hmax=1200)
                else:
                    hf.add_subtidal_flow_limiter(hmin=70, hmax=1000, lower_bound=-20,
                                                 upper_bound=1)
                nl += 1
        hf.flush()
        return hf.device_values.cpu().numpy()

    got = build(True)
    monkeypatch.setenv("OCSB_PASS_OVERLAP", "0")
    np.testing.assert_array_equal(got, build(False))


def _windows_case(ctx, h, w, chunk, seed, ops_of, n_reps=1):
    """ocsb_hfun_pass_windows over the window list of a chunked raster vs the
    same op list applied window by window through ocsb_hfun_pass."""
    import torch
    z, r, pts, pts2 = make_inputs(h, w, seed)
    wins_py = list(ob.get_iter_windows(w, h, chunk, 0))
    bounds = lambda wn: O.OracleRaster(z, S.dem_transform(h)).get_window_bounds(wn)
    zd = torch.from_numpy(z).cuda()
    ptd, pt2d = torch.from_numpy(pts).cuda(), torch.from_numpy(pts2).cuda()
    ops_dev = ops_of(ptd, pt2d)
    ops_host = ops_of(pts, pts2)
    cws = [_lib.make_win(bounds(wn), wn.width, wn.height) for wn in wins_py]
    arr = (_lib.ocsb_win * len(cws))(*cws)
    off = np.array([wn.row_off * w + wn.col_off for wn in wins_py], dtype=np.int64)
    # reference: one synchronous single-window pass per window (host buffers)
    ref = np.full((h, w), -1.0, np.float32)
    for wn, cw in zip(wins_py, cws):
        sub = ref[wn.row_off:, wn.col_off:]
        ctx.hfun_pass(cw, z[wn.row_off:, wn.col_off:], _lib.OCSB_F32, w, None, w, sub, w,
                      10.0, 10.0, 100.0, 1500.0, ops_host)
    return z, zd, arr, off, ops_dev, ref


@pytest.mark.parametrize("chunk", [64, 150, 257])
def test_pass_windows_equals_per_window_passes(ctx, chunk):
    import torch
    h, w = 300, 517
    mk = lambda a, b: [("feature", a, len(a), 0.001, 100.0, True),
                       ("limiter", 50.0, 1500.0, -30.0, 0.0, True)]
    z, zd, arr, off, ops, ref = _windows_case(ctx, h, w, chunk, 5, mk)
    out = torch.full((h, w), -1.0, dtype=torch.float32, device="cuda")
    ctx.hfun_pass_windows(arr, off, off, zd, _lib.OCSB_F32, w, h, w, None, w, out, w,
                          10.0, 10.0, 100.0, 1500.0, ops)
    np.testing.assert_array_equal(out.cpu().numpy(), ref)
    # the op list the tiled sweep does not take (two limiters with different
    # clamp fusing) runs window by window inside the same call
    mk2 = lambda a, b: [("feature", a, len(a), 0.001, 100.0, True),
                        ("limiter", 50.0, 1500.0, -30.0, 0.0, False)]
    z, zd, arr, off, ops, ref = _windows_case(ctx, h, w, chunk, 5, mk2)
    out.fill_(-1.0)
    ctx.hfun_pass_windows(arr, off, off, zd, _lib.OCSB_F32, w, h, w, None, w, out, w,
                          10.0, 10.0, 100.0, 1500.0, ops)
    np.testing.assert_array_equal(out.cpu().numpy(), ref)


def test_pass_windows_async_status_is_sticky(ctx):
    """Asynchronous mode: with the status word clear the pass runs and leaves it
    clear; with the status word set (an earlier pass of the sequence failed) the
    pass must not touch the raster, and the word stays set."""
    import torch
    h, w = 200, 260
    mk = lambda a, b: [("feature", a, len(a), 0.001, 100.0, True),
                       ("limiter", 50.0, 1500.0, None, None, True)]
    z, zd, arr, off, ops, ref = _windows_case(ctx, h, w, 128, 7, mk)
    status = torch.zeros(2, dtype=torch.int32, device="cuda")
    out = torch.full((h, w), -1.0, dtype=torch.float32, device="cuda")
    ctx.hfun_pass_windows(arr, off, off, zd, _lib.OCSB_F32, w, h, w, None, w, out, w,
                          10.0, 10.0, 100.0, 1500.0, ops, status=status)
    torch.cuda.synchronize()
    assert status.tolist() == [0, 0]
    np.testing.assert_array_equal(out.cpu().numpy(), ref)
    status[0] = 1
    out.fill_(-2.0)
    for these in (ops, ops[1:]):            # with and without a feature op
        ctx.hfun_pass_windows(arr, off, off, zd, _lib.OCSB_F32, w, h, w, None, w, out, w,
                              10.0, 10.0, 100.0, 1500.0, these, status=status)
        torch.cuda.synchronize()
        assert int(status[0]) == 1
        assert bool((out == -2.0).all())


def test_deferred_hfun_async_flush_and_verify(ctx):
    """HfunRaster(deferred=True) on a resident DEM: flush() only enqueues; the
    values are final at get_values(); a set status word (simulated pool
    overflow) makes _verify() redo the sequence synchronously."""
    h, w = 240, 300
    z = S.synth_dem(h, w, seed=4)
    tr = S.dem_transform(h)
    r = ob.Raster.from_numpy(z, tr, crs=32618)
    r.device_band(1, 0)
    lines = S.synth_shoreline(h, w, seed=4, n_lines=2)
    pts = F.feature_points(lines, 100.0)
    eager = ob.HfunRaster(r, hmin=100.0, hmax=1500.0)
    eager.add_feature_points(pts, 0.001, 100.0)
    eager.add_subtidal_flow_limiter(50.0, 1500.0, -30.0, 0.0)
    want = eager.get_values()
    hf = ob.HfunRaster(r, hmin=100.0, hmax=1500.0, deferred=True)
    for fail in (False, True):
        hf.reset()
        hf.add_feature_points(pts, 0.001, 100.0)
        hf.add_subtidal_flow_limiter(50.0, 1500.0, -30.0, 0.0)
        hf.flush()
        assert len(hf._pending) == 1
        if fail:
            # as if a pool had overflowed: the device raster is then stale
            hf._status[0] = 1
            hf._hfun.fill_(-3.0)
        np.testing.assert_array_equal(hf.get_values(), want)
        assert not hf._pending and int(hf._status[0]) == 0
