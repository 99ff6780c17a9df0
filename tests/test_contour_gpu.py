"""ocsb_contour_segments (marching-squares sweep on the
device) vs the host sweep of ocsmesh_b200.features / the oracle's independent
marching squares (restating ocsmesh/raster.py:1426-1470)."""
import numpy as np
import pytest

import ocsmesh_b200 as ob
from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, features as F, synthetic as S

pytestmark = pytest.mark.gpu


def canon(lines):
    """Order-free fingerprint of a set of polylines: sorted segment set."""
    segs = []
    for ln in lines:
        a, b = ln[:-1], ln[1:]
        s = np.concatenate([a, b], axis=1)
        swap = (s[:, 0] > s[:, 2]) | ((s[:, 0] == s[:, 2]) & (s[:, 1] > s[:, 3]))
        s[swap] = s[swap][:, [2, 3, 0, 1]]
        segs.append(s)
    s = np.concatenate(segs) if segs else np.empty((0, 4))
    return s[np.lexsort(s.T[::-1])]


@pytest.mark.parametrize("shape", [(2, 2), (3, 7), (64, 64), (257, 515), (1000, 33)])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("level", [0.0, -2.5])
def test_segments_match_host_sweep(ctx, shape, dtype, level):
    h, w = shape
    z = S.synth_dem(max(h, 8), max(w, 8), seed=h + w)[:h, :w].astype(dtype)
    z[0, 0] = level              # a value exactly on the level: not "above"
    tr = S.dem_transform(h)
    r = ob.Raster.from_numpy(z, tr, crs=32618)
    got = r.get_contour(level)
    ref = F.marching_squares(r.get_x(), r.get_y(), z, level)
    assert len(got) == len(ref)
    for p, q in zip(got, ref):
        np.testing.assert_array_equal(p, q)       # vertex for vertex, bit-exact


def test_saddles_and_plateaus(ctx):
    rng = np.random.default_rng(5)
    z = rng.integers(-1, 2, size=(120, 150)).astype(np.float32)   # many saddles, ties
    r = ob.Raster.from_numpy(z, S.dem_transform(120), crs=32618)
    for level in (0.0, 0.5, -0.5):
        got = r.get_contour(level)
        ref = F.marching_squares(r.get_x(), r.get_y(), z, level)
        assert len(got) == len(ref)
        for p, q in zip(got, ref):
            np.testing.assert_array_equal(p, q)


def test_against_oracle_marching_squares(ctx):
    """Independent restatement in oracle/: same segment set (stitch order and
    ring start points are implementation choices)."""
    z = S.synth_dem(300, 420, seed=21)
    r = ob.Raster.from_numpy(z, S.dem_transform(300), crs=32618)
    got = r.get_contour(0.0)
    ref = O.marching_squares(r.get_x(), r.get_y(), z, 0.0)
    a, b = canon(got), canon(ref)
    assert a.shape == b.shape
    np.testing.assert_allclose(a, b, rtol=0, atol=1e-9)


def test_window_and_band(ctx):
    z = S.synth_dem(200, 260, seed=8)
    r = ob.Raster.from_numpy(z, S.dem_transform(200), crs=32618)
    win = ob.Window(30, 20, 100, 90)
    got = r.get_contour(0.0, window=win)
    ref = F.marching_squares(r.get_x(win), r.get_y(win),
                             z[20:110, 30:130], 0.0)
    assert len(got) == len(ref)
    for p, q in zip(got, ref):
        np.testing.assert_array_equal(p, q)
    # raw ABI with a row band: the segments of cell rows [40, 120) only
    import torch
    zd = torch.from_numpy(z).cuda()
    full = _lib.make_win(r.get_window_bounds(ob.Window(0, 0, 260, 200)), 260, 200)
    e_all, p_all = ctx.contour_segments(full, zd, _lib.OCSB_F32, 260, 0.0)
    band = _lib.make_win(r.get_window_bounds(ob.Window(0, 0, 260, 200)), 260, 200, 40, 120)
    e_b, p_b = ctx.contour_segments(band, zd, _lib.OCSB_F32, 260, 0.0)
    # cell row of a segment: from its horizontal / vertical edge ids
    def cell_row(e):
        top = np.where(e >= 0, e // 259, (-e - 1) // 260)
        return top
    rows = np.minimum(cell_row(e_all[:, 0]), cell_row(e_all[:, 1]))
    sel = (rows >= 40) & (rows < 120)
    np.testing.assert_array_equal(e_b, e_all[sel])
    np.testing.assert_array_equal(p_b, p_all[sel])


def test_overflow_and_count_only(ctx):
    import ctypes as C
    z = S.synth_dem(64, 64, seed=3)
    win = _lib.make_win((0.0, 0.0, 640.0, 640.0), 64, 64)
    n = C.c_int64(0)
    lib = ctx.lib
    assert lib.ocsb_contour_segments(ctx.handle, C.byref(win), z.ctypes.data,
                                     _lib.OCSB_F32, 64, 0.0, 0, None, None,
                                     C.byref(n)) == 0
    assert n.value > 1
    e = np.empty((1, 2), np.int64)
    p = np.empty((1, 4))
    rc = lib.ocsb_contour_segments(ctx.handle, C.byref(win), z.ctypes.data,
                                   _lib.OCSB_F32, 64, 0.0, 1, e.ctypes.data,
                                   p.ctypes.data, C.byref(n))
    assert rc != 0 and b"capacity" in lib.ocsb_last_error()


@pytest.mark.parametrize("shape", [(64, 64), (257, 515), (600, 800)])
def test_device_stitch_equals_host_stitch(ctx, shape):
    """ocsb_contour_lines (list ranking on the device) vs the host stitch of the
    same device segments: identical polylines, order and orientation included."""
    h, w = shape
    z = S.synth_dem(h, w, seed=h + 3 * w)
    r = ob.Raster.from_numpy(z, S.dem_transform(h), crs=32618)
    for level in (0.0, 3.0, -7.5):
        dev = r.get_contour(level)
        host = r.get_contour(level, stitch="host")
        assert len(dev) == len(host)
        for p, q in zip(dev, host):
            np.testing.assert_array_equal(p, q)


@pytest.mark.parametrize("target,max_verts", [(100.0, 200), (35.0, 7), (5000.0, 200), (0.7, 200), (12.0, 400)])
def test_contour_points_equal_host_prologue(ctx, target, max_verts):
    """ocsb_contour_points == feature_points(get_contour(l) for l in levels):
    repartition into max_verts chunks + arc-length stations, bit for bit."""
    z = S.synth_dem(500, 640, seed=11)
    r = ob.Raster.from_numpy(z, S.dem_transform(500), crs=32618)
    levels = [-5.0, 0.0, 5.0]
    lines = []
    for lv in levels:
        lines.extend(r.get_contour(lv, stitch="host"))
    want = F.feature_points(lines, target, max_verts=max_verts, native=False)
    got = r.get_contour_points(levels, target, max_verts).cpu().numpy()
    np.testing.assert_array_equal(got, want)
    # a level without any contour contributes nothing; only such levels: empty
    assert r.get_contour_points([1e6], target).shape == (0, 2)
    got2 = r.get_contour_points([1e6, 0.0], target, max_verts).cpu().numpy()
    want2 = F.feature_points(r.get_contour(0.0, stitch="host"), target,
                             max_verts=max_verts, native=False)
    np.testing.assert_array_equal(got2, want2)


def test_rings_and_tiny_lines(ctx):
    """Closed rings (bumps inside the raster), two-vertex lines at corners and
    many saddles: the device ranking must order and orient them like the host."""
    g = np.arange(40.0)
    zz = (np.sin(g[:, None] * 0.9) * np.cos(g[None, :] * 0.7)).astype(np.float32)
    zz[0, 0] = 1.0; zz[0, 1] = -1.0; zz[1, 0] = -1.0      # corner cut: 2-vertex line
    r = ob.Raster.from_numpy(zz, S.dem_transform(40), crs=32618)
    for level in (0.0, 0.4, -0.8):
        dev = r.get_contour(level)
        host = F.marching_squares(r.get_x(), r.get_y(), zz, level)
        assert len(dev) == len(host) and len(dev) > 3
        assert any(np.array_equal(p[0], p[-1]) for p in dev)       # rings present
        for p, q in zip(dev, host):
            np.testing.assert_array_equal(p, q)


def test_chunked_raster_traces_window_by_window(ctx):
    """Reference Raster.get_contour on a chunked raster
    (_get_raster_contour_feathered, ocsmesh/raster.py:1471-1575): every window
    is traced on its own coordinates, the pieces are concatenated."""
    z = S.synth_dem(300, 420, seed=17)
    r = ob.Raster.from_numpy(z, S.dem_transform(300), crs=32618, chunk_size=128, overlap=2)
    wins = list(r.iter_windows())
    assert len(wins) > 4
    got = r.get_contour(0.0)
    want = []
    for wn in wins:
        sub = z[wn.row_off:wn.row_off + wn.height, wn.col_off:wn.col_off + wn.width]
        want.extend(F.marching_squares(r.get_x(wn), r.get_y(wn), sub, 0.0))
    assert len(got) == len(want)
    for p, q in zip(got, want):
        np.testing.assert_array_equal(p, q)
    pts = r.get_contour_points([0.0, 5.0], 100.0).cpu().numpy()
    lines = r.get_contour(0.0) + r.get_contour(5.0)
    np.testing.assert_array_equal(pts, F.feature_points(lines, 100.0, native=False))
