"""SURVEY.md 8(f) next-4, CPU side: the oracle's restatement of the fast
collector's big raster (ocsmesh/hfun/collector.py:2296-2442) against
hand-derived known answers.  GDAL is absent: the nearest-warp rule itself
(destination pixel centre -> floor of the source pixel coordinate) is a
restatement, parity unpinned."""
import numpy as np

from oracle import ocsmesh_oracle as O


def loop_resample(dst, dtr, src, strf, mask=None):
    """The rule, pixel by pixel."""
    out = dst.copy()
    for r in range(dst.shape[0]):
        for c in range(dst.shape[1]):
            X = dtr[2] + (c + 0.5) * dtr[0]
            Y = dtr[5] + (r + 0.5) * dtr[4]
            u = np.floor((X - strf[2]) / strf[0] + 1e-10)
            v = np.floor((Y - strf[5]) / strf[4] + 1e-10)
            if 0 <= u < src.shape[1] and 0 <= v < src.shape[0]:
                if mask is not None and mask[int(v), int(u)]:
                    continue
                out[r, c] = np.float32(src[int(v), int(u)])
    return out


def test_identity_and_integer_decimation():
    src = np.arange(48, dtype=np.float32).reshape(6, 8)
    tr = (10.0, 0.0, 500.0, 0.0, -10.0, 1060.0)
    dst = np.full((6, 8), -99999, np.float32)
    O.resample_nearest(dst, tr, src, tr)
    np.testing.assert_array_equal(dst, src)
    # 2x coarser grid with the same origin: the centre of coarse pixel (r, c)
    # sits on the corner shared by fine pixels (2r..2r+1, 2c..2c+1); floor(u +
    # 1e-10) picks the lower-right one, (2r + 1, 2c + 1)
    tr2 = (20.0, 0.0, 500.0, 0.0, -20.0, 1060.0)
    dst = np.full((3, 4), -99999, np.float32)
    O.resample_nearest(dst, tr2, src, tr)
    np.testing.assert_array_equal(dst, src[1::2, 1::2])


def test_partial_cover_mask_and_priority():
    rng = np.random.default_rng(3)
    a = rng.normal(size=(17, 23)).astype(np.float32)
    b = rng.normal(size=(11, 19))            # float64 source
    ta = (7.0, 0.0, 100.0, 0.0, -7.0, 400.0)
    tb = (5.0, 0.0, 180.0, 0.0, -5.0, 350.0)
    mb = rng.random(b.shape) < 0.2
    dtr = (9.5, 0.0, 80.0, 0.0, -9.5, 420.0)
    dst = np.full((20, 30), -99999, np.float32)
    want = loop_resample(loop_resample(dst, dtr, a, ta), dtr, b, tb, mb)
    O.resample_nearest(dst, dtr, a, ta)
    O.resample_nearest(dst, dtr, b, tb, mb)
    np.testing.assert_array_equal(dst, want)
    assert (dst == -99999).any() and (dst != -99999).any()
    # a source entirely outside the destination changes nothing
    before = dst.copy()
    O.resample_nearest(dst, dtr, a, (7.0, 0.0, 1e6, 0.0, -7.0, 400.0))
    np.testing.assert_array_equal(dst, before)


def test_big_raster_geometry_known_answer():
    # two 100 x 200 rasters of 10 m pixels side by side with a 20-px overlap,
    # hmin = 50: res = max(25, 10) = 25; union bbox 3800 m x 1000 m
    z = np.zeros((100, 200), np.float32)
    r0 = O.OracleRaster(z, (10.0, 0.0, 500000.0, 0.0, -10.0, 4101000.0))
    r1 = O.OracleRaster(z, (10.0, 0.0, 501800.0, 0.0, -10.0, 4101000.0))
    tr, w, h, ws = O.big_raster_geometry([r0, r1], 50.0, nprocs=3)
    assert (w, h) == (152, 40)
    # from_origin(x0 - res/2, y1 + res/2, res, res)      collector.py:2383
    assert tr == (25.0, 0.0, 500000.0 - 12.5, 0.0, -25.0, 4101000.0 + 12.5)
    # n_cell_lim = 20000 * 3 / 3 -> approx = 141 < 152: chunked, min(152, 40, 141)
    assert ws == 40
    tr, w, h, ws = O.big_raster_geometry([r0, r1], 50.0, nprocs=6)
    assert ws is None                      # approx = 200 >= 152
    # no global hmin: the worst source resolution
    tr, w, h, ws = O.big_raster_geometry([r0, r1], None, nprocs=6)
    assert tr[0] == 10.0 and (w, h) == (380, 100)


def test_create_big_raster_selection():
    rng = np.random.default_rng(5)
    za = rng.normal(size=(40, 60)).astype(np.float32)
    zb = rng.normal(size=(40, 60)).astype(np.float32)
    ra = O.OracleRaster(za, (10.0, 0.0, 0.0, 0.0, -10.0, 400.0))
    rb = O.OracleRaster(zb, (10.0, 0.0, 400.0, 0.0, -10.0, 400.0))
    big = O.create_big_raster([ra, rb], 30.0, nprocs=64)
    assert big.bands.dtype == np.float32 and big.chunk_size is None
    v = big.bands[0]
    assert not (v == -99999).all()
    only_b = O.create_big_raster([ra, rb], 30.0, nprocs=64, warp=[False, True]).bands[0]
    # columns whose centres lie left of raster b keep the deep-ocean fill
    X = big.transform[2] + (np.arange(big.width) + 0.5) * big.transform[0]
    assert (only_b[:, X < 400.0] == -99999).all()
    np.testing.assert_array_equal(only_b[:, X >= 400.0], v[:, X >= 400.0])


# ---- pinned to the reference's own source (tests/golden/make_golden_bigraster.py) ----
import json
import os
from types import SimpleNamespace

import pytest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden",
                                   "reference_golden_bigraster.json")))


@pytest.mark.parametrize("g", GOLD, ids=[g["name"] for g in GOLD])
def test_geometry_and_selection_equal_reference_source(g):
    """Outputs of HfunCollector._create_big_raster (collector.py:2296-2442) run
    from its lifted source: profile, transform, window size, fill, warped inputs."""
    rasters = [O.OracleRaster(np.zeros(r["shape"], np.float32), r["transform"])
               for r in g["rasters"]]
    tr, w, h, ws = O.big_raster_geometry(rasters, g["hmin"], g["nprocs"])
    assert (w, h) == (g["width"], g["height"])
    assert list(tr) == g["transform"]
    assert ws == g["chunk_size"]
    assert g["dtype"] == "float32" and g["count"] == 1
    assert all(fw["value"] == -99999.0 and fw["dtype"] == "float32" for fw in g["fill_writes"])
    sel = O.big_raster_selection(len(rasters), g["const_val_src"], g["flow_lim_src"],
                                 g["constraint_src"])
    # warped in list order (the last one wins on overlaps)
    assert [k for k, f in enumerate(sel) if f] == g["warped_in_order"]


@pytest.mark.parametrize("g", GOLD, ids=[g["name"] for g in GOLD])
def test_product_collector_logic_equals_reference_source(g):
    """The product's host logic (no device work in these two methods)."""
    from ocsmesh_b200.hfun import HfunCollector
    from ocsmesh_b200.raster import Affine
    coll = HfunCollector.__new__(HfunCollector)
    coll._hfuns = [SimpleNamespace(raster=SimpleNamespace(src=SimpleNamespace(
        shape=tuple(r["shape"]), transform=Affine(*r["transform"])))) for r in g["rasters"]]
    coll._nprocs, coll._hmin = g["nprocs"], g["hmin"]
    coll._const_val_info = [(s, None, None, 1.0) for s in g["const_val_src"]]
    coll._flow_lim_info = [(s, None, None, None, None) for s in g["flow_lim_src"]]
    coll._constraint_info = [(s, "add_topo_bound_constraint", (), {}) for s in g["constraint_src"]]
    tr, w, h, ws = coll._big_raster_geometry()
    assert (w, h, ws) == (g["width"], g["height"], g["chunk_size"])
    assert list(tr) == g["transform"]
    assert [k for k in range(len(g["rasters"])) if coll._needs_warp(k)] == g["warped_in_order"]
