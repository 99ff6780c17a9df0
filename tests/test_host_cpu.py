"""Host-side logic of the product package (no GPU): the C-ABI library loads and
exports every symbol include/ocsb200.h declares, window / coordinate / prologue
code matches the reference-generated golden vectors, GeoTIFF round trips, and
the package fails loudly without a CUDA device."""
import os
import re

import numpy as np
import pytest

import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, features as F, geotiff, synthetic as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def header_functions():
    text = open(os.path.join(ROOT, "include", "ocsb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ocsb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(lib):
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"libocsb200.so does not export {n}"
    # ... and the ctypes binding covers the whole header
    assert sorted(_lib.SIGNATURES) == names
    assert lib.ocsb_abi_version() == 5
    assert lib.ocsb_kernel_class_name(6) == b"feature_eval"


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_lib.OcsbError, match="no CUDA device|CUDA"):
        _lib.Context(0)
    with pytest.raises(_lib.OcsbError, match="no CPU fallback"):
        _lib.default_context()
    r = ob.Raster.from_numpy(np.zeros((8, 8), np.float32), S.dem_transform(8))
    with pytest.raises((_lib.OcsbError, RuntimeError, AssertionError)):
        ob.Hfun(r, hmin=1, hmax=2)


def test_windows_from_c_library_match_reference(lib, gold):
    k = 0
    while f"win_{k}" in gold.files:
        w, h, c, o = (int(v) for v in gold[f"win_{k}_args"])
        np.testing.assert_array_equal(_lib.windows(w, h, c, o), gold[f"win_{k}"])
        got = [tuple(x) for x in ob.get_iter_windows(w, h, c, o)]
        assert got == [tuple(x) for x in gold[f"win_{k}"]]
        k += 1
    # chunk_size 0 -> one full window; negative -> error (raster.py:194-198)
    np.testing.assert_array_equal(_lib.windows(7, 5, 0, 0), [[0, 0, 7, 5]])
    with pytest.raises(_lib.OcsbError, match="chunk_size must be >= 0"):
        _lib.windows(7, 5, -1, 0)
    with pytest.raises(ValueError):
        list(ob.get_iter_windows(7, 5, -1))


def test_raster_coordinates_match_reference(lib, gold):
    r = ob.Raster.from_numpy(gold["dem"], gold["dem_transform"], crs=32618)
    np.testing.assert_array_equal(r.get_xy(), gold["xy_full"])
    win = ob.Window(*(int(v) for v in gold["xy_win_args"]))
    np.testing.assert_array_equal(r.get_x(win), gold["x_win"])
    np.testing.assert_array_equal(r.get_y(win), gold["y_win"])
    np.testing.assert_array_equal(r.get_window_bounds(win), gold["bounds_win"])
    assert (r.dx, r.dy) == (10.0, 10.0)
    assert r.shape == gold["dem"].shape
    wins = list(r.iter_windows(chunk_size=50, overlap=3))
    assert wins[0] == ob.Window(0, 0, 53, 53)
    for w_, b in ob.Raster.from_numpy(gold["dem"], gold["dem_transform"], chunk_size=50):
        assert b == r.get_window_bounds(w_)


def test_feature_prologue_matches_reference(gold):
    line = gold["pro_line"]
    chunks = F.repartition_features(line, 200)
    np.testing.assert_array_equal([len(c) for c in chunks], gold["pro_chunk_lens"])
    np.testing.assert_array_equal(F.feature_points([line], 100.0, 200), gold["pro_points"])
    np.testing.assert_array_equal(F.feature_points([line], 100.0, 200, native=False),
                                  gold["pro_points"])
    np.testing.assert_array_equal(F.transform_linestring(line[:3], 1e6), gold["pro_short"])
    # reference tests/api/test_utils.py:873-916
    res = F.repartition_features(np.array([[i, i] for i in range(100000)], float), 200)
    assert (len(res), len(res[0]), len(res[-1])) == (503, 200, 102)
    with pytest.raises(ValueError, match="must be an integer"):
        F.repartition_features(line, 2.5)


def test_native_feature_points_bit_identical_to_numpy_walk():
    """ocsb_feature_points (host code of the library) against the numpy
    restatement of utils.py:247-312 on ragged random polylines: single
    vertices, lines shorter than target_size, chunk cuts at every max_verts."""
    rng = np.random.default_rng(3)
    for _ in range(60):
        lines = []
        for _ in range(int(rng.integers(1, 6))):
            m = int(rng.integers(1, 700))
            step = rng.choice([0.5, 7.0, 40.0, 350.0])
            lines.append(np.cumsum(rng.normal(0, step, (m, 2)), axis=0)
                         + [500000.0, 4100000.0])
        tgt = float(rng.choice([1.0, 33.3, 100.0, 1000.0]))
        mv = int(rng.choice([2, 3, 50, 200]))
        np.testing.assert_array_equal(F.feature_points(lines, tgt, mv, native=True),
                                      F.feature_points(lines, tgt, mv, native=False))
    assert F.feature_points([], 100.0).shape == (0, 2)
    with pytest.raises(ValueError, match="must be an integer"):
        F.feature_points([np.zeros((3, 2))], 100.0, 2.5)
    with pytest.raises(ValueError, match="greater than zero"):
        F.feature_points([np.zeros((3, 2))], 0.0)


def test_feature_argument_types():
    class LS:
        def __init__(self, c):
            self.coords = c

    class MLS:
        def __init__(self, ls):
            self.geoms = ls
    a = np.array([[0.0, 0.0], [1.0, 2.0], [3.0, 4.0]])
    assert len(F.as_linestrings(a)) == 1
    assert len(F.as_linestrings([a, a])) == 2
    assert len(F.as_linestrings(MLS([LS(a.tolist()), LS(a.tolist())]))) == 2
    np.testing.assert_array_equal(F.as_linestrings(LS(a.tolist()))[0], a)
    for bad in (42, "line", np.zeros(5)):
        with pytest.raises(TypeError):      # hfun/raster.py:1155-1158
            F.as_linestrings(bad)


def test_marching_squares_closed_and_open_lines():
    x = np.linspace(0, 10, 41)
    y = np.linspace(10, 0, 41)
    X, Y = np.meshgrid(x, y)
    z = (X - 5) ** 2 + (Y - 5) ** 2 - 9.0          # circle of radius 3
    lines = F.marching_squares(x, y, z, 0.0, block_rows=7)
    assert len(lines) == 1
    ring = lines[0]
    np.testing.assert_array_equal(ring[0], ring[-1])
    rad = np.hypot(ring[:, 0] - 5, ring[:, 1] - 5)
    assert np.all(np.abs(rad - 3.0) < 0.02)
    z2 = Y - 0.3 * X - 2.0                          # a straight open line
    l2 = F.marching_squares(x, y, z2, 0.0)
    assert len(l2) == 1 and len(l2[0]) > 40
    assert np.allclose(l2[0][:, 1], 0.3 * l2[0][:, 0] + 2.0, atol=1e-9)
    assert F.marching_squares(x, y, z, 1e9) == []


def test_geotiff_round_trip(tmp_path):
    rng = np.random.default_rng(3)
    for dtype in (np.float32, np.float64, np.int16):
        data = (rng.normal(size=(2, 37, 53)) * 100).astype(dtype)
        p = tmp_path / f"t_{np.dtype(dtype).name}.tif"
        geotiff.write_geotiff(p, data, (10.0, 0, 5e5, 0, -10.0, 4.1e6), epsg=32618,
                              nodata=-99999.0)
        info, bands = geotiff.read_geotiff(p)
        assert (info.width, info.height, info.count) == (53, 37, 2)
        assert info.epsg == 32618 and not info.is_geographic
        assert info.nodata == -99999.0
        assert info.transform == (10.0, 0.0, 5e5, 0.0, -10.0, 4.1e6)
        np.testing.assert_array_equal(bands, data)
    # file-backed Raster + raster_from_numpy (ocsmesh/utils.py:1932-1971) with
    # the reference's mgrid known answer (tests/api/test_utils.py:674-700)
    mg = np.mgrid[1:3:0.1, -1:1:0.1]
    z = rng.random(mg[0].shape)
    p = tmp_path / "mg.tif"
    assert ob.raster_from_numpy(p, z, mg, 4326) is None
    r = ob.Raster(p)
    assert np.all(np.isclose(mg.transpose([2, 1, 0]).reshape(-1, 2), r.get_xy()))
    assert np.all(z == r.values)
    assert r.crs == ob.CRS.from_epsg(4326) and r.crs.is_geographic


def test_masked_raster_and_bands(tmp_path):
    z = np.arange(48, dtype=np.float32).reshape(6, 8)
    zm = np.ma.MaskedArray(z, mask=z % 5 == 0, fill_value=-99999.0)
    mg = np.mgrid[0:8:1.0, 0:6:1.0]
    p = tmp_path / "m.tif"
    ob.raster_from_numpy(p, zm, mg, 32618)
    r = ob.Raster(p)
    got = r.get_values(masked=True)
    np.testing.assert_array_equal(np.ma.getmaskarray(got), zm.mask)
    assert r.nodata == -99999.0


def test_meshdata_and_errors():
    md = ob.MeshData(np.zeros((5, 2)), crs=32618)
    assert md.values.shape == (5, 1) and md.num_nodes == 5
    md.values = np.arange(5)
    assert md.values.shape == (5, 1)
    with pytest.raises(ValueError):
        ob.MeshData(np.zeros((5, 3)))
    with pytest.raises(TypeError):
        ob.Mesh("nope")
    with pytest.raises(TypeError):
        ob.Hfun(42)
    with pytest.raises(ValueError):
        ob.Raster.from_numpy(np.zeros((4, 4)), S.dem_transform(4), chunk_size=-3)


def test_synthetic_dem_row_blocks_are_consistent():
    full = S.synth_dem(300, 200, seed=5)
    part = S.synth_dem(300, 200, seed=5, rows=(120, 180))
    np.testing.assert_array_equal(full[120:180], part)
    g0, g1 = np.gradient(full, 10.0, 10.0)
    assert not np.any((g0 == 0) & (g1 == 0))


def test_stitch_is_order_independent():
    """features._stitch canonicalises the segment order: the device emits
    segments in cell order, the host block sweep by case."""
    rng = np.random.default_rng(0)
    z = S.synth_dem(90, 120, seed=12)
    x = np.linspace(0.0, 1190.0, 120)
    y = np.linspace(890.0, 0.0, 90)
    ref = F.marching_squares(x, y, z, 0.0)
    ea, eb, ids, px, py = F._ms_block(x, y, z, 0.0, 0, 120)
    lut = {int(i): k for k, i in enumerate(ids)}
    perm = rng.permutation(len(ea))
    flip = rng.random(len(ea)) < 0.5
    a, b = ea[perm], eb[perm]
    a, b = np.where(flip, b, a), np.where(flip, a, b)
    edges = np.stack([a, b], axis=1)
    ka = np.array([lut[int(i)] for i in a])
    kb = np.array([lut[int(i)] for i in b])
    pts = np.stack([px[ka], py[ka], px[kb], py[kb]], axis=1)
    for native in (False, True):     # numpy walk and the library's host code
        got = F.stitch_segments(edges, pts, native=native)
        assert len(got) == len(ref) > 0
        for p, q in zip(got, ref):
            np.testing.assert_array_equal(p, q)
    assert F.stitch_segments(np.empty((0, 2), np.int64), np.empty((0, 4))) == []


def test_topo_const_constraint_value_shapes():
    """features/constraint.py:171 `return_values[mask] = self.value`: numpy takes
    a scalar or a one-element array there, nothing else."""
    from ocsmesh_b200.constraint import TopoConstConstraint
    assert TopoConstConstraint(np.array([5.0])).value == 5.0
    assert TopoConstConstraint(np.float32(7)).value == 7
    with pytest.raises(ValueError, match="scalar"):
        TopoConstConstraint(np.ones((3, 3)))


def test_exact_collector_honours_source_index():
    """hfun/collector.py:1783-1786, 1951-1954: a refinement with a source index
    only goes to the member rasters it names (host logic, no device work)."""
    from ocsmesh_b200.hfun import HfunCollector
    calls = []

    class H:
        def __init__(self, k):
            self.k = k

        def add_constant_value(self, *a):
            calls.append((self.k, "cv", a))

        def add_subtidal_flow_limiter(self, *a):
            calls.append((self.k, "fl", a))

    c = HfunCollector.__new__(HfunCollector)
    c._method = "exact"
    c._hfuns = [H(0), H(1), H(2)]
    c.add_constant_value(5.0, source_index=1)
    c.add_subtidal_flow_limiter(1, 2, 3, 4, source_index=[0, 2])
    c.add_constant_value(7.0)
    assert calls == [(1, "cv", (5.0, None, None)), (0, "fl", (1, 2, 3, 4)),
                     (2, "fl", (1, 2, 3, 4)), (0, "cv", (7.0, None, None)),
                     (1, "cv", (7.0, None, None)), (2, "cv", (7.0, None, None))]
