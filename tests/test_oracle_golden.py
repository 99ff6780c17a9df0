"""Pin the CPU oracle against (a) tests/golden/reference_golden.npz -- outputs of
the REFERENCE'S OWN SOURCE executed by tests/golden/make_golden.py -- and (b) the
known-answer tests the reference carries for this path (file:line cited at each
test).  CPU only."""
import os

import numpy as np
import pytest

from oracle import ocsmesh_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def opt(v):
    return None if np.isnan(v) else float(v)


def hfun_case_names(gold):
    return sorted(k[5:-5] for k in gold.files if k.startswith("hfun_") and k.endswith("_args"))


# ---- golden: windows -------------------------------------------------------
def test_windows_match_reference(gold):
    k = 0
    while f"win_{k}" in gold.files:
        w, h, c, o = (int(v) for v in gold[f"win_{k}_args"])
        got = np.array([tuple(x) for x in O.get_iter_windows(w, h, c, o)])
        np.testing.assert_array_equal(got, gold[f"win_{k}"])
        k += 1
    assert k >= 9


def test_windows_survey_cases():
    # 20000^2, chunk 5000 -> 16 windows; overlap 2 -> 5002 wide
    a = list(O.get_iter_windows(20000, 20000, 5000, 0))
    assert len(a) == 16 and a[0] == O.Window(0, 0, 5000, 5000)
    b = list(O.get_iter_windows(20000, 20000, 5000, 2))
    assert b[0] == O.Window(0, 0, 5002, 5002) and b[1].col_off == 5000
    assert b[-1] == O.Window(15000, 15000, 5000, 5000)


# ---- golden + known answers: feature prologue -------------------------------
def test_prologue_matches_reference(gold):
    line = gold["pro_line"]
    chunks = O.repartition_features(line, 200)
    np.testing.assert_array_equal([len(c) for c in chunks], gold["pro_chunk_lens"])
    pts = np.concatenate([O.transform_linestring(c, 100.0) for c in chunks])
    np.testing.assert_array_equal(pts, gold["pro_points"])
    np.testing.assert_array_equal(O.feature_points([line], 100.0, 200), gold["pro_points"])
    np.testing.assert_array_equal(O.transform_linestring(line[:3], 1e6), gold["pro_short"])


@pytest.mark.parametrize("n,max_verts,n_seg,fix_len,last_len", [
    (10, 2, 9, 2, 2), (10, 3, 5, 3, 2), (11, 3, 5, 3, 3), (47, 5, 12, 5, 3),
    (100, 7, 17, 7, 4), (100000, 200, 503, 200, 102), (100000, 201, 500, 201, 200)])
def test_repartition_known_answers(n, max_verts, n_seg, fix_len, last_len):
    """reference tests/api/test_utils.py:873-916"""
    line = np.array([[i, i] for i in range(n)], dtype=float)
    res = O.repartition_features(line, max_verts)
    assert len(res) == n_seg
    assert all(len(c) == fix_len for c in res[:-1])
    assert len(res[-1]) == last_len
    for a, b in zip(res[:-1], res[1:]):
        np.testing.assert_array_equal(a[-1], b[0])   # shared end node


@pytest.mark.parametrize("fix_dist", [3, 0.3])
def test_transform_linestring_known_answers(fix_dist):
    """reference tests/api/test_utils.py:970-979"""
    line = np.array([[i, i] for i in range(50)], dtype=float)
    res = O.transform_linestring(line, fix_dist)
    length = O.linestring_length(line)
    exp_nseg = length // fix_dist + 1 * bool(length % fix_dist)
    assert exp_nseg == len(res) - 1
    lens = ((res[1:] - res[:-1]) ** 2).sum(axis=1) ** 0.5
    np.testing.assert_allclose(lens[:-1], fix_dist)
    np.testing.assert_array_equal(res[0], line[0])
    np.testing.assert_array_equal(res[-1], line[-1])


# ---- golden + known answers: coordinates ------------------------------------
def test_coordinates_match_reference(gold):
    r = O.OracleRaster(gold["dem"], gold["dem_transform"])
    np.testing.assert_array_equal(r.get_xy(), gold["xy_full"])
    win = O.Window(*(int(v) for v in gold["xy_win_args"]))
    np.testing.assert_array_equal(r.get_x(win), gold["x_win"])
    np.testing.assert_array_equal(r.get_y(win), gold["y_win"])
    np.testing.assert_array_equal(r.get_window_bounds(win), gold["bounds_win"])


def test_get_xy_mgrid_known_answer():
    """reference tests/api/test_utils.py:674-700 (raster_from_numpy transform
    ocsmesh/utils.py:1938-1945 + get_xy ordering)."""
    mg = np.mgrid[1:3:0.1, -1:1:0.1]
    x, y = mg[0][:, 0], mg[1][0, :]
    z = np.random.default_rng(0).random(mg[0].shape)
    res_x = (x[-1] - x[0]) / z.shape[1]
    res_y = (y[-1] - y[0]) / z.shape[0]
    r = O.OracleRaster(z, (res_x, 0.0, x[0], 0.0, res_y, y[0]))
    assert np.all(np.isclose(mg.transpose([2, 1, 0]).reshape(-1, 2), r.get_xy()))
    np.testing.assert_array_equal(r.get_values(), z)


# ---- golden: hfun ops --------------------------------------------------------
def test_hfun_ops_match_reference(gold):
    lines = [gold[f"feat_line_{k}"] for k in range(int(gold["feat_n_lines"]))]
    names = hfun_case_names(gold)
    assert len(names) >= 5
    for name in names:
        chunk, ovl, hmin, hmax, rate, target = (opt(v) for v in gold[f"hfun_{name}_args"])
        r = O.OracleRaster(gold["dem"], gold["dem_transform"],
                           chunk_size=None if chunk is None else int(chunk),
                           overlap=None if ovl is None else int(ovl))
        oh = O.OracleHfun(r, hmin, hmax)
        oh.add_feature(lines, rate, target)
        np.testing.assert_array_equal(oh.values, gold[f"hfun_{name}_feature"], err_msg=name)
        oh.add_subtidal_flow_limiter(50.0 if hmin else None, hmax, -30.0, 0.0)
        np.testing.assert_array_equal(oh.values, gold[f"hfun_{name}_limiter"], err_msg=name)
        oh.add_subtidal_flow_limiter(None, None, None, None)
        np.testing.assert_array_equal(oh.values, gold[f"hfun_{name}_gradient"], err_msg=name)


def test_plateau_matches_reference(gold):
    oh = O.OracleHfun(O.OracleRaster(gold["dem_plateau"], gold["dem_transform"]),
                      100.0, 1500.0)
    oh.add_subtidal_flow_limiter(50.0, 1500.0, -30.0, 0.0)
    np.testing.assert_array_equal(oh.values, gold["hfun_plateau_limiter"])
    assert np.all(oh.values[21:39, 31:69] == 100.0)    # zero gradient -> hmin


def test_limiter_closed_form_equals_whole_array():
    """the per-pixel closed form the kernel implements vs the
    literal masked-array statement, incl. the residual mask."""
    rng = np.random.default_rng(4)
    for trial in range(12):
        z = (rng.normal(size=(23, 31)) * 10).astype(np.float32)
        z[5:9, 4:20] = 2.0
        old = (rng.random(z.shape) * 3000).astype(np.float32)
        b = [None if rng.random() < 0.3 else v for v in (50.0, 1500.0, -3.0, 1.0, 100.0, 1200.0)]
        ref = O.flow_limiter_window(z, old, 10.0, 10.0, *b)
        g0, g1 = np.gradient(z, np.float32(10.0), np.float32(10.0))
        with np.errstate(all="ignore"):
            dh = np.sqrt(g0 * g0 + g1 * g1)
            h = np.where(dh == 0, 1.0 / 3.0, np.abs((1.0 / 3.0) * (z / dh).astype(np.float64)))
        m = dh == 0
        hmin, hmax, lb, ub, ghmin, ghmax = b
        if hmin is not None:
            s = h < hmin; h[s] = hmin; m[s] = False
        if hmax is not None:
            s = h > hmax; h[s] = hmax; m[s] = False
        if ub is not None:
            s = z > ub; h[s] = old[s]; m[s] = False
        if lb is not None:
            s = z < lb; h[s] = old[s]; m[s] = False
        if ghmin is not None:
            s = h < ghmin; h[s] = ghmin; m[s] = False
        if ghmax is not None:
            s = h > ghmax; h[s] = ghmax; m[s] = False
        out = np.minimum(old.astype(np.float64), h).astype(np.float32)
        np.testing.assert_array_equal(np.ma.getmaskarray(ref), m)
        np.testing.assert_array_equal(np.ma.getdata(ref)[~m], out[~m])


# ---- golden + known answers: interpolation -----------------------------------
def test_interpolation_matches_reference(gold):
    xy = gold["itp_xy"]
    init = np.full(len(xy), -7.5)
    ra = O.OracleRaster(gold["itp_dem1"], gold["itp_tr1"])
    for method in ("linear", "nearest", "spline"):
        if method == "spline":
            rb = O.OracleRaster(gold["itp_dem2_nomask"], gold["itp_tr2"])
        else:
            rb = O.OracleRaster(gold["itp_dem2"], gold["itp_tr2"], mask=gold["itp_mask2"])
        for k, r in enumerate((ra, rb)):
            mask, vals = O.mesh_interpolate_worker(xy, r, None, method)
            np.testing.assert_array_equal(mask, gold[f"itp_{method}_mask{k}"])
            np.testing.assert_array_equal(vals, gold[f"itp_{method}_vals{k}"])
        out = O.mesh_interpolate(xy, init, [ra, rb], method=method)
        np.testing.assert_array_equal(out, gold[f"itp_{method}_values"])


def test_constant_raster_known_answer():
    """reference tests/api/test_mesh.py:361-379: constant 1.0 / 2.0 bands."""
    bands = np.stack([np.ones((11, 8)), np.full((11, 8), 2.0)])
    r = O.OracleRaster(bands, (0.1, 0, -1.0, 0, -0.1, 0.1))
    rng = np.random.default_rng(1)
    xy = np.stack([rng.uniform(-0.95, -0.25, 200), rng.uniform(-0.95, 0.05, 200)], axis=1)
    for method in ("spline", "linear", "nearest"):
        for band, val in ((1, 1.0), (2, 2.0)):
            out = O.mesh_interpolate(xy, np.zeros(len(xy)), r, method=method, band=band)
            assert np.all(np.isclose(out, val))


def test_apply_rate_known_answer():
    """reference tests/api/test_features.py:436-455 (RegionConstraint with
    rate: [0,0,0,0,10,...] -> [2,3,4,5,...]); the nearest-vertex INDEX is what
    Constraint._apply_rate consumes (features/constraint.py:104-109)."""
    old = np.array([0, 0, 0, 0, 10, 10, 10, 10, 20, 20, 20, 20])
    loc = np.array([(i, 0) for i in range(len(old))], dtype=float)
    within = (loc[:, 0] > 2.5) & (loc[:, 0] < 3.5)
    vals = old.copy()
    mask = within & np.logical_not(np.greater(vals, 5))
    vals[mask] = 5
    out = O.apply_rate(vals, loc, mask, 0.199, -1, np.greater)
    check = old.copy()
    check[:4] = [2, 3, 4, 5]
    np.testing.assert_array_equal(out, check)
