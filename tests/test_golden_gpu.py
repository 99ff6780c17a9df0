"""The CUDA path vs tests/golden/reference_golden.npz -- outputs of the
reference's own source (tests/golden/make_golden.py) -- through the drop-in
API and the C ABI.  Sizes are bit-exact (tolerance stated by BASELINE.json:
1e-5 relative in fp32), nearest picks bit-exact, spline within 1e-9."""
import os

import numpy as np
import pytest

import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def opt(v):
    return None if np.isnan(v) else float(v)


def test_device_window_coordinates(ctx, gold):
    """ocsb_window_coords == np.linspace of the reference's get_x / get_y."""
    cw, rw, w, h = (int(v) for v in gold["xy_win_args"])
    win = _lib.make_win(gold["bounds_win"], w, h)
    x, y = np.zeros(w), np.zeros(h)
    ctx.window_coords(win, x, y)
    np.testing.assert_array_equal(x, gold["x_win"])
    np.testing.assert_array_equal(y, gold["y_win"])


def test_hfun_ops_match_reference(ctx, gold):
    lines = [gold[f"feat_line_{k}"] for k in range(int(gold["feat_n_lines"]))]
    names = sorted(k[5:-5] for k in gold.files
                   if k.startswith("hfun_") and k.endswith("_args"))
    assert len(names) >= 5
    for name in names:
        chunk, ovl, hmin, hmax, rate, target = (opt(v) for v in gold[f"hfun_{name}_args"])
        r = ob.Raster.from_numpy(gold["dem"], gold["dem_transform"], crs=32618,
                                 chunk_size=None if chunk is None else int(chunk),
                                 overlap=None if ovl is None else int(ovl))
        hf = ob.Hfun(r, hmin=hmin, hmax=hmax)
        hf.add_feature(lines, rate, target)
        np.testing.assert_array_equal(hf.get_values(), gold[f"hfun_{name}_feature"],
                                      err_msg=name)
        hf.add_subtidal_flow_limiter(50.0 if hmin else None, hmax, -30.0, 0.0)
        np.testing.assert_array_equal(hf.get_values(), gold[f"hfun_{name}_limiter"],
                                      err_msg=name)
        hf.add_gradient_delimiter(None, None)
        np.testing.assert_array_equal(hf.get_values(), gold[f"hfun_{name}_gradient"],
                                      err_msg=name)


def test_plateau_matches_reference(ctx, gold):
    r = ob.Raster.from_numpy(gold["dem_plateau"], gold["dem_transform"], crs=32618)
    hf = ob.Hfun(r, hmin=100.0, hmax=1500.0)
    hf.add_subtidal_flow_limiter(50.0, 1500.0, -30.0, 0.0)
    np.testing.assert_array_equal(hf.get_values(), gold["hfun_plateau_limiter"])


@pytest.mark.parametrize("method", ["linear", "nearest", "spline"])
def test_interpolation_matches_reference(ctx, gold, method):
    xy = gold["itp_xy"]
    ra = ob.Raster.from_numpy(gold["itp_dem1"], gold["itp_tr1"], crs=32618)
    if method == "spline":
        rb = ob.Raster.from_numpy(gold["itp_dem2_nomask"], gold["itp_tr2"], crs=32618)
    else:
        rb = ob.Raster.from_numpy(gold["itp_dem2"], gold["itp_tr2"], crs=32618,
                                  mask=gold["itp_mask2"])
    mesh = ob.Mesh(ob.MeshData(xy, values=np.full(len(xy), -7.5), crs=32618))
    mesh.interpolate([ra, rb], method=method)
    ref = gold[f"itp_{method}_values"]
    if method == "spline":
        err = np.abs(mesh.value - ref) / np.maximum(np.abs(ref), 1.0)
        assert err.max() < 1e-9
    else:
        np.testing.assert_array_equal(mesh.value, ref)
    # per-raster worker outputs (mask + values) through the C ABI
    for k, r in enumerate((ra, rb)):
        import torch
        dem = r.device_band(1, 0)
        m = r.device_mask(1, 0) if method != "spline" else None
        win = _lib.make_win(r.get_window_bounds(ob.Window(0, 0, r.width, r.height)),
                            r.width, r.height)
        touched = np.zeros(len(xy), np.uint8)
        vals = np.zeros(len(xy))
        code = {"linear": 0, "nearest": 1, "spline": 2}[method]
        ctx.interp(win, dem, _lib.OCSB_F32, r.width, m, r.width,
                   np.ascontiguousarray(xy), len(xy), code, touched, vals)
        mask = touched.astype(bool)
        np.testing.assert_array_equal(mask, gold[f"itp_{method}_mask{k}"])
        gv = gold[f"itp_{method}_vals{k}"].astype(np.float64)
        if method == "spline":
            assert np.abs(vals[mask] - gv).max() < 1e-8
        else:
            np.testing.assert_array_equal(vals[mask], gv)
