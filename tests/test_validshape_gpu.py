"""Mesh.interpolate(filter_by_shape=True) (ocsmesh/mesh/mesh.py:2298-2301): only
the nodes inside the raster's valid-data region are interpolated.  Oracle: the
same interpolation on the subset the restated region rule selects."""
import numpy as np
import pytest

import ocsmesh_b200 as ob
from ocsmesh_b200 import synthetic as S
from oracle import ocsmesh_oracle as O

pytestmark = pytest.mark.gpu


def _raster(seed, with_mask=True):
    h, w = 90, 120
    z = S.synth_dem(h, w, seed=seed)
    tr = S.dem_transform(h)
    mask = None
    if with_mask:
        rng = np.random.default_rng(seed)
        mask = rng.random((h, w)) < 0.08
        mask[20:40, 30:70] = True            # a hole
    return z, tr, mask


@pytest.mark.parametrize("method", ["linear", "nearest", "spline"])
def test_filter_by_shape_equals_oracle_subset(ctx, method):
    z, tr, mask = _raster(41)
    r = ob.Raster.from_numpy(z, tr, crs=32618, nodata=-9999.0, mask=mask)
    orr = O.OracleRaster(z, tr, mask=mask)
    xy = S.synth_nodes(6000, r.get_bbox(), seed=5, outside_frac=0.05)
    start = np.full(len(xy), -1.0)
    mesh = ob.Mesh(ob.MeshData(xy, values=start.copy()))
    mesh.interpolate(r, method=method, filter_by_shape=True)
    sel = O.valid_region_contains(r.get_x(), r.get_y(), mask, xy)
    assert 0 < sel.sum() < len(xy)
    ref = start.copy()
    ref[sel] = O.mesh_interpolate(xy[sel], start[sel], orr, method=method).ravel()
    got = mesh.value.ravel()
    if method == "spline":
        np.testing.assert_allclose(got, ref, rtol=1e-9, atol=1e-9)
        # without the filter the spline also fills nodes over the hole
        plain = ob.Mesh(ob.MeshData(xy, values=start.copy()))
        plain.interpolate(r, method=method)
        assert np.any((plain.value.ravel() != -1.0) & (got == -1.0))
    else:
        np.testing.assert_array_equal(got, ref)


def test_filter_by_shape_without_nodata_and_errors(ctx):
    z, tr, _ = _raster(43, with_mask=False)
    r = ob.Raster.from_numpy(z, tr, crs=32618)
    xy = S.synth_nodes(3000, r.get_bbox(), seed=6, outside_frac=0.05)
    a = ob.Mesh(ob.MeshData(xy, values=np.full(len(xy), np.nan)))
    b = ob.Mesh(ob.MeshData(xy, values=np.full(len(xy), np.nan)))
    a.interpolate(r, method="linear", filter_by_shape=True)
    b.interpolate(r, method="linear")
    np.testing.assert_array_equal(a.value, b.value)
    rc = ob.Raster.from_numpy(z, tr, crs=32618, chunk_size=50)
    with pytest.raises(NotImplementedError, match="chunked"):
        a.interpolate(rc, method="linear", filter_by_shape=True)
    with pytest.raises(NotImplementedError, match="info_out_path"):
        a.interpolate(r, method="linear", info_out_path="/tmp/x.csv")
