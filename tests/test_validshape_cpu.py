"""Mesh.interpolate(filter_by_shape=True): the valid-data region of a raster
(ocsmesh/raster.py:560-640 + ocsmesh/mesh/mesh.py:2299-2301) -- the product's
vectorised rule against the oracle's point-by-point restatement and
hand-derived known answers.  matplotlib's contourf is absent: parity unpinned."""
import numpy as np

from ocsmesh_b200 import features as F
from oracle import ocsmesh_oracle as O


def test_known_answers_one_nodata_corner():
    x = np.array([0.0, 10.0, 20.0])
    y = np.array([20.0, 10.0, 0.0])           # north-up: rows run downwards
    mask = np.zeros((3, 3), bool)
    mask[0, 0] = True                         # pixel at (x=0, y=20)
    pts = np.array([
        [1.0, 19.0],     # next to the nodata corner: cut off
        [4.9, 20.0],     # on the top edge, before its midpoint: cut off
        [5.0, 20.0],     # the midpoint itself: boundary -> inside
        [3.0, 17.0],     # |du| + |dv| = 0.3 + 0.3 > 0.5 -> inside
        [2.0, 18.0],     # 0.2 + 0.2 < 0.5 -> outside
        [15.0, 5.0],     # a cell without nodata
        [-1.0, 5.0],     # outside the raster
        [20.0, 0.0],     # the far corner: on the bbox -> inside
    ])
    want = np.array([False, False, True, True, False, True, False, True])
    np.testing.assert_array_equal(F.valid_region_contains(x, y, mask, pts), want)
    np.testing.assert_array_equal(O.valid_region_contains(x, y, mask, pts), want)


def test_no_nodata_is_the_bbox_test():
    x = np.linspace(5.0, 95.0, 7)
    y = np.linspace(80.0, 20.0, 5)
    pts = np.random.default_rng(1).uniform(0, 100, (500, 2))
    bbox = (pts[:, 0] >= 5) & (pts[:, 0] <= 95) & (pts[:, 1] >= 20) & (pts[:, 1] <= 80)
    got = F.valid_region_contains(x, y, np.zeros((5, 7), bool), pts)
    np.testing.assert_array_equal(got, bbox)


def test_random_masks_equal_oracle():
    rng = np.random.default_rng(7)
    for frac, flipx in ((0.15, False), (0.5, True), (0.85, False)):
        x = np.linspace(100.0, 340.0, 25)
        if flipx:
            x = x[::-1].copy()
        y = np.linspace(900.0, 700.0, 21)
        mask = rng.random((21, 25)) < frac
        pts = rng.uniform([90, 690], [350, 910], (4000, 2))
        # plus points exactly on cell edges and midpoints
        gx, gy = np.meshgrid(np.linspace(100, 340, 49), np.linspace(700, 900, 41))
        pts = np.concatenate([pts, np.stack([gx.ravel(), gy.ravel()], axis=1)])
        np.testing.assert_array_equal(F.valid_region_contains(x, y, mask, pts),
                                      O.valid_region_contains(x, y, mask, pts))


def test_nodata_rule_of_linear_implies_the_shape_filter_nearest_does_not():
    """For method='linear' every point the shape filter drops is dropped by the
    nodata rule anyway (mesh/mesh.py:2318-2325: RGI of the mask > 0 as soon as
    one corner of the cell is nodata).  For 'nearest' (and 'spline', which has no
    nodata rule) the filter does change the result: in a cell with three nodata
    corners the points beyond the valid corner's triangle still have that valid
    pixel as their nearest one."""
    from scipy.interpolate import RegularGridInterpolator
    rng = np.random.default_rng(9)
    x = np.linspace(0.0, 240.0, 25)
    y = np.linspace(200.0, 0.0, 21)
    mask = rng.random((21, 25)) < 0.3
    pts = rng.uniform([0, 0], [240, 200], (6000, 2))
    shape_ok = F.valid_region_contains(x, y, mask, pts)
    dropped = {}
    for method in ("linear", "nearest"):
        f = RegularGridInterpolator((x, np.flip(y)), np.flipud(mask).T.astype(bool),
                                    method=method)
        dropped[method] = f(pts) > 0
    assert not np.any(~shape_ok & ~dropped["linear"])
    assert np.any(~shape_ok & ~dropped["nearest"])
