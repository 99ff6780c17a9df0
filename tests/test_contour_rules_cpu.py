"""Contour tracing rules, hand-derived known answers.

The reference traces iso-lines with matplotlib/contourpy and merges the pieces
with shapely (ocsmesh/raster.py:1426-1470); neither is importable here, so the
tracer of this repository (device sweep ocsb_contour_segments + host stitch,
mirrored by features.marching_squares) is PARITY-UNPINNED against them.  What
can be pinned without them:

* the rules contourpy documents for a quad grid -- linear interpolation of the
  crossing along a cell edge, a corner equal to the level is "not above", a
  saddle cell is resolved by the mean of its four corners -- as hand-derived
  answers on cells small enough to check by eye;
* that the quantity the accelerated path consumes is robust against the part
  that stays unpinned: polyline start vertex and direction only move the
  resampled vertices ALONG the line, so the size function changes by at most
  rate * target * target / 2 anywhere.
"""
import numpy as np
import pytest
from scipy.spatial import cKDTree

from ocsmesh_b200 import features as F
from oracle import ocsmesh_oracle as O

X2 = np.array([0.0, 1.0])
Y2 = np.array([0.0, 1.0])


def segments_of(lines):
    out = set()
    for ln in lines:
        for a, b in zip(ln[:-1], ln[1:]):
            out.add(frozenset([tuple(np.round(a, 6)), tuple(np.round(b, 6))]))
    return out


@pytest.mark.parametrize("tracer", [F.marching_squares, O.marching_squares])
def test_saddle_is_resolved_by_the_cell_mean(tracer):
    # corners (row, col): (0,0)=0 (0,1)=1 (1,0)=1 (1,1)=v; level 0.5
    # mean above the level -> the two ABOVE corners are joined: each BELOW
    # corner is cut off by its own segment
    z = np.array([[0.0, 1.0], [1.0, 0.2]])           # mean 0.55
    t = (1.0 - 0.5) / (1.0 - 0.2)                    # crossing on the edges into (1,1)
    want = {frozenset([(0.5, 0.0), (0.0, 0.5)]),
            frozenset([(1.0, round(t, 6)), (round(t, 6), 1.0)])}
    assert segments_of(tracer(X2, Y2, z, 0.5)) == want
    # mean below the level -> the two BELOW corners are joined: each ABOVE
    # corner is cut off
    z = np.array([[0.0, 1.0], [1.0, -0.2]])          # mean 0.45
    t = (1.0 - 0.5) / (1.0 + 0.2)
    want = {frozenset([(0.5, 0.0), (1.0, round(t, 6))]),
            frozenset([(0.0, 0.5), (round(t, 6), 1.0)])}
    assert segments_of(tracer(X2, Y2, z, 0.5)) == want


@pytest.mark.parametrize("tracer", [F.marching_squares, O.marching_squares])
def test_a_corner_on_the_level_is_not_above(tracer):
    # top row exactly on the level, bottom row above: the line runs along the
    # top row itself (t = 0 on both vertical edges)
    z = np.array([[0.5, 0.5], [1.0, 1.0]])
    assert segments_of(tracer(X2, Y2, z, 0.5)) == {frozenset([(0.0, 0.0), (1.0, 0.0)])}
    # everything on or below the level: no line at all
    z = np.array([[0.5, 0.5], [0.5, 0.2]])
    assert tracer(X2, Y2, z, 0.5) == []


@pytest.mark.parametrize("tracer", [F.marching_squares, O.marching_squares])
def test_crossings_are_linear_along_the_edge_and_stitched(tracer):
    # a plane z = x on a 3 x 4 grid, level 1.25: one open line at x = 1.25
    # through all rows, 3 vertices (one per row), end points on the raster edge
    x = np.arange(4.0)
    y = np.arange(3.0)
    z = np.tile(x, (3, 1))
    lines = tracer(x, y, z, 1.25)
    assert len(lines) == 1 and lines[0].shape == (3, 2)
    np.testing.assert_allclose(lines[0][:, 0], 1.25)
    assert sorted(lines[0][:, 1]) == [0.0, 1.0, 2.0]
    # a closed loop: a bump in the middle of a 5 x 5 grid
    g = np.arange(5.0)
    zz = -np.hypot(*np.meshgrid(g - 2.0, g - 2.0))
    loop = tracer(g, g, zz, -1.5)
    assert len(loop) == 1
    np.testing.assert_array_equal(loop[0][0], loop[0][-1])     # closed: first == last
    # the iso-line of radius 1.5 cuts the grid lines x, y = 1, 2, 3 twice each
    # (12 crossings; linear interpolation puts those on x, y = 2 at 0.5 / 3.5)
    assert len(loop[0]) == 13                                  # 12 crossings + closure
    on_axis = sorted(tuple(p) for p in loop[0][:-1] if p[0] == 2.0 or p[1] == 2.0)
    assert on_axis == [(0.5, 2.0), (2.0, 0.5), (2.0, 3.5), (3.5, 2.0)]


def test_size_function_is_insensitive_to_start_vertex_and_direction():
    """What stays unpinned in the tracer only decides WHERE along the line the
    resampled vertices sit.  Two resamplings of the same closed line (rolled
    start, reversed direction) differ by at most target/2 in nearest-vertex
    distance, i.e. rate * target * target / 2 in size."""
    rate, target = 0.001, 100.0
    t = np.linspace(0.0, 2 * np.pi, 400, endpoint=False)
    ring = np.stack([3000.0 + 1800.0 * np.cos(t) * (1 + 0.2 * np.sin(3 * t)),
                     2500.0 + 1500.0 * np.sin(t)], axis=1)
    ring = np.vstack([ring, ring[:1]])
    rolled = np.roll(ring[:-1], 137, axis=0)
    variants = [ring, np.vstack([rolled, rolled[:1]]), ring[::-1].copy()]
    gx, gy = np.meshgrid(np.linspace(0.0, 6000.0, 61), np.linspace(0.0, 5000.0, 51))
    q = np.stack([gx.ravel(), gy.ravel()], axis=1)
    sizes = []
    for v in variants:
        pts = F.feature_points([v], target, native=False)
        d = cKDTree(pts).query(q)[0]
        sizes.append((rate * target * d + target).astype(np.float32))
    bound = rate * target * target / 2 + 1e-3
    assert np.abs(sizes[0] - sizes[1]).max() <= bound
    assert np.abs(sizes[0] - sizes[2]).max() <= bound
    assert np.abs(sizes[0] - sizes[1]).max() > 0          # they do differ: unpinned
