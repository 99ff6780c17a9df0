"""Host-side feature prologue of ``HfunRaster.add_feature`` / ``add_contour``.

The reference resamples every feature line to ``target_size`` spacing before
building the kd-tree (``ocsmesh/hfun/raster.py:1183-1225`` calling
``utils.repartition_features`` ``ocsmesh/utils.py:247-277`` and
``utils.transform_linestring`` ``ocsmesh/utils.py:280-312``).  That work is
O(vertices), not O(pixels), so it stays on the host -- and is hoisted out of
the window loop: for a projected CRS the result does not depend on the window.

shapely is not a dependency; a line is an ``(n, 2)`` float64 array.  Objects
with ``.coords`` (LineString-like) or ``.geoms`` (MultiLineString-like) are
accepted wherever the reference accepts shapely geometries.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence

import numpy as np


def _is_linestring_like(obj) -> bool:
    return hasattr(obj, "coords") and not hasattr(obj, "geoms")


def _is_multilinestring_like(obj) -> bool:
    return hasattr(obj, "geoms")


def as_linestrings(feature) -> List[np.ndarray]:
    """Flatten a feature argument into a list of ``(n, 2)`` float64 arrays.

    Raises ``TypeError`` for anything that is not line-like -- the
    reference's check at ``ocsmesh/hfun/raster.py:1155-1158``.
    """
    if _is_multilinestring_like(feature):
        out = []
        for g in feature.geoms:
            out.extend(as_linestrings(g))
        return out
    if _is_linestring_like(feature):
        return [np.asarray(feature.coords, dtype=float)[:, :2]]
    if isinstance(feature, np.ndarray):
        if feature.ndim == 2 and feature.shape[1] >= 2:
            return [np.asarray(feature, dtype=float)[:, :2]]
        raise TypeError("Argument feature must be an (n, 2) array, a "
                        "LineString or a MultiLineString, not an array of "
                        f"shape {feature.shape}.")
    if isinstance(feature, (list, tuple)):
        if len(feature) == 0:
            return []
        if all(isinstance(f, (list, tuple)) and len(f) == 2
               and all(np.isscalar(v) for v in f) for f in feature):
            return [np.asarray(feature, dtype=float)]
        out = []
        for f in feature:
            out.extend(as_linestrings(f))
        return out
    raise TypeError(
        "Argument feature must be of type LineString or MultiLineString "
        f"(or (n, 2) arrays), not type {type(feature)}.")


def _is_polygon_like(obj) -> bool:
    return hasattr(obj, "exterior") and hasattr(obj, "interiors")


def polygon_rings(shape) -> List[np.ndarray]:
    """Rings of a polygon input as closed ``(n, 2)`` float64 arrays: a shapely
    style Polygon (``.exterior`` / ``.interiors``) or MultiPolygon (``.geoms``),
    one ``(n, 2)`` array (a polygon without holes), or a list of those / of
    ``(exterior, [holes])`` pairs.  The masks built from them use the even-odd
    rule over ALL rings, so holes need no marking."""
    rings: List[np.ndarray] = []

    def add(r):
        a = np.asarray(r, dtype=np.float64)
        if a.ndim != 2 or a.shape[1] < 2 or len(a) < 3:
            raise TypeError("a polygon ring must be an (n >= 3, 2) coordinate array")
        a = np.ascontiguousarray(a[:, :2])
        if not np.array_equal(a[0], a[-1]):
            a = np.vstack([a, a[:1]])
        rings.append(a)

    def walk(o):
        if _is_polygon_like(o):
            add(o.exterior.coords)
            for i in o.interiors:
                add(i.coords)
        elif hasattr(o, "geoms"):
            for g in o.geoms:
                walk(g)
        elif isinstance(o, np.ndarray):
            add(o)
        elif isinstance(o, (list, tuple)):
            def is_pair(p):
                return isinstance(p, (list, tuple)) and len(p) in (2, 3) and \
                    all(np.isscalar(v) for v in p)
            if len(o) == 2 and isinstance(o[0], np.ndarray) and \
                    isinstance(o[1], (list, tuple)):
                add(o[0])                      # (exterior, [holes])
                for hole in o[1]:
                    add(hole)
            elif len(o) and is_pair(o[0]):
                add(o)                         # a plain list of (x, y) pairs
            else:
                for g in o:
                    walk(g)
        else:
            raise TypeError(f"Wrong type \"{type(o)}\" for multipolygon input.")

    walk(shape)
    if not rings:
        raise TypeError("empty polygon input")
    return rings


def ring_edges(rings: Sequence[np.ndarray]) -> np.ndarray:
    """``(n, 4)`` float64 edges (x0, y0, x1, y1) of closed rings."""
    return np.ascontiguousarray(np.concatenate(
        [np.hstack([r[:-1], r[1:]]) for r in rings], axis=0))


def repartition_features(linestring: np.ndarray, max_verts: int) -> List[np.ndarray]:
    """``ocsmesh/utils.py:247-277``: cut a line into chunks of at most
    ``max_verts`` vertices that share their end vertex."""
    if not isinstance(max_verts, (int, np.integer)) or isinstance(max_verts, bool):
        raise ValueError("Maximum number of vertices must be an integer!")
    coords = np.asarray(linestring, dtype=float)
    n = len(coords)
    if n <= max_verts:
        return [coords]
    step = max_verts - 1
    list_lens = [max_verts] * ((n - 1) // step)
    if (n - 1) % step != 0:
        list_lens.append((n - 1) % step + 1)
    ends = np.cumsum(np.array(list_lens) - 1)
    out, last = [], 0
    for e in ends:
        out.append(coords[last:e + 1])
        last = e
    return out


def transform_linestring(linestring: np.ndarray, target_size: float) -> np.ndarray:
    """``ocsmesh/utils.py:280-312``: points every ``target_size`` along the
    line (plus both ends); a line shorter than ``target_size`` is unchanged."""
    c = np.asarray(linestring, dtype=float)
    if len(c) < 2:
        return c
    lengths = ((c[1:] - c[:-1]) ** 2).sum(axis=1) ** 0.5
    cum_len = np.cumsum(lengths)
    total = float(cum_len[-1])
    distances = np.cumsum(np.ones(int(total // target_size)) * target_size)
    if len(distances) == 0:
        return c
    if distances[-1] < total:
        distances = np.append(distances, total)
    # index of the segment holding each station: first k with
    # distances <= cum_len[k]  (== len - count(distances <= cum_len))
    idx = np.searchsorted(cum_len, distances, side="left")
    idx = np.minimum(idx, len(cum_len) - 1)
    ratio = ((cum_len[idx] - distances) / lengths[idx])[:, None]
    pts = c[idx] * ratio + c[idx + 1] * (1 - ratio)
    return np.vstack((c[0], pts))


def feature_points(features: Sequence[np.ndarray], target_size: float,
                   max_verts: int = 200, native: bool = True) -> np.ndarray:
    """The ``points`` array handed to the kd-tree at
    ``ocsmesh/hfun/raster.py:1230`` (projected CRS): repartition, resample each
    chunk, concatenate.  ``native`` runs the library's host code
    (``ocsb_feature_points``, h_stitch.cu: bit-identical, ~10x faster); the
    numpy walk below stays as its cross-check (tests/test_host_cpu.py)."""
    if native:
        if not isinstance(max_verts, (int, np.integer)) or isinstance(max_verts, bool):
            raise ValueError("Maximum number of vertices must be an integer!")
        if not float(target_size) > 0:
            raise ValueError("Argument target_size must be greater than zero.")
        from . import _lib
        return _lib.feature_points(features, float(target_size), int(max_verts))
    parts = []
    for ls in features:
        for chunk in repartition_features(ls, max_verts):
            parts.append(transform_linestring(chunk, target_size))
    if not parts:
        return np.empty((0, 2), dtype=float)
    return np.ascontiguousarray(np.concatenate(parts, axis=0))


# ---------------------------------------------------------------------------
# contour tracing (CPU step before the accelerated path)
# ---------------------------------------------------------------------------
def _ms_block(x, y, z, level, j0, w):
    """Marching-squares segments of the row block ``z`` (rows j0 .. j0+hb-1 of
    a raster ``w`` wide).  Returns the two edge-id arrays of the segments and
    the (id, px, py) crossing table; ids are global: horizontal edge
    (j, i)-(j, i+1) -> j*(w-1)+i, vertical edge (j, i)-(j+1, i) ->
    n_he_total + j*w + i (the caller adds n_he_total, here left symbolic by
    a negative encoding: vertical ids are returned as -(j*w+i) - 1)."""
    z = np.asarray(z, dtype=float)
    hb = z.shape[0]
    above = z > level
    # corner bits: 1 = (j,i), 2 = (j,i+1), 4 = (j+1,i+1), 8 = (j+1,i)
    code = (above[:-1, :-1].astype(np.uint8)
            | (above[:-1, 1:].astype(np.uint8) << 1)
            | (above[1:, 1:].astype(np.uint8) << 2)
            | (above[1:, :-1].astype(np.uint8) << 3))
    jj, ii = np.nonzero((code != 0) & (code != 15))
    if len(jj) == 0:
        e = np.empty(0, dtype=np.int64)
        return e, e, e, np.empty(0), np.empty(0)
    cc = code[jj, ii]
    gj = jj.astype(np.int64) + j0
    ii = ii.astype(np.int64)
    top = gj * (w - 1) + ii
    bot = (gj + 1) * (w - 1) + ii
    lef = -(gj * w + ii) - 1
    rig = -(gj * w + ii + 1) - 1
    cen = 0.25 * (z[jj, ii] + z[jj, ii + 1] + z[jj + 1, ii] + z[jj + 1, ii + 1]) > level
    ea, eb = [], []

    def emit(sel, a, b):
        if np.any(sel):
            ea.append(a[sel])
            eb.append(b[sel])

    for (c0, c1), a, b in (((1, 14), top, lef), ((2, 13), top, rig),
                           ((4, 11), rig, bot), ((8, 7), lef, bot),
                           ((3, 12), lef, rig), ((6, 9), top, bot)):
        emit((cc == c0) | (cc == c1), a, b)
    s5, s10 = cc == 5, cc == 10
    emit(s5 & cen, top, rig)
    emit(s5 & cen, lef, bot)
    emit(s5 & ~cen, top, lef)
    emit(s5 & ~cen, rig, bot)
    emit(s10 & cen, top, lef)
    emit(s10 & cen, rig, bot)
    emit(s10 & ~cen, top, rig)
    emit(s10 & ~cen, lef, bot)
    ea = np.concatenate(ea)
    eb = np.concatenate(eb)
    # crossing points of the edges used by this block
    ids = np.unique(np.concatenate([ea, eb]))
    px = np.empty(len(ids))
    py = np.empty(len(ids))
    is_h = ids >= 0
    hj, hi = np.divmod(ids[is_h], w - 1)
    lj = hj - j0
    t = (level - z[lj, hi]) / (z[lj, hi + 1] - z[lj, hi])
    px[is_h] = x[hi] + t * (x[hi + 1] - x[hi])
    py[is_h] = y[hj]
    vj, vi = np.divmod(-ids[~is_h] - 1, w)
    lj = vj - j0
    t = (level - z[lj, vi]) / (z[lj + 1, vi] - z[lj, vi])
    px[~is_h] = x[vi]
    py[~is_h] = y[vj] + t * (y[vj + 1] - y[vj])
    return ea, eb, ids, px, py


def marching_squares(x, y, z, level, block_rows: int = 2048) -> List[np.ndarray]:
    """Iso-lines of ``z`` (``(h, w)``; row ``j`` at ``y[j]``, column ``i`` at
    ``x[i]``) at ``level`` as stitched polylines.

    The reference obtains these from matplotlib/contourpy + shapely linemerge
    (``ocsmesh/raster.py:1426-1470``); neither is available here.  Crossings
    are placed by linear interpolation along cell edges, saddles are resolved
    with the cell mean, and segments are stitched through shared cell edges.
    The raster is swept in row blocks so a 20000 x 20000 DEM needs no
    full-size temporaries.
    """
    x = np.asarray(x, dtype=float)
    y = np.asarray(y, dtype=float)
    h, w = z.shape
    if h < 2 or w < 2:
        return []
    eas, ebs, idl, pxl, pyl = [], [], [], [], []
    for j0 in range(0, h - 1, block_rows):
        j1 = min(h, j0 + block_rows + 1)
        ea, eb, ids, px, py = _ms_block(x, y, z[j0:j1], level, j0, w)
        if len(ea):
            eas.append(ea); ebs.append(eb)
            idl.append(ids); pxl.append(px); pyl.append(py)
    if not eas:
        return []
    ea = np.concatenate(eas)
    eb = np.concatenate(ebs)
    all_ids = np.concatenate(idl)
    ids, first_idx = np.unique(all_ids, return_index=True)
    px = np.concatenate(pxl)[first_idx]
    py = np.concatenate(pyl)[first_idx]
    return _stitch(ea, eb, ids, px, py)


def stitch_segments(edges, pts, native: bool = True) -> List[np.ndarray]:
    """Polylines from the output of ``ocsb_contour_segments``: ``edges``
    ``(n, 2)`` int64 global edge ids, ``pts`` ``(n, 4)`` float64 crossing
    points ``(xa, ya, xb, yb)``.  Same stitching (and therefore the same
    polylines, vertex for vertex) as :func:`marching_squares`."""
    edges = np.asarray(edges, dtype=np.int64).reshape(-1, 2)
    pts = np.asarray(pts, dtype=np.float64).reshape(-1, 4)
    if len(edges) == 0:
        return []
    if native:
        # the library's host code (h_stitch.cu): same rules, ~50x faster than
        # the numpy + Python walk below, which stays as its cross-check
        from . import _lib
        return _lib.stitch_segments(edges, pts)
    ids, first_idx = np.unique(edges.ravel(), return_index=True)
    xy = pts.reshape(-1, 2)[first_idx]
    return _stitch(edges[:, 0], edges[:, 1], ids, xy[:, 0], xy[:, 1])


def _stitch(ea, eb, ids, px, py) -> List[np.ndarray]:
    """Join segments (edge-id pairs) that share a cell edge into polylines.
    The segment list is put into a canonical order first, so the result does
    not depend on the order the segments were produced in (host blocks by
    case, device in cell order)."""
    lo, hi = np.minimum(ea, eb), np.maximum(ea, eb)
    canon = np.lexsort((hi, lo))
    ea, eb = lo[canon], hi[canon]
    a = np.searchsorted(ids, ea)
    b = np.searchsorted(ids, eb)
    nn = len(ids)
    ends = np.concatenate([a, b])
    other = np.concatenate([b, a])
    order = np.argsort(ends, kind="stable")
    ends, other = ends[order], other[order]
    first = np.ones(len(ends), dtype=bool)
    first[1:] = ends[1:] != ends[:-1]
    nbr = np.full((nn, 2), -1, dtype=np.int64)
    nbr[ends[first], 0] = other[first]
    nbr[ends[~first], 1] = other[~first]
    deg = (nbr >= 0).sum(axis=1)

    visited = np.zeros(nn, dtype=bool)
    nbr_l = nbr.tolist()
    lines = []

    def walk(start):
        path = [start]
        visited[start] = True
        cur = start
        while True:
            n0, n1 = nbr_l[cur]
            if n0 >= 0 and not visited[n0]:
                nxt = n0
            elif n1 >= 0 and not visited[n1]:
                nxt = n1
            else:
                if len(path) > 2 and start in (n0, n1):
                    path.append(start)  # closed ring
                break
            path.append(nxt)
            visited[nxt] = True
            cur = nxt
        return path

    for s in np.nonzero(deg == 1)[0].tolist():
        if not visited[s]:
            lines.append(walk(s))
    for s in np.nonzero(deg == 2)[0].tolist():
        if not visited[s]:
            lines.append(walk(s))
    out = []
    for path in lines:
        if len(path) < 2:
            continue
        e = np.asarray(path)
        out.append(np.stack([px[e], py[e]], axis=1))
    return out


def valid_region_contains(x: np.ndarray, y: np.ndarray, mask: np.ndarray,
                          xy: np.ndarray) -> np.ndarray:
    """``points.intersects(Raster.get_multipolygon())`` for one unchunked raster
    (``ocsmesh/raster.py:560-640`` called at ``ocsmesh/mesh/mesh.py:2299-2301``):
    the multipolygon is matplotlib's filled contour between the levels 0 and 1
    of the field that is +1 on valid and -1 on nodata pixels, i.e. the region
    bounded by the piecewise-linear zero line of that field -- it cuts every
    cell edge between a valid and a nodata pixel at its midpoint.

    matplotlib is not importable here (**parity unpinned**); the rule is restated
    per cell: a point is outside when it lies strictly inside the corner
    triangle (midpoint to midpoint) of a nodata corner, or strictly on the
    nodata side of the mid line when two neighbouring corners are nodata; a
    saddle cell (its mean is 0, not below the level) keeps its centre.
    ``x`` (ascending or descending) and ``y`` are the raster's ``get_x`` /
    ``get_y`` axes, ``mask`` is ``(len(y), len(x))`` with True = nodata.
    Boundary points count as inside (``intersects``)."""
    xy = np.asarray(xy, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    mask = np.asarray(mask, dtype=bool)
    px, py = xy[:, 0], xy[:, 1]
    inside = ((px >= min(x[0], x[-1])) & (px <= max(x[0], x[-1])) &
              (py >= min(y[0], y[-1])) & (py <= max(y[0], y[-1])))
    if len(x) < 2 or len(y) < 2:
        return inside & ~mask.any()
    if not mask.any():
        return inside

    def locate(axis, p):
        asc = axis if axis[0] <= axis[-1] else axis[::-1]
        k = np.clip(np.searchsorted(asc, p, side="right") - 1, 0, len(asc) - 2)
        t = (p - asc[k]) / (asc[k + 1] - asc[k])
        if axis[0] > axis[-1]:                    # back to the stored direction
            k = len(axis) - 2 - k
            t = 1.0 - t
        return k, np.clip(t, 0.0, 1.0)

    idx = np.nonzero(inside)[0]
    i, u = locate(x, px[idx])
    j, v = locate(y, py[idx])
    m00, m01 = mask[j, i], mask[j, i + 1]
    m10, m11 = mask[j + 1, i], mask[j + 1, i + 1]
    out = np.zeros(idx.shape, dtype=bool)
    # one corner (or two diagonal ones: the saddle keeps its centre): its triangle
    out |= m00 & ~m01 & ~m10 & (u + v < 0.5)
    out |= m01 & ~m00 & ~m11 & ((1 - u) + v < 0.5)
    out |= m10 & ~m00 & ~m11 & (u + (1 - v) < 0.5)
    out |= m11 & ~m01 & ~m10 & ((1 - u) + (1 - v) < 0.5)
    # two neighbouring corners: the mid line
    out |= m00 & m01 & ~m10 & ~m11 & (v < 0.5)
    out |= m10 & m11 & ~m00 & ~m01 & (v > 0.5)
    out |= m00 & m10 & ~m01 & ~m11 & (u < 0.5)
    out |= m01 & m11 & ~m00 & ~m10 & (u > 0.5)
    # three corners: only the triangle of the valid corner is inside
    out |= m01 & m10 & m11 & ~m00 & (u + v > 0.5)
    out |= m00 & m10 & m11 & ~m01 & ((1 - u) + v > 0.5)
    out |= m00 & m01 & m11 & ~m10 & (u + (1 - v) > 0.5)
    out |= m00 & m01 & m10 & ~m11 & ((1 - u) + (1 - v) > 0.5)
    out |= m00 & m01 & m10 & m11
    res = inside.copy()
    res[idx] = ~out
    return res
