#!/usr/bin/env python
"""Generate tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN SOURCE.

The reference package cannot be imported in the build container (pyproj,
rasterio, shapely, geopandas, matplotlib are missing), but the functions on the
hot path are plain numpy/scipy statements.  This script lifts the *source text*
of those functions out of /root/reference with ``ast`` (nothing is copied into
the repository), compiles it in a sandbox namespace whose only stand-ins are the
I/O shells around the arithmetic:

  * ``windows.Window`` / ``windows.transform`` / ``array_bounds`` (rasterio's
    documented affine algebra), an in-memory raster ``src.read``, and a
    ``modifying_raster`` context manager that collects ``write_band`` calls;
  * ``LineString`` / ``MultiLineString`` with ``.coords`` / ``.length`` /
    ``.geoms`` (shapely; GEOS length = sequential sum of segment lengths);
  * a serial ``pool.starmap``; ``utils.estimate_bounds_utm -> None``
    (projected CRS, ocsmesh/utils.py:1361).

and runs them on small seeded inputs.  Inputs AND outputs are stored, so the
fixtures pin (a) the numpy/scipy oracle (``-m "not gpu"``) and (b) the CUDA
path (``-m gpu``) against what the reference's code computes with this image's
numpy / scipy.

Run from the repository root:  python tests/golden/make_golden.py
(needs /root/reference; the tests only need the committed .npz files).
"""
from __future__ import annotations

import ast
import contextlib
import itertools
import logging
import math
import os
import sys
import textwrap
import warnings
from time import time

import numpy as np
from scipy.interpolate import RectBivariateSpline, RegularGridInterpolator
from scipy.spatial import cKDTree

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("OCSMESH_REFERENCE", "/root/reference")


# ---------------------------------------------------------------------------
# lifting reference source
# ---------------------------------------------------------------------------
def ref_source(relpath: str, name: str, cls: str | None = None) -> str:
    """Source of function ``name`` (or method ``cls.name``) without decorators."""
    path = os.path.join(REF, relpath)
    text = open(path).read()
    tree = ast.parse(text)
    body = tree.body
    if cls is not None:
        body = next(n for n in body if isinstance(n, ast.ClassDef) and n.name == cls).body
    fn = next(n for n in body if isinstance(n, ast.FunctionDef) and n.name == name)
    lines = text.splitlines()
    src = "\n".join(lines[fn.lineno - 1:fn.end_lineno])   # starts at 'def'
    return textwrap.dedent(src)


def ref_function(relpath: str, name: str, namespace: dict, cls: str | None = None):
    src = ref_source(relpath, name, cls)
    # annotations (rasterio / shapely / typing names) are never evaluated
    src = "from __future__ import annotations\n" + src
    code = compile(src, f"{REF}/{relpath}:{name}", "exec")
    exec(code, namespace)
    return namespace[name]


# ---------------------------------------------------------------------------
# stand-ins for rasterio / shapely / multiprocessing shells
# ---------------------------------------------------------------------------
class Window:
    def __init__(self, col_off, row_off, width, height):
        self.col_off, self.row_off = col_off, row_off
        self.width, self.height = width, height

    def astuple(self):
        return (self.col_off, self.row_off, self.width, self.height)


class Affine(tuple):
    def __new__(cls, a, b, c, d, e, f):
        return super().__new__(cls, (a, b, c, d, e, f))


class _windows_mod:
    Window = Window

    @staticmethod
    def transform(window, transform):
        # rasterio.windows.transform: x, y = T * (col_off, row_off);
        # Affine.translation(x - T.c, y - T.f) * T
        a, b, c, d, e, f = transform
        x = a * window.col_off + b * window.row_off + c
        y = d * window.col_off + e * window.row_off + f
        return Affine(a, b, c + (x - c), d, e, f + (y - f))


class _rasterio_mod:
    windows = _windows_mod


def array_bounds(height, width, transform):
    # rasterio.transform.array_bounds -> (west, south, east, north)
    a, b, c, d, e, f = transform
    return (c, width * d + height * e + f, width * a + height * b + c, f)


class LineString:
    def __init__(self, coords):
        self.coords = [tuple(map(float, c)) for c in np.asarray(coords)]

    @property
    def length(self):
        c = np.asarray(self.coords)
        if len(c) < 2:
            return 0.0
        seg = ((c[1:] - c[:-1]) ** 2).sum(axis=1) ** 0.5
        tot = 0.0
        for s in seg.tolist():          # GEOS: sequential accumulation
            tot += s
        return tot


class MultiLineString:
    def __init__(self, lines):
        self.geoms = [l if isinstance(l, LineString) else LineString(l) for l in lines]


class SerialPool:
    _processes = 1

    def starmap(self, fn, args):
        return list(itertools.starmap(fn, args))


class _Crs:
    is_geographic = False

    def equals(self, other):
        return True


class _Src:
    def __init__(self, bands, mask, transform):
        self.bands, self.mask, self.transform = bands, mask, transform
        self.crs = _Crs()

    @property
    def shape(self):
        return self.bands.shape[1:]

    def read(self, i=1, window=None, masked=False, **_kw):
        sl = (slice(window.row_off, window.row_off + window.height),
              slice(window.col_off, window.col_off + window.width))
        vals = np.array(self.bands[i - 1][sl])
        if masked:
            m = (np.zeros(vals.shape, bool) if self.mask is None
                 else np.array(self.mask[i - 1][sl]))
            return np.ma.MaskedArray(vals, mask=m)
        return vals


def make_raster_class():
    """A Raster whose coordinate / window methods ARE the reference's."""
    ns = {"np": np, "windows": _windows_mod, "rasterio": _rasterio_mod,
          "array_bounds": array_bounds, "math": math}
    ns["get_iter_windows"] = ref_function("ocsmesh/raster.py", "get_iter_windows", ns)
    methods = {}
    for name in ("get_x", "get_y", "get_xy", "get_values", "iter_windows",
                 "get_window_bounds", "get_window_transform"):
        methods[name] = ref_function("ocsmesh/raster.py", name, ns, cls="Raster")

    class RefRaster:
        registry = {}

        def __init__(self, bands, transform, chunk_size=None, overlap=None,
                     mask=None):
            bands = np.asarray(bands)
            if bands.ndim == 2:
                bands = bands[None]
            if mask is not None and np.asarray(mask).ndim == 2:
                mask = np.asarray(mask)[None]
            self.src = _Src(bands, mask, Affine(*transform))
            self.chunk_size = chunk_size
            self.overlap = 0 if overlap is None else overlap
            self.crs = self.src.crs
            self.tmpfile = f"mem://{id(self)}"
            RefRaster.registry[self.tmpfile] = self

        transform = property(lambda s: s.src.transform)
        width = property(lambda s: s.src.shape[1])
        height = property(lambda s: s.src.shape[0])
        shape = property(lambda s: s.src.shape)
        dx = property(lambda s: s.transform[0])       # ocsmesh/raster.py:1798
        dy = property(lambda s: -s.transform[4])      # ocsmesh/raster.py:1803

    for k, v in methods.items():
        setattr(RefRaster, k, v)
    return RefRaster, ns["get_iter_windows"]


class _Utils:
    """ocsmesh.utils as far as the hot path calls it."""

    @staticmethod
    def estimate_bounds_utm(bounds, crs="EPSG:4326"):
        return None     # ocsmesh/utils.py:1361: only geographic CRSs get a UTM


class _Dst:
    def __init__(self, out):
        self.out = out

    def write_band(self, band, values, window=None):
        assert band == 1
        r0, c0 = window.row_off, window.col_off
        self.out[r0:r0 + window.height, c0:c0 + window.width] = np.ma.getdata(values)
        if isinstance(values, np.ma.MaskedArray) and np.ma.is_masked(values):
            self.masked_cells += int(np.ma.getmaskarray(values).sum())

    masked_cells = 0


def make_hfun_class(RefRaster):
    """An HfunRaster whose add_feature / add_subtidal_flow_limiter /
    apply_constraints bodies ARE the reference's (decorators stripped:
    apply_constraints_wrap is re-applied by hand in ``_wrap``)."""
    ns_u = {"np": np, "LineString": LineString}
    _Utils.repartition_features = staticmethod(
        ref_function("ocsmesh/utils.py", "repartition_features", ns_u))
    _Utils.transform_linestring = staticmethod(
        ref_function("ocsmesh/utils.py", "transform_linestring", ns_u))

    import functools
    import gc
    import operator
    logger = logging.getLogger("golden")
    ns = {"np": np, "utils": _Utils, "LineString": LineString,
          "MultiLineString": MultiLineString, "cKDTree": cKDTree,
          "functools": functools, "operator": operator, "time": time,
          "_logger": logger, "warnings": warnings, "Transformer": None,
          "ops": None, "gc": gc}
    m_feat = ref_function("ocsmesh/hfun/raster.py", "add_feature", ns, cls="HfunRaster")
    m_lim = ref_function("ocsmesh/hfun/raster.py", "add_subtidal_flow_limiter", ns,
                         cls="HfunRaster")
    m_con = ref_function("ocsmesh/hfun/raster.py", "apply_constraints", ns,
                         cls="HfunRaster")

    class RefHfun:
        """hfun/raster.py:57-86, 197-251 restated as an in-memory holder."""

        def __init__(self, raster, hmin=None, hmax=None):
            self.raster = raster
            self._hmin = float(hmin) if hmin is not None else hmin
            self._hmax = float(hmax) if hmax is not None else hmax
            self.values = np.full(raster.shape, np.finfo(np.float32).max, np.float32)
            self.chunk_size, self.overlap = raster.chunk_size, raster.overlap
            self.crs = raster.crs
            self.src = raster.src      # only .crs / .shape are touched
            self.masked_cells = 0

        hmin = property(lambda s: s._hmin)
        hmax = property(lambda s: s._hmax)
        dx = property(lambda s: s.raster.dx)
        dy = property(lambda s: s.raster.dy)
        width = property(lambda s: s.raster.width)
        height = property(lambda s: s.raster.height)
        transform = property(lambda s: s.raster.transform)

        def dtype(self, i):
            return np.float32

        # coordinates / windows: the reference Raster methods (HfunRaster
        # inherits them)
        iter_windows = RefRaster.iter_windows
        get_window_bounds = RefRaster.get_window_bounds
        get_window_transform = RefRaster.get_window_transform
        get_x, get_y, get_xy = RefRaster.get_x, RefRaster.get_y, RefRaster.get_xy

        def get_values(self, window=None, band=None, **kw):
            if window is None:
                return self.values.copy()
            return self.values[window.row_off:window.row_off + window.height,
                               window.col_off:window.col_off + window.width].copy()

        @contextlib.contextmanager
        def modifying_raster(self, **_kw):
            # raster.py:366-424: writes go to a NEW file (reads keep seeing the
            # old one), swapped in on exit
            dst = _Dst(self.values.copy())
            yield dst
            self.values = dst.out
            self.masked_cells += dst.masked_cells

        def _wrap(self, fn, *a, **k):
            # features/constraint.py:420-423
            rv = fn(self, *a, **k)
            m_con(self, [])
            return rv

        def add_feature(self, feature, expansion_rate, target_size=None,
                        max_verts=200):
            return self._wrap(m_feat, feature, expansion_rate, target_size,
                              max_verts, pool=SerialPool())

        def add_subtidal_flow_limiter(self, hmin=None, hmax=None,
                                      lower_bound=None, upper_bound=None):
            return self._wrap(m_lim, hmin, hmax, lower_bound, upper_bound)

    return RefHfun


def make_interp(RefRaster):
    ns = {"np": np, "Raster": lambda path: RefRaster.registry[path],
          "RectBivariateSpline": RectBivariateSpline,
          "RegularGridInterpolator": RegularGridInterpolator,
          "Transformer": None}
    worker = ref_function("ocsmesh/mesh/mesh.py", "_mesh_interpolate_worker", ns)

    def interpolate(coords, values, rasters, method, band=1):
        """EuclideanMesh2D.interpolate, serial branch (mesh/mesh.py:365-376, 425)
        -- restated: the method body is Pool / pandas glue around these lines."""
        res = [worker(coords, _Crs(), r.tmpfile, r.chunk_size, method, False, band)
               for r in rasters]
        values = np.array(values, dtype=float).flatten()
        for (mask, _values) in res:
            values[mask] = _values
        return values.reshape((values.shape[0], 1)), res

    return worker, interpolate


# ---------------------------------------------------------------------------
# cases
# ---------------------------------------------------------------------------
WINDOW_CASES = [(20000, 20000, 5000, 0), (20000, 20000, 5000, 2), (70, 50, 30, 0),
                (70, 50, 30, 5), (100, 100, 100, 0), (101, 99, 25, 3),
                (8, 11, 3, 1), (2000, 2000, 512, 2), (7, 7, 10, 0)]

HFUN_CASES = [
    # name, chunk, overlap, hmin, hmax, rate, target
    ("c1_full", None, None, 100.0, 1500.0, 0.001, 100.0),
    ("c1_chunk50", 50, 0, 100.0, 1500.0, 0.001, 100.0),
    ("c1_chunk64_ovl3", 64, 3, 100.0, 1500.0, 0.001, 100.0),
    ("small_radius", None, None, 20.0, 60.0, 0.05, 20.0),
    ("no_hmax", 40, 2, 30.0, None, 0.01, None),
]


def main():
    from ocsmesh_b200 import synthetic as S
    RefRaster, ref_get_iter_windows = make_raster_class()
    RefHfun = make_hfun_class(RefRaster)
    worker, interpolate = make_interp(RefRaster)
    out = {}

    # ---- windows (raster.py:1891-1939) -------------------------------------
    for k, (w, h, c, o) in enumerate(WINDOW_CASES):
        wins = [x.astuple() for x in ref_get_iter_windows(w, h, c, o)]
        out[f"win_{k}_args"] = np.array([w, h, c, o])
        out[f"win_{k}"] = np.array(wins, dtype=np.int64)

    # ---- feature prologue (utils.py:247-312) --------------------------------
    rng = np.random.default_rng(7)
    t = np.linspace(0, 1, 523)
    line = np.stack([S.X0 + 9000 * t + 150 * np.sin(37 * t),
                     S.Y0 + 2500 * np.sin(5 * t) + 80 * rng.normal(size=t.size).cumsum() / 20],
                    axis=1)
    chunks = _Utils.repartition_features(LineString(line), 200)
    out["pro_line"] = line
    out["pro_chunk_lens"] = np.array([len(c.coords) for c in chunks])
    res = [np.array(_Utils.transform_linestring(c, 100.0).coords) for c in chunks]
    out["pro_points"] = np.concatenate(res)
    short = LineString(line[:3])          # shorter than target: unchanged
    out["pro_short"] = np.array(_Utils.transform_linestring(short, 1e6).coords)

    # ---- coordinates (raster.py:428-500) ------------------------------------
    H, W = 96, 128
    z = S.synth_dem(H, W, seed=31)
    tr = S.dem_transform(H)
    r = RefRaster(z, tr)
    out["dem"] = z
    out["dem_transform"] = np.array(tr)
    out["xy_full"] = r.get_xy()
    win = Window(17, 9, 50, 33)
    out["xy_win_args"] = np.array(win.astuple())
    out["x_win"], out["y_win"] = r.get_x(win), r.get_y(win)
    out["bounds_win"] = np.array(r.get_window_bounds(win))

    # ---- hfun ops (hfun/raster.py:1084-1261, 1315-1436, 509-568) -----------
    lines = S.synth_shoreline(H, W, seed=31, n_lines=2, spacing=25.0)
    for k, ln in enumerate(lines):
        out[f"feat_line_{k}"] = ln
    out["feat_n_lines"] = np.array(len(lines))
    mls = MultiLineString(lines)
    for name, chunk, ovl, hmin, hmax, rate, target in HFUN_CASES:
        rr = RefRaster(z, tr, chunk_size=chunk, overlap=ovl)
        hf = RefHfun(rr, hmin, hmax)
        hf.add_feature(mls, rate, target)
        out[f"hfun_{name}_feature"] = hf.values.copy()
        hf.add_subtidal_flow_limiter(50.0 if hmin else None, hmax, -30.0, 0.0)
        out[f"hfun_{name}_limiter"] = hf.values.copy()
        hf.add_subtidal_flow_limiter(None, None, None, None)
        out[f"hfun_{name}_gradient"] = hf.values.copy()
        assert hf.masked_cells == 0, "masked cells written: outside the oracle"
        out[f"hfun_{name}_args"] = np.array(
            [np.nan if v is None else v for v in (chunk, ovl, hmin, hmax, rate, target)],
            dtype=float)
    # limiter on a DEM with a plateau (zero gradient -> hmin)
    zp = z.copy()
    zp[20:40, 30:70] = -3.0
    hf = RefHfun(RefRaster(zp, tr), 100.0, 1500.0)
    hf.add_subtidal_flow_limiter(50.0, 1500.0, -30.0, 0.0)
    assert hf.masked_cells == 0
    out["dem_plateau"] = zp
    out["hfun_plateau_limiter"] = hf.values.copy()

    # ---- interpolation (mesh/mesh.py:2232-2349, 365-376) --------------------
    h2, w2 = 60, 80
    z1 = S.synth_dem(h2, w2, seed=33)
    z2 = S.synth_dem(h2, w2, seed=34)
    tr1 = S.dem_transform(h2)
    tr2 = list(tr1)
    tr2[2] += 300.0            # 30 px east: overlaps z1
    m2 = np.random.default_rng(35).random((h2, w2)) < 0.1
    z2m = z2.copy()
    z2m[m2] = -99999.0
    ra = RefRaster(z1, tr1)
    rb = RefRaster(z2m, tr2, mask=m2)
    x0, y0, x1, y1 = ra.get_window_bounds(Window(0, 0, w2, h2))
    xy = S.synth_nodes(3000, (x0 - 50, y0 - 50, x1 + 350, y1 + 50), seed=36)
    xs, ys = ra.get_x(), ra.get_y()
    extra = np.array([(x0, y0), (x1, y1), (xs[3], ys[5]), (xs[-1], ys[0]),
                      (0.5 * (xs[3] + xs[4]), 0.5 * (ys[5] + ys[6])),
                      (xs[10], 0.5 * (ys[7] + ys[8]))])
    xy = np.concatenate([xy, extra])
    out["itp_dem1"], out["itp_dem2"], out["itp_mask2"] = z1, z2m, m2
    out["itp_tr1"], out["itp_tr2"] = np.array(tr1), np.array(tr2)
    out["itp_xy"] = xy
    init = np.full(len(xy), -7.5)
    for method in ("linear", "nearest", "spline"):
        rasters = [ra, rb] if method != "spline" else [ra, RefRaster(z2, tr2)]
        vals, res = interpolate(xy, init, rasters, method)
        out[f"itp_{method}_values"] = vals
        for k, (mask, v) in enumerate(res):
            out[f"itp_{method}_mask{k}"] = mask
            out[f"itp_{method}_vals{k}"] = np.asarray(v)
    out["itp_dem2_nomask"] = z2

    path = os.path.join(HERE, "reference_golden.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1e3:.0f} kB "
          f"(numpy {np.__version__}, scipy {__import__('scipy').__version__})")


if __name__ == "__main__":
    main()
