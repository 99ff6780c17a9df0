#!/usr/bin/env python
"""Golden fixtures for the fast collector's big raster,
produced by EXECUTING THE REFERENCE'S OWN SOURCE of
``HfunCollector._create_big_raster`` (ocsmesh/hfun/collector.py:2296-2442),
lifted with ``ast`` like tests/golden/make_golden.py does (nothing of the
reference is copied into the repository; only inputs and outputs are stored).

Stand-ins, all of them I/O shells or identities for rasters of ONE projected
CRS: ``Transformer.from_crs`` is the identity, ``utils.estimate_bounds_utm``
returns a token CRS, ``Polygon`` / ``ops.transform`` keep coordinate arrays
(``.bounds`` = min / max), ``from_origin`` is rasterio's documented affine,
``rasterio.open(..., 'w')`` records the profile and the fill writes,
``reproject`` records WHICH inputs are warped in which order (GDAL itself is
absent: the warp rule stays parity-unpinned), ``Raster`` records the chunk size
the result is opened with.  What the fixtures pin: bounds, resolution, shape,
transform, window size, fill value and the selection / order of warped inputs.

Run from the repository root:  python tests/golden/make_golden_bigraster.py
"""
from __future__ import annotations

import json
import logging
import math
import os
import sys
from pathlib import Path

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import Affine, Window, array_bounds, ref_function  # noqa: E402


class _Bounds:
    def __init__(self, b):
        self.bounds = tuple(float(v) for v in b)


class Polygon:
    def __init__(self, coords):
        self.coords = np.asarray(coords, dtype=np.float64)

    @property
    def bounds(self):
        c = self.coords
        return (c[:, 0].min(), c[:, 1].min(), c[:, 0].max(), c[:, 1].max())


class _ops:
    @staticmethod
    def transform(func, poly):
        x, y = func(poly.coords[:, 0], poly.coords[:, 1])
        return Polygon(np.stack([x, y], axis=1))


class Transformer:
    @classmethod
    def from_crs(cls, *_a, **_k):
        return cls()

    @staticmethod
    def transform(x, y):
        return x, y


class _utils:
    @staticmethod
    def estimate_bounds_utm(bounds, crs):
        return "the-one-projected-crs"


def from_origin(west, north, xsize, ysize):
    return Affine(xsize, 0.0, west, 0.0, -ysize, north)


class _Src:
    def __init__(self, shape, transform):
        self.shape = shape
        self.transform = Affine(*transform)


class _RasterIn:
    def __init__(self, shape, transform):
        self.src = _Src(shape, transform)


class HfunRaster:
    """What the method touches of an input: ``raster.src.shape / .transform`` and
    ``get_bbox(crs=)`` (``box(*raster.bounds)``: the pixel-edge bounds)."""

    def __init__(self, shape, transform):
        self.raster = _RasterIn(shape, transform)

    def get_bbox(self, crs=None):
        h, w = self.raster.src.shape
        return _Bounds(array_bounds(h, w, self.raster.src.transform))


class Recorder:
    def __init__(self):
        self.profile = None
        self.writes = []
        self.warped = []
        self.chunk_size = "unset"


def make_namespace(rec: Recorder):
    class _Dst:
        def __init__(self, profile):
            rec.profile = profile

        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

        def write(self, z, band, window=None):
            rec.writes.append((tuple(z.shape), float(z.flat[0]), str(z.dtype),
                               None if window is None else window.astuple()))

    class _rasterio:
        @staticmethod
        def open(path, mode, **profile):
            assert mode == "w"
            return _Dst(profile)

        @staticmethod
        def band(ds, i):
            return (ds, i)

    class Resampling:
        nearest = "nearest"

    def reproject(source, destination, resampling, init_dest_nodata, num_threads,
                  warp_mem_limit):
        assert resampling == "nearest" and init_dest_nodata is False
        rec.warped.append(source[0])

    class Raster:
        def __init__(self, path, chunk_size=None):
            rec.chunk_size = chunk_size
            self.path = path

    ns = {"np": np, "math": math, "Path": Path, "HfunRaster": HfunRaster, "utils": _utils,
          "Transformer": Transformer, "Polygon": Polygon, "ops": _ops,
          "from_origin": from_origin, "rasterio": _rasterio, "reproject": reproject,
          "Resampling": Resampling, "Raster": Raster, "windows": None,
          "_logger": logging.getLogger("golden")}

    class _windows_mod:
        pass
    _windows_mod.Window = Window
    ns["windows"] = _windows_mod
    ns["get_iter_windows"] = ref_function("ocsmesh/raster.py", "get_iter_windows", ns)
    ns["_create_big_raster"] = ref_function("ocsmesh/hfun/collector.py", "_create_big_raster",
                                            ns, cls="HfunCollector")
    return ns


class _Self:
    def __init__(self, hfuns, nprocs, hmin, const_val, flow_lim, constraints):
        self._hfun_list = hfuns
        self._nprocs = nprocs
        self._size_info = {"hmin": hmin, "hmax": None}
        self._const_val_contour_coll = const_val
        self._flow_lim_coll = flow_lim
        self._constraint_info_coll = constraints


CASES = [
    # name, [(shape, transform)], nprocs, hmin, const_val src_idx list, flow_lim src_idx list,
    # constraint src_idx list   (None = "all sources"; [] = no such refinement)
    ("two_side_by_side", [((100, 200), (10.0, 0, 500000.0, 0, -10.0, 4101000.0)),
                          ((100, 200), (10.0, 0, 501800.0, 0, -10.0, 4101000.0))],
     3, 50.0, [], [None], []),
    ("two_side_by_side_unchunked", [((100, 200), (10.0, 0, 500000.0, 0, -10.0, 4101000.0)),
                                    ((100, 200), (10.0, 0, 501800.0, 0, -10.0, 4101000.0))],
     6, 50.0, [[1]], [], []),
    ("mixed_resolution", [((260, 330), (10.0, 0, 500000.0, 0, -10.0, 4102600.0)),
                          ((130, 170), (25.0, 0, 502900.0, 0, -25.0, 4102300.0)),
                          ((80, 90), (7.5, 0, 499000.0, 0, -7.5, 4101000.0))],
     1, 30.0, [], [[0, 2]], [[1]]),
    ("hmin_below_pixel", [((64, 48), (30.0, 0, 1000.0, 0, -30.0, 9000.0))],
     64, 20.0, [], [], []),
    ("nothing_selected", [((40, 60), (10.0, 0, 0.0, 0, -10.0, 400.0)),
                          ((40, 60), (10.0, 0, 400.0, 0, -10.0, 400.0))],
     64, 30.0, [], [], []),
]


def run_case(case):
    name, rasters, nprocs, hmin, cv, fl, cs = case
    rec = Recorder()
    ns = make_namespace(rec)
    hfuns = [HfunRaster(s, t) for s, t in rasters]
    const_val = [((idx, None, None), 1.0) for idx in cv]
    flow_lim = [(idx, None, None, None, None) for idx in fl]
    constraints = [(idx, object()) for idx in cs]
    me = _Self(hfuns, nprocs, hmin, const_val, flow_lim, constraints)
    ns["_create_big_raster"](me, "/tmp/golden_bigraster")
    warped = [hfuns.index(next(h for h in hfuns if h.raster.src is src)) for src in rec.warped]
    return {
        "name": name,
        "rasters": [{"shape": list(s), "transform": [float(v) for v in t]} for s, t in rasters],
        "nprocs": nprocs, "hmin": hmin,
        "const_val_src": cv, "flow_lim_src": fl, "constraint_src": cs,
        "width": int(rec.profile["width"]), "height": int(rec.profile["height"]),
        "transform": [float(v) for v in rec.profile["transform"]],
        "dtype": np.dtype(rec.profile["dtype"]).name, "count": int(rec.profile["count"]),
        "chunk_size": None if rec.chunk_size is None else int(rec.chunk_size),
        "fill_writes": [{"shape": [int(v) for v in w[0]], "value": w[1], "dtype": w[2], "window": None if w[3] is None else [int(v) for v in w[3]]}
                        for w in rec.writes],
        "warped_in_order": warped,
    }


def main():
    out = [run_case(c) for c in CASES]
    path = os.path.join(HERE, "reference_golden_bigraster.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    for o in out:
        print(o["name"], o["width"], o["height"], o["transform"], o["chunk_size"],
              o["warped_in_order"], len(o["fill_writes"]))
    print("wrote", path)


if __name__ == "__main__":
    main()
