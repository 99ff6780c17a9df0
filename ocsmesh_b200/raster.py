"""``Raster`` -- host-side mirror of ``ocsmesh.Raster`` for the accelerated path
(reference ``ocsmesh/raster.py``: windows ``:1577-1608, 1891-1939``,
coordinates ``:428-500``, values ``:502-531``, window bounds/transform
``:1642-1683``).

Window bookkeeping and coordinate conventions are integer / f64 host code and
stay in Python (delegating the window list to ``ocsb_windows`` in the C-ABI
library); pixel data are uploaded to the GPU once per raster band and cached as
a device tensor that the kernels read in place.

rasterio / pyproj / shapely are not dependencies: ``Window``, ``Affine`` and
``CRS`` below are small stand-ins with the attributes the path touches.
"""
from __future__ import annotations

import os
import pathlib
import tempfile
from collections import namedtuple
from typing import Any, Generator, List, Optional, Tuple, Union

import numpy as np

from . import _lib
from . import geotiff
from . import features as _features

#: stand-in for ``rasterio.windows.Window``
Window = namedtuple("Window", ["col_off", "row_off", "width", "height"])


class Affine(tuple):
    """Six-coefficient affine ``(a, b, c, d, e, f)``:
    ``x = a*col + b*row + c``, ``y = d*col + e*row + f`` (rasterio order)."""

    def __new__(cls, a, b, c, d, e, f):
        return super().__new__(cls, (float(a), float(b), float(c), float(d),
                                     float(e), float(f)))

    a = property(lambda s: s[0])
    b = property(lambda s: s[1])
    c = property(lambda s: s[2])
    d = property(lambda s: s[3])
    e = property(lambda s: s[4])
    f = property(lambda s: s[5])

    def __mul__(self, colrow):
        col, row = colrow
        return (self[0] * col + self[1] * row + self[2],
                self[3] * col + self[4] * row + self[5])

    @classmethod
    def from_gdal(cls, c, a, b, f, d, e):
        return cls(a, b, c, d, e, f)


class CRS:
    """Minimal CRS descriptor (EPSG code + geographic flag).

    The path only needs ``is_geographic`` (``ocsmesh/utils.py:1357-1382``:
    projected rasters take no UTM detour) and equality."""

    def __init__(self, epsg: Optional[int] = None, is_geographic: Optional[bool] = None):
        self.epsg = None if epsg is None else int(epsg)
        if is_geographic is None:
            is_geographic = self.epsg in (4326, 4269, 4267, 4258, 4979)
        self.is_geographic = bool(is_geographic)

    @classmethod
    def from_epsg(cls, code: int) -> "CRS":
        return cls(int(code))

    @classmethod
    def from_user_input(cls, value) -> "CRS":
        if isinstance(value, CRS):
            return value
        if value is None:
            return cls(None, False)
        if isinstance(value, int):
            return cls(value)
        if isinstance(value, str):
            s = value.strip().lower()
            if s.startswith("epsg:"):
                return cls(int(s.split(":")[1]))
            if s.isdigit():
                return cls(int(s))
        if hasattr(value, "to_epsg") and hasattr(value, "is_geographic"):
            return cls(value.to_epsg(), value.is_geographic)
        raise ValueError(f"unsupported CRS specification {value!r} "
                         "(pyproj is not available: use an EPSG code)")

    def to_epsg(self):
        return self.epsg

    def to_string(self):
        return f"EPSG:{self.epsg}" if self.epsg is not None else "None"

    def equals(self, other) -> bool:
        return self == other

    def __eq__(self, other):
        if other is None:
            return self.epsg is None
        try:
            o = CRS.from_user_input(other)
        except ValueError:
            return False
        return self.epsg == o.epsg and self.is_geographic == o.is_geographic

    def __hash__(self):
        return hash((self.epsg, self.is_geographic))

    def __repr__(self):
        return f"<CRS {self.to_string()}>"


def get_iter_windows(width: int, height: int, chunk_size: int = 0,
                     overlap: int = 0, row_off: int = 0, col_off: int = 0
                     ) -> Generator[Window, None, None]:
    """Square chunk windows of a raster, row-major
    (reference ``ocsmesh/raster.py:1891-1939``; ``row_off``/``col_off`` are
    accepted and unused there as well).  The list comes from ``ocsb_windows``
    in libocsb200."""
    del row_off, col_off
    if chunk_size is None or chunk_size is False:
        chunk_size = 0
    if chunk_size < 0:
        raise ValueError("Argument chunk_size must be >= 0.")
    for c, r, w, h in _lib.windows(int(width), int(height), int(chunk_size),
                                   int(overlap or 0)):
        yield Window(int(c), int(r), int(w), int(h))


def window_transform(window: Window, transform: Affine) -> Affine:
    """``rasterio.windows.transform`` (used at ``ocsmesh/raster.py:1683``)."""
    a, b, c, d, e, f = transform
    x, y = transform * (window.col_off, window.row_off)
    return Affine(a, b, c + (x - c), d, e, f + (y - f))


def array_bounds(height: int, width: int, transform: Affine):
    """``rasterio.transform.array_bounds`` (``ocsmesh/raster.py:1659-1662``):
    ``(west, south, east, north)``."""
    a, b, c, d, e, f = transform
    return (c, width * d + height * e + f, width * a + height * b + c, f)


class _Source:
    """What ``Raster.src`` exposes of a rasterio dataset on this path."""

    def __init__(self, bands: np.ndarray, transform: Affine, crs: CRS,
                 nodata: Optional[float], mask: Optional[np.ndarray],
                 row_origin: int = 0, full_height: Optional[int] = None):
        self.bands = bands
        self.transform = transform
        self.crs = crs
        self.nodata = nodata
        self._mask = mask
        # a rank of a multi-GPU job may hold only rows [row_origin,
        # row_origin + bands.shape[1]) of a raster `full_height` rows tall
        self.row_origin = int(row_origin)
        self.full_height = bands.shape[1] if full_height is None else int(full_height)

    @property
    def count(self):
        return self.bands.shape[0]

    @property
    def height(self):
        return self.full_height

    @property
    def width(self):
        return self.bands.shape[2]

    @property
    def shape(self):
        return (self.height, self.width)

    def _rows(self, window):
        a = window.row_off - self.row_origin
        b = a + window.height
        if a < 0 or b > self.bands.shape[1]:
            raise IndexError(
                f"rows [{window.row_off}, {window.row_off + window.height}) are "
                f"not held by this band raster (rows [{self.row_origin}, "
                f"{self.row_origin + self.bands.shape[1]}))")
        return slice(a, b)

    @property
    def dtypes(self):
        return tuple(str(self.bands.dtype) for _ in range(self.count))

    @property
    def nodatavals(self):
        return tuple(self.nodata for _ in range(self.count))

    def band_mask(self, i: int) -> Optional[np.ndarray]:
        """Boolean (h, w) nodata mask of band ``i`` (1-based) or None."""
        if self._mask is not None:
            return self._mask[i - 1]
        if self.nodata is None:
            return None
        b = self.bands[i - 1]
        if np.isnan(self.nodata):
            return np.isnan(b)
        return b == self.bands.dtype.type(self.nodata)

    def read(self, i: Optional[int] = None, window: Optional[Window] = None,
             masked: bool = False, **_kwargs):
        if window is None:
            window = Window(0, 0, self.width, self.height)
        sl = (self._rows(window),
              slice(window.col_off, window.col_off + window.width))
        if i is None:
            vals = np.array(self.bands[(slice(None),) + sl])
            if masked:
                ms = [self.band_mask(k + 1) for k in range(self.count)]
                m = np.stack([np.zeros(vals.shape[1:], bool) if mk is None
                              else mk[sl] for mk in ms])
                return np.ma.MaskedArray(vals, mask=m)
            return vals
        vals = np.array(self.bands[i - 1][sl])
        if masked:
            mk = self.band_mask(i)
            m = np.zeros(vals.shape, bool) if mk is None else np.array(mk[sl])
            return np.ma.MaskedArray(vals, mask=m)
        return vals


class Raster:
    """Wrapper class for basic raster handling (mirror of
    ``ocsmesh.Raster``, ``ocsmesh/raster.py:213-357``).

    Parameters are the reference's: ``path``, ``crs``, ``chunk_size``,
    ``overlap``.  ``Raster.from_numpy`` builds one without a file.
    """

    def __init__(self, path: Union[str, os.PathLike],
                 crs: Union[str, CRS, int, None] = None,
                 chunk_size: Optional[int] = None,
                 overlap: Optional[int] = None):
        self.chunk_size = chunk_size
        self.overlap = overlap
        self._path = pathlib.Path(path)
        info, bands = geotiff.read_geotiff(self._path)
        file_crs = CRS(info.epsg, info.is_geographic)
        self._crs = CRS.from_user_input(crs) if crs is not None else file_crs
        self._src = _Source(bands, Affine(*info.transform), self._crs,
                            info.nodata, None)
        self._dev_cache = {}
        self._dev_origin = {}

    @classmethod
    def from_numpy(cls, values, transform, crs=None, chunk_size=None,
                   overlap=None, nodata=None, mask=None, row_origin: int = 0,
                   full_height: Optional[int] = None) -> "Raster":
        """In-memory raster: ``values`` is ``(h, w)`` or ``(bands, h, w)``,
        ``transform`` the affine six-tuple ``(a, b, c, d, e, f)`` of the FULL
        raster.  ``row_origin`` / ``full_height``: ``values`` holds only rows
        ``[row_origin, row_origin + h)`` of a raster ``full_height`` rows tall
        (a rank's band of a row-partitioned raster)."""
        self = cls.__new__(cls)
        self.chunk_size = chunk_size
        self.overlap = overlap
        self._path = None
        values = np.asarray(values)
        if isinstance(values, np.ma.MaskedArray):
            if mask is None and values.mask is not np.ma.nomask:
                mask = np.ma.getmaskarray(values)
            values = np.ma.getdata(values)
        if values.ndim == 2:
            values = values[None]
        if values.ndim != 3:
            raise ValueError("Invalid data dimensions!")
        if mask is not None:
            mask = np.asarray(mask, dtype=bool)
            if mask.ndim == 2:
                mask = mask[None]
        self._crs = CRS.from_user_input(crs)
        self._src = _Source(values, Affine(*transform), self._crs, nodata, mask,
                            row_origin, full_height)
        self._dev_cache = {}
        self._dev_origin = {}
        return self

    # ---- descriptors (ocsmesh/raster.py:80-212) ---------------------------
    @property
    def chunk_size(self):
        return self._chunk_size

    @chunk_size.setter
    def chunk_size(self, val):
        if val is not None:
            if not isinstance(val, (int, np.integer)) or isinstance(val, bool):
                raise TypeError("Argument chunk_size must be an int or None.")
            if val < 0:
                raise ValueError("Argument chunk_size must be >= 0.")
        self._chunk_size = val

    @property
    def overlap(self):
        return self._overlap

    @overlap.setter
    def overlap(self, val):
        self._overlap = 0 if val is None else int(val)

    @property
    def path(self):
        return self._path

    @property
    def tmpfile(self):
        return self._path

    @property
    def src(self) -> _Source:
        return self._src

    @property
    def crs(self) -> CRS:
        return self._crs

    @property
    def transform(self) -> Affine:
        return self._src.transform

    @property
    def width(self) -> int:
        return self._src.width

    @property
    def height(self) -> int:
        return self._src.height

    @property
    def shape(self) -> Tuple[int, int]:
        return self._src.shape

    @property
    def count(self) -> int:
        return self._src.count

    @property
    def dx(self) -> float:
        return self.transform[0]          # ocsmesh/raster.py:1795-1798

    @property
    def dy(self) -> float:
        return -self.transform[4]         # ocsmesh/raster.py:1800-1803

    @property
    def nodata(self):
        return self._src.nodata

    @property
    def nodatavals(self):
        return self._src.nodatavals

    @property
    def dtypes(self):
        return self._src.dtypes

    @property
    def x(self):
        return self.get_x()

    @property
    def y(self):
        return self.get_y()

    @property
    def values(self):
        return self.get_values()

    @property
    def bbox(self):
        return self.get_bbox()

    def get_bbox(self, crs=None):
        """(xmin, ymin, xmax, ymax) of the raster (``ocsmesh/raster.py``
        ``get_bbox`` returns a shapely box; a tuple here)."""
        if crs is not None and CRS.from_user_input(crs) != self.crs:
            raise NotImplementedError("CRS transformation needs pyproj")
        x0, y0, x1, y1 = self.get_window_bounds(
            Window(0, 0, self.width, self.height))
        return (min(x0, x1), min(y0, y1), max(x0, x1), max(y0, y1))

    # ---- windows / coordinates --------------------------------------------
    def __iter__(self, chunk_size: int = None, overlap: int = None):
        for window in self.iter_windows(chunk_size, overlap):
            yield window, self.get_window_bounds(window)

    def iter_windows(self, chunk_size: Optional[int] = None,
                     overlap: Optional[int] = None
                     ) -> Generator[Window, None, None]:
        """``ocsmesh/raster.py:1577-1608``."""
        chunk_size = self.chunk_size if chunk_size is None else chunk_size
        overlap = self.overlap if overlap is None else overlap
        if chunk_size in [0, None, False]:
            yield Window(0, 0, self.width, self.height)
            return
        yield from get_iter_windows(self.width, self.height, chunk_size, overlap)

    def get_window_transform(self, window: Optional[Window]):
        if window is None:
            return None
        return window_transform(window, self.transform)

    def get_window_bounds(self, window: Window):
        return array_bounds(window.height, window.width,
                            self.get_window_transform(window))

    def _full(self, window):
        return Window(0, 0, self.width, self.height) if window is None else window

    def get_x(self, window: Optional[Window] = None) -> np.ndarray:
        window = self._full(window)
        x0, y0, x1, y1 = self.get_window_bounds(window)
        return np.linspace(x0, x1, window.width)

    def get_y(self, window: Optional[Window] = None) -> np.ndarray:
        window = self._full(window)
        x0, y0, x1, y1 = self.get_window_bounds(window)
        return np.linspace(y1, y0, window.height)

    def get_xy(self, window: Optional[Window] = None) -> np.ndarray:
        x, y = np.meshgrid(self.get_x(window), self.get_y(window))
        return np.vstack([x.flatten(), y.flatten()]).T

    def get_values(self, window: Optional[Window] = None,
                   band: Optional[int] = None, **kwargs: Any) -> np.ndarray:
        i = 1 if band is None else band
        return self._src.read(i, window=self._full(window), **kwargs)

    def get_xyz(self, window: Optional[Window] = None, band: Optional[int] = None
                ) -> np.ndarray:
        xy = self.get_xy(window)
        values = self.get_values(window=window, band=band).reshape((xy.shape[0], 1))
        return np.hstack([xy, values])

    def get_window_data(self, window: Window, masked: bool = True,
                        band: Optional[int] = None):
        x0, y0, x1, y1 = self.get_window_bounds(window)
        x = np.linspace(x0, x1, window.width)
        y = np.linspace(y1, y0, window.height)
        data = self._src.read(band, masked=masked, window=window)
        return x, y, data

    def _contour_window(self, window, device):
        """(ctx, ocsb_win restricted to the cell rows held on the device, DEM
        address, dtype code) for the contour calls; None when the device holds
        no cell row of the window."""
        from . import _lib
        window = self._full(window)
        ctx = _lib.default_context(device)
        dev = ctx.device
        dem = self.device_band(1, dev)
        origin = self.device_band_origin(1, dev)
        if getattr(self, "_halo_pending", {}).get((1, dev)):
            # a band raster of a multi-rank run (distributed.band_raster) whose
            # halo rows have not been exchanged yet: tracing now would run the
            # level against rows of zeros along the band edges
            raise RuntimeError(
                "the DEM band's halo rows have not been exchanged yet: call "
                "distributed.exchange_dem_halo (or BandExchange.exchange) before "
                "tracing contours on a band raster")
        code = _lib.OCSB_F32 if dem.element_size() == 4 else _lib.OCSB_F64
        w = self.width
        rows = dem.shape[0]
        # the rows of the window held on this device (a band-local raster
        # contributes the segments of its own cell rows)
        rb = max(0, origin - window.row_off)
        re = min(window.height, origin + rows - window.row_off)
        if re - rb < 2 or window.width < 2:
            return None
        # cell row j needs DEM rows j and j + 1: a band that stops short of
        # the window's last row gives up its last row of cells
        cw = _lib.make_win(self.get_window_bounds(window), window.width,
                           window.height, rb,
                           re - 1 if re < window.height else 0)
        addr = dem.data_ptr() + ((window.row_off - origin) * w
                                 + window.col_off) * dem.element_size()
        return ctx, cw, addr, code

    def get_contour(self, level: float, window: Optional[Window] = None,
                    device: Optional[int] = None, stitch: str = "device") -> List[np.ndarray]:
        """Iso-lines of band 1 at ``level`` as a list of ``(n, 2)`` polylines
        (stands in for ``Raster.get_contour`` ``ocsmesh/raster.py:1340-1372``,
        which returns a shapely (Multi)LineString traced by matplotlib; the
        tracer restates contourpy's documented rules, parity with it is
        unpinned: neither library is importable here).  Sweep, stitching and
        ordering all run on the device (``ocsb_contour_lines``);
        ``stitch="host"`` joins the device segments on the host instead
        (``ocsb_stitch_segments``: the same polylines, vertex for vertex)."""
        if window is None:
            wins = list(self.iter_windows())
            if len(wins) > 1:
                # chunked raster: the reference traces every window on its own
                # (own np.linspace coordinates, no cell between two windows) and
                # hands the pieces to shapely's linemerge
                # (_get_raster_contour_feathered, ocsmesh/raster.py:1471-1575).
                # Pieces of different windows only merge where end points
                # coincide exactly, which per-window linspaces do not produce:
                # the result is the list of the windows' polylines, in window
                # order
                lines = []
                for wn in wins:
                    lines.extend(self.get_contour(level, wn, device, stitch))
                return lines
        cw = self._contour_window(window, device)
        if cw is None:
            return []
        ctx, win, addr, code = cw
        if stitch == "host":
            edges, pts = ctx.contour_segments(win, addr, code, self.width, level)
            return _features.stitch_segments(edges, pts)
        xy, off = ctx.contour_lines(win, addr, code, self.width, level)
        return [xy[off[i]:off[i + 1]] for i in range(len(off) - 1)]

    def get_contour_points(self, levels, target_size: float, max_verts: int = 200,
                           window: Optional[Window] = None, device: Optional[int] = None):
        """The resampled vertex array of the contours at ``levels`` -- what
        ``HfunRaster.add_contour`` hands to the nearest-vertex search
        (``ocsmesh/hfun/raster.py:991-1002, 1219-1230``) -- as an ``(N, 2)``
        float64 CUDA tensor that never leaves the device: trace, stitch,
        repartition and arc-length resampling are all kernels
        (``ocsb_contour_points``).  Equal to ``feature_points(get_contour(l) for
        l in levels)`` bit for bit."""
        import torch
        if window is None:
            wins = list(self.iter_windows())
            if len(wins) > 1:
                # chunked raster: window by window, like get_contour; the
                # reference's add_contour flattens level by level, so the
                # levels are the outer loop
                lv = list(levels) if isinstance(levels, (list, tuple, np.ndarray)) \
                    else [levels]
                parts = [self.get_contour_points([l], target_size, max_verts, wn, device)
                         for l in lv for wn in wins]
                return torch.cat(parts, dim=0)
        cw = self._contour_window(window, device)
        if cw is None:
            dev = torch.cuda.current_device() if device is None else int(device)
            return torch.empty((0, 2), dtype=torch.float64, device=f"cuda:{dev}")
        ctx, win, addr, code = cw
        lv = list(levels) if isinstance(levels, (list, tuple, np.ndarray)) else [levels]
        return ctx.contour_points(win, addr, code, self.width, lv, target_size, max_verts)

    def save(self, path):
        epsg = self.crs.to_epsg() if self.crs is not None else None
        geotiff.write_geotiff(path, np.asarray(self._src.bands), self.transform,
                              epsg=epsg, is_geographic=self.crs.is_geographic,
                              nodata=self.nodata)

    # ---- device residency ---------------------------------------------------
    def device_band(self, band: int = 1, device=None):
        """Band ``band`` as a contiguous CUDA tensor (f32 stays f32, anything
        else becomes f64), uploaded once and cached."""
        import torch
        if device is None:
            device = torch.cuda.current_device()
        key = (int(band), int(device) if not isinstance(device, torch.device)
               else device.index)
        t = self._dev_cache.get(key)
        if t is None:
            host = np.asarray(self._src.bands[band - 1])
            if host.dtype != np.float32:
                host = host.astype(np.float64, copy=False)
            host = np.ascontiguousarray(host)
            t = torch.from_numpy(host).to(f"cuda:{key[1]}", non_blocking=False)
            self._dev_cache[key] = t
            self._dev_origin[key] = self._src.row_origin
        return t

    def device_band_origin(self, band: int = 1, device=None) -> int:
        """Raster row held by row 0 of ``device_band(band, device)``."""
        import torch
        if device is None:
            device = torch.cuda.current_device()
        return self._dev_origin.get((int(band), int(device)), self._src.row_origin)

    def set_device_band(self, tensor, row_origin: int, band: int = 1, device=None):
        """Install an already device-resident band (rows starting at raster row
        ``row_origin``), e.g. a rank's band with its halo rows."""
        import torch
        if device is None:
            device = tensor.device.index
        key = (int(band), int(device))
        self._dev_cache[key] = tensor
        if not hasattr(self, "_halo_pending"):
            self._halo_pending = {}
        self._halo_pending[key] = False
        self._dev_origin[key] = int(row_origin)

    def device_mask(self, band: int = 1, device=None):
        """uint8 nodata mask of a band on the device, or None."""
        import torch
        if device is None:
            device = torch.cuda.current_device()
        key = ("mask", int(band), int(device))
        if key in self._dev_cache:
            return self._dev_cache[key]
        m = self._src.band_mask(band)
        if m is None or not np.any(m):
            t = None
        else:
            t = torch.from_numpy(np.ascontiguousarray(m).astype(np.uint8)).to(
                f"cuda:{int(device)}")
        self._dev_cache[key] = t
        return t

    def drop_device_cache(self):
        self._dev_cache = {}
        self._dev_origin = {}


def raster_from_numpy(filename, data, mgrid, crs=4326) -> None:
    """``ocsmesh.utils.raster_from_numpy`` (``ocsmesh/utils.py:1932-1971``):
    write ``data`` (``(ny, nx)`` or ``(ny, nx, bands)``) on the grid ``mgrid``
    (ij-indexed ``np.mgrid``) as a GeoTIFF."""
    x = mgrid[0][:, 0]
    y = mgrid[1][0, :]
    res_x = (x[-1] - x[0]) / data.shape[1]
    res_y = (y[-1] - y[0]) / data.shape[0]
    transform = Affine(res_x, 0.0, x[0], 0.0, res_y, y[0])
    crs = CRS.from_user_input(crs)
    if data.ndim == 3:
        bands = np.moveaxis(np.ma.getdata(data), 2, 0)
    elif data.ndim == 2:
        bands = np.ma.getdata(data)[None]
    else:
        raise ValueError("Invalid data dimensions!")
    nodata = None
    if isinstance(data, np.ma.MaskedArray):
        nodata = float(data.fill_value)
        bands = np.moveaxis(data.filled(data.fill_value).reshape(
            data.shape[0], data.shape[1], -1), 2, 0)
    # a south-up (positive res_y) grid is stored as is: the window algebra
    # above handles either sign
    geotiff.write_geotiff(filename, bands, transform, epsg=crs.to_epsg(),
                          is_geographic=crs.is_geographic, nodata=nodata)


def temp_tiff_path(suffix=".tif") -> str:
    d = os.path.join(tempfile.gettempdir(), "ocsmesh_b200")
    os.makedirs(d, exist_ok=True)
    fd, p = tempfile.mkstemp(suffix=suffix, dir=d)
    os.close(fd)
    return p
