"""Minimal GeoTIFF / BigTIFF reader and writer (uncompressed strips).

rasterio/GDAL are what the reference uses for raster I/O
(``ocsmesh/raster.py:502-531``, ``ocsmesh/utils.py:1932-1971``); they are not
available to this package, and raster I/O is off the accelerated path, so this
module implements just the subset the path needs: north-up, uncompressed,
strip-organised float32 / float64 / int16 / int32 / uint8 rasters with the
GeoTIFF georeferencing tags (ModelPixelScale 33550, ModelTiepoint 33922,
ModelTransformation 34264, GeoKeyDirectory 34735) and GDAL_NODATA (42113).
Files above 4 GB are written as BigTIFF.  Pixel data are exposed as
``numpy.memmap`` so a 40000 x 40000 DEM never has to be resident on the host
twice.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

_TYPES = {1: ("B", 1), 2: ("c", 1), 3: ("H", 2), 4: ("I", 4), 5: ("II", 8),
          6: ("b", 1), 7: ("B", 1), 8: ("h", 2), 9: ("i", 4), 10: ("ii", 8),
          11: ("f", 4), 12: ("d", 8), 16: ("Q", 8), 17: ("q", 8), 18: ("Q", 8)}

_SAMPLE_DTYPES = {  # (sample_format, bits) -> numpy dtype
    (1, 8): "u1", (1, 16): "u2", (1, 32): "u4", (2, 8): "i1", (2, 16): "i2",
    (2, 32): "i4", (3, 32): "f4", (3, 64): "f8",
}


@dataclass
class GeoTiffInfo:
    width: int
    height: int
    count: int
    dtype: np.dtype
    transform: Tuple[float, float, float, float, float, float]  # a,b,c,d,e,f
    nodata: Optional[float]
    epsg: Optional[int]
    is_geographic: bool


def _read_ifd(f, bo: str, big: bool):
    if big:
        (n,) = struct.unpack(bo + "Q", f.read(8))
        esz, cnt_fmt, inline = 20, "Q", 8
    else:
        (n,) = struct.unpack(bo + "H", f.read(2))
        esz, cnt_fmt, inline = 12, "I", 4
    raw = f.read(n * esz)
    tags = {}
    for i in range(n):
        e = raw[i * esz:(i + 1) * esz]
        tag, typ = struct.unpack(bo + "HH", e[:4])
        (count,) = struct.unpack(bo + cnt_fmt, e[4:4 + inline])
        val = e[4 + inline:4 + 2 * inline]
        if typ not in _TYPES:
            continue
        fmt, size = _TYPES[typ]
        nbytes = size * count
        if nbytes <= inline:
            data = val[:nbytes]
        else:
            (off,) = struct.unpack(bo + cnt_fmt, val)
            pos = f.tell()
            f.seek(off)
            data = f.read(nbytes)
            f.seek(pos)
        if typ == 2:
            tags[tag] = data.rstrip(b"\0").decode("ascii", "replace")
        elif typ in (5, 10):
            v = struct.unpack(bo + fmt[0] * (2 * count), data)
            tags[tag] = tuple(v[2 * k] / v[2 * k + 1] if v[2 * k + 1] else 0.0
                              for k in range(count))
        else:
            tags[tag] = struct.unpack(bo + fmt * count, data)
    return tags


def read_geotiff(path) -> Tuple[GeoTiffInfo, np.ndarray]:
    """Open ``path``; returns ``(info, bands)`` with ``bands`` a ``(count, h,
    w)`` array (a read-only memmap whenever the strips are contiguous)."""
    path = os.fspath(path)
    with open(path, "rb") as f:
        head = f.read(4)
        bo = {b"II": "<", b"MM": ">"}.get(head[:2])
        if bo is None:
            raise ValueError(f"{path}: not a TIFF file")
        (magic,) = struct.unpack(bo + "H", head[2:4])
        if magic == 42:
            big = False
            (ifd,) = struct.unpack(bo + "I", f.read(4))
        elif magic == 43:
            big = True
            f.read(4)
            (ifd,) = struct.unpack(bo + "Q", f.read(8))
        else:
            raise ValueError(f"{path}: bad TIFF magic {magic}")
        f.seek(ifd)
        t = _read_ifd(f, bo, big)
    width, height = int(t[256][0]), int(t[257][0])
    spp = int(t.get(277, (1,))[0])
    bits = int(t.get(258, (1,))[0])
    sfmt = int(t.get(339, (1,))[0])
    if int(t.get(259, (1,))[0]) != 1:
        raise NotImplementedError(f"{path}: compressed TIFFs are not supported")
    if 322 in t or 324 in t:
        raise NotImplementedError(f"{path}: tiled TIFFs are not supported")
    key = (sfmt, bits)
    if key not in _SAMPLE_DTYPES:
        raise NotImplementedError(f"{path}: sample format {key}")
    dtype = np.dtype(bo + _SAMPLE_DTYPES[key])
    planar = int(t.get(284, (1,))[0])
    rps = int(t.get(278, (height,))[0])
    offsets = [int(o) for o in t[273]]
    counts = [int(c) for c in t[279]]
    contiguous = all(offsets[i] + counts[i] == offsets[i + 1]
                     for i in range(len(offsets) - 1))
    if contiguous:
        mm = np.memmap(path, dtype=dtype, mode="r", offset=offsets[0],
                       shape=(sum(counts) // dtype.itemsize,))
        if planar == 2 or spp == 1:
            bands = mm.reshape(spp, height, width)
        else:
            bands = np.moveaxis(mm.reshape(height, width, spp), 2, 0)
    else:
        with open(path, "rb") as f:
            chunks = []
            for o, c in zip(offsets, counts):
                f.seek(o)
                chunks.append(f.read(c))
        flat = np.frombuffer(b"".join(chunks), dtype=dtype)
        if planar == 2 or spp == 1:
            bands = flat.reshape(spp, height, width)
        else:
            bands = np.moveaxis(flat.reshape(height, width, spp), 2, 0)
    del rps
    # georeferencing
    if 34264 in t:
        m = t[34264]
        transform = (m[0], m[1], m[3], m[4], m[5], m[7])
    elif 33550 in t and 33922 in t:
        sx, sy = t[33550][0], t[33550][1]
        i, j, _, x, y, _ = t[33922][:6]
        transform = (sx, 0.0, x - i * sx, 0.0, -sy, y + j * sy)
    else:
        transform = (1.0, 0.0, 0.0, 0.0, 1.0, 0.0)
    epsg, geographic = None, False
    if 34735 in t:
        g = t[34735]
        for k in range(1, g[3] + 1):
            kid, loc, _cnt, val = g[4 * k:4 * k + 4]
            if loc != 0:
                continue
            if kid == 1024:
                geographic = (val == 2)
            elif kid == 3072 and val != 32767:
                epsg = val
            elif kid == 2048 and epsg is None and val != 32767:
                epsg = val
    nodata = None
    if 42113 in t:
        try:
            nodata = float(t[42113])
        except ValueError:
            nodata = None
    info = GeoTiffInfo(width, height, spp, dtype.newbyteorder("="),
                       tuple(float(v) for v in transform), nodata, epsg,
                       geographic)
    return info, bands


def write_geotiff(path, data: np.ndarray, transform, epsg: Optional[int] = None,
                  is_geographic: bool = False, nodata: Optional[float] = None):
    """Write ``data`` (``(h, w)`` or ``(count, h, w)``) as an uncompressed,
    single-strip-per-band, band-sequential (Big)TIFF."""
    data = np.asarray(data)
    if data.ndim == 2:
        data = data[None]
    count, height, width = data.shape
    dt = data.dtype.newbyteorder("<") if data.dtype.byteorder == ">" else data.dtype
    kind = {"f": 3, "i": 2, "u": 1}[dt.kind]
    bits = dt.itemsize * 8
    if (kind, bits) not in _SAMPLE_DTYPES:
        raise NotImplementedError(f"dtype {dt}")
    band_bytes = height * width * dt.itemsize
    big = band_bytes * count + 4096 >= (1 << 32) - 1
    a, b, c, d, e, f_ = (float(v) for v in transform)
    if b != 0.0 or d != 0.0:
        raise NotImplementedError("rotated rasters are not supported")
    geokeys = [1, 1, 0, 0]
    nkeys = 0
    model = 2 if is_geographic else 1
    keys = [(1024, 0, 1, model), (1025, 0, 1, 1)]
    if epsg is not None:
        keys.append((2048 if is_geographic else 3072, 0, 1, int(epsg)))
    for k in keys:
        geokeys.extend(k)
        nkeys += 1
    geokeys[3] = nkeys

    entries = []  # (tag, type, values)
    entries.append((256, 4, [width]))
    entries.append((257, 4, [height]))
    entries.append((258, 3, [bits] * count))
    entries.append((259, 3, [1]))
    entries.append((262, 3, [1]))
    offs_type = 16 if big else 4
    entries.append((273, offs_type, None))  # filled below
    entries.append((277, 3, [count]))
    entries.append((278, 4, [height]))
    entries.append((279, offs_type, [band_bytes] * count))
    entries.append((284, 3, [2]))
    entries.append((339, 3, [kind] * count))
    if e < 0 and a > 0:
        entries.append((33550, 12, [a, -e, 0.0]))
        entries.append((33922, 12, [0.0, 0.0, 0.0, c, f_, 0.0]))
    else:  # south-up / mirrored grids need the full matrix
        entries.append((34264, 12, [a, b, 0.0, c, d, e, 0.0, f_,
                                    0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 1.0]))
    entries.append((34735, 3, geokeys))
    if nodata is not None:
        entries.append((42113, 2, repr(float(nodata)).encode("ascii") + b"\0"))
    if count > 1:
        entries.append((338, 3, [0] * (count - 1)))
    entries.sort(key=lambda t: t[0])

    inline = 8 if big else 4
    esz = 20 if big else 12
    head = 16 if big else 8
    ifd_size = (8 if big else 2) + esz * len(entries) + (8 if big else 4)

    def pack_vals(typ, vals):
        if typ == 2:
            return bytes(vals)
        fmt = _TYPES[typ][0]
        return struct.pack("<" + fmt * len(vals), *vals)

    # layout: header | IFD | overflow values | pixel data (64-byte aligned)
    overflow_pos = head + ifd_size
    blobs = {}
    pos = overflow_pos
    for tag, typ, vals in entries:
        n = count if vals is None else len(vals)
        size = _TYPES[typ][1] * n
        if size > inline:
            blobs[tag] = pos
            pos += size + (size & 1)
    data_pos = (pos + 63) // 64 * 64
    strip_offsets = [data_pos + i * band_bytes for i in range(count)]

    with open(os.fspath(path), "wb") as f:
        if big:
            f.write(struct.pack("<2sHHHQ", b"II", 43, 8, 0, head))
        else:
            f.write(struct.pack("<2sHI", b"II", 42, head))
        f.write(struct.pack("<Q" if big else "<H", len(entries)))
        tail = []
        for tag, typ, vals in entries:
            if vals is None:
                vals = strip_offsets
            n = len(vals)
            payload = pack_vals(typ, vals)
            f.write(struct.pack("<HH", tag, typ))
            f.write(struct.pack("<Q" if big else "<I", n))
            if len(payload) <= inline:
                f.write(payload.ljust(inline, b"\0"))
            else:
                f.write(struct.pack("<Q" if big else "<I", blobs[tag]))
                tail.append((blobs[tag], payload))
        f.write(struct.pack("<Q" if big else "<I", 0))
        for off, payload in tail:
            f.seek(off)
            f.write(payload)
        f.seek(data_pos)
        for i in range(count):
            np.ascontiguousarray(data[i], dtype=dt).tofile(f)
