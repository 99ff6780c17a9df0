"""GeoTIFF / BigTIFF reader and writer.

rasterio/GDAL are what the reference uses for raster I/O
(``ocsmesh/raster.py:502-531``, ``ocsmesh/utils.py:1932-1971``); they are not
available to this package, and raster I/O is off the accelerated path, so this
module implements the subset DEM files use: north-up float32 / float64 / int16
/ int32 / uint8 rasters, organised in strips or tiles, uncompressed or with
DEFLATE (8, 32946: zlib), LZW (5) or PackBits (32773) (both decoded by
``ocsb_tiff_decode``, host C), predictors 1 / 2 (horizontal differencing) / 3
(floating point), chunky or planar samples, with the GeoTIFF georeferencing
tags (ModelPixelScale 33550, ModelTiepoint 33922, ModelTransformation 34264,
GeoKeyDirectory 34735) and GDAL_NODATA (42113).  Files above 4 GB are written
as BigTIFF.  Uncompressed contiguous strips are exposed as ``numpy.memmap`` so
a 40000 x 40000 DEM never has to be resident on the host twice; everything
else is decoded segment by segment into one array.  The writer produces
uncompressed single-strip files by default and tiled and / or DEFLATE files on
request (what GDAL's ``TILED=YES COMPRESS=DEFLATE PREDICTOR=n`` creates).
"""
from __future__ import annotations

import os
import struct
import zlib
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


def _decode(raw: bytes, compression: int, cap: int) -> np.ndarray:
    """One strip / tile -> its bytes (uint8 array)."""
    if compression == 1:
        return np.frombuffer(raw, dtype=np.uint8)
    if compression in (8, 32946):
        return np.frombuffer(zlib.decompress(raw), dtype=np.uint8)
    from . import _lib        # LZW / PackBits: host C in libocsb200
    return _lib.tiff_decode(compression, raw, cap)


def _unpredict(seg: np.ndarray, predictor: int, rows: int, cols: int, spp: int,
               dtype: np.dtype) -> np.ndarray:
    """Bytes of one segment -> ``(rows, cols, spp)`` samples in native order."""
    isz = dtype.itemsize
    native = dtype.newbyteorder("=")
    if predictor == 3:
        # TIFF Technical Note 3: per row, the bytes of all samples are grouped by
        # significance (most significant first) and differenced with stride spp
        b = seg.reshape(rows, cols * isz, spp)
        b = np.cumsum(b, axis=1, dtype=np.uint8).reshape(rows, isz, cols * spp)
        b = np.ascontiguousarray(np.moveaxis(b, 1, 2))
        return b.view(np.dtype(">f%d" % isz)).reshape(rows, cols, spp).astype(native)
    a = seg.view(dtype).reshape(rows, cols, spp)
    if predictor == 2:
        # horizontal differencing, modulo 2^bits (floats are differenced as the
        # integers of their width, which is how GDAL writes them)
        u = np.dtype("u%d" % isz)
        ai = a.astype(native).view(u)
        return np.cumsum(ai, axis=1, dtype=u).view(native)
    return a.astype(native, copy=False)


def _read_segments(path, offsets, counts, width, height, spp, dtype, planar, seg_w, seg_h,
                   tiled, compression, predictor) -> np.ndarray:
    """Strips (``seg_w == width``) or tiles -> ``(spp, height, width)``."""
    across = -(-width // seg_w)
    down = -(-height // seg_h)
    per_plane = across * down
    planes = spp if planar == 2 else 1
    seg_spp = 1 if planar == 2 else spp
    if len(offsets) < per_plane * planes:
        raise ValueError(f"{path}: {len(offsets)} segments, {per_plane * planes} expected")
    out = np.empty((spp, height, width), dtype=dtype.newbyteorder("="))

    def one(i: int) -> None:
        # segments are independent and write disjoint parts of `out`; zlib and
        # the C decoders release the GIL, so a thread per core decodes them
        p, k = divmod(i, per_plane)
        r0, c0 = (k // across) * seg_h, (k % across) * seg_w
        # tiles are stored whole; the last strip only holds the rows left
        rows = seg_h if tiled else min(seg_h, height - r0)
        full = rows * seg_w * seg_spp * dtype.itemsize
        with open(path, "rb") as f:
            f.seek(offsets[i])
            raw = f.read(counts[i])
        seg = _decode(raw, compression, full)
        if seg.size < full:
            raise ValueError(f"{path}: segment {i} is short ({seg.size} of {full} bytes)")
        a = _unpredict(seg[:full], predictor, rows, seg_w, seg_spp, dtype)
        nr, nc = min(rows, height - r0), min(seg_w, width - c0)
        if planar == 2:
            out[p, r0:r0 + nr, c0:c0 + nc] = a[:nr, :nc, 0]
        else:
            out[:, r0:r0 + nr, c0:c0 + nc] = np.moveaxis(a[:nr, :nc], 2, 0)

    n_seg = per_plane * planes
    workers = min(n_seg, os.cpu_count() or 1, 32)
    if workers <= 1 or compression == 1:
        for i in range(n_seg):
            one(i)
    else:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=workers) as ex:
            for _ in ex.map(one, range(n_seg)):
                pass
    return out


def read_geotiff(path) -> Tuple[GeoTiffInfo, np.ndarray]:
    """Open ``path``; returns ``(info, bands)`` with ``bands`` a ``(count, h,
    w)`` array (a read-only memmap whenever the file is uncompressed with
    contiguous strips)."""
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
    compression = int(t.get(259, (1,))[0])
    if compression not in (1, 5, 8, 32946, 32773):
        raise NotImplementedError(f"{path}: TIFF compression {compression} is not supported")
    predictor = int(t.get(317, (1,))[0])
    if predictor not in (1, 2, 3):
        raise NotImplementedError(f"{path}: TIFF predictor {predictor}")
    key = (sfmt, bits)
    if key not in _SAMPLE_DTYPES:
        raise NotImplementedError(f"{path}: sample format {key}")
    dtype = np.dtype(bo + _SAMPLE_DTYPES[key])
    if predictor == 3 and dtype.kind != "f":
        raise ValueError(f"{path}: floating point predictor on {dtype} samples")
    planar = int(t.get(284, (1,))[0])
    tiled = 322 in t or 324 in t
    if not tiled:
        rps = min(int(t.get(278, (height,))[0]), height)
        offsets = [int(o) for o in t[273]]
        counts = [int(c) for c in t[279]]
        contiguous = all(offsets[i] + counts[i] == offsets[i + 1]
                         for i in range(len(offsets) - 1))
    if not tiled and compression == 1 and predictor == 1 and contiguous:
        mm = np.memmap(path, dtype=dtype, mode="r", offset=offsets[0],
                       shape=(sum(counts) // dtype.itemsize,))
        if planar == 2 or spp == 1:
            bands = mm.reshape(spp, height, width)
        else:
            bands = np.moveaxis(mm.reshape(height, width, spp), 2, 0)
    else:
        if tiled:
            seg_w, seg_h = int(t[322][0]), int(t[323][0])
            offsets = [int(o) for o in t[324]]
            counts = [int(c) for c in t[325]]
        else:
            seg_w, seg_h = width, rps
        bands = _read_segments(path, offsets, counts, width, height, spp, dtype,
                               planar, seg_w, seg_h, tiled, compression, predictor)
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


def _predict(a: np.ndarray, predictor: int) -> bytes:
    """Inverse of :func:`_unpredict` for one ``(rows, cols)`` little-endian
    segment of one band."""
    rows, cols = a.shape
    isz = a.dtype.itemsize
    if predictor == 2:
        u = a.view(np.dtype("<u%d" % isz))
        d = u.copy()
        d[:, 1:] = u[:, 1:] - u[:, :-1]
        return d.tobytes()
    if predictor == 3:
        b = a.astype(np.dtype(">f%d" % isz)).view(np.uint8).reshape(rows, cols, isz)
        b = np.ascontiguousarray(np.moveaxis(b, 2, 1)).reshape(rows, isz * cols)
        d = b.copy()
        d[:, 1:] = b[:, 1:] - b[:, :-1]
        return d.tobytes()
    return a.tobytes()


def write_geotiff(path, data: np.ndarray, transform, epsg: Optional[int] = None,
                  is_geographic: bool = False, nodata: Optional[float] = None,
                  tile: Optional[int] = None, compress: Optional[str] = None,
                  predictor: int = 1):
    """Write ``data`` (``(h, w)`` or ``(count, h, w)``) as a band-sequential
    (Big)TIFF.  Default: uncompressed, one strip per band (the file can be
    memory-mapped back).  ``tile`` (a multiple of 16): square tiles of that
    size; ``compress='deflate'``: zlib per strip / tile, with ``predictor`` 1, 2
    or 3 (floating point) -- GDAL's ``TILED=YES BLOCKXSIZE=tile
    COMPRESS=DEFLATE PREDICTOR=n``."""
    data = np.asarray(data)
    if data.ndim == 2:
        data = data[None]
    count, height, width = data.shape
    dt = data.dtype.newbyteorder("<") if data.dtype.byteorder == ">" else data.dtype
    kind = {"f": 3, "i": 2, "u": 1}[dt.kind]
    bits = dt.itemsize * 8
    if (kind, bits) not in _SAMPLE_DTYPES:
        raise NotImplementedError(f"dtype {dt}")
    if compress not in (None, "deflate"):
        raise NotImplementedError(f"compression {compress!r}")
    if predictor not in (1, 2, 3) or (predictor == 3 and dt.kind != "f"):
        raise ValueError(f"predictor {predictor} on {dt} samples")
    if predictor != 1 and compress is None:
        raise ValueError("a predictor needs compress='deflate'")
    if tile is not None and (tile <= 0 or tile % 16):
        raise ValueError("tile must be a positive multiple of 16")
    plain = tile is None and compress is None
    band_bytes = height * width * dt.itemsize
    # segments of the non-plain layouts are built in memory (they are meant for
    # files of ordinary size; the plain layout streams band by band)
    segments = []
    if not plain:
        th, tw = (tile, tile) if tile else (min(height, max(1, (1 << 20) // max(
            1, width * dt.itemsize))), width)
        bands_c = [np.ascontiguousarray(data[b], dtype=dt) for b in range(count)]
        jobs = [(b, r0, c0) for b in range(count) for r0 in range(0, height, th)
                for c0 in range(0, width, tw)]

        def encode(job):
            b, r0, c0 = job
            blk = bands_c[b][r0:r0 + th, c0:c0 + tw]
            if tile and blk.shape != (th, tw):      # tiles are stored whole
                pad = np.zeros((th, tw), dtype=dt)
                pad[:blk.shape[0], :blk.shape[1]] = blk
                blk = pad
            raw = _predict(np.ascontiguousarray(blk), predictor)
            return zlib.compress(raw, 6) if compress else raw

        workers = min(len(jobs), os.cpu_count() or 1, 32)
        if workers > 1 and compress:
            from concurrent.futures import ThreadPoolExecutor
            with ThreadPoolExecutor(max_workers=workers) as ex:
                segments = list(ex.map(encode, jobs))      # in job order
        else:
            segments = [encode(j) for j in jobs]
        total = sum(len(sg) for sg in segments)
    else:
        th, tw = height, width
        total = band_bytes * count
    big = total + 4096 + 16 * len(segments) >= (1 << 32) - 1
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

    nseg = count if plain else len(segments)
    seg_sizes = [band_bytes] * count if plain else [len(sg) for sg in segments]
    entries = []  # (tag, type, values)
    entries.append((256, 4, [width]))
    entries.append((257, 4, [height]))
    entries.append((258, 3, [bits] * count))
    entries.append((259, 3, [8 if compress else 1]))
    entries.append((262, 3, [1]))
    offs_type = 16 if big else 4
    if tile:
        entries.append((322, 4, [tw]))
        entries.append((323, 4, [th]))
        entries.append((324, offs_type, None))  # filled below
        entries.append((325, offs_type, seg_sizes))
    else:
        entries.append((273, offs_type, None))  # filled below
        entries.append((278, 4, [th]))
        entries.append((279, offs_type, seg_sizes))
    entries.append((277, 3, [count]))
    entries.append((284, 3, [2]))
    if predictor != 1:
        entries.append((317, 3, [predictor]))
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
        n = nseg if vals is None else len(vals)
        size = _TYPES[typ][1] * n
        if size > inline:
            blobs[tag] = pos
            pos += size + (size & 1)
    data_pos = (pos + 63) // 64 * 64
    seg_offsets = []
    p = data_pos
    for sz in seg_sizes:
        seg_offsets.append(p)
        p += sz

    with open(os.fspath(path), "wb") as f:
        if big:
            f.write(struct.pack("<2sHHHQ", b"II", 43, 8, 0, head))
        else:
            f.write(struct.pack("<2sHI", b"II", 42, head))
        f.write(struct.pack("<Q" if big else "<H", len(entries)))
        tail = []
        for tag, typ, vals in entries:
            if vals is None:
                vals = seg_offsets
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
        if plain:
            for i in range(count):
                np.ascontiguousarray(data[i], dtype=dt).tofile(f)
        else:
            for sg in segments:
                f.write(sg)
