"""ctypes binding of libocsb200.so (the C ABI declared in include/ocsb200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device
is usable, the first call raises.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import threading
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
#: OCSB_LIB points the tools/ sweeps at an alternative build of the SAME library
LIB_PATH = os.environ.get("OCSB_LIB") or os.path.join(_HERE, "libocsb200.so")

OCSB_OK = 0
OCSB_EINVAL, OCSB_ENOMEM, OCSB_ECUDA, OCSB_ENCCL, OCSB_EOVERFLOW = -1, -2, -3, -4, -5
OCSB_F32, OCSB_F64 = 0, 1
INTERP_LINEAR, INTERP_NEAREST, INTERP_SPLINE = 0, 1, 2
KC_COUNT = 18
VT_MIN, VT_MAX = 0, 1
NONE = float("nan")

c_i32, c_i64, c_f64, c_f32, c_vp = C.c_int32, C.c_int64, C.c_double, C.c_float, C.c_void_p


class ocsb_win(C.Structure):
    _fields_ = [("x0", c_f64), ("y0", c_f64), ("x1", c_f64), ("y1", c_f64),
                ("width", c_i32), ("height", c_i32),
                ("row_begin", c_i32), ("row_end", c_i32)]


class ocsb_interp_raster(C.Structure):
    _fields_ = [("win", ocsb_win), ("dem", c_vp), ("dem_dtype", c_i32),
                ("dem_stride", c_i64), ("nodata_mask", c_vp),
                ("mask_stride", c_i64), ("spline_coef", c_vp)]


class ocsb_hfun_op(C.Structure):
    _fields_ = [("kind", c_i32), ("fuse_clamp", c_i32), ("verts_xy", c_vp),
                ("nv", c_i64), ("expansion_rate", c_f64), ("target_size", c_f64),
                ("hmin", c_f64), ("hmax", c_f64), ("lower_bound", c_f64),
                ("upper_bound", c_f64)]


OP_FEATURE, OP_LIMITER = 0, 1


class OcsbError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"libocsb200 error {status}: {message}")
        self.status = status
        self.message = message


#: every symbol include/ocsb200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "ocsb_abi_version": (c_i32, []),
    "ocsb_last_error": (C.c_char_p, []),
    "ocsb_ctx_create": (c_i32, [c_i32, c_vp, C.POINTER(c_vp)]),
    "ocsb_ctx_destroy": (c_i32, [c_vp]),
    "ocsb_ctx_set_stream": (c_i32, [c_vp, c_vp]),
    "ocsb_ctx_sync": (c_i32, [c_vp]),
    "ocsb_ctx_launch_count": (c_i64, [c_vp]),
    "ocsb_ctx_trim": (c_i32, [c_vp]),
    "ocsb_ctx_set_timing": (c_i32, [c_vp, c_i32]),
    "ocsb_ctx_get_timing": (c_i32, [c_vp, c_i32, C.POINTER(c_f64),
                                    C.POINTER(c_i64)]),
    "ocsb_ctx_reset_timing": (c_i32, [c_vp]),
    "ocsb_kernel_class_name": (C.c_char_p, [c_i32]),
    "ocsb_windows": (c_i32, [c_i32, c_i32, c_i32, c_i32, c_vp, c_i64,
                             C.POINTER(c_i64)]),
    "ocsb_window_coords": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_vp]),
    "ocsb_fill_f32": (c_i32, [c_vp, c_vp, c_i64, c_f32]),
    "ocsb_hfun_clamp": (c_i32, [c_vp, c_vp, c_i64, c_i32, c_i32, c_f64, c_f64]),
    "ocsb_hfun_feature": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i64,
                                  c_vp, c_i64, c_vp, c_i64, c_f64, c_f64,
                                  c_f64, c_f64, c_i32, c_vp, c_vp]),
    "ocsb_nearest_vertex": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_i64, c_f64,
                                    c_vp, c_vp]),
    "ocsb_nearest_vertex_static": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_i64, c_f64,
                                           c_vp, c_vp]),
    "ocsb_hfun_flow_limiter": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32,
                                       c_i64, c_vp, c_i64, c_vp, c_i64,
                                       c_f64, c_f64, c_f64, c_f64, c_f64,
                                       c_f64, c_f64, c_f64, c_i32]),
    "ocsb_hfun_pass": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32, c_i64,
                               c_vp, c_i64, c_vp, c_i64, c_f64, c_f64, c_f64,
                               c_f64, C.POINTER(ocsb_hfun_op), c_i32, c_i32]),
    "ocsb_hfun_op_windows": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_vp, c_i32,
                                     c_vp, c_i32, c_i64, c_vp, c_i64, c_vp, c_i64,
                                     c_f64, c_f64, c_f64, c_f64,
                                     C.POINTER(ocsb_hfun_op), c_i32]),
    "ocsb_hfun_pass_windows": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_vp, c_vp,
                                       c_i32, c_vp, c_i32, c_i64, c_i32, c_i32,
                                       c_vp, c_i64, c_vp, c_i64,
                                       c_f64, c_f64, c_f64, c_f64,
                                       C.POINTER(ocsb_hfun_op), c_i32, c_vp]),
    "ocsb_hfun_constant_value": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32,
                                         c_i64, c_vp, c_i64, c_vp, c_i64,
                                         c_f64, c_f64, c_f64, c_f64, c_f64,
                                         c_i32]),
    "ocsb_hfun_topo_constraint": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32,
                                          c_i64, c_vp, c_i64, c_vp, c_i32, c_i64,
                                          c_f64, c_f64, c_f64, c_i32, c_f64]),
    "ocsb_hfun_topo_mask_band": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32,
                                         c_i64, c_vp, c_i64, c_vp, c_i32, c_i64,
                                         c_f64, c_f64, c_f64, c_i32, c_vp, c_vp,
                                         c_vp, c_i64, C.POINTER(c_i64)]),
    "ocsb_hfun_apply_rate": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i64, c_vp,
                                     c_vp, c_vp, c_i64, c_f64, c_i32]),
    "ocsb_hfun_courant_constraint": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp,
                                             c_i32, c_i64, c_vp, c_i64, c_f64,
                                             c_f64, c_f64, c_i32]),
    "ocsb_contour_segments": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32,
                                      c_i64, c_f64, c_i64, c_vp, c_vp,
                                      C.POINTER(c_i64)]),
    "ocsb_band_pack": (c_i32, [c_vp, c_vp, c_i64, c_i64, c_vp, c_vp, c_i32, c_vp, c_f64]),
    "ocsb_band_unpack": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i64, c_i32, c_vp, c_vp, c_vp]),
    "ocsb_region_mask": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i64, c_i32, c_vp,
                                 c_i64]),
    "ocsb_contour_lines": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32, c_i64,
                                   c_f64, C.POINTER(c_i64), C.POINTER(c_i64)]),
    "ocsb_contour_points": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32, c_i64,
                                    c_vp, c_i32, c_f64, c_i32, C.POINTER(c_i64)]),
    "ocsb_take_result": (c_i32, [c_vp, c_i32, c_vp, c_i64]),
    "ocsb_hfun_sample_source": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i64,
                                        c_i32, c_f64, c_f64, c_vp, c_vp]),
    "ocsb_hfun_sample_nearest": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i64,
                                         c_i32, c_f64, c_f64, c_vp, c_i64, c_vp]),
    "ocsb_resample_nearest": (c_i32, [c_vp, c_vp, c_i32, c_i64, c_i32, c_i32, c_vp, c_vp,
                                      c_i64, c_vp, c_i64, c_i32, c_i32, c_vp]),
    "ocsb_band_p2p_bytes": (c_i64, [c_i32, c_i64, c_i32]),
    "ocsb_p2p_alloc": (c_i32, [c_vp, c_i64, C.POINTER(c_vp), c_vp]),
    "ocsb_p2p_open": (c_i32, [c_vp, c_vp, C.POINTER(c_vp)]),
    "ocsb_p2p_close": (c_i32, [c_vp, c_vp]),
    "ocsb_p2p_free": (c_i32, [c_vp, c_vp]),
    "ocsb_band_push": (c_i32, [c_vp, c_vp, c_i64, c_i64, c_vp, c_vp, c_i32, c_vp, c_i32,
                               c_i32, c_i64, c_f64, c_vp]),
    "ocsb_band_pull": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i64, c_i32, c_i64, c_vp, c_vp,
                               c_vp, c_vp]),
    "ocsb_tiff_decode": (c_i32, [c_i32, c_vp, c_i64, c_vp, c_i64, C.POINTER(c_i64)]),
    "ocsb_stitch_segments": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_i64, c_vp, c_i64,
                                     C.POINTER(c_i64), C.POINTER(c_i64)]),
    "ocsb_feature_points": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_i32, c_vp, c_i64,
                                    C.POINTER(c_i64)]),
    "ocsb_min_combine": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_i64, c_i32, c_i32]),
    "ocsb_interp": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32, c_i64,
                            c_vp, c_i64, c_vp, c_i64, c_i32, c_vp, c_vp]),
    "ocsb_interp_multi": (c_i32, [c_vp, C.POINTER(ocsb_interp_raster), c_i32,
                                  c_vp, c_i64, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "ocsb_bucket_nodes": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_f64, c_vp, c_vp]),
    "ocsb_spline_prefilter": (c_i32, [c_vp, C.POINTER(ocsb_win), c_vp, c_i32,
                                      c_i64, c_vp]),
}

_lib = None
_lib_lock = threading.Lock()


def load() -> C.CDLL:
    """dlopen libocsb200.so and bind every declared symbol (fails loudly)."""
    global _lib
    with _lib_lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise OcsbError(
                OCSB_ECUDA,
                f"{LIB_PATH} is missing: build it with "
                "`python -m ocsmesh_b200.build` (there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is absent
            fn.restype = res
            fn.argtypes = args
        _lib = lib
        return lib


def check(status: int) -> None:
    if status != OCSB_OK:
        msg = load().ocsb_last_error()
        raise OcsbError(status, msg.decode("utf-8", "replace") if msg else "")


def opt(v: Optional[float]) -> float:
    """None -> NaN (the ABI's encoding of an absent optional double)."""
    return NONE if v is None else float(v)


def ptr(buf) -> Optional[int]:
    """Address of a numpy array, torch tensor, raw int address, or None."""
    if buf is None:
        return None
    if isinstance(buf, int):
        return buf
    if isinstance(buf, np.ndarray):
        return buf.ctypes.data
    if hasattr(buf, "data_ptr"):
        return buf.data_ptr()
    raise TypeError(f"cannot take the address of {type(buf)!r}")


def make_win(bounds, width: int, height: int, row_begin: int = 0,
             row_end: int = 0) -> ocsb_win:
    """``bounds`` = Raster.get_window_bounds(window); rows [row_begin, row_end)
    select the device tile (0, 0 = the whole window)."""
    x0, y0, x1, y1 = (float(b) for b in bounds)
    return ocsb_win(x0, y0, x1, y1, int(width), int(height), int(row_begin),
                    int(row_end))


def windows(width: int, height: int, chunk_size: int, overlap: int) -> np.ndarray:
    """ocsb_windows -> (n, 4) int32 array of (col_off, row_off, w, h)."""
    lib = load()
    n = c_i64(0)
    check(lib.ocsb_windows(width, height, chunk_size, overlap, None, 0, C.byref(n)))
    out = np.empty((n.value, 4), dtype=np.int32)
    check(lib.ocsb_windows(width, height, chunk_size, overlap, out.ctypes.data,
                           n.value, C.byref(n)))
    return out


def tiff_decode(compression: int, raw: bytes, cap: int) -> np.ndarray:
    """One LZW (5) / PackBits (32773) strip or tile -> uint8 array (<= cap bytes)."""
    lib = load()
    src = np.frombuffer(raw, dtype=np.uint8)
    dst = np.empty(int(cap), dtype=np.uint8)
    n = c_i64(0)
    check(lib.ocsb_tiff_decode(int(compression), src.ctypes.data, src.size,
                               dst.ctypes.data, dst.size, C.byref(n)))
    return dst[:n.value]


def stitch_segments(edges: np.ndarray, pts: np.ndarray):
    """ocsb_stitch_segments (host code of the library) -> list of (k, 2)
    float64 polylines."""
    edges = np.ascontiguousarray(edges, dtype=np.int64).reshape(-1, 2)
    pts = np.ascontiguousarray(pts, dtype=np.float64).reshape(-1, 4)
    n = len(edges)
    if n == 0:
        return []
    xy = np.empty((2 * n, 2), dtype=np.float64)
    off = np.empty(n + 1, dtype=np.int64)
    nxy, nl = c_i64(0), c_i64(0)
    check(load().ocsb_stitch_segments(edges.ctypes.data, pts.ctypes.data, n,
                                      xy.ctypes.data, 2 * n, off.ctypes.data, n,
                                      C.byref(nxy), C.byref(nl)))
    return [xy[off[k]:off[k + 1]].copy() for k in range(nl.value)]


def feature_points(lines, target_size: float, max_verts: int = 200) -> np.ndarray:
    """ocsb_feature_points (host code of the library): the (N, 2) vertex array
    add_feature hands to the nearest-vertex search."""
    lines = [np.ascontiguousarray(np.asarray(l, dtype=np.float64)).reshape(-1, 2)
             for l in lines]
    if not lines:
        return np.empty((0, 2), dtype=np.float64)
    off = np.zeros(len(lines) + 1, dtype=np.int64)
    np.cumsum([len(l) for l in lines], out=off[1:])
    xy = np.ascontiguousarray(np.concatenate(lines, axis=0))
    n = c_i64(0)
    lib = load()
    check(lib.ocsb_feature_points(xy.ctypes.data, off.ctypes.data, len(lines),
                                  float(target_size), int(max_verts), None, 0,
                                  C.byref(n)))
    out = np.empty((n.value, 2), dtype=np.float64)
    check(lib.ocsb_feature_points(xy.ctypes.data, off.ctypes.data, len(lines),
                                  float(target_size), int(max_verts),
                                  out.ctypes.data, n.value, C.byref(n)))
    return out


class Context:
    """One ocsb_ctx (one CUDA device, one stream)."""

    def __init__(self, device: int = 0, stream: Optional[int] = None):
        self.lib = load()
        h = c_vp()
        check(self.lib.ocsb_ctx_create(int(device), stream, C.byref(h)))
        self.handle = h
        self.device = int(device)

    def close(self):
        if getattr(self, "handle", None):
            self.lib.ocsb_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(self.lib.ocsb_ctx_sync(self.handle))

    def set_stream(self, stream: Optional[int]):
        check(self.lib.ocsb_ctx_set_stream(self.handle, stream))

    @property
    def launch_count(self) -> int:
        return int(self.lib.ocsb_ctx_launch_count(self.handle))

    def trim(self):
        check(self.lib.ocsb_ctx_trim(self.handle))

    def set_timing(self, enabled: bool):
        check(self.lib.ocsb_ctx_set_timing(self.handle, int(bool(enabled))))

    def reset_timing(self):
        check(self.lib.ocsb_ctx_reset_timing(self.handle))

    def get_timing(self) -> dict:
        """{kernel class name: (total_ms, launches)} since the last reset."""
        out = {}
        for kc in range(KC_COUNT):
            ms, n = c_f64(0.0), c_i64(0)
            check(self.lib.ocsb_ctx_get_timing(self.handle, kc, C.byref(ms),
                                               C.byref(n)))
            if n.value:
                name = self.lib.ocsb_kernel_class_name(kc).decode()
                out[name] = (ms.value, n.value)
        return out

    # ---- thin typed wrappers (buffers: numpy arrays or torch tensors) ------
    def window_coords(self, win: ocsb_win, x_out, y_out):
        check(self.lib.ocsb_window_coords(self.handle, C.byref(win), ptr(x_out),
                                          ptr(y_out)))

    def fill_f32(self, dst, n: int, value: float):
        check(self.lib.ocsb_fill_f32(self.handle, ptr(dst), int(n), value))

    def hfun_clamp(self, hfun, stride: int, width: int, height: int, hmin, hmax):
        check(self.lib.ocsb_hfun_clamp(self.handle, ptr(hfun), int(stride),
                                       int(width), int(height), opt(hmin),
                                       opt(hmax)))

    def hfun_feature(self, win, verts, nv, old, old_stride, new, new_stride,
                     expansion_rate, target_size, hmin, hmax, fuse_clamp=False,
                     dist_out=None, idx_out=None):
        check(self.lib.ocsb_hfun_feature(
            self.handle, C.byref(win), ptr(verts), int(nv), ptr(old),
            int(old_stride), ptr(new), int(new_stride), float(expansion_rate),
            float(target_size), opt(hmin), opt(hmax), int(bool(fuse_clamp)),
            ptr(dist_out), ptr(idx_out)))

    def nearest_vertex(self, verts, nv, query, nq, r, dist_out, idx_out):
        check(self.lib.ocsb_nearest_vertex(
            self.handle, ptr(verts), int(nv), ptr(query), int(nq), opt(r),
            ptr(dist_out), ptr(idx_out)))

    def nearest_vertex_static(self, verts, nv, query, nq, r, dist_out, idx_out):
        """ocsb_nearest_vertex without any allocation (device buffers only,
        nv <= OCSB_STATIC_MAX_VERTS; calls share one fixed device array)."""
        check(self.lib.ocsb_nearest_vertex_static(
            self.handle, ptr(verts), int(nv), ptr(query), int(nq), opt(r),
            ptr(dist_out), ptr(idx_out)))

    def hfun_flow_limiter(self, win, dem, dem_dtype, dem_stride, old, old_stride,
                          new, new_stride, dx, dy, hmin, hmax, lower_bound,
                          upper_bound, ghmin, ghmax, fuse_clamp=False):
        check(self.lib.ocsb_hfun_flow_limiter(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            ptr(old), int(old_stride), ptr(new), int(new_stride), float(dx),
            float(dy), opt(hmin), opt(hmax), opt(lower_bound), opt(upper_bound),
            opt(ghmin), opt(ghmax), int(bool(fuse_clamp))))

    def hfun_op_windows(self, wins, dem_off, hfun_off, dem, dem_dtype, dem_stride,
                        hfun_old, old_stride, hfun_new, new_stride, dx, dy, ghmin,
                        ghmax, op, in_order=False):
        """One op tuple (see :meth:`hfun_pass`) on a list of windows of
        device-resident buffers: ``wins`` an ``ocsb_win`` ctypes array,
        ``dem_off`` / ``hfun_off`` int64 numpy arrays of element offsets."""
        arr = self._op_array([op])
        check(self.lib.ocsb_hfun_op_windows(
            self.handle, wins, ptr(dem_off), ptr(hfun_off), len(wins), ptr(dem),
            int(dem_dtype), int(dem_stride), ptr(hfun_old), int(old_stride),
            ptr(hfun_new), int(new_stride), float(dx), float(dy), opt(ghmin),
            opt(ghmax), arr, int(bool(in_order))))

    def hfun_pass_windows(self, wins, dem_off, hfun_off, dem, dem_dtype, dem_stride,
                          dem_rows, dem_cols, hfun_in, in_stride, hfun_out,
                          out_stride, dx, dy, ghmin, ghmax, ops, keep_wh=None,
                          status=None, op_array=None):
        """The op list (tuples as in :meth:`hfun_pass`) over a list of windows
        of device-resident buffers in one tiled sweep.  ``status``: None =
        synchronous; a device int32 tensor = asynchronous with the sticky
        failure flag (see ocsb200.h).  ``op_array``: a prebuilt ctypes op array
        (``_op_array(ops)``) to skip the marshalling on a hot path."""
        arr = op_array if op_array is not None else self._op_array(ops)
        check(self.lib.ocsb_hfun_pass_windows(
            self.handle, wins, ptr(dem_off), ptr(hfun_off), ptr(keep_wh), len(wins),
            ptr(dem), int(dem_dtype), int(dem_stride), int(dem_rows), int(dem_cols),
            ptr(hfun_in), int(in_stride), ptr(hfun_out), int(out_stride),
            float(dx), float(dy), opt(ghmin), opt(ghmax), arr, len(arr),
            ptr(status)))

    def hfun_pass(self, win, dem, dem_dtype, dem_stride, hfun_in, in_stride,
                  hfun_out, out_stride, dx, dy, ghmin, ghmax, ops, band_rows=0):
        """``ops``: list of ("feature", verts, nv, rate, target, fuse) /
        ("limiter", hmin, hmax, lower_bound, upper_bound, fuse) tuples."""
        arr = self._op_array(ops)
        check(self.lib.ocsb_hfun_pass(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            ptr(hfun_in), int(in_stride), ptr(hfun_out), int(out_stride),
            float(dx), float(dy), opt(ghmin), opt(ghmax), arr, len(ops),
            int(band_rows)))

    @staticmethod
    def _op_array(ops):
        arr = (ocsb_hfun_op * len(ops))()
        for o, t in zip(arr, ops):
            if t[0] == "feature":
                _, verts, nv, rate, target, fuse = t
                o.kind, o.fuse_clamp = OP_FEATURE, int(bool(fuse))
                o.verts_xy, o.nv = ptr(verts), int(nv)
                o.expansion_rate, o.target_size = float(rate), float(target)
                o.hmin = o.hmax = o.lower_bound = o.upper_bound = NONE
            else:
                _, hmin, hmax, lb, ub, fuse = t
                o.kind, o.fuse_clamp = OP_LIMITER, int(bool(fuse))
                o.verts_xy, o.nv = None, 0
                o.expansion_rate = o.target_size = NONE
                o.hmin, o.hmax = opt(hmin), opt(hmax)
                o.lower_bound, o.upper_bound = opt(lb), opt(ub)
        return arr

    def hfun_constant_value(self, win, dem, dem_dtype, dem_stride, old,
                            old_stride, new, new_stride, value, lower_bound,
                            upper_bound, ghmin, ghmax, fuse_clamp=False):
        check(self.lib.ocsb_hfun_constant_value(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            ptr(old), int(old_stride), ptr(new), int(new_stride), float(value),
            float(lower_bound), float(upper_bound), opt(ghmin), opt(ghmax),
            int(bool(fuse_clamp))))

    def hfun_topo_constraint(self, win, dem, dem_dtype, dem_stride, hfun,
                             hfun_stride, values, values_dtype, values_stride,
                             value, lower_bound, upper_bound, value_type, rate):
        """``value_type``: VT_MIN / VT_MAX; ``rate`` None = no relaxation."""
        check(self.lib.ocsb_hfun_topo_constraint(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            ptr(hfun), int(hfun_stride), ptr(values), int(values_dtype),
            int(values_stride), opt(value), float(lower_bound),
            float(upper_bound), int(value_type), opt(rate)))

    def hfun_topo_mask_band(self, win, dem, dem_dtype, dem_stride, hfun,
                            hfun_stride, values, values_dtype, values_stride,
                            value, lower_bound, upper_bound, value_type, mask_out,
                            verts_out=None, vals_out=None, cap=0) -> int:
        """Step 1 of a rate constraint on a row band; returns the number of
        mask-boundary pixels of the band."""
        n = c_i64(0)
        check(self.lib.ocsb_hfun_topo_mask_band(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            ptr(hfun), int(hfun_stride), ptr(values), int(values_dtype),
            int(values_stride), opt(value), float(lower_bound),
            float(upper_bound), int(value_type), ptr(mask_out), ptr(verts_out),
            ptr(vals_out), int(cap), C.byref(n)))
        return n.value

    def hfun_apply_rate(self, win, hfun, hfun_stride, mask, verts, vals, nv, rate,
                        value_type):
        check(self.lib.ocsb_hfun_apply_rate(
            self.handle, C.byref(win), ptr(hfun), int(hfun_stride), ptr(mask),
            ptr(verts), ptr(vals), int(nv), float(rate), int(value_type)))

    def hfun_courant_constraint(self, win, dem, dem_dtype, dem_stride, hfun,
                                hfun_stride, courant, timestep, wave_amplitude,
                                value_type):
        check(self.lib.ocsb_hfun_courant_constraint(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            ptr(hfun), int(hfun_stride), float(courant), float(timestep),
            float(wave_amplitude), int(value_type)))

    def contour_segments(self, win, dem, dem_dtype, dem_stride, level,
                         cap: Optional[int] = None):
        """-> (edges (n, 2) int64, pts (n, 4) float64) host arrays.  One sweep
        when ``cap`` (default: 16 segments per row + column of the window)
        holds the result; on OCSB_EOVERFLOW the call reports the exact count
        and is repeated once with buffers of that size."""
        if cap is None:
            cap = 16 * (win.width + win.height) + 1024
        n = c_i64(0)
        for _ in range(2):
            edges = np.empty((cap, 2), dtype=np.int64)
            pts = np.empty((cap, 4), dtype=np.float64)
            rc = self.lib.ocsb_contour_segments(
                self.handle, C.byref(win), ptr(dem), int(dem_dtype),
                int(dem_stride), float(level), cap, edges.ctypes.data,
                pts.ctypes.data, C.byref(n))
            if rc == OCSB_EOVERFLOW and n.value > cap:
                cap = n.value
                continue
            check(rc)
            break
        return edges[:n.value], pts[:n.value]

    def band_pack(self, verts, n, cap, row_first, row_last, width, send, far):
        check(self.lib.ocsb_band_pack(self.handle, ptr(verts), int(n), int(cap),
                                      ptr(row_first), ptr(row_last), int(width),
                                      ptr(send), float(far)))

    def band_unpack(self, recv, world, rank, cap, width, verts_out, halo_above, halo_below):
        check(self.lib.ocsb_band_unpack(self.handle, ptr(recv), int(world), int(rank),
                                        int(cap), int(width), ptr(verts_out),
                                        ptr(halo_above), ptr(halo_below)))

    # ---- band exchange over peer memory (CUDA IPC buffers) -------------------
    def p2p_alloc(self, nbytes: int):
        """-> (device address, 64-byte IPC handle) of a zeroed exchange buffer."""
        p = c_vp(0)
        h = (C.c_uint8 * 64)()
        check(self.lib.ocsb_p2p_alloc(self.handle, int(nbytes), C.byref(p), C.cast(h, c_vp)))
        return int(p.value), bytes(h)

    def p2p_open(self, handle: bytes) -> int:
        p = c_vp(0)
        h = (C.c_uint8 * 64).from_buffer_copy(handle)
        check(self.lib.ocsb_p2p_open(self.handle, C.cast(h, c_vp), C.byref(p)))
        return int(p.value)

    def p2p_close(self, addr: int):
        check(self.lib.ocsb_p2p_close(self.handle, c_vp(addr)))

    def p2p_free(self, addr: int):
        check(self.lib.ocsb_p2p_free(self.handle, c_vp(addr)))

    def band_push(self, verts, n, cap, row_first, row_last, width, peers_dev, world, rank,
                  epoch, far, done_counter):
        check(self.lib.ocsb_band_push(
            self.handle, ptr(verts) if n else None, int(n), int(cap), ptr(row_first),
            ptr(row_last), int(width), ptr(peers_dev), int(world), int(rank), int(epoch),
            float(far), ptr(done_counter)))

    def band_pull(self, own_addr, world, rank, cap, width, epoch, verts_out, halo_above,
                  halo_below, status):
        check(self.lib.ocsb_band_pull(
            self.handle, c_vp(own_addr), int(world), int(rank), int(cap), int(width),
            int(epoch), ptr(verts_out), ptr(halo_above), ptr(halo_below), ptr(status)))

    def region_mask(self, win, edges, n_edges, all_touched, mask_out, mask_stride):
        """0 / 1 float32 raster of a polygon set (``edges``: (n, 4) float64)."""
        check(self.lib.ocsb_region_mask(self.handle, C.byref(win), ptr(edges),
                                        int(n_edges), int(bool(all_touched)),
                                        ptr(mask_out), int(mask_stride)))

    def contour_lines(self, win, dem, dem_dtype, dem_stride, level):
        """Polylines of one level, traced, stitched and ordered on the device
        -> (xy (n, 2) float64, offsets (n_lines + 1,) int64) host arrays."""
        n_xy, n_lines = c_i64(0), c_i64(0)
        check(self.lib.ocsb_contour_lines(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            float(level), C.byref(n_xy), C.byref(n_lines)))
        xy = np.empty((n_xy.value, 2), dtype=np.float64)
        off = np.zeros(n_lines.value + 1 if n_lines.value else 0, dtype=np.int64)
        check(self.lib.ocsb_take_result(self.handle, 0, xy.ctypes.data, xy.nbytes))
        check(self.lib.ocsb_take_result(self.handle, 1, off.ctypes.data, off.nbytes))
        return xy, off

    def contour_points(self, win, dem, dem_dtype, dem_stride, levels, target_size,
                       max_verts=200, device=None):
        """The resampled contour vertices of ``levels`` (the array of
        hfun/raster.py:1230) as an (N, 2) float64 CUDA tensor: trace, stitch
        and resample all run on the device."""
        import torch
        lv = np.ascontiguousarray(np.asarray(levels, dtype=np.float64).ravel())
        n = c_i64(0)
        check(self.lib.ocsb_contour_points(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            lv.ctypes.data, len(lv), float(target_size), int(max_verts), C.byref(n)))
        dev = self.device if device is None else device
        out = torch.empty((n.value, 2), dtype=torch.float64, device=f"cuda:{dev}")
        check(self.lib.ocsb_take_result(self.handle, 0, ptr(out) if n.value else None,
                                        out.numel() * 8))
        return out

    def hfun_sample_source(self, win, hfun, hfun_stride, stride, hmin, hmax,
                           coords_out, values_out):
        check(self.lib.ocsb_hfun_sample_source(
            self.handle, C.byref(win), ptr(hfun), int(hfun_stride), int(stride),
            opt(hmin), opt(hmax), ptr(coords_out), ptr(values_out)))

    def hfun_sample_nearest(self, win, hfun, hfun_stride, stride, hmin, hmax,
                            xy, K, values_out):
        check(self.lib.ocsb_hfun_sample_nearest(
            self.handle, C.byref(win), ptr(hfun), int(hfun_stride), int(stride),
            opt(hmin), opt(hmax), ptr(xy), int(K), ptr(values_out)))

    def resample_nearest(self, src, src_dtype, src_stride, src_w, src_h, src_transform,
                         src_mask, mask_stride, dst, dst_stride, dst_w, dst_h,
                         dst_transform):
        """Nearest warp of a north-up raster onto another of the same CRS
        (transforms: affine six-tuples ``(a, b, c, d, e, f)``)."""
        sgt = (c_f64 * 6)(*[float(v) for v in tuple(src_transform)[:6]])
        dgt = (c_f64 * 6)(*[float(v) for v in tuple(dst_transform)[:6]])
        check(self.lib.ocsb_resample_nearest(
            self.handle, ptr(src), int(src_dtype), int(src_stride), int(src_w),
            int(src_h), C.cast(sgt, c_vp), ptr(src_mask), int(mask_stride), ptr(dst),
            int(dst_stride), int(dst_w), int(dst_h), C.cast(dgt, c_vp)))

    def min_combine(self, dst, dst_stride, src, src_stride, width, height):
        check(self.lib.ocsb_min_combine(self.handle, ptr(dst), int(dst_stride),
                                        ptr(src), int(src_stride), int(width),
                                        int(height)))

    def interp(self, win, dem, dem_dtype, dem_stride, mask, mask_stride, xy, K,
               method, touched, values):
        check(self.lib.ocsb_interp(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            ptr(mask), int(mask_stride), ptr(xy), int(K), int(method),
            ptr(touched), ptr(values)))

    def bucket_nodes(self, xy, K, bbox, cell_size, perm_out, xy_sorted_out):
        b = (c_f64 * 4)(*[float(v) for v in bbox])
        check(self.lib.ocsb_bucket_nodes(self.handle, ptr(xy), int(K), b,
                                         float(cell_size), ptr(perm_out),
                                         ptr(xy_sorted_out)))

    def interp_multi(self, entries, xy, K, method, touched, values, source=None,
                     perm=None):
        """``entries``: list of (win, dem, dem_dtype, dem_stride, mask,
        mask_stride[, spline_coef]) in raster-list order (last valid hit wins)."""
        arr = (ocsb_interp_raster * len(entries))()
        for a, e in zip(arr, entries):
            win, dem, dt, ds, mask, ms = e[:6]
            a.win = win
            a.dem = ptr(dem)
            a.dem_dtype = int(dt)
            a.dem_stride = int(ds)
            a.nodata_mask = ptr(mask)
            a.mask_stride = int(ms)
            a.spline_coef = ptr(e[6]) if len(e) > 6 else None
        check(self.lib.ocsb_interp_multi(self.handle, arr, len(entries), ptr(xy),
                                         int(K), int(method), ptr(touched),
                                         ptr(values), ptr(source), ptr(perm)))

    def spline_prefilter(self, win, dem, dem_dtype, dem_stride, coef):
        check(self.lib.ocsb_spline_prefilter(
            self.handle, C.byref(win), ptr(dem), int(dem_dtype), int(dem_stride),
            ptr(coef)))


_default_ctx = {}


def default_context(device: Optional[int] = None) -> Context:
    """Per-device context bound to torch's current stream on that device."""
    import torch
    if not torch.cuda.is_available():
        raise OcsbError(OCSB_ECUDA, "no CUDA device available; ocsmesh_b200 has "
                                    "no CPU fallback")
    if device is None:
        device = torch.cuda.current_device()
    ctx = _default_ctx.get(device)
    stream = torch.cuda.current_stream(device).cuda_stream
    if ctx is None:
        ctx = Context(device, None)
        ctx._stream = None
        _default_ctx[device] = ctx
    if ctx._stream != stream:
        # 0 is the legacy default stream, which is what torch uses by default
        ctx.set_stream(stream)
        ctx._stream = stream
    return ctx


def isnan(v) -> bool:
    return isinstance(v, float) and math.isnan(v)
