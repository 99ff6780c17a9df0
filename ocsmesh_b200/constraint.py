"""Size-function constraints on the raster path.

Mirror of ``ocsmesh/features/constraint.py``: same class names, constructor
arguments and ``value_type`` / ``rate`` semantics.  ``HfunRaster`` applies them
window by window on the device (``apply_constraints``,
``ocsmesh/hfun/raster.py:509-568``) through ``ocsb_hfun_topo_constraint`` /
``ocsb_hfun_courant_constraint``; there is no host implementation here.
``RegionConstraint`` is a ``TopoConstConstraint`` on the 0 / 1 raster of its
polygon (``ocsb_region_mask``: even-odd point-in-polygon of the pixel
coordinates; shapely is absent here, so that predicate is a restatement of
``within`` and its parity with the reference is unpinned).
"""
from __future__ import annotations

from enum import Enum
from typing import Callable

import numpy as np

from . import _lib

ConstraintValueType = Enum("ConstraintValueType", "MIN MAX")


class Constraint:
    """``features/constraint.py:17-80`` (value type, rate sign)."""

    def __init__(self, value_type: str = 'min', rate: float = 0.1):
        self._rate = rate
        if isinstance(value_type, ConstraintValueType):
            self._val_type = value_type
        elif isinstance(value_type, str):
            self._val_type = ConstraintValueType[value_type.upper()]
        else:
            raise ValueError("Invalid input for value type!")

    @property
    def type(self):
        return type(self)

    @property
    def value_type(self):
        return self._val_type

    @property
    def satisfies(self):
        if self.value_type == ConstraintValueType.MIN:
            return np.greater
        if self.value_type == ConstraintValueType.MAX:
            return np.less
        raise ValueError("Invalid value type for constraint!")

    @property
    def rate_sign(self):
        if self.value_type == ConstraintValueType.MIN:
            return -1
        if self.value_type == ConstraintValueType.MAX:
            return +1
        raise ValueError("Invalid value type for constraint!")

    @property
    def _vt_code(self) -> int:
        return (_lib.VT_MIN if self.value_type == ConstraintValueType.MIN
                else _lib.VT_MAX)

    def apply(self, *args, **kwargs):
        raise NotImplementedError(
            "constraints run on the device through HfunRaster."
            "apply_constraints; ocsmesh_b200 has no host implementation")

    # ---- a rate on a row band of the window (multi-GPU) ----------------------
    def _apply_rate_band(self, ctx, cwin, dem, dem_code, dem_stride, hfun,
                         hfun_stride, values, values_code, values_stride, value,
                         device, gather):
        """``ocsb_hfun_topo_mask_band`` on this band, ``gather`` of every band's
        mask-boundary pixels (rank order = row order), ``ocsb_hfun_apply_rate``
        on this band: bit-identical to the whole window on one device."""
        import torch
        rb = cwin.row_begin
        re = cwin.row_end if cwin.row_end > 0 else cwin.height
        mask = torch.empty((re - rb, cwin.width), dtype=torch.uint8, device=device)
        args = (cwin, dem, dem_code, dem_stride, hfun, hfun_stride, values,
                values_code, values_stride, value, float(self._lb), float(self._ub),
                self._vt_code, mask)
        n = ctx.hfun_topo_mask_band(*args)
        verts = torch.empty((n, 2), dtype=torch.float64, device=device)
        vals = torch.empty((n,), dtype=torch.float32, device=device)
        if n:
            ctx.hfun_topo_mask_band(*args, verts, vals, n)
        verts, vals = gather(verts), gather(vals)
        ctx.hfun_apply_rate(cwin, hfun, hfun_stride, mask, verts, vals,
                            int(verts.shape[0]), float(self._rate), self._vt_code)
        ctx.sync()      # mask / verts / vals are freed on return

    def _join_rate_band(self, device, gather) -> None:
        """A rank that owns no row of a window still takes part in the two
        gathers of :meth:`_apply_rate_band` (with nothing to contribute)."""
        import torch
        if self._rate is None:
            return
        gather(torch.empty((0, 2), dtype=torch.float64, device=device))
        gather(torch.empty((0,), dtype=torch.float32, device=device))

    # one (semantic window, device tile) of an HfunRaster; ``dem`` / ``hfun``
    # are (address of the window's top-left pixel, row stride in elements)
    def _apply_device(self, ctx, cwin, dem, dem_code, dem_stride, hfun,
                      hfun_stride, dem_view, gather=None):
        raise NotImplementedError


class TopoConstConstraint(Constraint):
    """``features/constraint.py:135-177``."""

    def __init__(self, value, upper_bound=np.inf, lower_bound=-np.inf,
                 value_type: str = 'min', rate=None):
        super().__init__(value_type, rate)
        # ``return_values[mask] = self.value`` (features/constraint.py:171):
        # numpy only takes a scalar or a one-element array there -- a full
        # (h, w) matrix fails with "boolean array indexing assignment cannot
        # assign ..." unless its size happens to equal the mask count
        if np.ndim(value) != 0:
            if np.size(value) != 1:
                raise ValueError(
                    "NumPy boolean array indexing assignment cannot assign "
                    f"{np.size(value)} input values to the output values where "
                    "the mask is true: TopoConstConstraint takes a scalar value")
            value = np.asarray(value).reshape(()).item()
        self._lb = lower_bound
        self._ub = upper_bound
        self._val = value

    @property
    def value(self):
        return self._val

    @property
    def topo_bounds(self):
        return self._lb, self._ub

    def _apply_device(self, ctx, cwin, dem, dem_code, dem_stride, hfun,
                      hfun_stride, dem_view, gather=None):
        if self._rate is not None and gather is not None:
            self._apply_rate_band(ctx, cwin, dem, dem_code, dem_stride, hfun,
                                  hfun_stride, None, _lib.OCSB_F32, 0,
                                  float(self._val), dem_view.device, gather)
            return
        ctx.hfun_topo_constraint(
            cwin, dem, dem_code, dem_stride, hfun, hfun_stride, None,
            _lib.OCSB_F32, 0, float(self._val), float(self._lb),
            float(self._ub), self._vt_code, self._rate)


class TopoFuncConstraint(Constraint):
    """``features/constraint.py:180-224``.  ``function`` is the caller's
    Python: it is offered the DEM window as a CUDA tensor first (arithmetic
    lambdas such as the default ``i / 2.0`` stay on the device) and, if it
    cannot take one, the window as a numpy array."""

    def __init__(self, function: Callable = lambda i: i / 2.0,
                 upper_bound=np.inf, lower_bound=-np.inf,
                 value_type: str = 'min', rate=None):
        super().__init__(value_type, rate)
        self._lb = lower_bound
        self._ub = upper_bound
        self._func = lambda i: i / 2.0
        if callable(function):
            self._func = function

    @property
    def topo_bounds(self):
        return self._lb, self._ub

    def _values(self, dem_view):
        import torch
        try:
            vals = self._func(dem_view)
            if not isinstance(vals, torch.Tensor):
                raise TypeError("function did not return a tensor")
        except Exception:  # pylint: disable=broad-except
            host = self._func(dem_view.cpu().numpy())
            vals = torch.from_numpy(np.ascontiguousarray(host)).to(dem_view.device)
        if vals.dtype not in (torch.float32, torch.float64):
            vals = vals.to(torch.float64)
        if tuple(vals.shape) != tuple(dem_view.shape):
            vals = vals.expand(dem_view.shape)
        return vals.contiguous()

    def _apply_device(self, ctx, cwin, dem, dem_code, dem_stride, hfun,
                      hfun_stride, dem_view, gather=None):
        import torch
        vals = self._values(dem_view)
        code = _lib.OCSB_F32 if vals.dtype == torch.float32 else _lib.OCSB_F64
        # `vals` covers the device tile only; the ABI wants window row 0
        rb = cwin.row_begin
        base = vals.data_ptr() - rb * vals.shape[1] * vals.element_size()
        if self._rate is not None and gather is not None:
            self._apply_rate_band(ctx, cwin, dem, dem_code, dem_stride, hfun,
                                  hfun_stride, base, code, vals.shape[1], None,
                                  dem_view.device, gather)
            return
        ctx.hfun_topo_constraint(
            cwin, dem, dem_code, dem_stride, hfun, hfun_stride, base, code,
            vals.shape[1], None, float(self._lb), float(self._ub),
            self._vt_code, self._rate)
        ctx.sync()   # `vals` is freed on return


class CourantNumConstraint(Constraint):
    """``features/constraint.py:227-313``."""

    def __init__(self, value: float, timestep: float = 150,
                 wave_amplitude: float = 2.0, value_type: str = 'max'):
        super().__init__(value_type, rate=None)
        self._value = value
        self._dt = timestep
        self._nu = wave_amplitude

    def _apply_device(self, ctx, cwin, dem, dem_code, dem_stride, hfun,
                      hfun_stride, dem_view, gather=None):
        ctx.hfun_courant_constraint(
            cwin, dem, dem_code, dem_stride, hfun, hfun_stride,
            float(self._value), float(self._dt), float(self._nu),
            self._vt_code)


class RegionConstraint(Constraint):
    """``features/constraint.py:318-395``: a fixed value where the pixel
    coordinate lies within ``shape`` and the size does not satisfy it, relaxed
    outside with ``rate``.  ``shape``: a shapely-style (Multi)Polygon or ring
    arrays (``features.polygon_rings``), in the raster's CRS (no reprojection
    here: pyproj is absent)."""

    def __init__(self, value: float, shape, crs=None, value_type: str = 'max',
                 rate: float = 0.1):
        super().__init__(value_type, rate)
        from . import features as _features
        try:
            self._rings = _features.polygon_rings(shape)
        except TypeError as e:
            raise ValueError("Invalid input shape type for region constraint: "
                             f"{type(shape)}") from e
        self._edges = _features.ring_edges(self._rings)
        self._val = value
        self._region = shape
        self._crs = crs
        # the region raster is 1 inside: a topo-const constraint with 0.5 < "z"
        self._lb, self._ub = 0.5, np.inf

    @property
    def value(self):
        return self._val

    def _apply_device(self, ctx, cwin, dem, dem_code, dem_stride, hfun,
                      hfun_stride, dem_view, gather=None):
        import torch
        rb = cwin.row_begin
        re = cwin.row_end if cwin.row_end > 0 else cwin.height
        w = cwin.width
        reg = torch.empty((re - rb, w), dtype=torch.float32, device=dem_view.device)
        base = reg.data_ptr() - rb * w * 4          # (virtual) window row 0
        ctx.region_mask(cwin, self._edges, len(self._edges), False, base, w)
        if self._rate is not None and gather is not None:
            self._apply_rate_band(ctx, cwin, base, _lib.OCSB_F32, w, hfun, hfun_stride,
                                  None, _lib.OCSB_F32, 0, float(self._val),
                                  dem_view.device, gather)
            return
        ctx.hfun_topo_constraint(
            cwin, base, _lib.OCSB_F32, w, hfun, hfun_stride, None, _lib.OCSB_F32, 0,
            float(self._val), self._lb, self._ub, self._vt_code, self._rate)
        ctx.sync()      # `reg` is freed on return
