"""Deterministic synthetic inputs (DEMs, shoreline features, node clouds).

Used by the tests, ``bench.py`` and ``__graft_entry__.smoke()`` so that the
CUDA path, the CPU oracle and the timed CPU baseline all see the *same*
inputs.  numpy only.

DEM(H, W, seed): projected (metre) CRS, 10 m pixels, north-up, float32,
    z = 40 sin(2 pi x / 7.3 km) cos(2 pi y / 5.1 km)
        + 0.004 (x - xc) - 0.003 (y - yc)
        + 4 octaves of random plane waves (amplitude 5 m, halving)
        + N(0, sigma) white noise
so the shoreline z = 0 crosses every part of the raster and the gradient is
non-zero everywhere.  Every term except the noise is separable in (x, y), so a
block of rows is one small matmul.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np

PIXEL = 10.0
X0 = 500000.0
Y0 = 4100000.0


def dem_transform(height: int, pixel: float = PIXEL, x0: float = X0,
                  y0: float = Y0) -> Tuple[float, ...]:
    """Affine (a, b, c, d, e, f) of a north-up raster whose south edge is y0."""
    return (pixel, 0.0, x0, 0.0, -pixel, y0 + pixel * height)


def _wave_terms(seed: int):
    rng = np.random.default_rng(seed)
    terms = []
    amp, lam = 5.0, 3000.0
    for _ in range(4):
        for _ in range(3):
            th = rng.uniform(0, 2 * np.pi)
            ph = rng.uniform(0, 2 * np.pi)
            k = 2 * np.pi / (lam * rng.uniform(0.8, 1.25))
            terms.append((amp / 3.0, k * np.cos(th), k * np.sin(th), ph))
        amp *= 0.5
        lam *= 0.5
    return terms


def synth_dem(height: int, width: int, seed: int = 1, noise: float = 0.01,
              pixel: float = PIXEL, x0: float = X0, y0: float = Y0,
              rows: Optional[Tuple[int, int]] = None,
              dtype=np.float32) -> np.ndarray:
    """(height, width) DEM (or the row block ``rows=(r0, r1)`` of it).

    Pixel (row j, col i) is evaluated at the pixel centre
    x = x0 + (i + 0.5) pixel, y = y0 + (height - j - 0.5) pixel.
    """
    r0, r1 = (0, height) if rows is None else rows
    x = x0 + (np.arange(width) + 0.5) * pixel
    xc = x0 + 0.5 * width * pixel
    yc = y0 + 0.5 * height * pixel
    terms = _wave_terms(seed)
    # separable expansion  z = sum_t Y_t(y) X_t(x)
    X = [np.sin(2 * np.pi * x / 7300.0), 0.004 * (x - xc), np.ones_like(x)]
    for a, kx, ky, ph in terms:
        # a sin(kx x + ky y + ph) = a [sin(kx x + ph) cos(ky y) + cos(kx x + ph) sin(ky y)]
        X.append(np.sin(kx * (x - xc) + ph))
        X.append(np.cos(kx * (x - xc) + ph))
    Xm = np.stack(X, axis=0).astype(np.float32)

    def y_terms(ra, rb):
        y = y0 + (height - np.arange(ra, rb) - 0.5) * pixel
        Y = [40.0 * np.cos(2 * np.pi * y / 5100.0), np.ones_like(y), -0.003 * (y - yc)]
        for a, kx, ky, ph in terms:
            Y.append(a * np.cos(ky * (y - yc)))
            Y.append(a * np.sin(ky * (y - yc)))
        return np.stack(Y, axis=1).astype(np.float32)

    out = np.empty((r1 - r0, width), dtype=dtype)
    # The matmul runs over blocks of ABSOLUTE rows [k B, (k + 1) B): whatever
    # `rows` is asked for, a raster row always comes out of the same call with
    # the same operands, so row blocks computed by different processes (the
    # ranks of a multi-GPU run, a rank's halo row and its neighbour's band row)
    # are bit-identical -- a BLAS result may depend on the block's row count.
    B = 256
    for k in range(r0 // B, (r1 + B - 1) // B):
        a, b = k * B, min(height, (k + 1) * B)
        blk = y_terms(a, b) @ Xm
        lo, hi = max(a, r0), min(b, r1)
        blk = blk[lo - a:hi - a]
        if noise:
            # one stream per (seed, absolute row) -> independent of `rows`
            for rr in range(lo, hi):
                g = np.random.default_rng([seed, 7919, rr])
                blk[rr - lo] += noise * g.standard_normal(width, dtype=np.float32)
        out[lo - r0:hi - r0] = blk
    return out


def synth_shoreline(height: int, width: int, seed: int = 1, n_lines: int = 3,
                    spacing: float = 25.0, pixel: float = PIXEL,
                    x0: float = X0, y0: float = Y0) -> List[np.ndarray]:
    """Smooth meandering polylines crossing the raster west-east (a synthetic
    shoreline to feed ``add_feature`` directly), vertices every ~``spacing`` m."""
    rng = np.random.default_rng(seed + 1000)
    ext_x, ext_y = width * pixel, height * pixel
    n = max(8, int(ext_x / spacing))
    lines = []
    for k in range(n_lines):
        t = np.linspace(0.0, 1.0, n)
        base = y0 + ext_y * (k + 0.5) / n_lines
        y = base + np.zeros_like(t)
        amp = 0.35 * ext_y / n_lines
        for o in range(1, 6):
            y += amp / o * np.sin(2 * np.pi * (o * rng.uniform(0.7, 1.6) * t
                                               + rng.uniform()))
        x = x0 + ext_x * (t + 0.01 * np.sin(2 * np.pi * 7 * t + rng.uniform()))
        lines.append(np.stack([x, y], axis=1))
    return lines


def synth_random_walk_features(n_lines: int, n_pts: int, step: float,
                               bbox, seed: int = 5) -> np.ndarray:
    """Config-5 style vertex set: ``n_lines`` random-walk polylines of
    ``n_pts`` points, ``step`` metres apart, clipped to ``bbox`` =
    (xmin, ymin, xmax, ymax); returns the (N, 2) vertex array."""
    rng = np.random.default_rng(seed)
    xmin, ymin, xmax, ymax = bbox
    start = np.stack([rng.uniform(xmin, xmax, n_lines),
                      rng.uniform(ymin, ymax, n_lines)], axis=1)
    heading = rng.uniform(0, 2 * np.pi, n_lines)
    dth = rng.normal(0.0, 0.15, size=(n_lines, n_pts - 1))
    th = heading[:, None] + np.cumsum(dth, axis=1)
    steps = step * np.stack([np.cos(th), np.sin(th)], axis=2)
    pts = np.concatenate([start[:, None, :],
                          start[:, None, :] + np.cumsum(steps, axis=1)], axis=1)
    pts = pts.reshape(-1, 2)
    keep = ((pts[:, 0] >= xmin) & (pts[:, 0] <= xmax) &
            (pts[:, 1] >= ymin) & (pts[:, 1] <= ymax))
    return np.ascontiguousarray(pts[keep])


def synth_nodes(n: int, bbox, seed: int = 100, outside_frac: float = 0.0
                ) -> np.ndarray:
    """(n, 2) float64 node cloud uniform in ``bbox``; a fraction is pushed
    outside (those nodes must stay untouched by interpolate)."""
    rng = np.random.default_rng(seed)
    xmin, ymin, xmax, ymax = bbox
    xy = np.empty((n, 2))
    xy[:, 0] = xmin + rng.random(n) * (xmax - xmin)
    xy[:, 1] = ymin + rng.random(n) * (ymax - ymin)
    n_out = int(n * outside_frac)
    if n_out:
        xy[:n_out, 0] += (xmax - xmin) * 1.5
    return xy
