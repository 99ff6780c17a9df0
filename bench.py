#!/usr/bin/env python
"""bench.py -- headline benchmark of the raster hot path (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Workload (BASELINE.json configs[1]): ONE synthetic 20000 x 20000 float32 DEM
(10 m pixels, projected CRS); one "step" = one full HfunRaster pass: FLT_MAX
init + add_contour at 3 levels (-5, 0, 5; expansion_rate 0.001, target_size
100, hmin 100, hmax 1500) + add_gradient_delimiter (hmin 50, hmax 1500), run as
ONE tiled sweep over the raster (HfunRaster(deferred=True): the add_* calls are
queued, flush() enqueues the sweep).  The contour polylines are traced and
resampled once on the host before the timed region: the accelerated path starts
at the resampled vertex array (reference ocsmesh/hfun/raster.py:1230), contour
tracing is the CPU step before it.

  value : Mpx/s, whole job, DEM + vertices + hfun resident in HBM
  e2e   : Mpx/s through the public API (ocsmesh_b200.Raster / Hfun) with HOST
          buffers: DEM and vertices H2D from pinned memory and the size raster
          D2H inside the timed region
  secondary.interp  : Mesh.interpolate of a 10 M-node cloud from 16 DEM tiles
          (BASELINE.json configs[3]) in Mnodes/s, linear / nearest / spline
  secondary.windows : the same pass on the raster chunked into semantic windows
  secondary.weak    : (N > 1) the round-1 weak-scaling run, 20000 rows per GPU

N > 1 is STRONG scaling: the same 20000 x 20000 raster split into N row bands
(2500 rows per GPU at N = 8; one semantic window = the whole raster, so
coordinates and radius law are global).  Every step the ranks all-gather the
contour vertices they traced on their own bands and their boundary DEM rows
over NCCL (one collective into persistent buffers), then run the sweep on
their band.  After the timed region every rank checks sampled rows of its band
bit-exactly against the oracle (cKDTree + numpy stencil): "parity_checked".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LEVELS = [-5.0, 0.0, 5.0]
RATE, TARGET, HMIN, HMAX = 0.001, 100.0, 100.0, 1500.0
LIM_HMIN, LIM_HMAX = 50.0, 1500.0
SEED = 2
# algorithmic bytes per pixel
BYTES_PER_PX = {"feature_eval": 8.0, "flow_limiter": 12.0, "fill": 4.0,
                "clamp": 8.0, "pass_tile": 12.0,
                # split pass: the limiter sweep of a FRESH raster (the bench step)
                # reads the DEM and writes the sizes; there is no pre-op raster to read
                "pass_limiter": 8.0}
PASS_BYTES_PER_PX = 12.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=20000,
                    help="DEM rows and columns (the whole raster)")
    ap.add_argument("--nodes", type=int, default=10_000_000)
    ap.add_argument("--tile", type=int, default=5000, help="interp DEM tile size")
    ap.add_argument("--cpu-sample", type=int, default=8000,
                    help="edge of the square crop timed on the CPU (>= 8000: "
                         "SURVEY.md 8(d) asks for a >= 64 Mpx crop)")
    ap.add_argument("--no-interp", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-windows", action="store_true")
    ap.add_argument("--no-weak", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-numa", action="store_true",
                    help="do not bind ranks to their GPU's local CPUs")
    ap.add_argument("--eager", action="store_true",
                    help="device-resident leg through the eager add_* calls "
                         "(one sweep per op) instead of the queued op list")
    return ap.parse_args()


# ---------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------
def host_dem(height, width, seed, rows=None, pin=True, full_height=None):
    """Synthetic DEM block in (pinned) host memory."""
    import torch
    from ocsmesh_b200 import synthetic as S
    r0, r1 = (0, height) if rows is None else rows
    t = torch.empty((r1 - r0, width), dtype=torch.float32)
    if pin and torch.cuda.is_available():
        t = t.pin_memory()
    a = t.numpy()
    blk = 2000
    for s in range(r0, r1, blk):
        e = min(r1, s + blk)
        a[s - r0:e - r0] = S.synth_dem(full_height or height, width, seed=seed,
                                       rows=(s, e))
    return t


def contour_points(z, transform, full_height, row0=0, native=True):
    """Contours at LEVELS of the DEM rows z = raster rows [row0, row0 + len(z))
    of a raster full_height rows tall, placed with the coordinates of the
    WHOLE raster (Raster.get_x / get_y of the full window), resampled to TARGET
    spacing: the (N, 2) vertex array.  A rank that owns rows [b0, b1) passes
    rows [b0, b1 + 1) (one halo row), so the cell rows of the bands partition
    the cell rows of the raster: the union of the ranks' polylines is the
    whole-raster contour cut at the band boundaries (each piece is resampled
    on its own, which is the vertex-set definition of an N-rank run)."""
    from ocsmesh_b200 import features as F
    h, w = z.shape
    a, _, c, _, e, f = transform
    x = np.linspace(c, c + a * w, w)
    y = np.linspace(f, f + e * full_height, full_height)[row0:row0 + h]
    lines = []
    for lv in LEVELS:
        lines.extend(F.marching_squares(x, y, z, lv))
    return F.feature_points(lines, TARGET, native=native)


class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,"
             "clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}",
                 "--format=csv,noheader,nounits", "-lms", "50", "-i",
                 str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        rows = [l for (t, l) in self.lines if t0 <= t <= t1 + 0.06] or \
               [l for (_, l) in self.lines]
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                 "sw_power_cap"]
        for l in rows:
            p = [v.strip() for v in l.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                smax = float(p[1])
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def ncu_traffic(kclass, px_per_launch):
    """DRAM bytes per launch of the dominant kernel from the committed
    ``ncu --set full`` capture (profiles/traffic.json: dram__bytes_read.sum +
    dram__bytes_write.sum per pixel of that capture), scaled to this launch's
    pixel count; None when no capture is committed for the kernel."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles",
                        "traffic.json")
    try:
        with open(path) as f:
            rec = json.load(f).get(kclass)
        if not rec:
            return None
        per = rec.get("dram_bytes_per_px", rec.get("dram_bytes_per_node"))
        return float(per) * px_per_launch
    except (OSError, ValueError, KeyError, TypeError):
        return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------
# CPU arm: the oracle restatement on the host cores
# ---------------------------------------------------------------------------
def cpu_hfun_sample(z, transform, points, sample, repeats=1):
    """Time the oracle's full pass on the top-left sample x sample crop.
    Returns (Mpx/s, seconds per pass, cores)."""
    from oracle import ocsmesh_oracle as O
    n = min(sample, z.shape[0], z.shape[1])
    crop = np.ascontiguousarray(z[:n, :n])
    r = O.OracleRaster(crop, transform)
    cores = os.cpu_count() or 1
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        oh = O.OracleHfun(r, hmin=HMIN, hmax=HMAX)
        oh.add_feature_points(points, RATE, TARGET, workers=-1)
        oh._clamp_pass()  # add_contour's outer decorator layer
        oh.add_subtidal_flow_limiter(LIM_HMIN, LIM_HMAX, None, None)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n * n / best / 1e6, best, cores, oh.values


def cpu_sample_text(n, nv, dt=None):
    s = (f"top-left {n}x{n} crop of the DEM as its own raster, the FULL "
         f"{nv}-vertex contour set of the 20000^2 raster, "
         "scipy cKDTree.query(workers=-1) + numpy stencil")
    if dt is not None:
        s += f", one pass {dt:.2f} s"
    return s


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port: the same
    scipy / numpy calls with the same arguments) on the host cores; under
    torchrun rank 0 alone runs it.  Nothing here loads libocsb200.so: the
    vertex set comes from the numpy prologue."""
    if rank != 0:
        return
    size = args.size
    from ocsmesh_b200 import synthetic as S
    tr = S.dem_transform(size)
    n = min(args.cpu_sample, size)
    z = host_dem(size, size, SEED, pin=False).numpy()
    pts = contour_points(z, tr, size, native=False)
    z = np.ascontiguousarray(z[:n, :n])
    times = []
    for i in range(args.warmup + args.steps):
        mpx, dt, cores, _ = cpu_hfun_sample(z, tr, pts, n)
        if i >= args.warmup:
            times.append(dt)
    t = float(np.mean(times))
    val = n * n / t / 1e6
    line = {
        "impl": "reference", "metric": "hfun_raster_pass_throughput",
        "value": val, "unit": "Mpx/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None,
        "dtype": "f64 distance / f32 size", "data": "synthetic",
        "config": workload_config(args, world),
        "n_vertices": int(len(pts)),
        "cpu_baseline": {"value": val, "unit": "Mpx/s", "cores": cores,
                         "kind": "port", "sample": cpu_sample_text(n, len(pts))},
        "e2e": {"value": val, "unit": "Mpx/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    """Identical in both arms (the reference arm times a bounded sample of the
    same workload; the sample is described in its cpu_baseline)."""
    return {
        "workload": (f"BASELINE.json configs[1]: synthetic {args.size}x"
                     f"{args.size} f32 DEM, HfunRaster full pass = init fill + "
                     f"add_contour(levels {LEVELS}, rate {RATE}, target {TARGET}, "
                     f"hmin {HMIN}, hmax {HMAX}) + add_gradient_delimiter(hmin "
                     f"{LIM_HMIN}, hmax {LIM_HMAX})"),
        "px": args.size * args.size,
        "l2": "inputs (DEM + hfun, 3.2 GB) exceed the 126 MB L2",
        "contours": "traced + resampled on the host once, outside the timed "
                    "region (parity boundary hfun/raster.py:1230)",
    }


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
class BandRun:
    """One rank's share of an H x W raster split into `world` row bands: host
    and device inputs, the exchange buffers, the queued-op HfunRaster."""

    def __init__(self, args, H, W, rank, world, dev, eager=False, chunk=None,
                 overlap=None):
        import torch
        import ocsmesh_b200 as ob
        from ocsmesh_b200 import synthetic as S, distributed as D
        self.H, self.W, self.rank, self.world, self.dev = H, W, rank, world, dev
        self.tr = S.dem_transform(H)
        self.part = D.RowBands(H, world)
        self.b0, self.b1 = self.part.band(rank)
        self.s0, self.s1 = self.part.stored(rank)
        # host: the band plus one halo row per side (what a rank reading its
        # rows of the GeoTIFF would over-read)
        self.z_host = host_dem(H, W, SEED, rows=(self.s0, self.s1), full_height=H)
        zb = self.z_host.numpy()
        c1 = min(self.b1 + 1, H)
        self.pts_local = contour_points(zb[self.b0 - self.s0:c1 - self.s0], self.tr,
                                        H, row0=self.b0)
        self.pts_pinned = torch.from_numpy(self.pts_local).pin_memory()
        self.raster = D.band_raster(zb[self.b0 - self.s0:self.b1 - self.s0], self.tr,
                                    H, self.part, rank, crs=32618, device=dev)
        if chunk is not None:
            self.raster.chunk_size = chunk
            self.raster.overlap = overlap
        self.hfun = ob.HfunRaster(self.raster, hmin=HMIN, hmax=HMAX, device=dev,
                                  row_band=(self.b0, self.b1), deferred=not eager)
        self.verts_local = self.pts_pinned.to(f"cuda:{dev}")
        self.exch = D.BandExchange(len(self.pts_local), W, torch.float32,
                                   f"cuda:{dev}", self.part, rank)
        self.nv_total = int(sum(self.exch.sizes))

    def close(self):
        """Release the peer-memory exchange buffers (every rank calls it)."""
        self.exch.close()

    def step(self):
        """A4 init, vertex all-gather, DEM halo rows, A6/A7 + A8(A9): one sweep."""
        h = self.hfun
        h.reset()
        verts = self.exch.exchange(self.verts_local,
                                   self.raster.device_band(1, self.dev),
                                   self.raster.device_band_origin(1, self.dev))
        h.add_feature_points(verts, RATE, TARGET)
        h.add_gradient_delimiter(hmin=LIM_HMIN, hmax=LIM_HMAX)
        h.flush()

    def check_parity(self, n_rows=4, seed=0):
        """Sampled rows of this rank's band against the oracle, bit-exact: the
        nearest-vertex distance of every pixel of the row from an exact
        cKDTree query over the REAL gathered vertices, the size law and clamps
        restated from hfun/raster.py:1240-1257, then the numpy stencil on the
        3-row strip (np.gradient is local).  Returns the number of pixels."""
        import torch
        from oracle import ocsmesh_oracle as O
        self.hfun.sync()
        got_all = self.hfun.device_values
        verts = self.exch.gather_vertices(self.verts_local)
        pts = self.exch.real_vertices(verts).cpu().numpy()
        H, W = self.H, self.W
        b = O.array_bounds(H, W, self.tr)
        x = np.linspace(b[0], b[2], W)
        y = np.linspace(b[3], b[1], H)
        rng = np.random.default_rng(seed + 17 * self.rank)
        lo, hi = max(self.b0, 1), min(self.b1, H - 1)
        rows = sorted(set(rng.integers(lo, hi, n_rows).tolist()) |
                      {lo, hi - 1})          # incl. the rows next to the band edges
        zb = self.z_host.numpy()
        n_px = 0
        for r in rows:
            strip = zb[r - 1 - self.s0:r + 2 - self.s0]
            yy = y[r - 1:r + 2]
            xy = np.stack([np.tile(x, 3), np.repeat(yy, W)], axis=1)
            old = np.full((3, W), np.finfo(np.float32).max, np.float32)
            feat = O.add_feature_window(old, pts, xy, 3, W, RATE, TARGET, HMIN, HMAX)
            feat = O.apply_constraints_clamp(feat, HMIN, HMAX)
            ref = O.flow_limiter_window(strip, feat, self.raster.dx, self.raster.dy,
                                        LIM_HMIN, LIM_HMAX, None, None, HMIN, HMAX)
            ref = O.apply_constraints_clamp(np.ma.getdata(ref), HMIN, HMAX)[1]
            got = got_all[r - self.b0].cpu().numpy()
            if not np.array_equal(got, ref):
                bad = int(np.flatnonzero(got != ref)[0])
                raise AssertionError(
                    f"rank {self.rank}: row {r} differs from the oracle at column "
                    f"{bad}: {got[bad]!r} vs {ref[bad]!r}")
            n_px += W
        return n_px


def time_steps(fn, n, dev, world):
    """ms per step: CUDA events around n steps, barrier + synchronize on both
    sides, max over ranks."""
    import torch
    import torch.distributed as dist

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    # host time to enqueue the steps: if it is close to the device time the
    # launches, not the kernels, set the pace
    time_steps.host_ms = (time.perf_counter() - t0) * 1e3 / n
    barrier()
    t1 = time.perf_counter()
    ms = e0.elapsed_time(e1)
    if world > 1:
        tt = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{dev}")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    return ms / n, t0, t1


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import ocsmesh_b200 as ob
    from ocsmesh_b200 import _lib, synthetic as S, distributed as D

    torch.cuda.set_device(local_rank)
    dev = local_rank
    # multi-rank: pinned staging buffers on the NUMA node of the rank's GPU
    numa = D.bind_to_gpu_numa(dev) if (world > 1 and not args.no_numa) else None
    ctx = _lib.default_context(dev)
    size = args.size
    H = W = size

    note("inputs")
    run = BandRun(args, H, W, rank, world, dev, eager=args.eager)
    nv_total = run.nv_total
    note(f"warm-up ({len(run.pts_local)} local / {nv_total} vertices)")
    for _ in range(max(args.warmup, 3)):
        run.step()
    torch.cuda.synchronize(dev)
    note("timed steps")

    # ---- timed: device-resident -------------------------------------------
    sampler = ClockSampler(dev)
    sampler.start()
    time.sleep(0.15)
    # headline region: K steps, nothing but the step's own work on the stream
    launches0 = ctx.launch_count
    ms_step, t0, t1 = time_steps(run.step, args.steps, dev, world)
    host_ms_step = time_steps.host_ms
    clocks = sampler.stop(t0, t1)
    launches = ctx.launch_count - launches0
    # per-kernel region: the same K steps again with an event pair around every
    # launch of the library (ocsb_ctx_set_timing).  The ~18 extra stream
    # operations per step cost ~0.025 ms (tools/host_profile.py), which a 0.3 ms
    # multi-GPU step notices: they stay out of the headline region
    ctx.reset_timing()
    ctx.set_timing(True)
    ms_step_instr, _, _ = time_steps(run.step, args.steps, dev, world)
    ctx.set_timing(False)
    ktimes = ctx.get_timing()
    px_total = float(H) * W
    px_rank = float(run.b1 - run.b0) * W
    value = px_total / (ms_step * 1e-3) / 1e6

    note(f"timed: {ms_step:.4f} ms per step (host enqueue {host_ms_step:.4f} ms per step)")
    # ---- parity at this rank count (outside the timed region) ---------------
    parity = None
    if not args.no_parity:
        err = None
        try:
            n_px = run.check_parity()
        except Exception as e:          # every rank must reach the reduction
            n_px, err = 0, f"{type(e).__name__}: {e}"
            print(f"[bench rank {rank}] parity check FAILED: {err}", file=sys.stderr,
                  flush=True)
        tot = torch.tensor([n_px, 1 if err else 0], dtype=torch.int64,
                           device=f"cuda:{dev}")
        if world > 1:
            dist.all_reduce(tot)
        parity = {"checked": True, "ok": int(tot[1].item()) == 0,
                  "failed_ranks": int(tot[1].item()), "pixels": int(tot[0].item()),
                  "how": "per rank: 4 random rows + the 2 rows next to its band "
                         "edges, every pixel bit-exact against cKDTree + size law "
                         "+ numpy stencil on the 3-row strip"}

    note("parity done")
    # ---- roofline of the dominant kernel -----------------------------------
    peak, peak_src = measured_peaks()
    # the HBM sweeps only: in the split pass the pyramid's launches (feature_refine)
    # queue behind the limiter sweep, so their event times include that wait
    sweeps = {k: v for k, v in ktimes.items() if k in BYTES_PER_PX}
    dom = max(sweeps.items(), key=lambda kv: kv[1][0])[0] if sweeps else None
    roofline = None
    if dom is not None:
        ms, n = ktimes[dom]
        avg = ms / n
        bpp = BYTES_PER_PX.get(dom, 8.0)
        px_launch = px_rank * (args.steps / n) if n else 0.0
        ach = bpp * px_launch / (avg * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak,
                    "unit": "GB/s", "frac": ach / peak,
                    "traffic": ncu_traffic(dom, px_launch),
                    "peak_source": peak_src,
                    "algorithmic_bytes_per_px": bpp,
                    "avg_launch_ms": avg,
                    # the north-star figure: the WHOLE pass (every kernel, gaps
                    # and collectives of a step) at 12 B/px per GPU
                    "pass_frac": (PASS_BYTES_PER_PX * px_rank / (ms_step * 1e-3)
                                  / 1e9 / peak),
                    "pass_achieved": PASS_BYTES_PER_PX * px_rank / (ms_step * 1e-3) / 1e9,
                    "kernel_ms_per_step": {k: v[0] / args.steps for k, v in ktimes.items()},
                    "split_pass": "pass_limiter" in ktimes,
                    "kernel_times_from": "a second timed region of the same K steps with "
                                         "an event pair around every launch "
                                         f"({ms_step_instr:.4f} ms per step there)",
                    "note": (("split pass (k_pass.cu): pass_limiter = limiter sweep of all "
                              "tiles on a second stream, pass_feature = feature op on the "
                              "tiles with candidates; feature_refine (the candidate pyramid) "
                              "is timed by events on its stream and includes its wait for "
                              "SM slots behind the sweep. ") if "pass_limiter" in ktimes else "")
                            + "algorithmic bytes per SURVEY.md 8(d): fused pass = DEM "
                              "read + hfun read + hfun write; the first op of a fresh "
                              "raster takes the initial state as a NULL pointer, so "
                              "the real traffic (`traffic`) is below it"}

    # ---- e2e: public API, host buffers -------------------------------------
    e2e = None
    if not args.no_e2e:
        e2e = bench_e2e(args, run, rank, world, dev, numa, px_total)

    # ---- CPU baseline (rank 0, N == 1) --------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        n = min(args.cpu_sample, size)
        zc = np.ascontiguousarray(run.z_host.numpy()[:n, :n])
        mpx, dt, cores, _ = cpu_hfun_sample(zc, run.tr, run.pts_local, n)
        # the GPU result on the same crop is NOT comparable pixel-for-pixel (a
        # crop is its own semantic window), so this leg only times the oracle
        cpu = {"value": mpx, "unit": "Mpx/s", "cores": cores, "kind": "port",
               "sample": cpu_sample_text(n, len(run.pts_local), dt)}

    # ---- secondary legs -----------------------------------------------------
    secondary = {}
    if not args.no_windows:
        secondary["windows"] = bench_windows(args, run, rank, world, dev, ms_step)
    if world == 1 and not args.no_windows:
        secondary["add_contour_api"] = bench_add_contour(args, run, dev)
    exchange_mode = run.exch.mode if world > 1 else "none (one rank)"
    exchange_failed = run.exch.p2p_failed()
    if world > 1 and not args.no_weak:
        run.close()
        del run
        torch.cuda.empty_cache()
        secondary["weak"] = bench_weak(args, rank, world, dev)
        run = None
    if not args.no_interp:
        secondary["interp"] = bench_interp(args, rank, world, dev, ctx, peak)

    if rank == 0:
        cfg = workload_config(args, world)
        line = {
            "metric": "hfun_raster_pass_throughput", "value": value,
            "unit": "Mpx/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64 distance / f32 size", "data": "synthetic",
            "config": cfg,
            "parallelism": (f"row bands x{world} of ONE {H}x{W} raster "
                            f"({H // world} rows per GPU); per step: vertex "
                            "all-gather + boundary DEM rows in one exchange (see "
                            "`exchange`), then the sweep of the band"),
            "exchange": {"mode": exchange_mode,
                         "p2p": "ocsb_band_push / ocsb_band_pull: stores into the "
                                "peers' CUDA-IPC buffers over NVLink + epoch flags, "
                                "no collective inside the step",
                         "nccl": "ocsb_band_pack + one all_gather_into_tensor + "
                                 "ocsb_band_unpack",
                         "gave_up_waiting": bool(exchange_failed)},
            "vertex_set": ("contours of each rank's band (cell rows partition the "
                           "raster's), each piece resampled on its own; all-"
                           "gathered every step"),
            "n_vertices": nv_total,
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "parity_checked": bool(parity and parity["ok"]), "parity": parity,
            "gpu_launches": int(launches), "clocks": clocks,
            "host_enqueue_ms_per_step": round(host_ms_step, 4),
            "secondary": secondary,
        }
        print(json.dumps(line), flush=True)


def bench_e2e(args, run, rank, world, dev, numa, px_total):
    """The same pass through the public API with HOST buffers: per step the
    rank's DEM rows and vertices go up from pinned memory, the size band comes
    down (ocsb_hfun_pass streams row bands: H2D, kernels and D2H overlap)."""
    import torch
    import torch.distributed as dist
    import ocsmesh_b200 as ob
    H, W = run.H, run.W
    rows = run.b1 - run.b0
    out_pinned = torch.empty((rows, W), dtype=torch.float32).pin_memory()
    z_e2e = run.z_host
    pts_pinned = run.pts_pinned

    def e2e_step():
        r = ob.Raster.from_numpy(z_e2e.numpy(), run.tr, crs=32618,
                                 row_origin=run.s0, full_height=H)
        h = ob.HfunRaster(r, hmin=HMIN, hmax=HMAX, device=dev,
                          row_band=(run.b0, run.b1), deferred=True)
        v = pts_pinned                                         # host vertices
        if world > 1:
            v = run.exch.gather_vertices(
                pts_pinned.to(f"cuda:{dev}", non_blocking=True))
        h.add_feature_points(v, RATE, TARGET)
        h.add_gradient_delimiter(hmin=LIM_HMIN, hmax=LIM_HMAX)
        h.get_values(out=out_pinned)             # DEM H2D + pass + hfun D2H
        return float(out_pinned[0, 0])

    e2e_step()
    e2e_steps = max(2, min(args.steps, 5))
    ems, _, _ = time_steps(e2e_step, e2e_steps, dev, world)
    res = {"value": px_total / (ems * 1e-3) / 1e6, "unit": "Mpx/s",
           "h2d_bytes_per_step": int(z_e2e.numel() * 4 + pts_pinned.numel() * 8),
           "d2h_bytes_per_step": int(out_pinned.numel() * 4),
           "path": "HfunRaster(deferred=True).get_values(out=pinned) -> "
                   "ocsb_hfun_pass: row bands streamed host->device->host, copies "
                   "overlapped with the kernels",
           "steps": e2e_steps, "ms_per_step": ems,
           "bound": "PCIe on one GPU (1.6 GB up + 1.6 GB down per 400 Mpx); beyond "
                    "two ranks the host's copy bandwidth (DESIGN.md 6)",
           "host_placement": numa}
    del out_pinned
    return res


def bench_windows(args, run, rank, world, dev, ms_single):
    """configs[1] "tiled by raster windows": the same pass with the raster cut
    into semantic windows (own linspace coordinates, radius and border rules
    per window); the window table is consumed by the same launches."""
    import torch
    import ocsmesh_b200 as ob
    out = {"single_window_ms": ms_single}
    steps = max(3, min(args.steps, 10))
    cases = [("chunk5000_overlap0", 5000, 0), ("chunk1000_overlap0", 1000, 0),
             ("chunk5000_overlap2", 5000, 2)]
    for name, chunk, ovl in cases:
        if chunk >= run.W:
            continue
        raster = run.raster
        raster.chunk_size, raster.overlap = chunk, ovl
        hf = ob.HfunRaster(raster, hmin=HMIN, hmax=HMAX, device=dev,
                           row_band=(run.b0, run.b1), deferred=True)
        verts = run.exch.gather_vertices(run.verts_local)

        def step():
            hf.reset()
            hf.add_feature_points(verts, RATE, TARGET)
            hf.add_gradient_delimiter(hmin=LIM_HMIN, hmax=LIM_HMAX)
            hf.flush()
        for _ in range(2):
            step()
        ms, _, _ = time_steps(step, steps, dev, world)
        hf.sync()
        n_win = len([w for w in hf._windows() if hf._band_of(w) is not None])
        out[name] = {"ms_per_step": ms, "windows_on_rank0": n_win,
                     "vs_single_window": ms / ms_single,
                     "mode": "one sweep for the op list" if hf._can_defer()
                             else "one sweep per op (overlapping windows: every "
                                  "window must see the pre-op raster)"}
        del hf
    run.raster.chunk_size, run.raster.overlap = None, None
    return out


def bench_add_contour(args, run, dev):
    """The pass through the public call that starts at the DEM:
    Hfun.add_contour(levels, ...) traces the contours (sweep, stitch by list
    ranking, repartition, arc-length resampling -- all on the device) and feeds
    the vertex array to the sweep; nothing is traced outside the timed region."""
    import torch
    import ocsmesh_b200 as ob
    hf = ob.HfunRaster(run.raster, hmin=HMIN, hmax=HMAX, device=dev,
                       row_band=(run.b0, run.b1), deferred=True)

    def step():
        hf.reset()
        hf.add_contour(LEVELS, RATE, TARGET)
        hf.add_gradient_delimiter(hmin=LIM_HMIN, hmax=LIM_HMAX)
        hf.flush()
    for _ in range(2):
        step()
    steps = max(3, min(args.steps, 10))
    ms, _, _ = time_steps(step, steps, dev, 1)
    hf.sync()

    pts = run.raster.get_contour_points(LEVELS, TARGET, 200, device=dev)
    same = bool(torch.equal(pts, run.verts_local))
    return {"ms_per_step": ms,
            "n_vertices": int(pts.shape[0]),
            "equals_host_prologue": same,
            "path": "HfunRaster.add_contour(levels) + add_gradient_delimiter + flush: "
                    "per level count + emit sweeps (4 B/px each), radix sort of the "
                    "segment ends, pointer-jumping list ranking, chunk-wise "
                    "resampling; 3 host syncs per level for the array sizes",
            "parity": "polylines identical to the host stitch (tests/"
                      "test_contour_gpu.py); parity with contourpy / shapely "
                      "unpinned (not importable here)"}


def bench_weak(args, rank, world, dev):
    """Round-1 layout kept as a secondary: the raster grows with N (20000 rows
    per GPU), same exchanges per step."""
    import torch
    H, W = args.size * world, args.size
    run = BandRun(args, H, W, rank, world, dev)
    for _ in range(3):
        run.step()
    steps = max(3, min(args.steps, 10))
    ms, _, _ = time_steps(run.step, steps, dev, world)
    run.hfun.sync()
    res = {"scaling": "weak", "raster": f"{H}x{W}", "ms_per_step": ms,
           "value": float(H) * W / (ms * 1e-3) / 1e6, "unit": "Mpx/s",
           "n_vertices": run.nv_total, "exchange": run.exch.mode}
    run.close()
    del run
    torch.cuda.empty_cache()
    return res


def bench_interp(args, rank, world, dev, ctx, peak):
    """BASELINE.json configs[3]: 10 M nodes x 16 DEM tiles, linear + nearest.
    Nodes are sharded over ranks (contiguous chunks), tiles replicated."""
    import torch
    import torch.distributed as dist
    import ocsmesh_b200 as ob
    from ocsmesh_b200 import synthetic as S
    from ocsmesh_b200.mesh import EuclideanMesh2D
    T, ovl, grid = args.tile, 100, 4
    rasters = []
    for k in range(grid * grid):
        ty, tx = divmod(k, grid)
        z = S.synth_dem(T, T, seed=20 + k)
        tr = list(S.dem_transform(T))
        tr[2] += (T - ovl) * S.PIXEL * tx
        tr[5] += (T - ovl) * S.PIXEL * ty
        r = ob.Raster.from_numpy(z, tr, crs=32618)
        r.device_band(1, dev)
        rasters.append(r)
    ext = ((T - ovl) * (grid - 1) + T) * S.PIXEL
    bbox = (S.X0, S.Y0, S.X0 + ext, S.Y0 + ext)
    K = args.nodes * world        # weak scaling: nodes per GPU fixed
    k0, k1 = rank * args.nodes, (rank + 1) * args.nodes
    xy_host = torch.from_numpy(S.synth_nodes(K, bbox, seed=400, outside_frac=0.02)[k0:k1].copy()).pin_memory()
    n = k1 - k0
    xy_raw = xy_host.to(f"cuda:{dev}")
    # resident layout: nodes bucketed into 64x64-px cells once (the bucketing
    # pass itself is timed separately and is inside the e2e figure)
    torch.cuda.synchronize(dev)
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    xy, perm = EuclideanMesh2D.bucket_nodes(ctx, rasters, xy_raw, dev)
    b0.record()
    xy, perm = EuclideanMesh2D.bucket_nodes(ctx, rasters, xy_raw, dev)
    b1.record()
    torch.cuda.synchronize(dev)
    bucket_ms = b0.elapsed_time(b1)
    vals = torch.full((n,), float("nan"), dtype=torch.float64, device=f"cuda:{dev}")
    touched = torch.zeros(n, dtype=torch.uint8, device=f"cuda:{dev}")
    out = {}
    steps = max(3, min(args.steps, 10))
    for method, bpn in (("linear", 40.0), ("nearest", 28.0)):
        def timed(p):
            for _ in range(3):
                EuclideanMesh2D._interp_device(ctx, rasters, xy, vals, touched, method, 1, dev, perm=p)
            torch.cuda.synchronize(dev)
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                EuclideanMesh2D._interp_device(ctx, rasters, xy, vals, touched, method, 1, dev, perm=p)
            e1.record()
            torch.cuda.synchronize(dev)
            return e0.elapsed_time(e1) / steps
        # bucket-resident layout: nodes AND values live in bucket order; with
        # `perm` the kernel scatters the values back to mesh order; the headline
        # is what the public Mesh.interpolate runs: a gather in mesh order
        ms_bucket = timed(None)
        ms_unpermute = timed(perm)
        xy_save, xy = xy, xy_raw
        ms = timed(None)
        xy = xy_save
        # e2e: coords H2D + values D2H through Mesh.interpolate
        vh = torch.empty(n, dtype=torch.float64).pin_memory()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(2):
            x2 = xy_host.to(f"cuda:{dev}", non_blocking=True)
            v2 = torch.full((n,), float("nan"), dtype=torch.float64, device=f"cuda:{dev}")
            t2 = torch.zeros(n, dtype=torch.uint8, device=f"cuda:{dev}")
            EuclideanMesh2D._interp_device(ctx, rasters, x2, v2, t2, method, 1, dev)
            vh.copy_(v2)
        a1.record()
        torch.cuda.synchronize(dev)
        ems = a0.elapsed_time(a1) / 2
        if world > 1:
            tt = torch.tensor([ms, ems], dtype=torch.float64, device=f"cuda:{dev}")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms, ems = float(tt[0]), float(tt[1])
        # one launch serves all tiles: coordinates read once, 4 (1) texels and
        # one value per node
        ach = (bpn * n) / (ms * 1e-3) / 1e9      # headline: the mesh-order path
        # what the gather really moves (committed ncu capture): a touched DEM
        # row costs a 64-byte DRAM access whatever part of it is used
        traffic = ncu_traffic("interp_" + method, n)
        out[method] = {"Mnodes_per_s": K / (ms * 1e-3) / 1e6, "ms_per_step": ms,
                       "path": "mesh-order gather (what Mesh.interpolate runs), "
                               "device-timed",
                       "roofline_frac": ach / peak,
                       "roofline_bound": "sectors: one node per ~40 px, so every "
                                         "texel row of a node costs its own 64-byte "
                                         "DRAM access (profiles/r1h_interp_ncu_"
                                         "summary.txt: 151 / 110 B per node moved "
                                         "for 40 / 28 algorithmic)",
                       "algorithmic_bytes_per_node": bpn,
                       "bucket_resident": {
                           "ms_per_step": ms_bucket,
                           "roofline_frac": (bpn * n) / (ms_bucket * 1e-3) / 1e9 / peak,
                           "traffic": traffic,
                           "traffic_frac_of_peak": (traffic / (ms_bucket * 1e-3) / 1e9
                                                    / peak if traffic else None),
                           "ms_with_unpermuted_outputs": ms_unpermute,
                           "bucket_pass_ms": bucket_ms},
                       "e2e_Mnodes_per_s": K / (ems * 1e-3) / 1e6}
    # spline (the reference's default method): prefilter of every tile + one
    # evaluation launch, nodes in mesh order; no bucket / scatter variants
    def spline_call():
        EuclideanMesh2D._interp_device(ctx, rasters, xy_raw, vals, touched, "spline", 1, dev)
    for r in rasters:                       # forget cached spline coefficients
        for k in [k for k in r._dev_cache if isinstance(k, tuple) and k and
                  k[0] == "spline_coef"]:
            del r._dev_cache[k]
    torch.cuda.synchronize(dev)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    spline_call()                           # first call: prefilter of all tiles
    f1.record()
    torch.cuda.synchronize(dev)
    first_ms = f0.elapsed_time(f1)
    ctx.reset_timing()
    ctx.set_timing(True)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(3):
        spline_call()
    s1.record()
    torch.cuda.synchronize(dev)
    ctx.set_timing(False)
    sms = s0.elapsed_time(s1) / 3
    skt = {k: v[0] / 3 for k, v in ctx.get_timing().items() if k.startswith("spline")}
    if world > 1:
        tt = torch.tensor([sms], dtype=torch.float64, device=f"cuda:{dev}")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sms = float(tt[0])
    out["spline"] = {"Mnodes_per_s": K / (sms * 1e-3) / 1e6, "ms_per_step": sms,
                     "first_call_ms": first_ms,
                     "roofline_frac": 88.0 * n / (sms * 1e-3) / 1e9 / peak,
                     "algorithmic_bytes_per_node": 88.0,
                     "kernel_ms_per_step": skt,
                     "note": "the B-spline coefficients of a raster window are "
                             "computed by the first call (prefilter of all 16 "
                             "tiles, 16 B/px: first_call_ms) and kept with the "
                             "raster's device band; later calls only evaluate "
                             "(88 B/node), nodes in mesh order"}
    out["config"] = (f"BASELINE.json configs[3]: {K} nodes ({args.nodes}/GPU), "
                     f"16 tiles {T}x{T} (4x4, {ovl}-px overlap), last wins; "
                     "nodes (and their values) resident in spatial-bucket order "
                     "(ocsb_bucket_nodes, 64-px cells); e2e = what Mesh.interpolate "
                     "does: H2D of mesh-order coordinates + gather in mesh order "
                     "+ D2H of the values")
    return out


def note(msg):
    """Progress marker on stderr (BENCH_VERBOSE=1)."""
    if os.environ.get("BENCH_VERBOSE"):
        print(f"[bench rank {os.environ.get('RANK', '0')} +{time.perf_counter() - T_START:.1f}s] "
              f"{msg}", file=sys.stderr, flush=True)


T_START = time.perf_counter()


def main():
    args = parse_args()
    wd = os.environ.get("BENCH_WATCHDOG")
    if wd:
        # a hung run prints where every thread of every rank stands, then exits
        import faulthandler
        faulthandler.dump_traceback_later(float(wd), exit=True)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU "
                         "fallback (use --impl reference for the CPU arm)")
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
