"""Multi-GPU plumbing: one process per GPU, ``torch.distributed`` (NCCL over
NVLink on the GPU box, gloo in the CPU tests).

The raster hot path shards by raster rows:

* **feature / contour distance** -- pixels are independent given the full
  vertex set: each rank owns a row band of the (single) semantic window and
  the ranks ``all_gather`` the contour vertices they traced on their own bands
  (:func:`all_gather_vertices`).
* **flow limiter / gradient stencil** -- one exchange: a rank needs the DEM row
  above and below its band (:func:`exchange_dem_halo`, pairwise send/recv of
  ``W`` floats).  Rows on the border of the *semantic window* get no halo: the
  kernel switches to numpy's one-sided difference there.
* **multi-raster min-combine** -- raster ``r`` lives on rank ``r % world``;
  every overlapping pair swaps its overlap strips and takes the min
  (:func:`min_combine_overlaps`); a strip shared by all ranks can also be
  min-reduced in one ``all_reduce(MIN)`` (:func:`min_combine_strip`).
* **Mesh.interpolate** -- nodes are independent: contiguous node shards,
  rasters replicated (:func:`node_shard`); results are all-gathered.

Nothing here touches pixel data on the host.
"""
from __future__ import annotations

import logging
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

_logger = logging.getLogger(__name__)


class RowBands:
    """Contiguous, near-equal row bands of a raster ``height`` rows tall."""

    def __init__(self, height: int, world: int):
        if world < 1 or height < world:
            raise ValueError(f"cannot split {height} rows over {world} ranks")
        self.height = int(height)
        self.world = int(world)
        base, rem = divmod(self.height, self.world)
        self.starts = [r * base + min(r, rem) for r in range(self.world + 1)]

    def band(self, rank: int) -> Tuple[int, int]:
        return self.starts[rank], self.starts[rank + 1]

    def stored(self, rank: int) -> Tuple[int, int]:
        """Rows a rank keeps of the DEM: its band plus one halo row per side
        (clipped to the raster)."""
        r0, r1 = self.band(rank)
        return max(0, r0 - 1), min(self.height, r1 + 1)

    def owner(self, row: int) -> int:
        for r in range(self.world):
            if self.starts[r] <= row < self.starts[r + 1]:
                return r
        raise IndexError(row)


def node_shard(n_nodes: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous node range of a rank for Mesh.interpolate."""
    base, rem = divmod(int(n_nodes), int(world))
    k0 = rank * base + min(rank, rem)
    return k0, k0 + base + (1 if rank < rem else 0)


def band_raster(band_values: np.ndarray, transform, full_height: int,
                part: RowBands, rank: int, crs=None, device: Optional[int] = None):
    """A ``Raster`` that holds only this rank's band ``band_values`` (rows
    ``part.band(rank)``) of a raster ``full_height`` rows tall, with its DEM
    band uploaded to ``device`` into a buffer that has room for the halo rows
    (filled by :func:`exchange_dem_halo`)."""
    import torch
    from .raster import Raster
    r0, r1 = part.band(rank)
    if band_values.shape[0] != r1 - r0:
        raise ValueError("band_values must hold exactly the rank's band rows")
    r = Raster.from_numpy(band_values, transform, crs=crs, row_origin=r0,
                          full_height=full_height)
    if device is not None:
        s0, s1 = part.stored(rank)
        w = band_values.shape[1]
        dtype = torch.float32 if band_values.dtype == np.float32 else torch.float64
        buf = torch.zeros((s1 - s0, w), dtype=dtype, device=f"cuda:{device}")
        src = torch.from_numpy(np.ascontiguousarray(band_values))
        buf[r0 - s0:r1 - s0].copy_(src, non_blocking=True)
        r.set_device_band(buf, s0, band=1, device=device)
        # halo slots are zeros until the first exchange (get_contour refuses)
        r._halo_pending[(1, int(device))] = (s1 - s0) != (r1 - r0)
    return r


def exchange_halo_rows(buf, origin: int, part: RowBands, rank: int,
                       group=None) -> None:
    """``buf`` holds raster rows ``[origin, origin + len(buf))`` of this rank
    (its band plus halo slots).  Sends the first / last owned row to the
    previous / next rank and receives their boundary rows into the halo slots
    (``W`` elements each; the send/recv pairs are batched in one group)."""
    import torch.distributed as dist
    if part.world == 1:
        return
    r0, r1 = part.band(rank)
    ops = []
    if rank > 0:
        ops.append(dist.P2POp(dist.isend, buf[r0 - origin], rank - 1, group))
        ops.append(dist.P2POp(dist.irecv, buf[r0 - 1 - origin], rank - 1, group))
    if rank < part.world - 1:
        ops.append(dist.P2POp(dist.isend, buf[r1 - 1 - origin], rank + 1, group))
        ops.append(dist.P2POp(dist.irecv, buf[r1 - origin], rank + 1, group))
    for req in dist.batch_isend_irecv(ops):
        req.wait()


def exchange_dem_halo(raster, part: RowBands, rank: int, device: Optional[int] = None,
                      band: int = 1, group=None) -> None:
    """Fill the halo rows of this rank's device DEM band (see
    :func:`band_raster`) from its neighbours over NCCL."""
    import torch
    if device is None:
        device = torch.cuda.current_device()
    exchange_halo_rows(raster.device_band(band, device),
                       raster.device_band_origin(band, device), part, rank, group)
    if hasattr(raster, "_halo_pending"):
        raster._halo_pending[(int(band), int(device))] = False


def gather_sizes(n_local: int, device, group=None) -> List[int]:
    """Every rank's row count, on the host (one small all-gather + one host
    sync).  Vertex sets that do not change between passes exchange their sizes
    once and hand the list to :func:`all_gather_ragged` afterwards, which then
    never waits on the host."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n = torch.tensor([int(n_local)], dtype=torch.int64, device=device)
    sizes = torch.empty(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(sizes, n, group=group)
    return [int(v) for v in sizes.tolist()]


def all_gather_ragged(local, group=None, sizes: Optional[Sequence[int]] = None):
    """All ranks' ``(n_i, ...)`` tensors (same dtype and trailing shape, any
    ``n_i`` including 0) concatenated in rank order on every rank.  Ragged
    sizes are handled by padding to the maximum.  ``sizes`` (from
    :func:`gather_sizes`) skips the size exchange and its host sync."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if world == 1:
        return local
    if sizes is None:
        sizes_h = gather_sizes(local.shape[0], local.device, group)
    else:
        sizes_h = [int(v) for v in sizes]
        if len(sizes_h) != world or sizes_h[dist.get_rank(group)] != local.shape[0]:
            raise ValueError("sizes does not describe this gather")
    m = max(sizes_h)
    tail = tuple(local.shape[1:])
    if m == 0:
        return local.new_empty((0,) + tail)
    pad = torch.zeros((m,) + tail, dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    out = torch.empty((world * m,) + tail, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    parts = [out[r * m:r * m + sizes_h[r]] for r in range(world)]
    return torch.cat(parts, dim=0).contiguous()


def all_gather_vertices(verts_local, group=None, sizes: Optional[Sequence[int]] = None):
    """All ranks' ``(n_i, 2)`` float64 vertex arrays concatenated in rank
    order -> ``(sum n_i, 2)`` on every rank (the contour-vertex all-gather)."""
    return all_gather_ragged(verts_local, group, sizes)


class BandExchange:
    """Per-step exchanges of a row-band run with persistent buffers: nothing is
    allocated, padded or concatenated inside a step.

    * vertices: every rank contributes its ``(n_i, 2)`` float64 contour vertices
      in a fixed-capacity block (``cap`` = the largest ``n_i``, exchanged once);
      the unused tail of a block repeats the rank's last vertex -- duplicates
      cannot change a nearest-vertex distance -- or, for a rank without
      vertices, a far-away finite point that every candidate test drops.  ONE
      ``all_gather_into_tensor`` fills the ``(world * cap, 2)`` array the
      kernels read in place.
    * DEM halo rows: every rank contributes its first and last owned row
      (``2 x W`` elements) to ONE ``all_gather_into_tensor``; the two rows a rank
      needs are then copied into the halo slots of its stored band.
    """

    FAR = 1.0e13    # |coordinate| of the filler vertex of an empty rank (< 1e14)

    def __init__(self, n_local: int, width: int, dem_dtype, device, part: RowBands,
                 rank: int, group=None):
        import torch
        import torch.distributed as dist
        self.group = group
        self.part, self.rank = part, int(rank)
        self.world = part.world
        self.device = device
        self.sizes = gather_sizes(n_local, device, group) if self.world > 1 \
            else [int(n_local)]
        self.cap = max(max(self.sizes), 1)
        self.block = torch.empty((self.cap, 2), dtype=torch.float64, device=device)
        self.verts = torch.empty((self.world * self.cap, 2), dtype=torch.float64,
                                 device=device)
        self.rows_out = torch.empty((2, width), dtype=dem_dtype, device=device)
        self.rows_all = torch.empty((self.world * 2, width), dtype=dem_dtype,
                                    device=device)
        # packed form (one collective per step for both): per rank
        # [first / last owned DEM row | vertex block] as raw bytes
        es = torch.empty((), dtype=dem_dtype).element_size()
        self._row_bytes = 2 * int(width) * es
        self._row_bytes += (-self._row_bytes) % 16           # float64 pairs behind it
        blk = self._row_bytes + self.cap * 16
        self._send = torch.zeros(blk, dtype=torch.uint8, device=device)
        self._recv = torch.zeros((self.world, blk), dtype=torch.uint8, device=device)
        self._send_rows = self._send[:2 * int(width) * es].view(dem_dtype).view(2, width)
        self._send_verts = self._send[self._row_bytes:].view(torch.float64).view(self.cap, 2)
        self._recv_rows = self._recv[:, :2 * int(width) * es].view(dem_dtype)   # (world, 2 W)
        self._recv_verts = self._recv[:, self._row_bytes:].view(torch.float64)  # (world, 2 cap)
        self._width = int(width)
        # peer-memory form of exchange() (ocsb_band_push / ocsb_band_pull): set up
        # when every rank can map every other rank's buffer, else the collective
        self.mode = "nccl"
        self._p2p = None
        if (self.world > 1 and dem_dtype == torch.float32 and torch.device(device).type == "cuda"
                and os.environ.get("OCSB_BAND_P2P", "1") != "0"):
            self._init_p2p()

    def _init_p2p(self) -> None:
        """Allocate this rank's exchange buffer, swap CUDA IPC handles, map the
        peers' buffers.  All ranks agree on the outcome (one all-gather of a
        success flag): anything short of everybody succeeding keeps the
        collective path."""
        import torch
        import torch.distributed as dist
        from . import _lib
        dev = torch.device(self.device)
        ctx = _lib.default_context(dev.index)
        own = None
        opened = {}
        ok = True
        try:
            nbytes = _lib.load().ocsb_band_p2p_bytes(self.world, self.cap, self._width)
            if nbytes <= 0:
                raise RuntimeError("ocsb_band_p2p_bytes refused the geometry")
            own, handle = ctx.p2p_alloc(nbytes)
        except Exception as e:      # pylint: disable=broad-except
            _logger.info("band exchange: no peer-memory buffer (%s)", e)
            ok, handle = False, b""
        handles = [None] * self.world
        dist.all_gather_object(handles, handle if ok else None, group=self.group)
        if ok and all(h is not None for h in handles):
            try:
                for r, h in enumerate(handles):
                    if r != self.rank:
                        opened[r] = ctx.p2p_open(h)
            except Exception as e:  # pylint: disable=broad-except
                _logger.info("band exchange: cannot map a peer buffer (%s)", e)
                ok = False
        else:
            ok = False
        flags = [None] * self.world
        dist.all_gather_object(flags, bool(ok), group=self.group)
        if not all(flags):
            for a in opened.values():
                ctx.p2p_close(a)
            if own is not None:
                dist.barrier(group=self.group)      # nobody maps it any more
                ctx.p2p_free(own)
            return
        addrs = [own if r == self.rank else opened[r] for r in range(self.world)]
        self._p2p = {
            "own": own, "opened": opened,
            "peers": torch.tensor(addrs, dtype=torch.int64, device=dev),
            "done": torch.zeros(1, dtype=torch.int64, device=dev),
            "status": torch.zeros(1, dtype=torch.int32, device=dev),
            "epoch": 0,
        }
        self.mode = "p2p"

    def close(self) -> None:
        """Unmap the peers' buffers and free this rank's (collective: every rank
        calls it)."""
        import torch
        import torch.distributed as dist
        if self._p2p is None:
            return
        from . import _lib
        ctx = _lib.default_context(torch.device(self.device).index)
        torch.cuda.synchronize(torch.device(self.device))
        p, self._p2p, self.mode = self._p2p, None, "nccl"
        for a in p["opened"].values():
            ctx.p2p_close(a)
        dist.barrier(group=self.group)
        ctx.p2p_free(p["own"])

    def p2p_failed(self) -> bool:
        """True when a pull gave up waiting for a peer (host sync)."""
        return self._p2p is not None and bool(self._p2p["status"].item())

    def gather_vertices(self, verts_local):
        """All ranks' vertices, (world * cap, 2), rank order (padded blocks)."""
        import torch.distributed as dist
        n = int(verts_local.shape[0])
        if n != self.sizes[self.rank]:
            raise ValueError("vertex count differs from the exchanged size")
        if n:
            self.block[:n].copy_(verts_local)
            if n < self.cap:
                self.block[n:].copy_(verts_local[n - 1:n].expand(self.cap - n, 2))
        else:
            self.block.fill_(self.FAR)
        if self.world == 1:
            return self.block
        dist.all_gather_into_tensor(self.verts, self.block, group=self.group)
        return self.verts

    def exchange(self, verts_local, buf, origin: int, raster=None):
        """Both exchanges of a step in ONE collective: returns the gathered
        (padded) vertex array and fills the halo slots of ``buf`` (raster rows
        ``[origin, origin + len(buf))``: band + halo slots).  ``raster``: the
        band raster ``buf`` belongs to (its halo rows are then known valid)."""
        import torch
        import torch.distributed as dist
        if raster is not None and hasattr(raster, "_halo_pending"):
            for k in raster._halo_pending:
                raster._halo_pending[k] = False
        n = int(verts_local.shape[0])
        if n != self.sizes[self.rank]:
            raise ValueError("vertex count differs from the exchanged size")
        if self.world == 1:
            # nothing to exchange: the local vertices are the whole set
            return verts_local if n else self.gather_vertices(verts_local)
        r0, r1 = self.part.band(self.rank)
        if buf.is_cuda and buf.dtype == torch.float32:
            # two launches of the library around the collective instead of eight
            # small tensor ops (a 0.4 ms step notices every one of them)
            from . import _lib
            ctx = _lib.default_context(buf.device.index)
            if self._p2p is not None:
                # stores into the peers' buffers + epoch flags: no collective
                p = self._p2p
                p["epoch"] += 1
                ctx.band_push(verts_local, n, self.cap, buf[r0 - origin],
                              buf[r1 - 1 - origin], self._width, p["peers"], self.world,
                              self.rank, p["epoch"], self.FAR, p["done"])
                ctx.band_pull(p["own"], self.world, self.rank, self.cap, self._width,
                              p["epoch"], self.verts,
                              buf[r0 - 1 - origin] if self.rank > 0 else None,
                              buf[r1 - origin] if self.rank < self.world - 1 else None,
                              p["status"])
                return self.verts
            ctx.band_pack(verts_local, n, self.cap, buf[r0 - origin], buf[r1 - 1 - origin],
                          self._width, self._send, self.FAR)
            dist.all_gather_into_tensor(self._recv.view(-1), self._send, group=self.group)
            ctx.band_unpack(self._recv, self.world, self.rank, self.cap, self._width,
                            self.verts,
                            buf[r0 - 1 - origin] if self.rank > 0 else None,
                            buf[r1 - origin] if self.rank < self.world - 1 else None)
            return self.verts
        sv = self._send_verts
        if n:
            sv[:n].copy_(verts_local)
            if n < self.cap:
                sv[n:].copy_(verts_local[n - 1:n].expand(self.cap - n, 2))
        else:
            sv.fill_(self.FAR)
        self._send_rows[0].copy_(buf[r0 - origin])
        self._send_rows[1].copy_(buf[r1 - 1 - origin])
        dist.all_gather_into_tensor(self._recv.view(-1), self._send, group=self.group)
        self.verts.view(self.world, 2 * self.cap).copy_(self._recv_verts)
        w = self._width
        if self.rank > 0:                       # last row of the rank above
            buf[r0 - 1 - origin].copy_(self._recv_rows[self.rank - 1, w:])
        if self.rank < self.world - 1:          # first row of the rank below
            buf[r1 - origin].copy_(self._recv_rows[self.rank + 1, :w])
        return self.verts

    def real_vertices(self, gathered):
        """The gathered array without the filler entries (host-side checks)."""
        import torch
        if self.world == 1:
            return gathered[:self.sizes[0]]
        return torch.cat([gathered[r * self.cap:r * self.cap + self.sizes[r]]
                          for r in range(self.world)], dim=0)

    def exchange_halo(self, buf, origin: int) -> None:
        """``buf`` = raster rows ``[origin, origin + len(buf))`` of this rank
        (band + halo slots): fill the halo slots from the neighbours."""
        import torch.distributed as dist
        if self.world == 1:
            return
        r0, r1 = self.part.band(self.rank)
        self.rows_out[0].copy_(buf[r0 - origin])
        self.rows_out[1].copy_(buf[r1 - 1 - origin])
        dist.all_gather_into_tensor(self.rows_all, self.rows_out, group=self.group)
        if self.rank > 0:                       # last row of the rank above
            buf[r0 - 1 - origin].copy_(self.rows_all[2 * (self.rank - 1) + 1])
        if self.rank < self.world - 1:          # first row of the rank below
            buf[r1 - origin].copy_(self.rows_all[2 * (self.rank + 1)])


def min_combine_strip(strip, group=None):
    """Pixel-wise min of a co-registered overlap strip over all ranks of the
    group, in place (multi-raster combine)."""
    import torch.distributed as dist
    dist.all_reduce(strip, op=dist.ReduceOp.MIN, group=group)
    return strip


def raster_owner(index: int, world: int) -> int:
    """Multi-raster runs (BASELINE.json configs[2]): raster ``index`` lives on
    rank ``index % world``."""
    return int(index) % int(world)


def min_combine_overlaps(geoms, owners: Sequence[int], rank: int, get_window,
                         combine, group=None) -> int:
    """Multi-rank form of ``HfunCollector.min_combine_overlaps``: ``geoms[i]`` =
    ``HfunRaster.geometry()`` of raster ``i`` (known to every rank), raster ``i``
    lives on rank ``owners[i]``.  Three phases: (1) every overlap window of an
    owned raster is snapshotted, (2) the strips of pairs split over two ranks
    are swapped in ONE batched isend/irecv, in the same pair order on every
    rank, (3) every owned window takes the min with its partner's snapshot.
    Because the partners' PRE-combine values are used, a pixel covered by any
    number of rasters ends up with the min over all of them.

    ``get_window(i, window)`` -> contiguous tensor of an owned raster's window,
    ``combine(i, window, tensor)`` -> raster ``i``'s window = min(window, tensor).
    Returns the number of overlapping pairs."""
    import torch
    import torch.distributed as dist
    from .hfun import overlap_windows
    n = 0
    todo = []        # (raster, window, partner's strip)
    p2p, pending = [], []
    for i in range(len(geoms)):
        for j in range(i + 1, len(geoms)):
            ov = overlap_windows(geoms[i], geoms[j])
            if ov is None:
                continue
            n += 1
            wi, wj = ov
            oi, oj = owners[i], owners[j]
            if rank != oi and rank != oj:
                continue
            if oi == oj:
                si, sj = get_window(i, wi), get_window(j, wj)
                todo.append((i, wi, sj))
                todo.append((j, wj, si))
                continue
            mine, win, peer = (i, wi, oj) if rank == oi else (j, wj, oi)
            send = get_window(mine, win)
            recv = torch.empty_like(send)
            p2p.append(dist.P2POp(dist.isend, send, peer, group))
            p2p.append(dist.P2POp(dist.irecv, recv, peer, group))
            pending.append((mine, win, recv))
    if p2p:
        for req in dist.batch_isend_irecv(p2p):
            req.wait()
    for i, w, t in todo + pending:
        combine(i, w, t)
    return n


def all_gather_node_values(values_local, n_nodes: int, group=None):
    """Concatenate the ranks' contiguous node-value shards (``node_shard``)
    into the full ``(n_nodes,)`` array on every rank."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    if world == 1:
        return values_local
    sizes = [node_shard(n_nodes, world, r) for r in range(world)]
    m = max(b - a for a, b in sizes)
    pad = torch.zeros(m, dtype=values_local.dtype, device=values_local.device)
    pad[:values_local.shape[0]] = values_local
    out = torch.empty(world * m, dtype=values_local.dtype,
                      device=values_local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * m:r * m + (b - a)]
                      for r, (a, b) in enumerate(sizes)]).contiguous()


# ---------------------------------------------------------------------------
# host placement: the streamed pass (ocsb_hfun_pass) moves 8 B/px over PCIe,
# so on a multi-socket box each rank's pinned buffers must live on the NUMA
# node its GPU hangs off
# ---------------------------------------------------------------------------
def parse_cpulist(text: str) -> List[int]:
    """``"0-3,8,10-11"`` (sysfs cpulist format) -> ``[0, 1, 2, 3, 8, 10, 11]``."""
    cpus: List[int] = []
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-", 1)
            cpus.extend(range(int(a), int(b) + 1))
        else:
            cpus.append(int(part))
    return cpus


def gpu_pci_address(device: int) -> Optional[str]:
    """``domain:bus:device.0`` of a CUDA device as sysfs spells it."""
    import torch
    try:
        p = torch.cuda.get_device_properties(device)
        return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except (AttributeError, RuntimeError):
        return None


def bind_to_gpu_numa(device: int, sysfs: str = "/sys/bus/pci/devices") -> dict:
    """Pin the calling thread (and the threads it creates later) to the CPUs
    local to ``device``, so that memory it allocates and pins afterwards --
    the DEM / hfun staging buffers of the streamed pass -- is node-local to the
    GPU's PCIe root.  No-op (and says so) when sysfs has no locality to offer.
    Call it before the first pinned allocation."""
    import os
    info = {"device": int(device), "bound": False}
    addr = gpu_pci_address(device)
    info["pci"] = addr
    if addr is None or not hasattr(os, "sched_setaffinity"):
        return info
    try:
        with open(os.path.join(sysfs, addr, "local_cpulist")) as f:
            local = set(parse_cpulist(f.read()))
        try:
            with open(os.path.join(sysfs, addr, "numa_node")) as f:
                info["numa_node"] = int(f.read().strip())
        except (OSError, ValueError):
            pass
    except (OSError, ValueError):
        return info
    allowed = os.sched_getaffinity(0)
    cpus = sorted(local & allowed)
    info["local_cpus"] = len(local)
    if not cpus or len(cpus) == len(allowed):
        return info                       # no locality information / already local
    os.sched_setaffinity(0, cpus)
    info["bound"] = True
    info["cpus"] = len(cpus)
    return info
