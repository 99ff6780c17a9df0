"""N > 1 host logic on CPU: world_size-2 (and 3) gloo process groups exercise
the same collectives the GPU ranks run over NCCL (vertex all-gather, DEM halo
rows, min-combine, node shards)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ocsmesh_b200 import distributed as D, synthetic as S


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, fn, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def run_ranks(fn, world=2):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), fn, ret), nprocs=world, join=True)
    return [ret[r] for r in range(world)]


def test_row_bands_and_node_shards():
    p = D.RowBands(20000, 8)
    assert [p.band(r) for r in range(8)][:2] == [(0, 2500), (2500, 5000)]
    assert p.stored(0) == (0, 2501) and p.stored(3) == (7499, 10001)
    assert p.stored(7) == (17499, 20000)
    p = D.RowBands(10, 3)
    assert [p.band(r) for r in range(3)] == [(0, 4), (4, 7), (7, 10)]
    assert [p.owner(r) for r in (0, 3, 4, 9)] == [0, 0, 1, 2]
    with pytest.raises(ValueError):
        D.RowBands(2, 3)
    shards = [D.node_shard(10, 4, r) for r in range(4)]
    assert shards == [(0, 3), (3, 6), (6, 8), (8, 10)]


def _gather_case(rank, world):
    rng = np.random.default_rng(100 + rank)
    n = [5, 0, 9][rank % 3] + rank          # ragged, one rank may be empty
    v = torch.from_numpy(rng.random((n, 2)))
    out = D.all_gather_vertices(v)
    return v.numpy(), out.numpy()


@pytest.mark.parametrize("world", [2, 3])
def test_all_gather_vertices_ragged(world):
    res = run_ranks(_gather_case, world)
    expect = np.concatenate([r[0] for r in res])
    for _, got in res:
        np.testing.assert_array_equal(got, expect)


def _halo_case(rank, world):
    H, W = 41, 37
    z = S.synth_dem(H, W, seed=3)
    part = D.RowBands(H, world)
    r0, r1 = part.band(rank)
    s0, s1 = part.stored(rank)
    buf = torch.zeros((s1 - s0, W), dtype=torch.float32)
    buf[r0 - s0:r1 - s0] = torch.from_numpy(z[r0:r1])
    D.exchange_halo_rows(buf, s0, part, rank)
    # after the exchange the stored rows equal the raster rows, so the band's
    # np.gradient (with halos) equals the full-window gradient on owned rows
    ok = np.array_equal(buf.numpy(), z[s0:s1])
    g_full = np.gradient(z, 10.0, 10.0)[0][r0:r1]
    g_band = np.gradient(buf.numpy(), 10.0, 10.0)[0][r0 - s0:r1 - s0]
    interior = slice(1 if rank == 0 else 0, (r1 - r0) - (1 if rank == world - 1 else 0))
    return ok and np.array_equal(g_full[interior], g_band[interior])


@pytest.mark.parametrize("world", [2, 3])
def test_dem_halo_exchange(world):
    assert all(run_ranks(_halo_case, world))


def _min_case(rank, world):
    rng = np.random.default_rng(rank)
    strip = torch.from_numpy(rng.random((6, 9)).astype(np.float32))
    mine = strip.clone()
    D.min_combine_strip(strip)
    return mine.numpy(), strip.numpy()


def test_min_combine_strip():
    res = run_ranks(_min_case, 2)
    expect = np.minimum(res[0][0], res[1][0])
    for _, got in res:
        np.testing.assert_array_equal(got, expect)


def _nodes_case(rank, world):
    n = 11
    k0, k1 = D.node_shard(n, world, rank)
    vals = torch.arange(k0, k1, dtype=torch.float64) * 1.5
    return D.all_gather_node_values(vals, n).numpy()


def test_all_gather_node_values():
    for got in run_ranks(_nodes_case, 3):
        np.testing.assert_array_equal(got, np.arange(11) * 1.5)


def test_cpulist_parsing_and_numa_binding_without_gpu(tmp_path, monkeypatch):
    """Host placement helper: sysfs cpulist format, and binding is a no-op
    (reported as such) when the device has no sysfs entry."""
    from ocsmesh_b200 import distributed as D
    assert D.parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert D.parse_cpulist("") == []
    import os
    before = os.sched_getaffinity(0)
    monkeypatch.setattr(D, "gpu_pci_address", lambda d: "0000:1b:00.0")
    info = D.bind_to_gpu_numa(0, sysfs=str(tmp_path))       # no such device dir
    assert info["bound"] is False and os.sched_getaffinity(0) == before
    # a device whose local CPUs are a strict subset of ours: bound to them
    if len(before) > 1:
        d = tmp_path / "0000:1b:00.0"
        d.mkdir()
        keep = sorted(before)[:1]
        (d / "local_cpulist").write_text(",".join(map(str, keep)) + "\n")
        (d / "numa_node").write_text("1\n")
        try:
            info = D.bind_to_gpu_numa(0, sysfs=str(tmp_path))
            assert info["bound"] and info["numa_node"] == 1
            assert os.sched_getaffinity(0) == set(keep)
        finally:
            os.sched_setaffinity(0, before)


# ---- configs[2]: rasters on rank r % world, pairwise overlap strips -----------
def _c3_layout(tile_h=40, tile_w=50, ovl=10, nx=4, ny=2):
    geoms = []
    for k in range(nx * ny):
        ty, tx = divmod(k, nx)
        tr = list(S.dem_transform(tile_h))
        tr[2] += (tile_w - ovl) * S.PIXEL * tx
        tr[5] -= (tile_h - ovl) * S.PIXEL * ty      # next row of tiles lies south
        geoms.append((tuple(tr), tile_w, tile_h))
    return geoms


def _c3_values(k, h, w):
    return np.random.default_rng(100 + k).random((h, w)).astype(np.float32)


def _c3_expected(geoms):
    """Global mosaic min, then each raster's view of it."""
    base = geoms[0]
    # offsets of every raster in the pixel grid of raster 0
    offs = []
    for (tr, w, h) in geoms:
        oc = int(round((tr[2] - base[0][2]) / base[0][0]))
        orow = int(round((tr[5] - base[0][5]) / base[0][4]))
        offs.append((orow, oc))
    H = max(o[0] + g[2] for o, g in zip(offs, geoms))
    W = max(o[1] + g[1] for o, g in zip(offs, geoms))
    G = np.full((H, W), np.inf, dtype=np.float32)
    for k, ((r, c), (_, w, h)) in enumerate(zip(offs, geoms)):
        G[r:r + h, c:c + w] = np.minimum(G[r:r + h, c:c + w], _c3_values(k, h, w))
    return [G[r:r + h, c:c + w].copy() for (r, c), (_, w, h) in zip(offs, geoms)]


def _c3_case(rank, world):
    geoms = _c3_layout()
    owners = [D.raster_owner(k, world) for k in range(len(geoms))]
    owned = {k: torch.from_numpy(_c3_values(k, g[2], g[1]))
             for k, g in enumerate(geoms) if owners[k] == rank}

    def get_window(i, w):
        return owned[i][w.row_off:w.row_off + w.height,
                        w.col_off:w.col_off + w.width].contiguous()

    def combine(i, w, t):
        v = owned[i][w.row_off:w.row_off + w.height, w.col_off:w.col_off + w.width]
        torch.minimum(v, t, out=v)

    n = D.min_combine_overlaps(geoms, owners, rank, get_window, combine)
    return n, {k: v.numpy() for k, v in owned.items()}


@pytest.mark.parametrize("world", [2, 3])
def test_min_combine_overlaps_pairwise_exchange(world):
    """8 rasters 4x2 with overlaps (corners covered by 4 rasters) spread over
    the ranks: after the pairwise strip swaps every raster equals its view of
    the global mosaic min -- whatever the rank count."""
    geoms = _c3_layout()
    want = _c3_expected(geoms)
    res = run_ranks(_c3_case, world)
    seen = set()
    for n, owned in res:
        assert n == 4 * 1 + 3 * 2 + 2 * 3      # E-W, N-S and diagonal neighbours
        for k, v in owned.items():
            np.testing.assert_array_equal(v, want[k])
            seen.add(k)
    assert seen == set(range(len(geoms)))


def test_overlap_windows_geometry():
    from ocsmesh_b200.hfun import overlap_windows
    from ocsmesh_b200 import Window
    g = _c3_layout()
    wa, wb = overlap_windows(g[0], g[1])          # east neighbour
    assert wa == Window(40, 0, 10, 40) and wb == Window(0, 0, 10, 40)
    wa, wb = overlap_windows(g[0], g[4])          # south neighbour
    assert wa == Window(0, 30, 50, 10) and wb == Window(0, 0, 50, 10)
    wa, wb = overlap_windows(g[5], g[0])          # diagonal, reversed order
    assert wa == Window(0, 0, 10, 10) and wb == Window(40, 30, 10, 10)
    assert overlap_windows(g[0], g[2]) is None
    bad = (tuple(np.array(g[1][0]) + [0, 0, 3.0, 0, 0, 0]), 50, 40)
    with pytest.raises(ValueError, match="co-registered"):
        overlap_windows(g[0], bad)


def _ragged_case(rank, world):
    # rank r contributes r values (rank 0: none) and (2 r + 1, 2) vertices
    vals = torch.arange(rank, dtype=torch.float32) + 10.0 * rank
    verts = torch.full((2 * rank + 1, 2), float(rank), dtype=torch.float64)
    none = torch.empty((0, 2), dtype=torch.float64)
    return (D.all_gather_ragged(vals).numpy(), D.all_gather_ragged(verts).numpy(),
            D.all_gather_ragged(none).numpy())


def test_all_gather_ragged_any_rank_and_empty_contributions():
    """The exchange behind a rate constraint on row bands: 1-D values and (n, 2)
    vertices, ranks with nothing to contribute, nobody contributing."""
    world = 3
    want_vals = np.concatenate([np.arange(r, dtype=np.float32) + 10.0 * r
                                for r in range(world)])
    want_verts = np.concatenate([np.full((2 * r + 1, 2), float(r)) for r in range(world)])
    for vals, verts, none in run_ranks(_ragged_case, world):
        np.testing.assert_array_equal(vals, want_vals)
        np.testing.assert_array_equal(verts, want_verts)
        assert none.shape == (0, 2)


def _ragged_plan_case(rank, world):
    verts = torch.full((3 * rank + 2, 2), float(rank), dtype=torch.float64)
    sizes = D.gather_sizes(verts.shape[0], verts.device)
    a = D.all_gather_vertices(verts).numpy()
    b = D.all_gather_vertices(verts, sizes=sizes).numpy()
    bad = False
    try:
        D.all_gather_vertices(verts, sizes=[1] * world)
    except ValueError:
        bad = True
    return sizes, a, b, bad


def test_all_gather_with_pre_exchanged_sizes():
    for sizes, a, b, bad in run_ranks(_ragged_plan_case, 2):
        assert sizes == [2, 5] and bad
        np.testing.assert_array_equal(a, b)


def _band_exchange_case(rank, world):
    """BandExchange: padded-block vertex all-gather (the filler entries repeat a
    real vertex or are far away) and the boundary-row all-gather that fills the
    DEM halo slots, both into persistent buffers."""
    from scipy.spatial import cKDTree
    H, W = 41, 37
    z = S.synth_dem(H, W, seed=3)
    part = D.RowBands(H, world)
    r0, r1 = part.band(rank)
    s0, s1 = part.stored(rank)
    rng = np.random.default_rng(7 + rank)
    n = [5, 0, 9][rank % 3]               # ragged; rank 1 contributes nothing
    v = torch.from_numpy(rng.random((n, 2)) * 100.0)
    ex = D.BandExchange(n, W, torch.float32, "cpu", part, rank)
    # host tensors: the collective form (the peer-memory form needs CUDA IPC)
    assert ex.mode == "nccl" and not ex.p2p_failed()
    ex.close()                            # nothing to release: a no-op
    out = {}
    for rep in range(2):                  # persistent buffers: a second step too
        g = ex.gather_vertices(v)
        real = ex.real_vertices(g).numpy()
        buf = torch.zeros((s1 - s0, W), dtype=torch.float32)
        buf[r0 - s0:r1 - s0] = torch.from_numpy(z[r0:r1])
        ex.exchange_halo(buf, s0)
        # the packed form: both exchanges in one collective
        buf2 = torch.zeros((s1 - s0, W), dtype=torch.float32)
        buf2[r0 - s0:r1 - s0] = torch.from_numpy(z[r0:r1])
        g2 = ex.exchange(v, buf2, s0)
        same = bool(torch.equal(g2, g)) and bool(torch.equal(buf2, buf))
        out[rep] = (np.array_equal(buf.numpy(), z[s0:s1]) and same, real.copy(),
                    g.numpy().copy())
    # nearest-vertex distances over the padded array == over the real vertices
    q = rng.random((200, 2)) * 100.0
    d_pad = cKDTree(out[1][2]).query(q)[0]
    d_real = cKDTree(out[1][1]).query(q)[0]
    return (out[0][0] and out[1][0], v.numpy(), out[1][1],
            bool(np.array_equal(d_pad, d_real)), list(ex.sizes))


@pytest.mark.parametrize("world", [2, 3])
def test_band_exchange_persistent_buffers(world):
    res = run_ranks(_band_exchange_case, world)
    expect = np.concatenate([r[1] for r in res])
    for ok, _, real, same_d, sizes in res:
        assert ok
        np.testing.assert_array_equal(real, expect)
        assert same_d
        assert sizes == [len(r[1]) for r in res]


def test_band_raster_refuses_contours_before_the_halo_exchange(monkeypatch):
    """get_contour on a band raster whose halo slots are still zeros
    would trace the level along the band edges; the raster knows its halo state
    (set by band_raster, cleared by exchange_dem_halo / BandExchange.exchange)."""
    from ocsmesh_b200.raster import Raster
    r = Raster.from_numpy(S.synth_dem(8, 9, seed=1), S.dem_transform(20), crs=32618,
                          row_origin=4, full_height=20)
    r._halo_pending = {(1, 0): True}
    r._dev_cache[(1, 0)] = torch.zeros((10, 9))
    r._dev_origin = {(1, 0): 3}

    class Ctx:
        device = 0
    import ocsmesh_b200._lib as L
    monkeypatch.setattr(L, "default_context", lambda d=None: Ctx())
    with pytest.raises(RuntimeError, match="halo rows have not been exchanged"):
        r.get_contour(0.0)
