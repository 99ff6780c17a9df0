"""Row-band sharding: a raster split into per-rank bands with
band-local storage (+1 DEM halo row) gives bit-identical results to the
un-sharded object.  The ranks are emulated one after the other on one GPU (the
collectives themselves are covered by tests/test_distributed_cpu.py over gloo,
and by the 2-GPU torchrun test below when two devices are present)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import ocsmesh_b200 as ob
from ocsmesh_b200 import _lib, distributed as D, features as F, synthetic as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 3, 5])
def test_banded_hfun_equals_full(ctx, world):
    import torch
    H, W = 301, 417
    z = S.synth_dem(H, W, seed=8)
    tr = S.dem_transform(H)
    full_r = ob.Raster.from_numpy(z, tr, crs=32618)
    lines = S.synth_shoreline(H, W, seed=8, n_lines=2)
    pts = F.feature_points(lines, 100.0)
    full = ob.HfunRaster(full_r, hmin=100, hmax=1500)
    full.add_feature_points(pts, 0.001, 100.0)
    full.add_subtidal_flow_limiter(50, 1500, -30, 0)
    full.add_gradient_delimiter(50, 1500)
    ref = full.get_values()

    part = D.RowBands(H, world)
    rasters = [D.band_raster(z[slice(*part.band(r))], tr, H, part, r, crs=32618,
                             device=0) for r in range(world)]
    # what exchange_dem_halo does over NCCL: neighbours' boundary rows
    for r in range(world):
        buf, org = rasters[r].device_band(1, 0), rasters[r].device_band_origin(1, 0)
        r0, r1 = part.band(r)
        if r > 0:
            buf[r0 - 1 - org].copy_(torch.from_numpy(z[r0 - 1]))
        if r < world - 1:
            buf[r1 - org].copy_(torch.from_numpy(z[r1]))
    got = np.empty_like(ref)
    for r in range(world):
        h = ob.HfunRaster(rasters[r], hmin=100, hmax=1500, row_band=part.band(r))
        assert h.device_values.shape == (part.band(r)[1] - part.band(r)[0], W)
        h.add_feature_points(pts, 0.001, 100.0)
        h.add_subtidal_flow_limiter(50, 1500, -30, 0)
        h.add_gradient_delimiter(50, 1500)
        got[slice(*part.band(r))] = h.get_values()
    np.testing.assert_array_equal(got, ref)


def test_sharded_nodes_equal_full(ctx):
    rasters = []
    for k in range(2):
        z = S.synth_dem(90, 120, seed=40 + k)
        tr = list(S.dem_transform(90))
        tr[2] += 700.0 * k
        rasters.append(ob.Raster.from_numpy(z, tr, crs=32618))
    x0, y0, _, y1 = rasters[0].get_bbox()
    xy = S.synth_nodes(250000, (x0, y0, rasters[1].get_bbox()[2], y1), seed=1,
                       outside_frac=0.03)
    for method in ("linear", "nearest"):
        full = ob.Mesh(ob.MeshData(xy, values=np.full(len(xy), np.nan)))
        full.interpolate(rasters, method=method)       # bucketed path (>= 200k)
        shard = ob.Mesh(ob.MeshData(xy, values=np.full(len(xy), np.nan)))
        for r in range(3):
            shard.interpolate(rasters, method=method,
                              node_range=D.node_shard(len(xy), 3, r))
        np.testing.assert_array_equal(full.value, shard.value)


def test_two_gpu_bench_smoke():
    """torchrun x2 over NCCL (only where two devices are visible)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
           "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "bench.py"), "--gpus", "2",
           "--steps", "2", "--warmup", "3", "--size", "2048", "--nodes", "200000",
           "--tile", "512", "--no-cpu"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    assert '"n_gpus": 2' in out.stdout


def test_band_pack_unpack_blocks(ctx):
    """ocsb_band_pack / ocsb_band_unpack: the blocks of the one-collective
    exchange, emulated for 3 ranks on one device (the collective is a torch.cat)."""
    import torch
    W, cap, world = 37, 9, 3
    rng = np.random.default_rng(3)
    rows = [torch.from_numpy(rng.random((2, W)).astype(np.float32)).cuda() for _ in range(world)]
    verts = [torch.from_numpy(rng.random((n, 2))).cuda() for n in (5, 0, 9)]
    row_bytes = (2 * W * 4 + 15) // 16 * 16
    blk = row_bytes + cap * 16
    blocks = []
    for r in range(world):
        send = torch.zeros(blk, dtype=torch.uint8, device="cuda")
        ctx.band_pack(verts[r], len(verts[r]), cap, rows[r][0], rows[r][1], W, send, 1e13)
        blocks.append(send)
    recv = torch.stack(blocks)
    for rank in range(world):
        out = torch.empty((world * cap, 2), dtype=torch.float64, device="cuda")
        above = torch.full((W,), -1.0, dtype=torch.float32, device="cuda")
        below = torch.full((W,), -1.0, dtype=torch.float32, device="cuda")
        ctx.band_unpack(recv, world, rank, cap, W, out,
                        above if rank > 0 else None, below if rank < world - 1 else None)
        torch.cuda.synchronize()
        for r in range(world):
            n = len(verts[r])
            got = out[r * cap:(r + 1) * cap].cpu().numpy()
            if n:
                np.testing.assert_array_equal(got[:n], verts[r].cpu().numpy())
                assert (got[n:] == verts[r][-1].cpu().numpy()).all()
            else:
                assert (got == 1e13).all()
        if rank > 0:
            assert torch.equal(above, rows[rank - 1][1])
        else:
            assert bool((above == -1.0).all())
        if rank < world - 1:
            assert torch.equal(below, rows[rank + 1][0])


def test_band_push_pull_peer_memory_emulated(ctx):
    """ocsb_band_push / ocsb_band_pull (the exchange as stores into the peers'
    buffers + epoch flags), 3 ranks emulated on one device: the 'peer' buffers
    are three local ocsb_p2p_alloc buffers.  Two steps check the two slot
    generations; a pull before a peer's push gives up (bounded wait) and says so."""
    import torch
    W, cap, world = 41, 7, 3
    lib = _lib.load()
    nbytes = lib.ocsb_band_p2p_bytes(world, cap, W)
    assert nbytes == 2 * world * ((2 * W * 4 + 15) // 16 * 16 + cap * 16) + 256
    bufs = [ctx.p2p_alloc(nbytes) for _ in range(world)]
    assert all(len(h) == 64 for _, h in bufs)
    peers = torch.tensor([a for a, _ in bufs], dtype=torch.int64, device="cuda")
    done = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    rng = np.random.default_rng(5)
    try:
        for epoch in (1, 2, 3):
            rows = [torch.from_numpy(rng.random((2, W)).astype(np.float32)).cuda()
                    for _ in range(world)]
            verts = [torch.from_numpy(rng.random((n, 2))).cuda() for n in (4, 0, 7)]
            for r in range(world):
                ctx.band_push(verts[r], len(verts[r]), cap, rows[r][0], rows[r][1], W, peers,
                              world, r, epoch, 1e13, done[r])
            for rank in range(world):
                out = torch.empty((world * cap, 2), dtype=torch.float64, device="cuda")
                above = torch.full((W,), -1.0, dtype=torch.float32, device="cuda")
                below = torch.full((W,), -1.0, dtype=torch.float32, device="cuda")
                ctx.band_pull(bufs[rank][0], world, rank, cap, W, epoch, out,
                              above if rank > 0 else None,
                              below if rank < world - 1 else None, status)
                torch.cuda.synchronize()
                for r in range(world):
                    n = len(verts[r])
                    got = out[r * cap:(r + 1) * cap].cpu().numpy()
                    if n:
                        np.testing.assert_array_equal(got[:n], verts[r].cpu().numpy())
                        assert (got[n:] == verts[r][-1].cpu().numpy()).all()
                    else:
                        assert (got == 1e13).all()
                if rank > 0:
                    assert torch.equal(above, rows[rank - 1][1])
                if rank < world - 1:
                    assert torch.equal(below, rows[rank + 1][0])
            assert int(status.item()) == 0
        # epoch 4: rank 1 never pushes -> rank 0's pull gives up after ~2 s
        ctx.band_push(verts[0], len(verts[0]), cap, rows[0][0], rows[0][1], W, peers, world, 0,
                      4, 1e13, done[0])
        out = torch.empty((world * cap, 2), dtype=torch.float64, device="cuda")
        ctx.band_pull(bufs[0][0], world, 0, cap, W, 4, out, None,
                      torch.empty(W, dtype=torch.float32, device="cuda"), status)
        torch.cuda.synchronize()
        assert int(status.item()) == 1
        with pytest.raises(_lib.OcsbError):
            ctx.band_pull(bufs[0][0], world, 0, cap, W, 4, out,
                          torch.empty(W, dtype=torch.float32, device="cuda"), None, status)
    finally:
        torch.cuda.synchronize()
        for a, _ in bufs:
            ctx.p2p_free(a)
