"""ocsb_nearest_vertex_static: the allocation-free variant of the nearest-vertex
query (no counterpart in the reference; include/ocsb200.h).  Collected last on
purpose.

The scratch of the query lives in one fixed device array that every call
shares and that keeps the previous call's contents, so the checks run several
calls in a row with vertex sets of different size and extent, each against
ocsb_nearest_vertex on the same inputs -- bit-exact, the tolerance of every
index / distance test in this suite -- and each also against brute force.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_VERTS = 262144      # OCSB_STATIC_MAX_VERTS


def _case(torch, seed, nv, nq, scale, shift):
    g = torch.Generator(device="cpu").manual_seed(seed)
    v = (torch.rand((nv, 2), generator=g, dtype=torch.float64) * scale + shift).cuda()
    q = (torch.rand((nq, 2), generator=g, dtype=torch.float64) * (scale * 1.2)
         + (shift - 0.1 * scale)).cuda()
    return v, q


def _both(torch, ctx, v, q, r):
    nq = q.shape[0]
    out = []
    for fn in (ctx.nearest_vertex, ctx.nearest_vertex_static):
        d = torch.full((nq,), -7.0, dtype=torch.float64, device="cuda")
        i = torch.full((nq,), -7, dtype=torch.int64, device="cuda")
        fn(v, v.shape[0], q, nq, r, d, i)
        torch.cuda.synchronize()
        out.append((d, i))
    return out


def test_static_matches_allocating_entry_point_on_calls_in_a_row(ctx):
    import torch
    # (seed, nv, nq, scale, shift, r): a large set first, then smaller ones whose
    # grid covers only a part of what the previous call left behind
    cases = [
        (1, 200000, 300000, 50000.0, 1.0e5, None),
        (2, 777, 50000, 10.0, -3.0, None),
        (3, 40000, 100000, 2000.0, 4.0e6, 35.0),
        (1, 200000, 300000, 50000.0, 1.0e5, None),
        (4, 1, 1000, 1.0, 0.0, None),
        (5, MAX_VERTS, 20000, 1.0e4, 0.0, 120.0),
    ]
    for seed, nv, nq, scale, shift, r in cases:
        v, q = _case(torch, seed, nv, nq, scale, shift)
        (d0, i0), (d1, i1) = _both(torch, ctx, v, q, r)
        assert torch.equal(i0, i1), (seed, nv)
        assert torch.equal(d0.view(torch.int64), d1.view(torch.int64)), (seed, nv)
        # and against brute force on a sample (same FMA-free arithmetic in f64)
        k = min(nq, 512)
        dd = ((q[:k, None, :] - v[None, :, :]) ** 2)
        d2 = dd[..., 0] + dd[..., 1]
        best, arg = d2.min(dim=1)
        hit = torch.ones_like(best, dtype=torch.bool) if r is None else best < r * r
        assert torch.equal(i1[:k][hit], arg[hit])
        assert torch.equal(d1[:k][hit], best[hit].sqrt())
        assert bool((i1[:k][~hit] == nv).all()) and bool(torch.isinf(d1[:k][~hit]).all())


def test_static_allocates_nothing(ctx):
    import torch
    v, q = _case(torch, 9, 30000, 60000, 500.0, 10.0)
    d = torch.empty((60000,), dtype=torch.float64, device="cuda")
    i = torch.empty((60000,), dtype=torch.int64, device="cuda")
    ctx.nearest_vertex_static(v, 30000, q, 60000, None, d, i)      # module load, first use
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    tm0 = torch.cuda.memory_allocated()
    for _ in range(3):
        ctx.nearest_vertex_static(v, 30000, q, 60000, None, d, i)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0, "device memory was taken during the call"
    assert torch.cuda.memory_allocated() == tm0


def test_static_rejects_what_it_cannot_hold(ctx, lib):
    import torch
    from ocsmesh_b200 import _lib
    nv = MAX_VERTS + 1
    v = torch.zeros((nv, 2), dtype=torch.float64, device="cuda")
    q = torch.zeros((4, 2), dtype=torch.float64, device="cuda")
    d = torch.empty((4,), dtype=torch.float64, device="cuda")
    i = torch.empty((4,), dtype=torch.int64, device="cuda")
    with pytest.raises(_lib.OcsbError, match="exceed the fixed scratch array"):
        ctx.nearest_vertex_static(v, nv, q, 4, None, d, i)
    # host buffers would need staging, which allocates
    hv = np.zeros((8, 2))
    with pytest.raises(_lib.OcsbError, match="device memory"):
        ctx.nearest_vertex_static(hv, 8, q, 4, None, d, i)
    with pytest.raises(_lib.OcsbError, match="empty vertex set"):
        ctx.nearest_vertex_static(v, 0, q, 4, None, d, i)
    # the symbol is there under its C name
    assert isinstance(lib.ocsb_nearest_vertex_static, C._CFuncPtr)
