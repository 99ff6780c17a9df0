"""K1 parity: ocsb_hfun_feature vs scipy.spatial.cKDTree through the oracle
(reference ocsmesh/hfun/raster.py:1230-1257)."""
import numpy as np
import pytest

from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, features as F, synthetic as S

pytestmark = pytest.mark.gpu

FLT_MAX = np.finfo(np.float32).max


def make_case(h, w, seed, n_lines=2, target=100.0):
    tr = S.dem_transform(h)
    r = O.OracleRaster(np.zeros((h, w), np.float32), tr)
    lines = S.synth_shoreline(h, w, seed=seed, n_lines=n_lines)
    pts = F.feature_points(lines, target)
    return r, pts


def run_gpu(ctx, r, pts, old, rate, target, hmin, hmax, fuse=False,
            want=False, band=(0, 0), device=True):
    import torch
    h, w = old.shape
    win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h, *band)
    dist = idx = None
    if device:
        od = torch.from_numpy(old).cuda()
        nd = od.clone()
        vd = torch.from_numpy(np.ascontiguousarray(pts)).cuda()
        if want:
            dist = torch.full((h, w), -1.0, dtype=torch.float64, device="cuda")
            idx = torch.full((h, w), -1, dtype=torch.int64, device="cuda")
        ctx.hfun_feature(win, vd, len(pts), od, w, nd, w, rate, target, hmin,
                         hmax, fuse, dist, idx)
        ctx.sync()
        if want:
            return nd.cpu().numpy(), dist.cpu().numpy(), idx.cpu().numpy()
        return nd.cpu().numpy()
    new = old.copy()
    if want:
        dist = np.full((h, w), -1.0)
        idx = np.full((h, w), -1, np.int64)
    ctx.hfun_feature(win, np.ascontiguousarray(pts), len(pts), old, w, new, w,
                     rate, target, hmin, hmax, fuse, dist, idx)
    return (new, dist, idx) if want else new


def check_indices(idx, ref_idx, pts, xy, miss):
    """Bit-exact nearest-vertex indices; where several vertices attain the
    bit-identical minimal distance (duplicates are guaranteed by
    repartition_features) any of them is accepted."""
    idx = idx.ravel()
    ref_idx = ref_idx.ravel()
    bad = np.nonzero(idx != ref_idx)[0]
    if len(bad) == 0:
        return
    assert not np.any(miss.ravel()[bad])
    da = ((xy[bad] - pts[idx[bad]]) ** 2).sum(axis=1)
    db = ((xy[bad] - pts[ref_idx[bad]]) ** 2).sum(axis=1)
    np.testing.assert_array_equal(da, db)


@pytest.mark.parametrize("shape,seed", [((64, 64), 1), ((200, 333), 2),
                                        ((515, 257), 3), ((1000, 1000), 4)])
@pytest.mark.parametrize("hmax", [1500.0, None])
def test_distance_index_and_size(ctx, shape, seed, hmax):
    h, w = shape
    r, pts = make_case(h, w, seed)
    rng = np.random.default_rng(seed)
    old = np.full(shape, FLT_MAX, np.float32)
    old[rng.random(shape) < 0.2] = 321.0
    xy = r.get_xy()
    ref, ref_d, ref_i = O.add_feature_window(old, pts, xy, h, w, 0.001, 100.0,
                                             100.0, hmax, return_dist=True)
    out, d, i = run_gpu(ctx, r, pts, old, 0.001, 100.0, 100.0, hmax, want=True)
    # distances: the kernel evaluates sqrt(dx*dx + dy*dy) in f64 like scipy
    np.testing.assert_array_equal(d, ref_d)
    miss = ref_i == len(pts)
    np.testing.assert_array_equal(i == len(pts), miss)
    check_indices(i, ref_i, pts, xy, miss)
    # sizes: 1e-5 relative in fp32 is the stated tolerance; same op order ->
    # bit-exact
    np.testing.assert_array_equal(out, ref)


def test_bounded_equals_unbounded_after_clamp(ctx):
    r, pts = make_case(300, 400, 7)
    old = np.full((300, 400), FLT_MAX, np.float32)
    # r = (300 - 50) / (0.05 * 50) = 100 m: most pixels are beyond the radius
    a = run_gpu(ctx, r, pts, old, 0.05, 50.0, 50.0, 300.0)
    win = r.full_window()
    xy = r.get_xy(win)
    ref = O.add_feature_window(old, pts, xy, 300, 400, 0.05, 50.0, 50.0, 300.0)
    np.testing.assert_array_equal(a, ref)
    assert a.max() == 300.0 and a.min() >= 50.0


def test_duplicates_and_single_vertex(ctx):
    r, pts = make_case(128, 160, 5)
    pts = np.concatenate([pts, pts[::3], pts[:1]])      # duplicated vertices
    old = np.full((128, 160), FLT_MAX, np.float32)
    xy = r.get_xy()
    ref, ref_d, ref_i = O.add_feature_window(old, pts, xy, 128, 160, 0.001, 100.0,
                                             100.0, None, return_dist=True)
    out, d, i = run_gpu(ctx, r, pts, old, 0.001, 100.0, 100.0, None, want=True)
    np.testing.assert_array_equal(d, ref_d)
    check_indices(i, ref_i, pts, xy, ref_i == len(pts))
    np.testing.assert_array_equal(out, ref)
    one = pts[:1]
    ref = O.add_feature_window(old, one, xy, 128, 160, 0.001, 100.0, 100.0, 1500.0)
    out = run_gpu(ctx, r, one, old, 0.001, 100.0, 100.0, 1500.0)
    np.testing.assert_array_equal(out, ref)


def test_random_vertex_cloud(ctx):
    # not a polyline at all: uniformly random vertices (worst case for the
    # dominance pruning)
    h, w = 256, 384
    r, _ = make_case(h, w, 1)
    rng = np.random.default_rng(5)
    x0, y0, x1, y1 = r.get_window_bounds(r.full_window())
    pts = np.stack([rng.uniform(x0 - 500, x1 + 500, 3000),
                    rng.uniform(y0 - 500, y1 + 500, 3000)], axis=1)
    old = np.full((h, w), FLT_MAX, np.float32)
    xy = r.get_xy()
    for hmax in (None, 150.0):
        ref, ref_d, ref_i = O.add_feature_window(old, pts, xy, h, w, 0.01, 100.0,
                                                 10.0, hmax, return_dist=True)
        out, d, i = run_gpu(ctx, r, pts, old, 0.01, 100.0, 10.0, hmax, want=True)
        np.testing.assert_array_equal(d, ref_d)
        check_indices(i, ref_i, pts, xy, ref_i == len(pts))
        np.testing.assert_array_equal(out, ref)


def test_host_buffers_bands_and_fused_clamp(ctx):
    h, w = 300, 260
    r, pts = make_case(h, w, 8)
    old = np.full((h, w), FLT_MAX, np.float32)
    xy = r.get_xy()
    ref = O.add_feature_window(old, pts, xy, h, w, 0.001, 100.0, 100.0, 1500.0)
    out = run_gpu(ctx, r, pts, old, 0.001, 100.0, 100.0, 1500.0, device=False)
    np.testing.assert_array_equal(out, ref)
    acc = old.copy()
    for rb, re in [(0, 7), (7, 130), (130, 300)]:
        part = run_gpu(ctx, r, pts, old, 0.001, 100.0, 100.0, 1500.0, band=(rb, re))
        assert np.array_equal(part[:rb], old[:rb]) and np.array_equal(part[re:], old[re:])
        acc[rb:re] = part[rb:re]
    np.testing.assert_array_equal(acc, ref)
    # hmax None + global clamp folded in
    ref2 = O.add_feature_window(old, pts, xy, h, w, 0.001, 100.0, None, None)
    ref2 = O.apply_constraints_clamp(ref2, None, None)
    out2 = run_gpu(ctx, r, pts, old, 0.001, 100.0, None, None, fuse=True)
    np.testing.assert_array_equal(out2, ref2)


def test_errors(ctx):
    r, pts = make_case(32, 32, 1)
    old = np.full((32, 32), FLT_MAX, np.float32)
    with pytest.raises(_lib.OcsbError):
        run_gpu(ctx, r, pts, old, 0.001, -1.0, None, None, device=False)
    with pytest.raises(_lib.OcsbError):
        run_gpu(ctx, r, pts[:0], old, 0.001, 100.0, None, None, device=False)


@pytest.mark.parametrize("rate,target,hmin,hmax", [
    (0.001, 100.0, 100.0, 1500.0), (0.0137, 73.1, None, None),
    (0.3, 5.0, 1.0, 4000.0)])
def test_fast_size_law_is_bit_exact(ctx, rate, target, hmin, hmax):
    """Without dist/idx outputs the epilogue uses the checked fast sqrt
    (k_feature_eval.cuh); its float32 sizes must equal the f64 reference
    path bit for bit on a few million pixels."""
    h, w = 1500, 1400
    r, pts = make_case(h, w, 12, n_lines=3, target=target)
    old = np.full((h, w), FLT_MAX, np.float32)
    ref = O.add_feature_window(old, pts, r.get_xy(), h, w, rate, target, hmin, hmax)
    out = run_gpu(ctx, r, pts, old, rate, target, hmin, hmax)
    np.testing.assert_array_equal(out, ref)
    # and the exact (dist/idx) variant agrees with the fast one
    out2, _, _ = run_gpu(ctx, r, pts, old, rate, target, hmin, hmax, want=True)
    np.testing.assert_array_equal(out2, out)


@pytest.mark.parametrize("fuse", [False, True])
@pytest.mark.parametrize("hmax", [1500.0, None])
def test_null_old_is_initial_state(ctx, fuse, hmax):
    """hfun_old == NULL means "the raster is still finfo(float32).max
    everywhere" (A4): same bits as passing that buffer, device and host."""
    import torch
    h, w = 515, 389
    r, pts = make_case(h, w, 21, n_lines=3)
    old = np.full((h, w), FLT_MAX, np.float32)
    ref = run_gpu(ctx, r, pts, old, 0.001, 100.0, 100.0, hmax, fuse=fuse)
    win = _lib.make_win(r.get_window_bounds(r.full_window()), w, h, 0, 0)
    nd = torch.full((h, w), -7.0, dtype=torch.float32, device="cuda")
    vd = torch.from_numpy(np.ascontiguousarray(pts)).cuda()
    ctx.hfun_feature(win, vd, len(pts), None, w, nd, w, 0.001, 100.0, 100.0,
                     hmax, fuse)
    ctx.sync()
    np.testing.assert_array_equal(nd.cpu().numpy(), ref)
    new = np.full((h, w), -7.0, np.float32)
    ctx.hfun_feature(win, np.ascontiguousarray(pts), len(pts), None, w, new, w,
                     0.001, 100.0, 100.0, hmax, fuse)
    np.testing.assert_array_equal(new, ref)


def test_fast_size_law_edge_distances(ctx):
    """Pixels that coincide with a vertex (d = 0), sit beyond the radius, or
    make the size law land next to a float32 rounding tie all leave the fast
    path; a dense vertex lattice makes the first two common."""
    h, w = 257, 300
    r, _ = make_case(h, w, 1)
    xy = r.get_xy()
    pts = np.ascontiguousarray(xy.reshape(h, w, 2)[::16, ::16].reshape(-1, 2))
    old = np.full((h, w), FLT_MAX, np.float32)
    for rate, target, hmin, hmax in [(0.01, 10.0, 1.0, 10.5), (0.5, 3.0, None, None)]:
        ref = O.add_feature_window(old, pts, xy, h, w, rate, target, hmin, hmax)
        out = run_gpu(ctx, r, pts, old, rate, target, hmin, hmax)
        np.testing.assert_array_equal(out, ref)
