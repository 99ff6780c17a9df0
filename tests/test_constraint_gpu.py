"""SURVEY.md 8(f) next-2 parity: add_constant_value, Topo / Courant constraints,
Constraint._apply_rate and the arbitrary-point nearest-vertex query vs the
numpy/scipy oracle (reference ocsmesh/hfun/raster.py:509-736, 1438-1502,
ocsmesh/features/constraint.py:83-313)."""
import numpy as np
import pytest
from scipy.spatial import cKDTree

import ocsmesh_b200 as ob
from oracle import ocsmesh_oracle as O
from ocsmesh_b200 import _lib, synthetic as S

pytestmark = pytest.mark.gpu

FLT_MAX = np.finfo(np.float32).max


def both(h, w, seed, chunk_size=None, overlap=None, dtype=np.float32):
    z = S.synth_dem(h, w, seed=seed).astype(dtype)
    tr = S.dem_transform(h)
    gr = ob.Raster.from_numpy(z, tr, crs=32618, chunk_size=chunk_size, overlap=overlap)
    orr = O.OracleRaster(z, tr, chunk_size=chunk_size, overlap=overlap)
    return gr, orr


def seeded(gr, orr, hmin=50.0, hmax=2000.0, seed=5):
    """Both size functions after one contour pass (non-trivial old values)."""
    hf = ob.Hfun(gr, hmin=hmin, hmax=hmax)
    oh = O.OracleHfun(orr, hmin=hmin, hmax=hmax)
    lines = S.synth_shoreline(gr.height, gr.width, seed=seed, n_lines=2)
    hf.add_feature(lines, 0.01, 60.0)
    oh.add_feature(lines, 0.01, 60.0)
    np.testing.assert_array_equal(hf.get_values(), oh.values)
    return hf, oh


# ---- ocsb_nearest_vertex (arbitrary query points) ----------------------------
@pytest.mark.parametrize("nv,nq", [(1, 50), (7, 1000), (5000, 20000), (200000, 100000)])
@pytest.mark.parametrize("bounded", [False, True])
def test_nearest_vertex_vs_ckdtree(ctx, nv, nq, bounded):
    rng = np.random.default_rng(nv + nq)
    verts = np.stack([5e5 + rng.random(nv) * 4e4, 4.1e6 + rng.random(nv) * 2e4], axis=1)
    if nv > 100:      # duplicates: exact ties (repartition shares end nodes)
        verts[10:20] = verts[30:40]
    q = np.stack([5e5 - 5e3 + rng.random(nq) * 5e4, 4.1e6 - 5e3 + rng.random(nq) * 3e4],
                 axis=1)
    r = 900.0 if bounded else None
    dist = np.empty(nq)
    idx = np.empty(nq, dtype=np.int64)
    ctx.nearest_vertex(verts, nv, q, nq, r, dist, idx)
    tree = cKDTree(verts)
    if bounded:
        d_ref, i_ref = tree.query(q, distance_upper_bound=r)
    else:
        d_ref, i_ref = tree.query(q)
    np.testing.assert_array_equal(dist, d_ref)        # bit-exact f64 (inf on miss)
    miss = np.isinf(d_ref)
    assert np.all(idx[miss] == nv)
    hit = ~miss
    # index equal modulo exact ties: the vertex we name attains the minimum
    d_ours = np.sqrt(((q[hit] - verts[idx[hit]]) ** 2).sum(axis=1))
    np.testing.assert_array_equal(d_ours, d_ref[hit])
    differs = idx[hit] != i_ref[hit]
    if differs.any():
        a, b = verts[idx[hit][differs]], verts[i_ref[hit][differs]]
        d_a = ((q[hit][differs] - a) ** 2).sum(axis=1)
        d_b = ((q[hit][differs] - b) ** 2).sum(axis=1)
        np.testing.assert_array_equal(d_a, d_b)


def test_nearest_vertex_device_buffers(ctx):
    import torch
    rng = np.random.default_rng(2)
    verts = rng.random((3000, 2)) * 1000
    q = rng.random((4000, 2)) * 1200 - 100
    vd, qd = torch.from_numpy(verts).cuda(), torch.from_numpy(q).cuda()
    dist = torch.empty(4000, dtype=torch.float64, device="cuda")
    idx = torch.empty(4000, dtype=torch.int64, device="cuda")
    ctx.nearest_vertex(vd, 3000, qd, 4000, None, dist, idx)
    ctx.sync()
    d_ref, i_ref = cKDTree(verts).query(q)
    np.testing.assert_array_equal(dist.cpu().numpy(), d_ref)
    np.testing.assert_array_equal(idx.cpu().numpy(), i_ref)


# ---- add_constant_value --------------------------------------------------------
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("chunk,overlap", [(None, None), (100, 0), (96, 5)])
@pytest.mark.parametrize("lb,ub", [(None, None), (-3.0, 2.5), (None, 0.1), (0.1, None)])
def test_add_constant_value(ctx, dtype, chunk, overlap, lb, ub):
    gr, orr = both(230, 310, 4, chunk, overlap, dtype)
    hf, oh = seeded(gr, orr)
    hf.add_constant_value(123.456, lb, ub)
    oh.add_constant_value(123.456, lb, ub)
    np.testing.assert_array_equal(hf.get_values(), oh.values)


def test_add_constant_value_first_op(ctx):
    gr, orr = both(64, 100, 9)
    hf = ob.Hfun(gr, hmin=10.0, hmax=500.0)
    oh = O.OracleHfun(orr, hmin=10.0, hmax=500.0)
    hf.add_constant_value(77.7, -1.0, 4.0)
    oh.add_constant_value(77.7, -1.0, 4.0)
    np.testing.assert_array_equal(hf.get_values(), oh.values)


# ---- topo constraints ------------------------------------------------------------
@pytest.mark.parametrize("value_type", ["min", "max"])
@pytest.mark.parametrize("rate", [None, 0.01, 0.2])
@pytest.mark.parametrize("chunk,overlap", [(None, None), (128, 0), (100, 6)])
def test_topo_bound_constraint(ctx, value_type, rate, chunk, overlap):
    gr, orr = both(220, 300, 6, chunk, overlap)
    hf, oh = seeded(gr, orr)
    value = 400.0 if value_type == "min" else 90.0
    hf.add_topo_bound_constraint(value, upper_bound=3.0, lower_bound=-8.0,
                                 value_type=value_type, rate=rate)
    oh.add_topo_bound_constraint(value, upper_bound=3.0, lower_bound=-8.0,
                                 value_type=value_type, rate=rate)
    np.testing.assert_array_equal(hf.get_values(), oh.values)
    # constraints stay registered: the next op re-applies them
    hf.add_subtidal_flow_limiter(hmin=30.0, hmax=900.0)
    oh.add_subtidal_flow_limiter(30.0, 900.0, None, None)
    np.testing.assert_array_equal(hf.get_values(), oh.values)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rate", [None, 0.05])
def test_topo_func_constraint(ctx, dtype, rate):
    gr, orr = both(180, 260, 7, None, None, dtype)
    hf, oh = seeded(gr, orr)
    hf.add_topo_func_constraint(upper_bound=-1.0, value_type="min", rate=rate)
    oh.add_topo_func_constraint(upper_bound=-1.0, value_type="min", rate=rate)
    got, ref = hf.get_values(), oh.values
    if rate is None:
        np.testing.assert_array_equal(got, ref)
    else:
        # per-vertex values: exact distance ties may pick another mask pixel
        # (cKDTree's choice is traversal dependent)
        bad = got != ref
        assert bad.mean() < 1e-4
    # numpy-only callable (falls back to the host array for the callback)
    f = lambda z: np.abs(z) * 3.0 + 20.0      # noqa: E731
    hf.add_topo_func_constraint(f, lower_bound=0.5, value_type="max", rate=None)
    oh.add_topo_func_constraint(f, lower_bound=0.5, value_type="max", rate=None)
    if rate is None:
        np.testing.assert_array_equal(hf.get_values(), oh.values)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("ub,lb", [(0.9, None), (0.8, 0.3), (None, 0.2)])
def test_courant_constraint(ctx, dtype, ub, lb):
    gr, orr = both(200, 280, 8, 90, 3, dtype)
    hf, oh = seeded(gr, orr)
    hf.add_courant_num_constraint(upper_bound=ub, lower_bound=lb, timestep=20,
                                  wave_amplitude=1.5)
    oh.add_courant_num_constraint(upper_bound=ub, lower_bound=lb, timestep=20,
                                  wave_amplitude=1.5)
    np.testing.assert_array_equal(hf.get_values(), oh.values)


def test_courant_bounds_required(ctx):
    gr, _ = both(32, 32, 1)
    hf = ob.Hfun(gr, hmin=10.0, hmax=100.0)
    with pytest.raises(ValueError, match="can NOT be None"):
        hf.add_courant_num_constraint(upper_bound=None, lower_bound=None)


def test_hold_then_apply_once(ctx):
    gr, orr = both(150, 200, 11)
    hf, oh = seeded(gr, orr)
    with hf.hold_applying_added_constratins():
        hf.add_topo_bound_constraint(300.0, upper_bound=0.0, value_type="min", rate=0.02)
        hf.add_constant_value(80.0, 1.0, 2.0)
    # the reference applies the list once on exit (hfun/raster.py:476-485)
    oh._constraints = [O.TopoConstConstraint(300.0, 0.0, -np.inf, "min", 0.02)]
    new = oh.values.copy()
    z = orr.get_values(band=1)
    new = O.add_constant_value_window(oh.values, z, 80.0, 1.0, 2.0)
    oh.values = new
    oh._clamp_pass()
    np.testing.assert_array_equal(hf.get_values(), oh.values)


def test_apply_rate_known_answer(ctx):
    """tests/api/test_features.py:436-455 restated on a 1 x 12 raster: value 5
    at x = 3, rate 0.199, 'min' -> [2.015, 3.01, 4.005, 5, 10, ...]."""
    import torch
    w = 12
    z = np.zeros((2, w), np.float32)
    z[:, 3] = 1.0                        # the "region": topo band (0.5, 1.5)
    old = np.array([[0, 0, 0, 0, 10, 10, 10, 10, 20, 20, 20, 20]] * 2, np.float32)
    # window whose linspace axes are x = 0..11, y = 0 and 1e6 (rows far apart)
    win = _lib.make_win((0.0, 0.0, 11.0, 1e6), w, 2)
    zd, hd = torch.from_numpy(z).cuda(), torch.from_numpy(old).cuda()
    ctx.hfun_topo_constraint(win, zd, _lib.OCSB_F32, w, hd, w, None,
                             _lib.OCSB_F32, 0, 5.0, 0.5, 1.5, _lib.VT_MIN, 0.199)
    ctx.sync()
    got = hd.cpu().numpy()
    expect = np.array([5 * (1 - 3 * 0.199), 5 * (1 - 2 * 0.199), 5 * (1 - 0.199), 5,
                       10, 10, 10, 10, 20, 20, 20, 20])
    np.testing.assert_array_equal(got[0], expect.astype(np.float32))
    np.testing.assert_array_equal(np.trunc(got[0]), [2, 3, 4, 5, 10, 10, 10, 10,
                                                     20, 20, 20, 20])


# ---- golden: outputs of the reference's own source ------------------------------
def test_golden_constraints(ctx):
    import os
    from test_oracle_constraints import GOLD, OPS, case_names, lines_of, opt_int
    gold = np.load(GOLD)
    for name in case_names(gold):
        chunk, ovl = (opt_int(v) for v in gold[f"{name}_args"])

        def fresh():
            r = ob.Raster.from_numpy(gold["dem"], tuple(gold["dem_transform"]),
                                     crs=32618, chunk_size=chunk, overlap=ovl)
            hf = ob.Hfun(r, hmin=40.0, hmax=2000.0)
            hf.add_feature(lines_of(gold), 0.01, 60.0)
            return hf
        np.testing.assert_array_equal(fresh().get_values(), gold[f"{name}_seed"])
        for op, fn in OPS.items():
            hf = fresh()
            fn(hf)
            np.testing.assert_array_equal(hf.get_values(), gold[f"{name}_{op}"],
                                          err_msg=f"{name}/{op}")
            if op == "topo_min_rate":
                hf.add_subtidal_flow_limiter(30.0, 900.0, None, None)
                np.testing.assert_array_equal(
                    hf.get_values(), gold[f"{name}_topo_min_rate_then_limiter"])


# ---- a rate constraint on row bands (multi-GPU split of one window) -----------
@pytest.mark.parametrize("value_type,value", [("min", 400.0), ("max", 90.0)])
@pytest.mark.parametrize("cuts", [(0, 165, 330), (0, 97, 98, 250, 330)])
def test_rate_on_row_bands_equals_whole_window(ctx, value_type, value, cuts):
    """ocsb_hfun_topo_mask_band on every band, the bands' mask-boundary pixels
    concatenated in band order (what the ranks all-gather), ocsb_hfun_apply_rate
    on every band == ocsb_hfun_topo_constraint on the whole window, bit for bit
    (the nearest mask pixel of a band pixel often lies in another band)."""
    import torch
    gr, orr = both(330, 290, 8)
    hf, _ = seeded(gr, orr)
    H, W = gr.height, gr.width
    base = hf.device_values.clone()
    dem = gr.device_band(1, 0)
    bounds = gr.get_window_bounds(ob.Window(0, 0, W, H))
    vt = _lib.VT_MIN if value_type == "min" else _lib.VT_MAX
    lb, ub, rate = -8.0, 3.0, 0.02
    want = base.clone()
    ctx.hfun_topo_constraint(_lib.make_win(bounds, W, H), dem, _lib.OCSB_F32, W,
                             want, W, None, _lib.OCSB_F32, 0, value, lb, ub, vt, rate)
    got = base.clone()
    wins, masks, verts, vals = [], [], [], []
    for a, b in zip(cuts[:-1], cuts[1:]):
        cw = _lib.make_win(bounds, W, H, a, b)
        mask = torch.empty((b - a, W), dtype=torch.uint8, device="cuda")
        args = (cw, dem, _lib.OCSB_F32, W, got, W, None, _lib.OCSB_F32, 0, value,
                lb, ub, vt, mask)
        n = ctx.hfun_topo_mask_band(*args)
        v = torch.empty((n, 2), dtype=torch.float64, device="cuda")
        l = torch.empty((n,), dtype=torch.float32, device="cuda")
        if n:
            assert ctx.hfun_topo_mask_band(*args, v, l, n) == n
            with pytest.raises(_lib.OcsbError):      # capacity protocol
                ctx.hfun_topo_mask_band(*args, v, l, n - 1)
        wins.append(cw); masks.append(mask); verts.append(v); vals.append(l)
    allv, alll = torch.cat(verts), torch.cat(vals)
    assert len(allv) > 100
    for cw, mask in zip(wins, masks):
        ctx.hfun_apply_rate(cw, got, W, mask, allv, alll, len(allv), rate, vt)
    ctx.sync()
    assert torch.equal(got, want)
    assert not torch.equal(got, base)
    # no mask pixel anywhere: nothing happens (constraint.py:87-88)
    ctx.hfun_apply_rate(wins[0], got, W, masks[0], allv[:0], alll[:0], 0, rate, vt)
    assert torch.equal(got, want)
