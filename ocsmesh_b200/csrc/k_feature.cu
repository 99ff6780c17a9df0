// K1 -- nearest-vertex distance sizing (reference ocsmesh/hfun/raster.py:1230-1257:
// cKDTree(points).query(xy[, distance_upper_bound=r]) + size law + clamps + min).
//
// Design (B200-first, no kd-tree):
//
//  * The window is covered by a pyramid of pixel tiles (..., 4096, 1024, 256,
//    64 px square).  Every tile carries a CANDIDATE LIST of vertex ids that is
//    a superset of {nearest vertex of p : p pixel of the tile}.  A child list
//    is filtered from its parent list by two exact, conservative tests:
//      - radius : drop v if mindist(v, tile rect) > r          (bounded query)
//      - dominance: drop v if some kept vertex u is at least as close as v at
//        all four rect corners -- then u is at least as close on the whole
//        rect (|p-v|^2 - |p-u|^2 is linear in p), so v can never be the
//        strict nearest neighbour of a pixel in the tile.  Dominators are
//        the parent candidates nearest to the rect centre and to its corners.
//    For a shoreline sampled every `target_size` metres this leaves
//    ~(tile width / target_size + 2) candidates per tile *independently of
//    the distance to the shoreline*, where a bounding-box test leaves
//    O(sqrt(distance * tile)).
//  * The last filtering step (64x64 tile -> 32x8 leaf) is fused into the
//    evaluation kernel: one warp per leaf, lane = candidate for the filter
//    (redux.sync min + ballot), then lane = 4x2 pixels for the evaluation
//    with candidates broadcast by shuffle.
//  * Everything that decides a result is float64 and FMA-free:
//    d2 = dx*dx + dy*dy with separately rounded products, d = sqrt_rn(d2),
//    which is bit-identical to what scipy's cKDTree returns;
//    pixel coordinates follow numpy.linspace bit for bit.  B200 runs FP64 at
//    half the FP32 rate, which the pruning makes affordable.
//  * Epilogue fused: size law in f64 -> f32, [hmin,hmax] clamp, min(old, .),
//    optional A10 re-clamp, 16-byte hfun loads/stores.
//
// Algorithmic HBM bytes: 8 B/px (hfun read + write) + 16 B x N once.
#include "ocsb_common.cuh"

#include <limits>
#include <algorithm>
#include <cstdlib>

namespace {

#include "k_geom.cuh"

// ---------------------------------------------------------------------------
// level refinement
// ---------------------------------------------------------------------------
struct LevelDev {
    int tw, th, ntx, nty;   // tile size (px) and tile grid
    int *off;               // per tile: offset into ids
    int *cnt;               // per tile: list length
    int *ids;               // pool
};

struct RefineArgs {
    Geom g;
    LevelDev child;
    LevelDev parent;        // parent.tw == 0 -> root (ids = 0..n_root-1)
    int n_root;
    const double2 *verts;
    unsigned long long *pool_ctr;   // child pool bump counter
    unsigned long long pool_cap;
    int *overflow;
    double r2lim;           // r^2 (1+eps), +inf if unbounded
};

__device__ __forceinline__ void child_px(const RefineArgs &A, int t, int *c0,
                                         int *c1, int *r0, int *r1) {
    int tx = t % A.child.ntx, ty = t / A.child.ntx;
    *c0 = tx * A.child.tw;
    *c1 = min(A.g.w, *c0 + A.child.tw) - 1;
    *r0 = ty * A.child.th;
    *r1 = min(A.g.h, *r0 + A.child.th) - 1;
}

__device__ __forceinline__ void parent_list(const RefineArgs &A, int c0, int r0,
                                            int *off, int *cnt,
                                            const int **ids) {
    if (A.parent.tw == 0) {
        *off = 0; *cnt = A.n_root; *ids = nullptr;
    } else {
        int pt = (r0 / A.parent.th) * A.parent.ntx + (c0 / A.parent.tw);
        *off = A.parent.off[pt]; *cnt = A.parent.cnt[pt]; *ids = A.parent.ids;
    }
}

// block-wide argmin of 5 (d2,id) slots; result valid in all threads via smem
template <int BT>
__device__ __forceinline__ void block_best5(Best b[5], Best *s_out /*[5]*/,
                                            Best *s_tmp /*[BT/32][5]*/) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        Best r = best_warp_min(b[j]);
        if (lane == 0) s_tmp[wid * 5 + j] = r;
    }
    __syncthreads();
    if (threadIdx.x < 5) {
        Best r = s_tmp[threadIdx.x];
        for (int k = 1; k < BT / 32; ++k) {
            Best o = s_tmp[k * 5 + threadIdx.x];
            best_upd(r, o.d2, o.id);
        }
        s_out[threadIdx.x] = r;
    }
    __syncthreads();
}

__device__ __forceinline__ void scan_best5(const Rect &q, const double2 *verts,
                                           const int *ids, int off, int lo,
                                           int hi, int stride, Best b[5]) {
    for (int i = lo + (int)threadIdx.x; i < hi; i += stride) {
        int id = ids ? ids[off + i] : i;
        double2 v = verts[id];
        CornerD2 c = corner_d2(q, v);
        best_upd(b[0], center_d2(q, v), id);
        best_upd(b[1], c.d00, id);
        best_upd(b[2], c.d01, id);
        best_upd(b[3], c.d10, id);
        best_upd(b[4], c.d11, id);
    }
}

__device__ __forceinline__ void build_dom(const Rect &q, const double2 *verts,
                                          const Best *s_best, Dom *s_dom) {
    if (threadIdx.x < 5) {
        int id = s_best[threadIdx.x].id;
        CornerD2 c = corner_d2(q, verts[id]);
        const double m = 1.0 + kDomEps;
        s_dom->c[threadIdx.x][0] = c.d00 * m;
        s_dom->c[threadIdx.x][1] = c.d01 * m;
        s_dom->c[threadIdx.x][2] = c.d10 * m;
        s_dom->c[threadIdx.x][3] = c.d11 * m;
    }
    __syncthreads();
}

__device__ __forceinline__ bool keep_vertex(const Rect &q, const Dom &D,
                                            double r2lim, double2 v) {
    if (rect_d2(q, v) > r2lim) return false;
    CornerD2 c = corner_d2(q, v);
    return !dominated(D, c);
}

// fused single-CTA-per-tile refinement (levels with many tiles)
template <int BT>
__global__ void __launch_bounds__(BT) refine_fused_kernel(RefineArgs A) {
    __shared__ Best s_best[5];
    __shared__ Best s_tmp[(BT / 32) * 5];
    __shared__ Dom s_dom;
    __shared__ int s_count, s_pos;
    __shared__ unsigned long long s_base;
    if (*A.overflow) return;
    const int t = blockIdx.x;
    int c0, c1, r0, r1;
    child_px(A, t, &c0, &c1, &r0, &r1);
    const Rect q = tile_rect(A.g, c0, c1, r0, r1);
    int off, cnt;
    const int *ids;
    parent_list(A, c0, r0, &off, &cnt, &ids);
    if (cnt == 0) {
        if (threadIdx.x == 0) { A.child.off[t] = 0; A.child.cnt[t] = 0; }
        return;
    }
    Best b[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) { b[j].d2 = INFINITY; b[j].id = 0x7fffffff; }
    scan_best5(q, A.verts, ids, off, 0, cnt, BT, b);
    block_best5<BT>(b, s_best, s_tmp);
    build_dom(q, A.verts, s_best, &s_dom);
    if (threadIdx.x == 0) { s_count = 0; s_pos = 0; }
    __syncthreads();
    int local = 0;
    for (int i = threadIdx.x; i < cnt; i += BT) {
        int id = ids ? ids[off + i] : i;
        local += keep_vertex(q, s_dom, A.r2lim, A.verts[id]) ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&s_count, local);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long base = 0;
        int total = s_count;
        if (total) {
            base = atomicAdd(A.pool_ctr, (unsigned long long)total);
            if (base + total > A.pool_cap) { *A.overflow = 1; total = 0; }
        }
        s_base = base;
        s_count = total;
        A.child.off[t] = (int)base;
        A.child.cnt[t] = total;
    }
    __syncthreads();
    if (s_count == 0) return;
    const unsigned long long base = s_base;
    for (int i = threadIdx.x; i < cnt; i += BT) {
        int id = ids ? ids[off + i] : i;
        if (keep_vertex(q, s_dom, A.r2lim, A.verts[id])) {
            int pos = atomicAdd(&s_pos, 1);
            A.child.ids[base + pos] = id;
        }
    }
}

// warp-per-tile refinement for the finest levels (short parent lists, very
// many tiles): a one-warp CTA per tile is capped at 32 resident tiles per SM by
// the CTA limit, WPB tiles per CTA fill all 64 warp slots.  No block-level
// synchronisation; survivors are compacted in list order (ballot + popc).
template <int WPB>
__global__ void __launch_bounds__(WPB * 32) refine_warp_kernel(RefineArgs A, int ntiles) {
    __shared__ Dom s_dom[WPB];
    if (*A.overflow) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int t = blockIdx.x * WPB + wid;
    if (t >= ntiles) return;
    int c0, c1, r0, r1;
    child_px(A, t, &c0, &c1, &r0, &r1);
    const Rect q = tile_rect(A.g, c0, c1, r0, r1);
    int off, cnt;
    const int *ids;
    parent_list(A, c0, r0, &off, &cnt, &ids);
    if (cnt == 0) {
        if (lane == 0) { A.child.off[t] = 0; A.child.cnt[t] = 0; }
        return;
    }
    Best b[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) { b[j].d2 = INFINITY; b[j].id = 0x7fffffff; }
    for (int i = lane; i < cnt; i += 32) {
        const int id = ids ? ids[off + i] : i;
        const double2 v = A.verts[id];
        const CornerD2 c = corner_d2(q, v);
        best_upd(b[0], center_d2(q, v), id);
        best_upd(b[1], c.d00, id);
        best_upd(b[2], c.d01, id);
        best_upd(b[3], c.d10, id);
        best_upd(b[4], c.d11, id);
    }
    Dom &D = s_dom[wid];
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const Best r = best_warp_min(b[j]);
        if (lane == j) {
            const CornerD2 c = corner_d2(q, A.verts[r.id]);
            const double m = 1.0 + kDomEps;
            D.c[j][0] = c.d00 * m; D.c[j][1] = c.d01 * m;
            D.c[j][2] = c.d10 * m; D.c[j][3] = c.d11 * m;
        }
    }
    __syncwarp();
    // keep flags of the first 4 chunks are remembered, longer lists re-test
    unsigned keepm[4] = {0u, 0u, 0u, 0u};
    int total = 0;
    for (int b0 = 0, ch = 0; b0 < cnt; b0 += 32, ++ch) {
        const int i = b0 + lane;
        bool k = false;
        if (i < cnt) {
            const int id = ids ? ids[off + i] : i;
            k = keep_vertex(q, D, A.r2lim, A.verts[id]);
        }
        const unsigned m = __ballot_sync(0xffffffffu, k);
        if (ch < 4) keepm[ch] = m;
        total += __popc(m);
    }
    unsigned long long base = 0;
    if (lane == 0) {
        if (total) {
            base = atomicAdd(A.pool_ctr, (unsigned long long)total);
            if (base + total > A.pool_cap) { *A.overflow = 1; total = 0; }
        }
        A.child.off[t] = (int)base;
        A.child.cnt[t] = total;
    }
    total = __shfl_sync(0xffffffffu, total, 0);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (total == 0) return;
    int run = 0;
    for (int b0 = 0, ch = 0; b0 < cnt; b0 += 32, ++ch) {
        const int i = b0 + lane;
        int id = 0;
        if (i < cnt) id = ids ? ids[off + i] : i;
        unsigned m;
        if (ch < 4) {
            m = keepm[ch];
        } else {
            bool k = false;
            if (i < cnt) k = keep_vertex(q, D, A.r2lim, A.verts[id]);
            m = __ballot_sync(0xffffffffu, k);
        }
        if ((m >> lane) & 1u)
            A.child.ids[base + run + __popc(m & ((1u << lane) - 1u))] = id;
        run += __popc(m);
    }
}

// ---- split path for levels with few tiles and long parent lists ------------
struct SplitScratch {
    Best *part;     // [ntiles][nsplit][5]
    int *counts;    // [ntiles][nsplit]  (count, then exclusive offset)
    int nsplit;
};

__device__ __forceinline__ void split_range(int cnt, int nsplit, int s, int *lo,
                                            int *hi) {
    int chunk = (cnt + nsplit - 1) / nsplit;
    *lo = min(cnt, s * chunk);
    *hi = min(cnt, *lo + chunk);
}

__global__ void __launch_bounds__(256)
split_p1_kernel(RefineArgs A, SplitScratch S) {
    __shared__ Best s_best[5];
    __shared__ Best s_tmp[8 * 5];
    const int t = blockIdx.x / S.nsplit, s = blockIdx.x % S.nsplit;
    int c0, c1, r0, r1;
    child_px(A, t, &c0, &c1, &r0, &r1);
    const Rect q = tile_rect(A.g, c0, c1, r0, r1);
    int off, cnt;
    const int *ids;
    parent_list(A, c0, r0, &off, &cnt, &ids);
    int lo, hi;
    split_range(cnt, S.nsplit, s, &lo, &hi);
    Best b[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) { b[j].d2 = INFINITY; b[j].id = 0x7fffffff; }
    scan_best5(q, A.verts, ids, off, lo, hi, 256, b);
    block_best5<256>(b, s_best, s_tmp);
    if (threadIdx.x < 5)
        S.part[((int64_t)t * S.nsplit + s) * 5 + threadIdx.x] = s_best[threadIdx.x];
}

__device__ __forceinline__ bool split_dom(const RefineArgs &A,
                                          const SplitScratch &S, int t,
                                          const Rect &q, Best *s_best,
                                          Dom *s_dom) {
    if (threadIdx.x < 5) {
        Best r; r.d2 = INFINITY; r.id = 0x7fffffff;
        for (int k = 0; k < S.nsplit; ++k) {
            Best o = S.part[((int64_t)t * S.nsplit + k) * 5 + threadIdx.x];
            best_upd(r, o.d2, o.id);
        }
        s_best[threadIdx.x] = r;
    }
    __syncthreads();
    if (s_best[0].id == 0x7fffffff) return false;  // empty parent list
    build_dom(q, A.verts, s_best, s_dom);
    return true;
}

__global__ void __launch_bounds__(256)
split_p2_kernel(RefineArgs A, SplitScratch S) {
    __shared__ Best s_best[5];
    __shared__ Dom s_dom;
    __shared__ int s_count;
    const int t = blockIdx.x / S.nsplit, s = blockIdx.x % S.nsplit;
    int c0, c1, r0, r1;
    child_px(A, t, &c0, &c1, &r0, &r1);
    const Rect q = tile_rect(A.g, c0, c1, r0, r1);
    int off, cnt;
    const int *ids;
    parent_list(A, c0, r0, &off, &cnt, &ids);
    if (threadIdx.x == 0) s_count = 0;
    if (!split_dom(A, S, t, q, s_best, &s_dom)) {
        if (threadIdx.x == 0) S.counts[(int64_t)t * S.nsplit + s] = 0;
        return;
    }
    int lo, hi;
    split_range(cnt, S.nsplit, s, &lo, &hi);
    int local = 0;
    for (int i = lo + threadIdx.x; i < hi; i += 256) {
        int id = ids ? ids[off + i] : i;
        local += keep_vertex(q, s_dom, A.r2lim, A.verts[id]) ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&s_count, local);
    __syncthreads();
    if (threadIdx.x == 0) S.counts[(int64_t)t * S.nsplit + s] = s_count;
}

__global__ void split_p3_kernel(RefineArgs A, SplitScratch S, int ntiles) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    int total = 0;
    for (int k = 0; k < S.nsplit; ++k) {
        int c = S.counts[(int64_t)t * S.nsplit + k];
        S.counts[(int64_t)t * S.nsplit + k] = total;
        total += c;
    }
    unsigned long long base = 0;
    if (total) {
        base = atomicAdd(A.pool_ctr, (unsigned long long)total);
        if (base + total > A.pool_cap) { *A.overflow = 1; total = 0; }
    }
    A.child.off[t] = (int)base;
    A.child.cnt[t] = total;
}

__global__ void __launch_bounds__(256)
split_p4_kernel(RefineArgs A, SplitScratch S) {
    __shared__ Best s_best[5];
    __shared__ Dom s_dom;
    __shared__ int s_pos;
    const int t = blockIdx.x / S.nsplit, s = blockIdx.x % S.nsplit;
    if (A.child.cnt[t] == 0) return;
    int c0, c1, r0, r1;
    child_px(A, t, &c0, &c1, &r0, &r1);
    const Rect q = tile_rect(A.g, c0, c1, r0, r1);
    int off, cnt;
    const int *ids;
    parent_list(A, c0, r0, &off, &cnt, &ids);
    if (threadIdx.x == 0) s_pos = 0;
    if (!split_dom(A, S, t, q, s_best, &s_dom)) return;
    int lo, hi;
    split_range(cnt, S.nsplit, s, &lo, &hi);
    const int64_t base = (int64_t)A.child.off[t] + S.counts[(int64_t)t * S.nsplit + s];
    for (int i = lo + threadIdx.x; i < hi; i += 256) {
        int id = ids ? ids[off + i] : i;
        if (keep_vertex(q, s_dom, A.r2lim, A.verts[id])) {
            int pos = atomicAdd(&s_pos, 1);
            A.child.ids[base + pos] = id;
        }
    }
}

#include "k_feature_eval.cuh"

// ---------------------------------------------------------------------------
// host orchestration
// ---------------------------------------------------------------------------
struct LevelHost {
    int tw, th, ntx, nty;
    int64_t ntiles;
};

template <int BT>
void launch_fused(ocsb_ctx *ctx, const RefineArgs &A, int64_t ntiles) {
    OCSB_TIME(ctx, KC_FEATURE_REFINE);
    refine_fused_kernel<BT><<<(unsigned)ntiles, BT, 0, ctx->stream>>>(A);
    OCSB_LAUNCHED(ctx);
}

void launch_warp(ocsb_ctx *ctx, const RefineArgs &A, int64_t ntiles) {
    constexpr int WPB = 8;
    OCSB_TIME(ctx, KC_FEATURE_REFINE);
    refine_warp_kernel<WPB><<<(unsigned)((ntiles + WPB - 1) / WPB), WPB * 32, 0,
                              ctx->stream>>>(A, (int)ntiles);
    OCSB_LAUNCHED(ctx);
}

// development knobs (tools/sweep_pass.py); defaults are the measured optimum
int knob(const char *name, int dflt) {
    const char *e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}

}  // namespace

int32_t k_hfun_feature(ocsb_ctx *ctx, const FeatureArgs &a, const double *verts,
                       int64_t nv, const float *old_, int64_t old_stride,
                       float *new_, int64_t new_stride, double *dist_out,
                       int64_t *idx_out) {
    OCSB_REQUIRE(((uintptr_t)verts & 15) == 0,
                 "ocsb_hfun_feature: verts_xy must be 16-byte aligned");
    const int w = a.ax.n, h = a.row_end - a.row_begin;
    Geom g{a.ax, a.ay, w, h, a.row_begin};

    // ---- level plan --------------------------------------------------------
    std::vector<LevelHost> lv;
    for (int s = kTile0; (int)lv.size() < kMaxLevels; s *= 4) {
        LevelHost L;
        L.tw = L.th = s;
        L.ntx = (w + s - 1) / s;
        L.nty = (h + s - 1) / s;
        L.ntiles = (int64_t)L.ntx * L.nty;
        lv.push_back(L);
        if (L.ntiles == 1) break;
        if (s > (1 << 28)) break;
    }
    // the coarsest level actually built: the finest one with <= 4 tiles or, if
    // the parent list is the whole vertex set, any level is fine -- keep all
    // levels down from the first with ntiles <= 16 so upper lists stay short
    // ... but the top level scans the WHOLE vertex set once per tile, and each
    // further level costs a launch chain: start at the finest level whose
    // root scan (ntiles x nv vertex tests) stays within a fixed budget
    int top = 0;
    for (int i = 0; i < (int)lv.size(); ++i) {
        top = i;
        if (lv[i].ntiles <= 16) break;
    }
    {
        static const double budget = (double)knob("OCSB_REFINE_BUDGET_M", 4) * 1e6;
        for (int i = 0; i < top; ++i)
            if ((double)lv[i].ntiles * (double)nv <= budget) { top = i; break; }
    }

    const double inf = std::numeric_limits<double>::infinity();
    const bool bounded = a.r < inf;
    const double r2 = bounded ? a.r * a.r : inf;
    const double r2lim = bounded ? (a.r > 0 ? r2 * (1.0 + kDomEps) : -1.0) : inf;

    ScratchGuard sg(ctx);
    {
        // counters + per-level offset / count tables in one allocation
        size_t fixed = ScratchGuard::piece((kMaxLevels + 2) * sizeof(unsigned long long));
        for (int i = 0; i <= top; ++i)
            fixed += 2 * ScratchGuard::piece((size_t)lv[i].ntiles * sizeof(int));
        OCSB_TRY(sg.reserve(fixed));
    }
    unsigned long long *ctrs = nullptr;   // [kMaxLevels] + overflow flag behind
    OCSB_TRY(sg.get(&ctrs, kMaxLevels + 2));
    int *overflow = reinterpret_cast<int *>(ctrs + kMaxLevels);
    const bool async_mode = a.ext_overflow != nullptr;
    if (async_mode) overflow = a.ext_overflow;

    std::vector<LevelDev> dev(lv.size());
    for (int i = 0; i <= top; ++i) {
        dev[i].tw = lv[i].tw; dev[i].th = lv[i].th;
        dev[i].ntx = lv[i].ntx; dev[i].nty = lv[i].nty;
        OCSB_TRY(sg.get(&dev[i].off, (size_t)lv[i].ntiles));
        OCSB_TRY(sg.get(&dev[i].cnt, (size_t)lv[i].ntiles));
        dev[i].ids = nullptr;
    }

    bool hook_done = false;
    for (int attempt = 0; attempt < 6; ++attempt) {
        ScratchGuard pools(ctx);
        OCSB_CUDA(cudaMemsetAsync(ctrs, 0, (kMaxLevels + 2) * sizeof(*ctrs),
                                  ctx->stream));
        // fork: the refinement launches below go to a.refine_stream
        struct StreamSwap {
            ocsb_ctx *c; cudaStream_t saved; bool on = false;
            void swap(cudaStream_t s) { saved = c->stream; c->stream = s; on = true; }
            void restore() { if (on) { c->stream = saved; on = false; } }
            ~StreamSwap() { restore(); }
        } sw{ctx, nullptr};
        const bool fork = a.refine_stream && a.between && !hook_done;
        if (fork) {
            OCSB_CUDA(cudaEventRecord(a.ev_fork, ctx->stream));
            OCSB_CUDA(cudaStreamWaitEvent(a.refine_stream, a.ev_fork, 0));
            sw.swap(a.refine_stream);
        }
        std::vector<unsigned long long> cap(top + 1);
        size_t pool_bytes = 0;
        for (int i = 0; i <= top; ++i) {
            unsigned long long c = 4ull * (unsigned long long)nv +
                                   48ull * (unsigned long long)lv[i].ntiles +
                                   (1ull << 16);
            c <<= (2 * attempt);
            if (async_mode) c <<= 1;   // one shot: start with doubled pools
            if (c > 0x7fffffffull) c = 0x7fffffffull;  // int offsets
            cap[i] = c;
            pool_bytes += ScratchGuard::piece((size_t)c * sizeof(*dev[i].ids));
        }
        // all candidate pools of this attempt in one allocation (+ room for
        // the split scratch of a typical level)
        OCSB_TRY(pools.reserve(pool_bytes + (1u << 20)));
        for (int i = 0; i <= top; ++i)
            OCSB_TRY(pools.get(&dev[i].ids, (size_t)cap[i]));
        for (int i = top; i >= 0; --i) {
            RefineArgs A;
            A.g = g;
            A.child = dev[i];
            if (i == top) {
                A.parent = LevelDev{0, 0, 0, 0, nullptr, nullptr, nullptr};
            } else {
                A.parent = dev[i + 1];
            }
            A.n_root = (int)nv;
            A.verts = reinterpret_cast<const double2 *>(verts);
            A.pool_ctr = ctrs + i;
            A.pool_cap = cap[i];
            A.overflow = overflow;
            A.r2lim = r2lim;
            const int64_t ntiles = lv[i].ntiles;
            // long parent lists on few tiles: split each list over several CTAs
            int nsplit = 1;
            const int64_t want = 2 * (int64_t)ctx->num_sms;
            static const int topsplit = knob("OCSB_REFINE_TOPSPLIT", 0);
            if ((ntiles < want && (i == top ? nv : 4096) >= 4096) ||
                (topsplit && i == top && nv >= 8192)) {
                int64_t est = (i == top) ? nv : nv / 2;   // parent list scale
                nsplit = (int)((want + ntiles - 1) / ntiles);
                // a root scan is long whatever the tile count: split it too
                if (i == top && topsplit && nsplit < (int)((nv + 4095) / 4096))
                    nsplit = (int)((nv + 4095) / 4096);
                int64_t maxs = (est + 2047) / 2048;
                if (nsplit > maxs) nsplit = (int)maxs;
                if (nsplit > kMaxSplit) nsplit = kMaxSplit;
                if (nsplit < 1) nsplit = 1;
            }
            // dense vertex sets (configs[4]: 5 M vertices): the parent lists of
            // the upper levels hold ~ nv x (parent tile area / raster area)
            // entries whatever the tile count; one CTA per 8 k expected
            // entries keeps every scan short (empty splits exit at once)
            if (nv >= 262144) {
                double frac = 1.0;
                if (i != top)
                    frac = 4.0 * ((double)lv[i + 1].tw * (double)lv[i + 1].th) /
                           ((double)w * (double)h);
                if (frac > 1.0) frac = 1.0;
                const double est_len = frac * (double)nv;
                int by_len = (int)std::min<double>((double)kMaxSplit, est_len / 8192.0);
                if (est_len >= 65536.0 && by_len > nsplit) nsplit = by_len;
            }
            if (nsplit > 1) {
                SplitScratch S;
                S.nsplit = nsplit;
                OCSB_TRY(pools.get(&S.part, (size_t)ntiles * nsplit * 5));
                OCSB_TRY(pools.get(&S.counts, (size_t)ntiles * nsplit));
                const unsigned gridn = (unsigned)(ntiles * nsplit);
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                split_p1_kernel<<<gridn, 256, 0, ctx->stream>>>(A, S);
                OCSB_LAUNCHED(ctx);
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                split_p2_kernel<<<gridn, 256, 0, ctx->stream>>>(A, S);
                OCSB_LAUNCHED(ctx);
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                split_p3_kernel<<<(unsigned)((ntiles + 127) / 128), 128, 0,
                                  ctx->stream>>>(A, S, (int)ntiles);
                OCSB_LAUNCHED(ctx);
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                split_p4_kernel<<<gridn, 256, 0, ctx->stream>>>(A, S);
                OCSB_LAUNCHED(ctx);
            } else if (i == 0) {
                static const int warp0 = knob("OCSB_REFINE_WARP0", 1);
                if (i == top && nv > 1024) launch_fused<256>(ctx, A, ntiles);
                else if (warp0) launch_warp(ctx, A, ntiles);
                else launch_fused<32>(ctx, A, ntiles);
            } else if (i == 1 && i != top) {
                static const int bt1 = knob("OCSB_REFINE_BT1", 256);
                if (bt1 == 64) launch_fused<64>(ctx, A, ntiles);
                else if (bt1 == 128) launch_fused<128>(ctx, A, ntiles);
                else launch_fused<256>(ctx, A, ntiles);
            } else {
                launch_fused<256>(ctx, A, ntiles);
            }
            OCSB_CHECK_LAUNCH();
        }

        if (fork) {
            // join: other ops of the list first, then wait for the pyramid
            OCSB_CUDA(cudaEventRecord(a.ev_join, a.refine_stream));
            sw.restore();
            hook_done = true;
            OCSB_TRY((*a.between)());
            OCSB_CUDA(cudaStreamWaitEvent(ctx->stream, a.ev_join, 0));
        }
        EvalArgs E;
        E.g = g;
        E.lvl = dev[0];
        E.verts = reinterpret_cast<const double2 *>(verts);
        E.old_ = old_; E.os = old_stride;
        E.new_ = new_; E.ns = new_stride;
        E.dist = dist_out; E.idx = idx_out; E.nv = nv;
        E.rate_target = a.rate_target; E.target = a.target;
        E.r = a.r; E.r2 = (bounded && a.r <= 0) ? -1.0 : r2; E.r2lim = r2lim;
        {
            volatile double m = a.rate_target * a.r;   // no FMA contraction
            volatile double s = m + a.target;
            E.v_far = s;
        }
        const float finf = std::numeric_limits<float>::infinity();
        E.lo = a.has_hmin ? a.hmin_f : -finf;
        E.hi = a.has_hmax ? a.hmax_f : finf;
        // the fast size law needs v64 = rt*d + target to be a normal float32
        // with |error| <= 2^-39 v64: rt, target >= 0, target not tiny, no
        // overflow for d <= 1e15
        E.fast_ok = (a.rate_target >= 0.0 && a.rate_target <= 1e15 &&
                     a.target >= 1e-30 && a.target <= 1e30) ? 1 : 0;
        {
            auto hi_word = [](double v) {
                uint64_t b; memcpy(&b, &v, 8); return (uint32_t)(b >> 32);
            };
            const double d2max = (bounded && r2 < 1e30) ? r2 : 1e30;
            const uint32_t lo = hi_word(1e-30) + 1, hi = hi_word(d2max);
            E.hi_lo = lo;
            E.hi_span = (E.fast_ok && d2max > 1e-29 && hi > lo) ? hi - lo : 0u;
            E.v_far_f = (float)E.v_far;
            E.old_const = std::numeric_limits<float>::max();
        }
        E.fuse_clamp = a.fuse_clamp;
        E.vec_old = old_ && (((uintptr_t)old_ & 15) == 0) && (old_stride % 4 == 0);
        E.vec_new = (((uintptr_t)new_ & 15) == 0) && (new_stride % 4 == 0);
        E.overflow = overflow;
        E.rate_vals = a.rate_vals; E.rate_mask = a.rate_mask;
        E.rate = a.rate; E.rate_sign = a.rate_sign; E.rate_is_min = a.rate_is_min;
        OCSB_REQUIRE(!a.rate_vals || (old_ && a.rate_mask),
                     "k_hfun_feature: rate epilogue needs old values and a mask");
        OCSB_TIME(ctx, KC_FEATURE_EVAL);
        const dim3 egrid((unsigned)lv[0].ntx, (unsigned)lv[0].nty);
        if (idx_out || dist_out || a.rate_vals) {
            if (old_) eval_kernel<true, true><<<egrid, kEvalThreads, 0, ctx->stream>>>(E);
            else eval_kernel<true, false><<<egrid, kEvalThreads, 0, ctx->stream>>>(E);
        } else {
            if (old_) eval_kernel<false, true><<<egrid, kEvalThreads, 0, ctx->stream>>>(E);
            else eval_kernel<false, false><<<egrid, kEvalThreads, 0, ctx->stream>>>(E);
        }
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();

        // asynchronous mode: the caller inspects *ext_overflow later
        if (async_mode) return OCSB_OK;

        // candidate pools are bump-allocated with a size guess: read the
        // overflow flag back and retry with larger pools if it tripped (the
        // evaluation kernel writes nothing in that case)
        int *flag_h = reinterpret_cast<int *>(ctx->pinned);
        int flag_local = 0;
        if (!flag_h) flag_h = &flag_local;
        OCSB_CUDA(cudaMemcpyAsync(flag_h, overflow, sizeof(int),
                                  cudaMemcpyDeviceToHost, ctx->stream));
        OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
        if (*flag_h == 0) return OCSB_OK;
    }
    ocsb_set_error("ocsb_hfun_feature: candidate pools overflowed after 6 "
                   "attempts (degenerate vertex set?)");
    return OCSB_EOVERFLOW;
}
