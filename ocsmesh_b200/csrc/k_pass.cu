// K-pass -- the op list of a size-function pass in ONE sweep over the raster
// (reference: the window loops of HfunRaster.add_feature / add_contour,
// ocsmesh/hfun/raster.py:1212-1260, and add_subtidal_flow_limiter, :1393-1436,
// over Raster.iter_windows, ocsmesh/raster.py:1577-1608).
//
// Every op of the path has the form  new = clampG(min(old, f_op(pixel)))  with a
// per-pixel f_op, so a list of ops collapses into one read-modify-write of the
// size raster: 12 B/px (DEM read, hfun read, hfun write) for the whole list
// instead of 8 + 12 B/px per (feature, limiter) pair.  One CTA owns one 64x64
// pixel tile of one semantic window:
//
//   * the DEM tile (+ halo) is fetched by ONE cp.async.bulk.tensor.2d (TMA, one
//     elected thread, one mbarrier) at the very start and consumed last, so the
//     whole nearest-vertex stage hides its latency; window borders are index
//     tests (one-sided differences), the zero fill of out-of-bounds box parts is
//     never used;
//   * feature ops: the 256-px candidate list of the tile's parent is filtered
//     by the CTA to the tile (shared memory), by each warp to its 32x16 rect,
//     then every lane evaluates 4x4 pixels against the survivors in float64
//     without FMA contraction (bit-identical to cKDTree distances); the size
//     law epilogue rounds through a checked fast square root;
//   * limiter ops: float32 gradient / sqrt / divide sequences that are
//     bit-identical to IEEE (see k_limiter.cu), issued as packed f32x2
//     instructions (FFMA2 / FMUL2 / FADD2: half the issue slots of the scalar
//     forms -- the kernel is issue-bound, not FLOP-bound).
//
// The window table makes semantic windows units of work of the SAME launch:
// tile -> (window, tile-in-window) by a search in the table; coordinates,
// search radius and border rules are the window's own.
//
// Candidate lists come from a two-level pyramid per feature op (1024- and
// 256-px tiles).  The 1024-px lists are built from a uniform grid of the
// vertices (counting sort by cell) instead of a scan of the whole set per
// tile: ring search around the tile centre for dominators, then only the cells
// within min(r, dominator bound) of the tile are tested.
#include "ocsb_common.cuh"

#include <cuda.h>

#include <algorithm>
#include <limits>

namespace {

#include "k_geom.cuh"
#include "k_limiter_math.cuh"

constexpr int kT0 = 64, kT1 = 256, kT2 = 1024;   // tile sizes of the levels (px)
constexpr int kPassThreads = 256;
#ifndef OCSB_PASS_MIN_CTAS
#define OCSB_PASS_MIN_CTAS 3
#endif
constexpr int kListCap = 256;        // tile candidate list in shared memory
constexpr int kSurv = 64;            // survivors of a warp's 32x16 rect
constexpr int kDemPitch = 76;        // floats per staged DEM row: 4 + 64 + 4, +4 pad
constexpr int kDemRows = kT0 + 2;
constexpr int kMaxFeat = 4, kMaxLim = 4;
constexpr double kBig = 1e300;       // "no candidate" squared distance (finite)

#ifdef OCSB_TIMELINE
// tuning build only: first CTA start / last CTA end of each launch of a pass on
// the GPU's global timer (ns) -- which launches really overlap
__device__ unsigned long long g_timeline[8][2];
__device__ __forceinline__ unsigned long long tl_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
struct TlScope {
    int k;
    __device__ explicit TlScope(int k_) : k(k_) {
        if (threadIdx.x == 0) atomicMin(&g_timeline[k][0], tl_now());
    }
    __device__ ~TlScope() {
        if (threadIdx.x == 0) atomicMax(&g_timeline[k][1], tl_now());
    }
};
#define TL_SCOPE(k) TlScope tl_scope__(k)
#else
#define TL_SCOPE(k)
#endif

#ifdef OCSB_PASS_STATS
// tuning build only (tools/pass_lab.py): list lengths and per-phase clocks
// (one slot row per SM: no contention between SMs)
__device__ unsigned long long g_pass_stats[160][32];
__device__ __forceinline__ unsigned pstat_smid() {
    unsigned r;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(r));
    return r;
}
#define PSTAT_ADD(i, v) atomicAdd(&g_pass_stats[pstat_smid()][i], (unsigned long long)(v))
#define PSTAT_CLK(var) const long long var = clock64()
#else
#define PSTAT_ADD(i, v)
#define PSTAT_CLK(var)
#endif

// ---------------------------------------------------------------------------
// window table
// ---------------------------------------------------------------------------
struct __align__(16) PWin {     // 144 bytes: bulk-copied into shared memory
    LinAxis ax, ay;
    int rb, re;             // rows [rb, re) of the window live on this device
    int keep_w, keep_h;     // window pixels (r < keep_h, c < keep_w) are written
    long long hoff;         // element offset of window pixel (0, 0) in the hfun buffers
    int dem_c, dem_r;       // DEM buffer column / row of window pixel (0, 0)
    int n0x, n0y, base0;    // 64-px tiles over [0, w) x [rb, re); first global index
    int n1x, n1y, base1;
    int n2x, n2y, base2;
    int vec;                // 16-byte hfun accesses are aligned in this window
    int pad_[2];
};
static_assert(sizeof(PWin) % 16 == 0, "PWin is moved by cp.async.bulk");

__device__ __forceinline__ int find_win(const PWin *wins, int n, int t, int level) {
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        const int b = level == 0 ? wins[mid].base0 : level == 1 ? wins[mid].base1
                                                                : wins[mid].base2;
        if (b <= t) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ Rect win_rect(const PWin &W, int c0, int c1, int r0, int r1) {
    // pixel rect (band-relative rows) -> coordinates; axes may run either way
    const double xa = lin_at(W.ax, c0), xb = lin_at(W.ax, c1);
    const double ya = lin_at(W.ay, W.rb + r0), yb = lin_at(W.ay, W.rb + r1);
    Rect q;
    q.xlo = xa < xb ? xa : xb; q.xhi = xa < xb ? xb : xa;
    q.ylo = ya < yb ? ya : yb; q.yhi = ya < yb ? yb : ya;
    return q;
}

// ---------------------------------------------------------------------------
// vertex grid (counting sort of the vertices by cell, row-major cell order)
// ---------------------------------------------------------------------------
struct VGrid {
    double x0, y0, inv, cell;
    int ncx, ncy;           // interior cells; stored grid is (ncx+2) x (ncy+2):
                            // the border ring collects everything outside
    int *start;             // (ncx+2)*(ncy+2) + 1 offsets
    double2 *sv;            // vertices in cell order
    int *sid;               // their original indices
};

__device__ __forceinline__ int vg_cell1(double t, int n) {
    // monotone in t: a coordinate interval maps onto a cell index interval
    return t < 0.0 ? 0 : (t >= (double)n ? n + 1 : (int)t + 1);
}
__device__ __forceinline__ int vg_cx(const VGrid &G, double x) {
    return vg_cell1((x - G.x0) * G.inv, G.ncx);
}
__device__ __forceinline__ int vg_cy(const VGrid &G, double y) {
    return vg_cell1((y - G.y0) * G.inv, G.ncy);
}

__global__ void vg_hist_kernel(VGrid G, const double2 *verts, int nv, int *counts,
                               int *far_flag) {
    TL_SCOPE(0);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nv) return;
    const double2 v = verts[i];
    atomicAdd(&counts[vg_cy(G, v.y) * (G.ncx + 2) + vg_cx(G, v.x)], 1);
    // the fast size law bounds d2 < 1e30 through |coordinates| <= 1e14
    if (!(fabs(v.x) <= 1e14 && fabs(v.y) <= 1e14)) *far_flag = 1;
}

// exclusive scan of counts[0..n) into start[0..n], cursor = copy of start
__global__ void __launch_bounds__(1024) vg_scan_kernel(const int *counts, int n,
                                                        int *start, int *cursor) {
    __shared__ int s_part[1024];
    const int per = (n + 1023) / 1024;
    const int lo = min(n, (int)threadIdx.x * per), hi = min(n, lo + per);
    int sum = 0;
    for (int i = lo; i < hi; ++i) sum += counts[i];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        int v = threadIdx.x >= o ? s_part[threadIdx.x - o] : 0;
        __syncthreads();
        s_part[threadIdx.x] += v;
        __syncthreads();
    }
    int run = s_part[threadIdx.x] - sum;
    for (int i = lo; i < hi; ++i) {
        start[i] = run; cursor[i] = run;
        run += counts[i];
    }
    if (threadIdx.x == 1023) start[n] = s_part[1023];
}

__global__ void vg_scatter_kernel(VGrid G, const double2 *verts, int nv, int *cursor) {
    TL_SCOPE(1);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nv) return;
    const double2 v = verts[i];
    const int pos = atomicAdd(&cursor[vg_cy(G, v.y) * (G.ncx + 2) + vg_cx(G, v.x)], 1);
    G.sv[pos] = v;
    G.sid[pos] = i;
}

// scan + scatter in one launch for grids of up to kVgFusedCells cells: every CTA
// redoes the exclusive scan of the (small) histogram in shared memory instead of
// waiting for a one-CTA scan launch of its own (that launch and the gaps around
// it were ~25 us of the chain every step, profiles/r2c_split_lab.txt run 5);
// positions inside a cell come from a zeroed fill counter
constexpr int kVgFusedCells = 4096;

// WITH_HIST: the histogram too, in the same launch -- the CTAs meet at a counter
// (`bar`, zeroed with the histogram).  Only launched with a grid that is resident
// as a whole (<= 2 CTAs per SM of 256 threads), so the spin cannot starve anybody.
template <bool WITH_HIST>
__global__ void __launch_bounds__(256) vg_scan_scatter_kernel(VGrid G, const double2 *verts,
                                                              int nv, int *counts, int *fill,
                                                              int *far_flag, int *bar,
                                                              int *overflow, int *sticky) {
    TL_SCOPE(1);
    __shared__ int s_start[kVgFusedCells + 1];
    __shared__ int s_warp[8];
    if (WITH_HIST) {
        const int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < nv) {
            const double2 v = verts[i];
            atomicAdd(&counts[vg_cy(G, v.y) * (G.ncx + 2) + vg_cx(G, v.x)], 1);
            if (!(fabs(v.x) <= 1e14 && fabs(v.y) <= 1e14)) *far_flag = 1;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(bar, 1);
            // bounded (~1 s): if the grid is ever not resident as a whole the pass
            // is flagged like a pool overflow (the host redoes it) instead of hanging
            long long spins = 0;
            while (*(volatile int *)bar < (int)gridDim.x) {
                if (++spins > 10000000ll) {
                    *overflow = 1;
                    if (sticky) *sticky = 1;
                    break;
                }
                __nanosleep(100);
            }
            __threadfence();
        }
        __syncthreads();
    }
    const int ncell = (G.ncx + 2) * (G.ncy + 2);
    const int per = (ncell + 255) / 256;
    const int lo = min(ncell, (int)threadIdx.x * per), hi = min(ncell, lo + per);
    int sum = 0;
    for (int i = lo; i < hi; ++i) sum += __ldcg(counts + i);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < wid; ++w) base += s_warp[w];
    int run = base + inc - sum;
    for (int i = lo; i < hi; ++i) {
        s_start[i] = run;
        run += __ldcg(counts + i);
    }
    if (threadIdx.x == 255) s_start[ncell] = run;
    __syncthreads();
    if (blockIdx.x == 0)
        for (int i = threadIdx.x; i <= ncell; i += 256) G.start[i] = s_start[i];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nv) return;
    const double2 v = verts[i];
    const int c = vg_cy(G, v.y) * (G.ncx + 2) + vg_cx(G, v.x);
    const int pos = s_start[c] + atomicAdd(&fill[c], 1);
    G.sv[pos] = v;
    G.sid[pos] = i;
}

// very small vertex sets: histogram, scan and scatter in ONE launch of one CTA
// -- the three launches above are pure launch latency at that size
constexpr int kVgSmallCells = 8192;
constexpr int kVgSmallVerts = 4096;   // beyond that one CTA's shared-memory atomics
                                      // on a few populated cells are slower (measured)

__global__ void __launch_bounds__(1024) vg_build_small_kernel(VGrid G, const double2 *verts,
                                                              int nv, int *far_flag) {
    __shared__ int s_cnt[kVgSmallCells];
    __shared__ int s_part[1024];
    const int NX = G.ncx + 2, ncell = NX * (G.ncy + 2);
    for (int i = threadIdx.x; i < ncell; i += 1024) s_cnt[i] = 0;
    __syncthreads();
    bool far = false;
    for (int i = threadIdx.x; i < nv; i += 1024) {
        const double2 v = verts[i];
        atomicAdd(&s_cnt[vg_cy(G, v.y) * NX + vg_cx(G, v.x)], 1);
        far = far || !(fabs(v.x) <= 1e14 && fabs(v.y) <= 1e14);
    }
    if (far) *far_flag = 1;
    __syncthreads();
    // exclusive scan: each thread owns a contiguous run of cells
    const int per = (ncell + 1023) / 1024;
    const int lo = min(ncell, (int)threadIdx.x * per), hi = min(ncell, lo + per);
    int sum = 0;
    for (int i = lo; i < hi; ++i) sum += s_cnt[i];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = threadIdx.x >= o ? s_part[threadIdx.x - o] : 0;
        __syncthreads();
        s_part[threadIdx.x] += v;
        __syncthreads();
    }
    int run = s_part[threadIdx.x] - sum;
    for (int i = lo; i < hi; ++i) {
        const int c = s_cnt[i];
        G.start[i] = run;
        s_cnt[i] = run;                 // becomes the scatter cursor
        run += c;
    }
    if (threadIdx.x == 1023) G.start[ncell] = s_part[1023];
    __syncthreads();
    for (int i = threadIdx.x; i < nv; i += 1024) {
        const double2 v = verts[i];
        const int pos = atomicAdd(&s_cnt[vg_cy(G, v.y) * NX + vg_cx(G, v.x)], 1);
        G.sv[pos] = v;
        G.sid[pos] = i;
    }
}

// ---------------------------------------------------------------------------
// dominators of a rect, found by the warps of a block (lane = candidate)
// ---------------------------------------------------------------------------
struct DomTable {
    double c[5][4];         // corner distances of the 5 dominators, x (1 + eps)
};

// one 32-candidate step of a warp's dominator search: `valid` lanes hold a
// candidate v; key[j] / tab track the warp's best so far (warp-uniform keys)
__device__ __forceinline__ void dom_step(const Rect &q, bool valid, double2 v,
                                         unsigned key[5], DomTable &tab, int lane) {
    const CornerD2 c = corner_d2(q, v);
    const double dd[5] = {center_d2(q, v), c.d00, c.d01, c.d10, c.d11};
    // the five warp minima first (independent reductions in flight together),
    // then the rare table updates
    unsigned k[5], m[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) k[j] = valid ? __float_as_uint((float)dd[j]) : 0xffffffffu;
#pragma unroll
    for (int j = 0; j < 5; ++j) m[j] = __reduce_min_sync(0xffffffffu, k[j]);
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        if (m[j] < key[j]) {   // warp-uniform
            key[j] = m[j];
            const int src = __ffs(__ballot_sync(0xffffffffu, k[j] == m[j])) - 1;
            if (lane == src) {
                const double mg = 1.0 + kDomEps;
                tab.c[j][0] = c.d00 * mg; tab.c[j][1] = c.d01 * mg;
                tab.c[j][2] = c.d10 * mg; tab.c[j][3] = c.d11 * mg;
            }
        }
    }
}

// The same search for LONG candidate runs: every lane keeps the minima of its
// own candidates (no warp reduction per 32 candidates -- that reduction is two
// thirds of dom_step's instructions) and the warp reduces once at the end.
// Which of several equally near vertices becomes a dominator may differ from
// dom_step; any vertex is a valid dominator, the lists stay supersets of the
// nearest vertices and the sizes do not change.
struct DomAcc {
    unsigned key[5];
    int idx[5];
};

__device__ __forceinline__ void dom_acc_init(DomAcc &a) {
#pragma unroll
    for (int j = 0; j < 5; ++j) { a.key[j] = 0xffffffffu; a.idx[j] = 0; }
}

__device__ __forceinline__ void dom_acc_step(const Rect &q, double2 v, int i, DomAcc &a) {
    const CornerD2 c = corner_d2(q, v);
    const double dd[5] = {center_d2(q, v), c.d00, c.d01, c.d10, c.d11};
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const unsigned k = __float_as_uint((float)dd[j]);
        if (k < a.key[j]) { a.key[j] = k; a.idx[j] = i; }
    }
}

// warp-level end of a run: key[] / tab as dom_step leaves them; fetch(i) returns
// candidate i of the run again
template <class Fetch>
__device__ __forceinline__ void dom_acc_finish(const Rect &q, const DomAcc &a, unsigned key[5],
                                               DomTable &tab, int lane, Fetch fetch) {
    unsigned m[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) m[j] = __reduce_min_sync(0xffffffffu, a.key[j]);
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        if (m[j] < key[j]) {   // warp-uniform
            key[j] = m[j];
            const int src = __ffs(__ballot_sync(0xffffffffu, a.key[j] == m[j])) - 1;
            if (lane == src) {
                const CornerD2 c = corner_d2(q, fetch(a.idx[j]));
                const double mg = 1.0 + kDomEps;
                tab.c[j][0] = c.d00 * mg; tab.c[j][1] = c.d01 * mg;
                tab.c[j][2] = c.d10 * mg; tab.c[j][3] = c.d11 * mg;
            }
        }
    }
}

__device__ __forceinline__ bool dominated_tab(const DomTable &D, const CornerD2 &v) {
    bool dom = false;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const double2 a = *reinterpret_cast<const double2 *>(&D.c[j][0]);
        const double2 b = *reinterpret_cast<const double2 *>(&D.c[j][2]);
        dom = dom || (v.d00 > a.x && v.d01 > a.y && v.d10 > b.x && v.d11 > b.y);
    }
    return dom;
}

// block-level combine: every warp has (key[5], tab) in s_tab[wid]; afterwards
// s_fin holds the block's 5 dominators and s_key0 the centre key
template <int NW>
__device__ __forceinline__ void dom_combine(const unsigned key[5], unsigned (*s_key)[5],
                                            DomTable *s_tab, DomTable *s_fin,
                                            unsigned *s_key0) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane < 5) {
        unsigned k = 0;
#pragma unroll
        for (int j = 0; j < 5; ++j) if (lane == j) k = key[j];
        s_key[wid][lane] = k;
    }
    __syncthreads();
    if (threadIdx.x < 5) {
        unsigned best = 0xffffffffu;
        int bw = 0;
        for (int w = 0; w < NW; ++w)
            if (s_key[w][threadIdx.x] < best) { best = s_key[w][threadIdx.x]; bw = w; }
#pragma unroll
        for (int c = 0; c < 4; ++c) s_fin->c[threadIdx.x][c] = s_tab[bw].c[threadIdx.x][c];
        if (threadIdx.x == 0) *s_key0 = best;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// list pools
// ---------------------------------------------------------------------------
struct ListDev {
    int *off, *cnt, *ids;
    double2 *xy;            // level 1 only: the listed vertices themselves
};

struct BuildArgs {
    const PWin *wins;
    int n_wins;
    const double2 *verts;
    VGrid grid;
    ListDev l2, l1;
    unsigned long long *ctr2, *ctr1;
    unsigned long long cap2, cap1;
    int *overflow;
    int *sticky;            // status word of a pass sequence (or nullptr): a pass is a
                            // no-op when it is set, and sets it when a pool overflows
    double r, r2lim;        // search radius (+inf: unbounded) and r^2 (1 + eps)
};

// pool overflow: this pass is void (and, in a sequence, every later one)
__device__ __forceinline__ void flag_overflow(int *overflow, int *sticky) {
    *overflow = 1;
    if (sticky) *sticky = 1;
}

// survivors of a block are collected in shared memory first (one pass over the
// parent), the pool chunk is reserved with ONE atomic so a list is contiguous
template <int CAP>
struct Collect {
    int buf[CAP];
    int count, pos;
    unsigned long long base;
};

// ---------------------------------------------------------------------------
// level 2: 1024-px tiles from the vertex grid
// ---------------------------------------------------------------------------
#ifndef OCSB_ROOT_THREADS
#define OCSB_ROOT_THREADS 256
#endif
constexpr int kRootThreads = OCSB_ROOT_THREADS;
constexpr int kRootCap = 2048;

__global__ void __launch_bounds__(kRootThreads) root_kernel(BuildArgs A) {
    TL_SCOPE(2);
    constexpr int NW = kRootThreads / 32;
    __shared__ unsigned s_key[NW][5];
    __shared__ DomTable s_tab[NW];
    __shared__ DomTable s_fin;
    __shared__ unsigned s_key0;
    __shared__ Collect<kRootCap> s_col;
    if (A.sticky && *A.sticky) {
        // an earlier pass of the sequence failed: void this one (the kernels
        // after this one look at the overflow flag only)
        if (threadIdx.x == 0) *A.overflow = 1;
        return;
    }
    if (*A.overflow) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int t = blockIdx.x;
    const int wi = find_win(A.wins, A.n_wins, t, 2);
    const PWin &W = A.wins[wi];
    const int lt = t - W.base2;
    const int tx = lt % W.n2x, ty = lt / W.n2x;
    const int bw = W.ax.n, bh = W.re - W.rb;
    const int c0 = tx * kT2, r0 = ty * kT2;
    const int c1 = min(bw, c0 + kT2) - 1, r1 = min(bh, r0 + kT2) - 1;
    const Rect q = win_rect(W, c0, c1, r0, r1);
    const VGrid &G = A.grid;
    const int NX = G.ncx + 2, NY = G.ncy + 2;

    // ---- pass 1: ring search around the centre cell for the dominators -------
    const double cxm = 0.5 * (q.xlo + q.xhi), cym = 0.5 * (q.ylo + q.yhi);
    const int ccx = vg_cx(G, cxm), ccy = vg_cy(G, cym);
    const double hdx = 0.5 * (q.xhi - q.xlo), hdy = 0.5 * (q.yhi - q.ylo);
    const double halfdiag = sqrt(hdx * hdx + hdy * hdy);
    unsigned key[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) key[j] = 0xffffffffu;
    const int kmax = max(max(ccx, NX - 1 - ccx), max(ccy, NY - 1 - ccy));
    bool found = false;
    DomAcc acc;
    dom_acc_init(acc);
    for (int k = 0; k <= kmax; ++k) {
        // cells of Chebyshev ring k, as contiguous slices of the sorted array
        const int j0 = max(0, ccy - k), j1 = min(NY - 1, ccy + k);
        for (int j = j0; j <= j1; ++j) {
            const bool full = (j == ccy - k) || (j == ccy + k);
            for (int side = 0; side < (full ? 1 : 2); ++side) {
                int i0, i1;
                if (full) { i0 = max(0, ccx - k); i1 = min(NX - 1, ccx + k); }
                else {
                    i0 = i1 = side == 0 ? ccx - k : ccx + k;
                    if (i0 < 0 || i0 > NX - 1) continue;
                }
                const int lo = G.start[j * NX + i0], hi = G.start[j * NX + i1 + 1];
                for (int i = lo + (int)threadIdx.x; i < hi; i += kRootThreads)
                    dom_acc_step(q, G.sv[i], i, acc);
            }
        }
        dom_acc_finish(q, acc, key, s_tab[wid], lane, [&](int i) { return G.sv[i]; });
        __syncwarp();
        dom_combine<NW>(key, s_key, s_tab, &s_fin, &s_key0);
        const unsigned k0 = s_key0;
        __syncthreads();     // s_key / s_key0 are rewritten by the next ring
        if (k0 != 0xffffffffu) {
            found = true;
            // every unvisited vertex is at least k cells from the centre
            const double reach = (double)k * G.cell;
            if ((double)__uint_as_float(k0) <= reach * reach) break;
        }
        // bounded query: nothing within the radius of any pixel of the tile
        if (!found && (double)k * G.cell > A.r + halfdiag + G.cell) break;
    }
    if (!found) {
        if (threadIdx.x == 0) { A.l2.off[t] = 0; A.l2.cnt[t] = 0; }
        return;
    }
    // ---- pass 2: cells within min(r, dominator bound) of the rect -------------
    // every pixel of the rect is at most D from the centre dominator (max of
    // its corner distances), so a nearest vertex lies within D of the rect
    const double D2 = fmax(fmax(s_fin.c[0][0], s_fin.c[0][1]),
                           fmax(s_fin.c[0][2], s_fin.c[0][3]));
    const double R2 = fmin(A.r2lim, D2 * (1.0 + 1e-9));
    const double R = sqrt(R2) * (1.0 + 1e-9);
    const int ix0 = vg_cx(G, q.xlo - R), ix1 = vg_cx(G, q.xhi + R);
    const int iy0 = vg_cy(G, q.ylo - R), iy1 = vg_cy(G, q.yhi + R);
    if (threadIdx.x == 0) { s_col.count = 0; s_col.pos = 0; }
    __syncthreads();
    auto sweep = [&](bool write) {
        for (int j = iy0; j <= iy1; ++j) {
            const int lo = G.start[j * NX + ix0], hi = G.start[j * NX + ix1 + 1];
            for (int i = lo + (int)threadIdx.x; i < hi; i += kRootThreads) {
                const double2 v = G.sv[i];
                if (rect_d2(q, v) > R2) continue;
                if (dominated_tab(s_fin, corner_d2(q, v))) continue;
                const int id = G.sid[i];
                if (!write) {
                    const int p = atomicAdd(&s_col.count, 1);
                    if (p < kRootCap) s_col.buf[p] = id;
                } else {
                    const int p = atomicAdd(&s_col.pos, 1);
                    A.l2.ids[s_col.base + p] = id;
                    A.l2.xy[s_col.base + p] = v;
                }
            }
        }
    };
    sweep(false);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long base = 0;
        int total = s_col.count;
        if (total) {
            base = atomicAdd(A.ctr2, (unsigned long long)total);
            if (base + total > A.cap2) { flag_overflow(A.overflow, A.sticky); total = 0; }
        }
        s_col.base = base;
        s_col.count = total;
        A.l2.off[t] = (int)base;
        A.l2.cnt[t] = total;
    }
    __syncthreads();
    const int total = s_col.count;
    if (total == 0) return;
    if (total <= kRootCap) {
        for (int i = threadIdx.x; i < total; i += kRootThreads) {
            const int id = s_col.buf[i];
            A.l2.ids[s_col.base + i] = id;
            A.l2.xy[s_col.base + i] = A.verts[id];
        }
    } else {
        sweep(true);
    }
}

// ---------------------------------------------------------------------------
// level 1: 256-px tiles from their 1024-px parent list
// ---------------------------------------------------------------------------
constexpr int kL1Threads = 256;
constexpr int kL1Cap = 1024;

__global__ void __launch_bounds__(kL1Threads) l1_kernel(BuildArgs A) {
    constexpr int NW = kL1Threads / 32;
    __shared__ unsigned s_key[NW][5];
    __shared__ DomTable s_tab[NW];
    __shared__ DomTable s_fin;
    __shared__ unsigned s_key0;
    __shared__ Collect<kL1Cap> s_col;
    if (*A.overflow) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int t = blockIdx.x;
    const int wi = find_win(A.wins, A.n_wins, t, 1);
    const PWin &W = A.wins[wi];
    const int lt = t - W.base1;
    const int tx = lt % W.n1x, ty = lt / W.n1x;
    const int pt = W.base2 + (ty >> 2) * W.n2x + (tx >> 2);
    const int pcnt = A.l2.cnt[pt], poff = A.l2.off[pt];
    if (pcnt == 0) {
        if (threadIdx.x == 0) { A.l1.off[t] = 0; A.l1.cnt[t] = 0; }
        return;
    }
    const int bw = W.ax.n, bh = W.re - W.rb;
    const int c0 = tx * kT1, r0 = ty * kT1;
    const int c1 = min(bw, c0 + kT1) - 1, r1 = min(bh, r0 + kT1) - 1;
    const Rect q = win_rect(W, c0, c1, r0, r1);
    unsigned key[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) key[j] = 0xffffffffu;
    for (int b0 = wid * 32; b0 < pcnt; b0 += kL1Threads) {
        const int i = b0 + lane;
        const bool valid = i < pcnt;
        double2 v = make_double2(0.0, 0.0);
        if (valid) v = A.l2.xy[poff + i];
        dom_step(q, valid, v, key, s_tab[wid], lane);
    }
    __syncwarp();
    dom_combine<NW>(key, s_key, s_tab, &s_fin, &s_key0);
    if (threadIdx.x == 0) { s_col.count = 0; s_col.pos = 0; }
    __syncthreads();
    auto sweep = [&](bool write) {
        for (int i = threadIdx.x; i < pcnt; i += kL1Threads) {
            const double2 v = A.l2.xy[poff + i];
            if (rect_d2(q, v) > A.r2lim) continue;
            if (dominated_tab(s_fin, corner_d2(q, v))) continue;
            const int id = A.l2.ids[poff + i];
            if (!write) {
                const int p = atomicAdd(&s_col.count, 1);
                if (p < kL1Cap) s_col.buf[p] = id;
            } else {
                const int p = atomicAdd(&s_col.pos, 1);
                A.l1.ids[s_col.base + p] = id;
                A.l1.xy[s_col.base + p] = v;
            }
        }
    };
    sweep(false);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long base = 0;
        int total = s_col.count;
        if (total) {
            base = atomicAdd(A.ctr1, (unsigned long long)total);
            if (base + total > A.cap1) { flag_overflow(A.overflow, A.sticky); total = 0; }
        }
        s_col.base = base;
        s_col.count = total;
        A.l1.off[t] = (int)base;
        A.l1.cnt[t] = total;
    }
    __syncthreads();
    const int total = s_col.count;
    if (total == 0) return;
    if (total <= kL1Cap) {
        for (int i = threadIdx.x; i < total; i += kL1Threads) {
            const int id = s_col.buf[i];
            A.l1.ids[s_col.base + i] = id;
            A.l1.xy[s_col.base + i] = A.verts[id];
        }
    } else {
        sweep(true);
    }
}

// ---------------------------------------------------------------------------
// the tile kernel
// ---------------------------------------------------------------------------
struct FeatDev {
    const double2 *verts;
    const int *off1, *cnt1, *ids1;
    double rt, target;      // rate * target, target
    double r2, r2lim;       // r^2 (+inf unbounded; -1 when r <= 0) and r^2 (1 + eps)
    float v_far_f;          // float32(fl(fl(rt * r) + target)), +inf unbounded
    int fuse;
    unsigned hi_lo;         // fast size law needs hi(d2) >= hi_lo; 0xffffffff: off
};

struct TileArgs {
    const PWin *wins;
    int n_wins;
    const float *in;        // nullptr: every pixel = finfo(float32).max (A4)
    float *out;
    long long in_stride, out_stride;
    int nf, nl;
    FeatDev feat[kMaxFeat];
    LimiterDev<float> lim[kMaxLim];
    int lim_band[kMaxLim];
    float lo, hi;           // global [hmin, hmax] as float32, -/+inf when None
    const float *dem;       // element (0, 0) of the stored DEM block
    long long dem_stride;
    int dem_rows, dem_cols;
    int use_tma;
    const int *overflow;    // int2 {pool overflow, a vertex beyond 1e14}
    const int *sticky;      // status word of a pass sequence (or nullptr)
};

struct WarpScr {
    DomTable dom;           // 160 B
    double2 sv[kSurv];      // survivors of the warp's rect
};

__device__ __forceinline__ float min3_nan(float a, float b, float c) {
    float r;
    asm("min.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float max3_nan(float a, float b, float c) {
    float r;
    asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ double rsqrt_approx_f64(double x) {
    double r;   // MUFU.RSQ64H: ~20-bit seed from the high word
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    return r;
}

// float32(fl(fl(rt * sqrt(d2)) + target)) for 8 squared distances (all f64 in
// the reference).  Fast path: sqrt to 2^-39.4 relative (MUFU.RSQ64H seed + one
// Newton step), accepted only if the 29 mantissa bits that the float32
// rounding drops are not within 2^15 ulp64 of the rounding tie -- then
// float32(v64) is provably the reference's float32 (see k_feature_eval.cuh).
// The test is accumulated over the lane's 8 pixels (one compare per lane, not
// per pixel); a lane that fails recomputes its 8 pixels with sqrt.rn.  Beyond
// the radius the law is monotone, so min(., v_far) needs no domain test; d2
// below 1e-30 (a pixel ON a vertex) takes the exact path through the same flag.
__device__ __forceinline__ void size_law8(const double (&best)[8], const FeatDev &F,
                                          unsigned hi_lo, float (&vf)[8]) {
    unsigned tmin = 0xffffffffu, hmin = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const double d2 = best[k];
        const double y = rsqrt_approx_f64(d2);
        const double g = d2 * y;
        const double e = fma(-g, g, d2);
        const double d = fma(e * y, 0.5, g);
        const double v64 = fma(F.rt, d, F.target);
        const unsigned t = ((unsigned)__double2loint(v64) << 3) - (0x80000000u - (1u << 18));
        tmin = min(tmin, t);
        hmin = min(hmin, (unsigned)__double2hiint(d2));
        vf[k] = (float)v64;
    }
    if (tmin < (1u << 19) || hmin < hi_lo) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const double d2 = best[k];
            float v = F.v_far_f;
            if (d2 < F.r2)
                v = (float)__dadd_rn(__dmul_rn(F.rt, __dsqrt_rn(d2)), F.target);
            vf[k] = v;
        }
    }
}

// ---- packed float32 pairs ---------------------------------------------------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void upk2(f32x2 v, float &a, float &b) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// NOTE (measured, tools/ubench/ubench2.cu): ptxas 12.9 contracts
// mul.rn.f32x2 + add.rn.f32x2 into FFMA2 although both carry .rn, so a packed
// product must never feed a packed add where numpy rounds twice (g0^2 + g1^2):
// those sums are scalar adds of the unpacked halves.

// One row of 4 interior pixels of the limiter law, packed (see header): returns
// the raw h = |(z / sqrt(g0^2 + g1^2)) / 3| and whether the group is inside
// the range where the unguarded sequences are IEEE-exact.
__device__ __forceinline__ bool limiter_raw4(const float4 up, const float4 zc,
                                             const float4 dn, float west, float east,
                                             const LimiterDev<float> &p, float (&h)[4]) {
    // axis 0 (rows): (dn - up) / (2 dx); axis 1 (cols): (zr - zl) / (2 dy)
    const f32x2 a0_01 = sub2(pk2(dn.x, dn.y), pk2(up.x, up.y));
    const f32x2 a0_23 = sub2(pk2(dn.z, dn.w), pk2(up.z, up.w));
    const f32x2 a1_12 = sub2(pk2(zc.z, zc.w), pk2(zc.x, zc.y));     // pixels 1, 2
    const f32x2 a1_03 = pk2(__fsub_rn(zc.y, west), __fsub_rn(east, zc.z));   // 0, 3
    const f32x2 r0 = pk2(p.d2_0.r, p.d2_0.r), nc0 = pk2(-p.d2_0.c, -p.d2_0.c);
    const f32x2 r1 = pk2(p.d2_1.r, p.d2_1.r), nc1 = pk2(-p.d2_1.c, -p.d2_1.c);
    // Markstein a / c:  q0 = a r;  g = fma(fma(q0, -c, a), r, q0)
    auto cdiv2 = [](f32x2 a, f32x2 r, f32x2 nc) {
        const f32x2 q0 = mul2(a, r);
        return fma2(fma2(q0, nc, a), r, q0);
    };
    const f32x2 g0_01 = cdiv2(a0_01, r0, nc0), g0_23 = cdiv2(a0_23, r0, nc0);
    const f32x2 g1_12 = cdiv2(a1_12, r1, nc1), g1_03 = cdiv2(a1_03, r1, nc1);
    float p0[4], p1[4];
    upk2(mul2(g0_01, g0_01), p0[0], p0[1]);
    upk2(mul2(g0_23, g0_23), p0[2], p0[3]);
    upk2(mul2(g1_12, g1_12), p1[1], p1[2]);
    upk2(mul2(g1_03, g1_03), p1[0], p1[3]);
    float s[4], rs[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        s[k] = __fadd_rn(p0[k], p1[k]);     // scalar: rounded twice like numpy
        rs[k] = mufu_rsq(s[k]);
    }
    const f32x2 half = pk2(0.5f, 0.5f), one = pk2(1.0f, 1.0f);
    const f32x2 y3 = pk2(1.0f / 3.0f, 1.0f / 3.0f), m3 = pk2(-3.0f, -3.0f);
    float q[4];
    float rc[4];
    f32x2 ndh[2], z2[2];
    z2[0] = pk2(zc.x, zc.y); z2[1] = pk2(zc.z, zc.w);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        // sqrt.rn fast sequence on the NEGATED chain (no packed negation
        // exists): ns = -s, G = ns rs = -g, e' = G^2 - s, -dh = G + e' (rs / 2)
        const f32x2 ns = pk2(-s[2 * i], -s[2 * i + 1]);
        const f32x2 rs2 = pk2(rs[2 * i], rs[2 * i + 1]);
        const f32x2 G = mul2(ns, rs2);
        const f32x2 hh = mul2(rs2, half);
        const f32x2 e = fma2(G, G, ns);
        ndh[i] = fma2(e, hh, G);
        float a, b;
        upk2(ndh[i], a, b);
        rc[2 * i] = mufu_rcp(-a);           // rcp(dh)
        rc[2 * i + 1] = mufu_rcp(-b);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        // div.rn fast sequence z / dh with -dh at hand
        const f32x2 rc2 = pk2(rc[2 * i], rc[2 * i + 1]);
        const f32x2 t = fma2(rc2, ndh[i], one);
        const f32x2 r1_ = fma2(rc2, t, rc2);
        const f32x2 qq = mul2(z2[i], r1_);
        const f32x2 rem = fma2(qq, ndh[i], z2[i]);
        const f32x2 qv = fma2(r1_, rem, qq);
        upk2(qv, q[2 * i], q[2 * i + 1]);
        // |q / 3| == float32(|(1/3) * (double)q|) for every finite q (exhaustive
        // check, tools/ubench/ubench2.cu): Markstein with y = RN(1/3)
        const f32x2 h0 = mul2(qv, y3);
        const f32x2 hv = fma2(fma2(h0, m3, qv), y3, h0);
        float a, b;
        upk2(hv, a, b);
        h[2 * i] = fabsf(a);
        h[2 * i + 1] = fabsf(b);
    }
    // NaN-propagating min / max: a NaN anywhere fails every compare below
    const float smax = max_nan_f32(max3_nan(s[0], s[1], s[2]), s[3]);
    const float smin = min_nan_f32(min3_nan(s[0], s[1], s[2]), s[3]);
    const float qmax = max_nan_f32(max3_nan(fabsf(q[0]), fabsf(q[1]), fabsf(q[2])), fabsf(q[3]));
    const float qmin = min_nan_f32(min3_nan(fabsf(q[0]), fabsf(q[1]), fabsf(q[2])), fabsf(q[3]));
    return smin >= 1e-30f && smax <= 1e30f && qmin >= 1e-15f && qmax <= 1e15f;
}

// limiter tails of one row of 4 pixels for one op (BAND / FUSE are uniform)
template <bool BAND, bool FUSE>
__device__ __forceinline__ void tails4(const float (&h)[4], const float (&zz)[4],
                                       float *o4, const LimiterDev<float> &p) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
        o4[k] = limiter_tail<BAND, FUSE, float>(h[k], zz[k], o4[k], p);
}
__device__ __forceinline__ void tails4_dyn(const float (&h)[4], const float (&zz)[4],
                                           float *o4, const LimiterDev<float> &p,
                                           int band) {
    if (p.fuse_clamp) {
        if (band) tails4<true, true>(h, zz, o4, p); else tails4<false, true>(h, zz, o4, p);
    } else {
        if (band) tails4<true, false>(h, zz, o4, p); else tails4<false, false>(h, zz, o4, p);
    }
}

// four staged DEM values of a row; `shift` (dem_c & 3, uniform) != 0 when the
// window does not start on a 16-byte boundary of the DEM rows
__device__ __forceinline__ float4 dem_row4(const float *p, int shift) {
    if (shift == 0) return *reinterpret_cast<const float4 *>(p);
    return make_float4(p[0], p[1], p[2], p[3]);
}

// Pixel ownership: warp = 32x16 rect of the tile (2 x 4 warps), lane = 4 columns
// x 4 rows of it, processed two rows ("half") at a time so that only 8 running
// size values live in registers.  Per feature op: CTA filter, warp filter, then
// per half  old -> nearest vertex -> size law -> (last op: limiter ops) -> store.
// With several feature ops the intermediate raster goes through `out` (the
// thread re-reads its own pixels: L1/L2 hits).
template <bool HAS_OLD>
__global__ void __launch_bounds__(kPassThreads, OCSB_PASS_MIN_CTAS)
tile_kernel(const __grid_constant__ CUtensorMap tmap, const TileArgs A) {
    __shared__ __align__(128) float s_dem[kDemRows * kDemPitch];
    __shared__ __align__(16) double s_x[kT0];
    __shared__ __align__(16) double s_y[kT0];
    __shared__ __align__(16) double2 s_v[kListCap];
    __shared__ __align__(16) WarpScr s_w[kPassThreads / 32];
    __shared__ __align__(16) PWin s_win;
    __shared__ unsigned s_key[kPassThreads / 32][5];
    __shared__ DomTable s_fin;
    __shared__ int s_n;
    __shared__ __align__(8) unsigned long long s_bar;

    PSTAT_CLK(st0);
    const int2 flags = *reinterpret_cast<const int2 *>(A.overflow);   // {overflow, far}
    if (flags.x || (A.sticky && *A.sticky)) return;
    const bool far = flags.y != 0;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int wi = 0, tx, ty;
    if (A.n_wins == 1) {
        tx = blockIdx.x; ty = blockIdx.y;
    } else {
        const int t = blockIdx.x;
        wi = find_win(A.wins, A.n_wins, t, 0);
        const int lt = t - A.wins[wi].base0, n0x = A.wins[wi].n0x;
        ty = lt / n0x; tx = lt - ty * n0x;
    }
    const PWin &G = A.wins[wi];         // global copy: prologue only
    const int c0t = tx * kT0, r0t = ty * kT0;

    // ---- DEM tile: one TMA box (cols c0t-4 .. c0t+71, rows r0t-1 .. r0t+64) --
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if (A.nl && A.use_tma && threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_dem);
        // the box starts on a 16-byte boundary of the DEM rows (TMA requirement)
        const int cx = (G.dem_c + c0t - 4) & ~3, cy = G.dem_r + G.rb + r0t - 1;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                     ::"r"(bar), "r"((uint32_t)(kDemRows * kDemPitch * 4)) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
            " [%0], [%1, {%2, %3}], [%4];"
            ::"r"(dst), "l"(&tmap), "r"(cx), "r"(cy), "r"(bar) : "memory");
    }
    // ---- window record + coordinates of the tile's columns / rows -------------
    if (threadIdx.x < kT0) {
        s_x[threadIdx.x] = lin_at(G.ax, min(c0t + (int)threadIdx.x, G.ax.n - 1));
    } else if (threadIdx.x < 2 * kT0) {
        const int j = threadIdx.x - kT0;
        s_y[j] = lin_at(G.ay, G.rb + min(r0t + j, G.re - G.rb - 1));
    } else if (threadIdx.x < 2 * kT0 + (int)(sizeof(PWin) / 4)) {
        const int i = threadIdx.x - 2 * kT0;
        reinterpret_cast<int *>(&s_win)[i] = reinterpret_cast<const int *>(&G)[i];
    }
    if (A.nl && !A.use_tma) {
        // unaligned DEM buffers: plain loads of the same box, zeros outside
        const int gr0 = G.dem_r + G.rb + r0t - 1, gc0 = (G.dem_c + c0t - 4) & ~3;
        for (int i = threadIdx.x; i < kDemRows * kDemPitch; i += kPassThreads) {
            const int rr = i / kDemPitch, cc = i - rr * kDemPitch;
            const int gr = gr0 + rr, gc = gc0 + cc;
            float v = 0.0f;
            if (gr >= 0 && gr < A.dem_rows && gc >= 0 && gc < A.dem_cols)
                v = A.dem[(long long)gr * A.dem_stride + gc];
            s_dem[i] = v;
        }
    }
    __syncthreads();
    PSTAT_CLK(st1);
    const PWin &W = s_win;
    const int bw = W.ax.n, rb = W.rb, bh = W.re - rb;

    // ---- this warp's 32x16 rect and this lane's 4x4 pixels ---------------------
    const int lcol = (wid & 1) * 32 + (lane & 7) * 4;      // tile-local
    const int lrow = (wid >> 1) * 16 + (lane >> 3) * 4;
    const int sc0 = c0t + (wid & 1) * 32, sr0 = r0t + (wid >> 1) * 16;
    const bool sub_valid = sc0 < bw && sr0 < bh;
    const int sc1 = min(bw, sc0 + 32) - 1, sr1 = min(bh, sr0 + 16) - 1;
    const int gcol = c0t + lcol, grow = r0t + lrow;        // band-relative rows
    const int wc1 = min(sc1, W.keep_w - 1);                // last written column
    const int wr1 = min(sr1, W.keep_h - 1 - rb);           // last written row
    const bool vec_r = W.vec && gcol + 3 <= sc1;
    const bool vec_w = W.vec && gcol + 3 <= wc1;
    float *const out0 = A.out + W.hoff + (long long)(rb + grow) * A.out_stride + gcol;
    const float *const in0 =
        HAS_OLD ? A.in + W.hoff + (long long)(rb + grow) * A.in_stride + gcol : nullptr;

    WarpScr &WS = s_w[wid];
    const int nf = A.nf;
    const int nit = nf > 0 ? nf : 1;
    for (int f = 0; f < nit; ++f) {
        const bool last = f == nit - 1;
        const FeatDev &F = A.feat[f];
        int ntile = 0, poff = 0, ns = 0;
        bool in_smem = true, refilter = false;
        PSTAT_CLK(st2);
        if (nf > 0) {
            const int t1 = W.base1 + (ty >> 2) * W.n1x + (tx >> 2);
            const int pcnt = F.cnt1[t1];
            poff = F.off1[t1];
            // ---- CTA: parent list -> tile list in shared memory ----------------------
            if (pcnt > 0) {
                Rect qT;
                {
                    const int tc1 = min(bw, c0t + kT0) - 1 - c0t;
                    const int tr1 = min(bh, r0t + kT0) - 1 - r0t;
                    const double xa = s_x[0], xb = s_x[tc1], ya = s_y[0], yb = s_y[tr1];
                    qT.xlo = xa < xb ? xa : xb; qT.xhi = xa < xb ? xb : xa;
                    qT.ylo = ya < yb ? ya : yb; qT.yhi = ya < yb ? yb : ya;
                }
                unsigned key[5];
#pragma unroll
                for (int j = 0; j < 5; ++j) key[j] = 0xffffffffu;
                for (int b0 = wid * 32; b0 < pcnt; b0 += kPassThreads) {
                    const int i = b0 + lane;
                    const bool valid = i < pcnt;
                    double2 v = make_double2(0.0, 0.0);
                    if (valid) v = F.verts[F.ids1[poff + i]];
                    dom_step(qT, valid, v, key, WS.dom, lane);
                }
                __syncwarp();
                if (threadIdx.x == 0) s_n = 0;
                // combine the warps' dominators (their tables live in s_w[].dom)
                if (lane < 5) {
                    unsigned k = 0;
#pragma unroll
                    for (int j = 0; j < 5; ++j) if (lane == j) k = key[j];
                    s_key[wid][lane] = k;
                }
                __syncthreads();
                if (threadIdx.x < 5) {
                    unsigned best = 0xffffffffu;
                    int bwp = 0;
                    for (int w = 0; w < kPassThreads / 32; ++w)
                        if (s_key[w][threadIdx.x] < best) { best = s_key[w][threadIdx.x]; bwp = w; }
#pragma unroll
                    for (int c = 0; c < 4; ++c)
                        s_fin.c[threadIdx.x][c] = s_w[bwp].dom.c[threadIdx.x][c];
                }
                __syncthreads();
                for (int b0 = wid * 32; b0 < pcnt; b0 += kPassThreads) {
                    const int i = b0 + lane;
                    const bool valid = i < pcnt;
                    double2 v = make_double2(0.0, 0.0);
                    if (valid) v = F.verts[F.ids1[poff + i]];
                    const bool keep = valid && !(rect_d2(qT, v) > F.r2lim) &&
                                      !dominated_tab(s_fin, corner_d2(qT, v));
                    const unsigned m = __ballot_sync(0xffffffffu, keep);
                    int base = 0;
                    if (lane == 0 && m) base = atomicAdd(&s_n, __popc(m));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    const int pos = base + __popc(m & ((1u << lane) - 1u));
                    if (keep && pos < kListCap) s_v[pos] = v;
                }
                __syncthreads();
                ntile = s_n;
                if (ntile > kListCap) { in_smem = false; ntile = pcnt; }
            }
            if (threadIdx.x == 0) {
                PSTAT_ADD(0, 1); PSTAT_ADD(1, pcnt); PSTAT_ADD(2, ntile);
                PSTAT_ADD(3, pcnt > 32); PSTAT_ADD(4, ntile > 32);
            }
        }
        PSTAT_CLK(st3);
        auto cand = [&](int i) -> double2 {
            return in_smem ? s_v[i] : F.verts[F.ids1[poff + i]];
        };
        // the warp's rect (recomputed where needed: not live across the evaluation)
        auto sub_rect = [&]() -> Rect {
            Rect q;
            const double xa = s_x[sc0 - c0t], xb = s_x[sc1 - c0t];
            const double ya = s_y[sr0 - r0t], yb = s_y[sr1 - r0t];
            q.xlo = xa < xb ? xa : xb; q.xhi = xa < xb ? xb : xa;
            q.ylo = ya < yb ? ya : yb; q.yhi = ya < yb ? yb : ya;
            return q;
        };
        auto filter_chunk = [&](const Rect &q, int b0, int base) -> int {
            const int i = b0 + lane;
            const bool valid = i < ntile;
            double2 v = make_double2(0.0, 0.0);
            if (valid) v = cand(i);
            const bool keep = valid && !(rect_d2(q, v) > F.r2lim) &&
                              !dominated_tab(WS.dom, corner_d2(q, v));
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            const int n = __popc(m);
            if (keep && base + n <= kSurv)
                WS.sv[base + __popc(m & ((1u << lane) - 1u))] = v;
            return n;
        };

        if (sub_valid) {
            // ---- warp: tile list -> survivors of the 32x16 rect -----------------------
            if (ntile > 0) {
                const Rect q = sub_rect();
                unsigned key[5];
#pragma unroll
                for (int j = 0; j < 5; ++j) key[j] = 0xffffffffu;
                for (int b0 = 0; b0 < ntile; b0 += 32) {
                    const int i = b0 + lane;
                    const bool valid = i < ntile;
                    double2 v = make_double2(0.0, 0.0);
                    if (valid) v = cand(i);
                    dom_step(q, valid, v, key, WS.dom, lane);
                }
                __syncwarp();
                for (int b0 = 0; b0 < ntile; b0 += 32) {
                    const int n = filter_chunk(q, b0, ns);
                    if (ns + n > kSurv) { refilter = true; break; }
                    ns += n;
                }
                __syncwarp();
            }
            PSTAT_CLK(st4);
            if (lane == 0) {
                PSTAT_ADD(5, 1); PSTAT_ADD(6, ns); PSTAT_ADD(7, refilter);
                PSTAT_ADD(8, ns == 0); PSTAT_ADD(9, ns == 1); PSTAT_ADD(10, ns == 2);
                PSTAT_ADD(11, ns >= 3 && ns < 5); PSTAT_ADD(12, ns >= 5);
                PSTAT_ADD(16, st1 - st0); PSTAT_ADD(17, st3 - st2); PSTAT_ADD(18, st4 - st3);
            }
            if (last && A.nl && A.use_tma) {
                // the DEM box has had the whole list stage to land
                uint32_t done = 0;
                while (!done) {
                    asm volatile(
                        "{\n\t.reg .pred p;\n\t"
                        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                        "selp.u32 %0, 1, 0, p;\n\t}"
                        : "=r"(done) : "r"(bar), "r"(0u) : "memory");
                }
            }

            // ---- lane: 4x4 pixels, two rows at a time ----------------------------------
            PSTAT_CLK(st5);
            if (lane == 0) PSTAT_ADD(19, st5 - st4);
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
                const int hr = grow + half * 2;            // first row of the pair
                // (not with the chunk-wise path: its warp-wide votes need every lane)
                if (!refilter && (hr > sr1 || gcol > sc1)) break;
                PSTAT_CLK(sh0);
                // ---- feature op: nearest vertex + size law -------------------------------
                float vf[8];
                if (nf > 0) {
                    if (ntile == 0) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) vf[k] = F.v_far_f;
                    } else {
                        double px[4];
                        {
                            const double2 a = *reinterpret_cast<const double2 *>(&s_x[lcol]);
                            const double2 b = *reinterpret_cast<const double2 *>(&s_x[lcol + 2]);
                            px[0] = a.x; px[1] = a.y; px[2] = b.x; px[3] = b.y;
                        }
                        const double2 py =
                            *reinterpret_cast<const double2 *>(&s_y[lrow + half * 2]);
                        double best[8];
                        auto eval_one = [&](const double2 sv, bool first) {
                            double dx2[4];
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const double d = __dsub_rn(px[k], sv.x);
                                dx2[k] = __dmul_rn(d, d);
                            }
                            const double e0 = __dsub_rn(py.x, sv.y), e1 = __dsub_rn(py.y, sv.y);
                            const double dy2[2] = {__dmul_rn(e0, e0), __dmul_rn(e1, e1)};
#pragma unroll
                            for (int j = 0; j < 2; ++j)
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const double d2 = __dadd_rn(dx2[k], dy2[j]);
                                    // finite coordinates: compare + select, not fmin
                                    best[j * 4 + k] = (first || d2 < best[j * 4 + k])
                                                          ? d2 : best[j * 4 + k];
                                }
                        };
                        if (!refilter) {
                            if (ns > 0) eval_one(WS.sv[0], true);
                            else {
#pragma unroll
                                for (int k = 0; k < 8; ++k) best[k] = kBig;
                            }
                            for (int s = 1; s < ns; ++s) eval_one(WS.sv[s], false);
                        } else {
                            // more than kSurv survivors: filter and evaluate chunk-wise
#pragma unroll
                            for (int k = 0; k < 8; ++k) best[k] = kBig;
                            const Rect q = sub_rect();
                            for (int b0 = 0; b0 < ntile; b0 += 32) {
                                const int n = filter_chunk(q, b0, 0);
                                __syncwarp();
                                for (int s = 0; s < n; ++s) eval_one(WS.sv[s], false);
                                __syncwarp();
                            }
                        }
                        size_law8(best, F, far ? 0xffffffffu : F.hi_lo, vf);
                    }
                }
                PSTAT_CLK(sh1);
                // ---- pre-op values (loaded late: not live during the search) ------------
                float o[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = 3.402823466e+38f;
                if (HAS_OLD || f > 0) {
                    const float *src = (f > 0)
                        ? out0 + (long long)(half * 2) * A.out_stride
                        : in0 + (long long)(half * 2) * A.in_stride;
                    const long long st = (f > 0) ? A.out_stride : A.in_stride;
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        if (hr + j > sr1) continue;
                        const float *orow = src + j * st;
                        if (vec_r) {
                            const float4 v4 = __ldcs(reinterpret_cast<const float4 *>(orow));
                            o[j * 4 + 0] = v4.x; o[j * 4 + 1] = v4.y;
                            o[j * 4 + 2] = v4.z; o[j * 4 + 3] = v4.w;
                        } else {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (gcol + k <= sc1) o[j * 4 + k] = orow[k];
                        }
                    }
                }
                if (nf > 0) {
                    // reference: clamp(vf); min(old, .); [A10: clamp again]; the clamp is
                    // monotone and idempotent, so fused clamp(min(old, vf)) is the same
                    if (F.fuse) {
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            o[k] = clamp_nan(min3_nan(o[k], vf[k], F.v_far_f), A.lo, A.hi);
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            o[k] = min_nan_f32(o[k], clamp_nan(fminf(vf[k], F.v_far_f), A.lo, A.hi));
                    }
                }
                PSTAT_CLK(sh2);
                // ---- limiter ops ---------------------------------------------------------
                if (last && A.nl) {
                    const int H = W.ay.n;
                    // staged row of tile row j is j + 1, staged column of tile column c is c + 4
                    const int shift = W.dem_c & 3;
                    const float *sd = s_dem + (lrow + half * 2) * kDemPitch + 4 + lcol + shift;
                    float4 rowv[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) rowv[j] = dem_row4(sd + j * kDemPitch, shift);
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        if (hr + j > sr1) continue;
                        const int wr = rb + hr + j;        // window row
                        const float4 up = rowv[j], zc = rowv[j + 1], dn = rowv[j + 2];
                        const float west = sd[(j + 1) * kDemPitch - 1];
                        const float east = sd[(j + 1) * kDemPitch + 4];
                        const bool interior = wr >= 1 && wr <= H - 2 && gcol >= 1 &&
                                              gcol + 4 <= bw - 1;
                        float h[4];
                        bool done = false;
                        if (interior) done = limiter_raw4(up, zc, dn, west, east, A.lim[0], h);
                        const float zz[4] = {zc.x, zc.y, zc.z, zc.w};
                        if (!done) {
                            // generic IEEE path: window borders, zero / subnormal / huge
                            // gradients, NaN (bit-identical, just slower)
                            const bool top = (wr == 0), bot = (wr == H - 1);
                            const float uu[4] = {up.x, up.y, up.z, up.w};
                            const float dd[4] = {dn.x, dn.y, dn.z, dn.w};
                            const LimiterDev<float> &p = A.lim[0];
                            const ConstDiv<float> d0 = (top || bot) ? p.d1_0 : p.d2_0;
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const int c = gcol + k;
                                const float u = top ? zz[k] : uu[k];
                                const float d = bot ? zz[k] : dd[k];
                                const float g0 = cdiv(__fsub_rn(d, u), d0);
                                const float zl0 = (k == 0) ? west : zz[k > 0 ? k - 1 : 0];
                                const float zr0 = (k == 3) ? east : zz[k < 3 ? k + 1 : 3];
                                const bool lft = (c == 0), rgt = (c >= bw - 1);
                                const float zl = lft ? zz[k] : zl0;
                                const float zr = rgt ? zz[k] : zr0;
                                const ConstDiv<float> d1 = (lft || rgt) ? p.d1_1 : p.d2_1;
                                const float g1 = cdiv(__fsub_rn(zr, zl), d1);
                                const float dh = __fsqrt_rn(
                                    __fadd_rn(__fmul_rn(g0, g0), __fmul_rn(g1, g1)));
                                const float qv = __fdiv_rn(zz[k], dh);
                                const float hv = (float)fabs(__dmul_rn(1.0 / 3.0, (double)qv));
                                h[k] = (dh == 0.0f) ? (float)(1.0 / 3.0) : hv;
                            }
                        }
                        if (A.nl == 1) {
                            tails4_dyn(h, zz, o + j * 4, A.lim[0], A.lim_band[0]);
                        } else {
                            for (int l = 0; l < A.nl; ++l)
                                tails4_dyn(h, zz, o + j * 4, A.lim[l], A.lim_band[l]);
                        }
                    }
                }
                PSTAT_CLK(sh3);
                // ---- store -----------------------------------------------------------------
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    if (hr + j > wr1 || gcol > wc1) continue;
                    float *nrow = out0 + (long long)(half * 2 + j) * A.out_stride;
                    if (vec_w) {
                        __stcs(reinterpret_cast<float4 *>(nrow),
                               make_float4(o[j * 4 + 0], o[j * 4 + 1], o[j * 4 + 2], o[j * 4 + 3]));
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (gcol + k <= wc1) nrow[k] = o[j * 4 + k];
                    }
                }
                PSTAT_CLK(sh4);
                if (lane == 0) {
                    PSTAT_ADD(20, sh1 - sh0); PSTAT_ADD(21, sh2 - sh1);
                    PSTAT_ADD(22, sh3 - sh2); PSTAT_ADD(23, sh4 - sh3);
                }
            }
        } else if (last && A.nl && A.use_tma) {
            // a CTA must not retire with its bulk copy in flight
            uint32_t done = 0;
            while (!done) {
                asm volatile(
                    "{\n\t.reg .pred p;\n\t"
                    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                    "selp.u32 %0, 1, 0, p;\n\t}"
                    : "=r"(done) : "r"(bar), "r"(0u) : "memory");
            }
        }
        if (!last) __syncthreads();   // s_v / s_w are reused by the next op
#ifdef OCSB_PASS_STATS
        if (last && lane == 0) {
            const long long dur = clock64() - st0;
            PSTAT_ADD(ntile > 0 ? 24 : 25, dur);
            PSTAT_ADD(ntile > 0 ? 26 : 27, 1);
        }
#endif
    }
}

#include "k_pass2.cuh"

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t,
                                  void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) !=
                cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// one instantiation of the persistent kernel: resident grid, opt-in shared memory
// `reserve` > 0: that many of the resident CTA slots are left free (the limiter
// sweep of a split pass leaves room for the pyramid's launches); LIST:
// P.active / P.n_active
template <bool HAS_OLD, bool HAS_FEAT, bool HAS_LIM, bool BAND, bool LIST>
int32_t launch_pass2_t(ocsb_ctx *ctx, const CUtensorMap &tmap, P2Args P, int reserve) {
    auto kern = pass2_kernel<HAS_OLD, HAS_FEAT, HAS_LIM, BAND, LIST>;
    static int resident[64] = {};     // per device: CTAs that fit the whole GPU
    int dev = 0;
    OCSB_CUDA(cudaGetDevice(&dev));
    OCSB_REQUIRE(dev >= 0 && dev < 64, "launch_pass2: device index out of range");
    if (resident[dev] == 0) {
        OCSB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(P2Smem)));
        int n = 0, sms = 0;
        OCSB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, kP2Threads,
                                                                sizeof(P2Smem)));
        OCSB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        OCSB_REQUIRE(n > 0 && sms > 0, "pass2_kernel does not fit an SM");
        resident[dev] = n * sms;
    }
    int res = resident[dev];
    if (reserve > 0) res = std::max(ctx->num_sms, res - reserve);
    // chunk: up to kChunkMax tiles per counter fetch, fewer when that would leave a CTA with
    // less than ~8 fetches (a band of a multi-GPU run): the tail of a launch is
    // one chunk long.  A list's length is only known on the device: a quarter
    // of the tiles is assumed
    int chunk = (int)((LIST ? P.n_tiles / 4 : P.n_tiles) / ((long long)res * 8));
    chunk = std::max(1, std::min(kChunkMax, chunk));
    static const int chunk_env = [] {
        const char *e = getenv("OCSB_PASS_CHUNK");      // tuning override
        return e ? atoi(e) : 0;
    }();
    if (chunk_env > 0) chunk = std::min(kChunkMax, chunk_env);
    P.chunk = chunk;
    const int grid = (int)std::min<long long>((P.n_tiles + chunk - 1) / chunk, (long long)res);
    kern<<<grid, kP2Threads, sizeof(P2Smem), ctx->stream>>>(tmap, P);
    return OCSB_OK;
}

int32_t launch_pass2(ocsb_ctx *ctx, const CUtensorMap &tmap, const P2Args &P, bool old,
                     bool feat, bool lim, bool band, int reserve = 0) {
    if (P.active) {
        // the feature half of a split pass: in place, tiles from the list
        OCSB_REQUIRE(old && feat && !lim, "launch_pass2: bad list variant");
        return launch_pass2_t<true, true, false, false, true>(ctx, tmap, P, reserve);
    }
#define P2_CASE(O, F, L, B)                                                        \
    if (old == O && feat == F && lim == L && band == B)                            \
        return launch_pass2_t<O, F, L, B, false>(ctx, tmap, P, reserve);
    P2_CASE(false, true, true, false)
    P2_CASE(false, true, true, true)
    P2_CASE(true, true, true, false)
    P2_CASE(true, true, true, true)
    P2_CASE(false, true, false, false)
    P2_CASE(true, true, false, false)
    P2_CASE(false, false, true, false)
    P2_CASE(false, false, true, true)
    P2_CASE(true, false, true, false)
    P2_CASE(true, false, true, true)
#undef P2_CASE
    ocsb_set_error("launch_pass2: no such variant");
    return OCSB_EINVAL;
}

}  // namespace

bool k_pass_supported(const PassDesc &d) {
    if (d.nf > kMaxFeat || d.nl > kMaxLim || d.nf + d.nl == 0) return false;
    if (d.nl && d.dem_dtype != OCSB_F32) return false;
    // one fuse_clamp setting: the ops then commute (see header)
    int fuse = -1;
    for (int i = 0; i < d.nf; ++i) {
        if (fuse >= 0 && d.feats[i].fuse_clamp != fuse) return false;
        fuse = d.feats[i].fuse_clamp;
        if (d.feats[i].nv <= 0 || d.feats[i].nv >= 0x7fffffffll) return false;
    }
    for (int i = 0; i < d.nl; ++i) {
        if (fuse >= 0 && d.lims[i].fuse_clamp != fuse) return false;
        fuse = d.lims[i].fuse_clamp;
    }
    for (int k = 0; k < d.n_wins; ++k)
        if (d.wins[k].ax.n < 1 || d.wins[k].re <= d.wins[k].rb) return false;
    return d.n_wins > 0;
}

int32_t k_pass_tiled(ocsb_ctx *ctx, const PassDesc &d) {
    OCSB_REQUIRE(k_pass_supported(d), "k_pass_tiled: unsupported op list");
    const double inf = std::numeric_limits<double>::infinity();
    const float finf = std::numeric_limits<float>::infinity();

    // ---- window table ---------------------------------------------------------
    std::vector<PWin> wins(d.n_wins);
    long long n0 = 0, n1 = 0, n2 = 0;
    double xmin = inf, xmax = -inf, ymin = inf, ymax = -inf, pix = 0.0;
    const bool vec_base = (((uintptr_t)d.out & 15) == 0) && (d.out_stride % 4 == 0) &&
                          (!d.in || ((((uintptr_t)d.in & 15) == 0) && (d.in_stride % 4 == 0)));
    for (int k = 0; k < d.n_wins; ++k) {
        const PassWin &s = d.wins[k];
        PWin &w = wins[k];
        w.ax = s.ax; w.ay = s.ay;
        w.rb = s.rb; w.re = s.re;
        w.keep_w = s.keep_w; w.keep_h = s.keep_h;
        w.hoff = s.hoff;
        w.dem_c = s.dem_c; w.dem_r = s.dem_r;
        const int bw = s.ax.n, bh = s.re - s.rb;
        w.n0x = (bw + kT0 - 1) / kT0; w.n0y = (bh + kT0 - 1) / kT0;
        w.n1x = (bw + kT1 - 1) / kT1; w.n1y = (bh + kT1 - 1) / kT1;
        w.n2x = (bw + kT2 - 1) / kT2; w.n2y = (bh + kT2 - 1) / kT2;
        w.base0 = (int)n0; w.base1 = (int)n1; w.base2 = (int)n2;
        n0 += (long long)w.n0x * w.n0y;
        n1 += (long long)w.n1x * w.n1y;
        n2 += (long long)w.n2x * w.n2y;
        w.vec = vec_base && (s.hoff % 4 == 0);
        xmin = std::min(xmin, std::min(s.ax.start, s.ax.stop));
        xmax = std::max(xmax, std::max(s.ax.start, s.ax.stop));
        ymin = std::min(ymin, std::min(s.ay.start, s.ay.stop));
        ymax = std::max(ymax, std::max(s.ay.start, s.ay.stop));
        pix = std::max(pix, std::max(fabs(s.ax.step), fabs(s.ay.step)));
    }
    OCSB_REQUIRE(n0 < 0x7fffffffll, "k_pass_tiled: too many tiles");
    const bool coords_ok = std::max(std::max(fabs(xmin), fabs(xmax)),
                                    std::max(fabs(ymin), fabs(ymax))) <= 1e14;

    // ---- counters / flags --------------------------------------------------------
    // [0..2*nf): pool counters; then per feature a far flag; then the overflow flag
    // ... followed by the cell histograms of the feature ops' vertex grids: ONE
    // memset per attempt clears everything
    const double ext_x = std::max(xmax - xmin, pix), ext_y = std::max(ymax - ymin, pix);
    std::vector<VGrid> grids(d.nf);
    std::vector<size_t> cnt_off(d.nf);
    size_t n_counts = 0;
    for (int f = 0; f < d.nf; ++f) {
        // grid: ~1024-px cells, finer for dense vertex sets (<= ~512 per cell)
        VGrid &G = grids[f];
        double cell = kT2 * pix;
        const double dense = sqrt(ext_x * ext_y * 512.0 / (double)d.feats[f].nv);
        if (dense < cell) cell = std::max(dense, 16.0 * pix);
        while ((ext_x / cell + 1) * (ext_y / cell + 1) > 1048576.0) cell *= 2.0;
        G.cell = cell; G.inv = 1.0 / cell;
        G.x0 = xmin; G.y0 = ymin;
        G.ncx = (int)(ext_x / cell) + 1; G.ncy = (int)(ext_y / cell) + 1;
        cnt_off[f] = n_counts;
        // histogram + (fused scan / scatter) the fill counters behind it
        n_counts += 2 * (size_t)(G.ncx + 2) * (G.ncy + 2);
    }
    ScratchGuard sg(ctx);
    const size_t nctr = 2 * kMaxFeat + 2;
    const size_t zero_words = nctr + 8 + (n_counts + 1) / 2;
    {
        // every piece of scratch of a first attempt from ONE kept arena (the sizes
        // restate the get() calls below; a later attempt's larger pools that do
        // not fit fall back to their own allocations): a band step of a multi-GPU
        // run is bound by host time, most of it stream-ordered malloc / free calls
        auto pc = [](size_t count, size_t size) { return ScratchGuard::piece(count * size); };
        size_t total = pc((size_t)d.n_wins, sizeof(PWin)) + pc(zero_words, 8) + 4096;
        const bool amode = d.ext_overflow != nullptr;
        for (int f = 0; f < d.nf; ++f) {
            const size_t nv = (size_t)d.feats[f].nv;
            const size_t ncell = (size_t)(grids[f].ncx + 2) * (grids[f].ncy + 2);
            size_t cap2 = 4 * nv + 48 * (size_t)n2 + (1 << 16);
            size_t cap1 = 4 * nv + 48 * (size_t)n1 + (1 << 16);
            size_t cap0 = 8 * nv + 8 * (size_t)n0 + 48 * (size_t)n1 + (1 << 17);
            if (amode) { cap2 <<= 1; cap1 <<= 1; cap0 <<= 1; }
            total += pc(ncell, 4) + pc(ncell + 1, 4) + pc(nv, 16) + pc(nv, 4);  // vertex grid
            total += 2 * pc((size_t)n2, 4) + pc(cap2, 4) + 2 * pc((size_t)n1, 4) + pc(cap1, 4) +
                     pc(cap1, 16) + pc(cap2, 16);                                // 1024 / 256 px
            total += pc((size_t)n0 * kSlot, 16) + pc(cap0, 16) + pc((size_t)n0, 4);  // 64 px
        }
        OCSB_TRY(sg.reserve_kept(total));
    }
    PWin *d_wins = nullptr;
    OCSB_TRY(sg.get(&d_wins, (size_t)d.n_wins));
    {
        const size_t wb = sizeof(PWin) * (size_t)d.n_wins;
        const bool at_head = (void *)d_wins == ctx->keep_arena;
        if (at_head && ctx->wins_cache.size() == wb &&
            memcmp(ctx->wins_cache.data(), wins.data(), wb) == 0) {
            // the same table is already there (uploaded by an earlier pass of this stream)
        } else {
            OCSB_TRY(ocsb_stage_upload(ctx, d_wins, wins.data(), wb));
            if (at_head)
                ctx->wins_cache.assign((const unsigned char *)wins.data(),
                                       (const unsigned char *)wins.data() + wb);
        }
    }
    unsigned long long *ctrs = nullptr;
    OCSB_TRY(sg.get(&ctrs, zero_words));
    int *counts_all = reinterpret_cast<int *>(ctrs + nctr + 8);
    // {pool overflow, far vertex}: one 8-byte word the tile kernel reads at once
    int *overflow = reinterpret_cast<int *>(ctrs + nctr);
    int *far_flag = overflow + 1;
    const bool async_mode = d.ext_overflow != nullptr;

    // ---- tensor map of the DEM block ------------------------------------------------
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof(tmap));
    int use_tma = 0;
    if (d.nl) {
        EncodeTiledFn enc = encode_tiled_fn();
        const bool aligned = (((uintptr_t)d.dem & 15) == 0) && ((d.dem_stride * 4) % 16 == 0);
        if (enc && aligned && d.dem_rows > 0 && d.dem_cols > 0) {
            const cuuint64_t dims[2] = {(cuuint64_t)d.dem_cols, (cuuint64_t)d.dem_rows};
            const cuuint64_t strides[1] = {(cuuint64_t)d.dem_stride * 4};
            const cuuint32_t box[2] = {(cuuint32_t)kDemPitch, (cuuint32_t)kDemRows};
            const cuuint32_t es[2] = {1, 1};
            const CUresult r = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                                   const_cast<void *>(d.dem), dims, strides, box, es,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            use_tma = (r == CUDA_SUCCESS) ? 1 : 0;
        }
    }


    // ---- persistent pipeline (k_pass2.cuh) for the common shape of a pass: at
    //      most one feature and one limiter op, fused clamps, TMA-able DEM
    static const bool v2_on = [] {
        const char *e = getenv("OCSB_PASS_V2");
        return !(e && atoi(e) == 0);
    }();
    bool v2 = v2_on && d.nf <= 1 && d.nl <= 1 && (d.nl == 0 || use_tma) &&
              n0 + 4096 < 0x7fffffffll;
    for (int f = 0; f < d.nf; ++f) v2 = v2 && d.feats[f].fuse_clamp;
    for (int l = 0; l < d.nl; ++l) v2 = v2 && d.lims[l].fuse_clamp;

    // ---- split form of a (feature, limiter) pass ------------------------------------
    // The ops commute bit for bit (one fuse_clamp setting), so the limiter may go
    // first: its sweep needs no candidate lists and runs on a second stream
    // BESIDE the pyramid's chain of latency-bound launches; the feature op then
    // only visits the tiles that ended with candidates (l10_kernel lists them),
    // in place.  Skipping a tile without candidates is exact when the op's miss
    // value cannot lower a clamped pixel: min(x, v_far) = x for every x <= ghmax.
    static const int split_env = [] {
        const char *e = getenv("OCSB_PASS_SPLIT");
        return e ? atoi(e) : 1;
    }();
    bool split = split_env != 0 && v2 && d.nf == 1 && d.nl == 1 &&
                 (d.in == nullptr || (d.in == d.out && d.in_stride == d.out_stride));
    if (split) {
        const PassFeat &pf = d.feats[0];
        volatile double m = pf.rate_target * pf.r;       // as F.v_far_f below
        volatile double sfar = m + pf.target;
        const float v_far = pf.r < inf ? (float)sfar : finf;
        const float hi = ocsb_is_none(d.ghmax) ? finf : (float)d.ghmax;
        split = v_far >= hi;
        for (int k = 0; k < d.n_wins; ++k)      // every window writes all of its pixels
            split = split && d.wins[k].keep_w >= d.wins[k].ax.n &&
                    d.wins[k].keep_h >= d.wins[k].re;
    }
    if (split && !ctx->s_split) {
        // an SM runs ONE shared-memory carveout at a time: the pyramid's launches
        // only fit beside the sweep's resident CTAs (which take the largest one)
        // when they ask for the same split of L1 / shared memory
        // (measured, profiles/r2c_split_lab.txt: with the preference set the root
        //  launch runs slower beside the sweep and nothing is gained -- opt-in)
        const char *cv = getenv("OCSB_PASS_SPLIT_CARVEOUT");
        if (cv && atoi(cv) != 0) {
            OCSB_CUDA(cudaFuncSetAttribute(root_kernel,
                                           cudaFuncAttributePreferredSharedMemoryCarveout,
                                           (int)cudaSharedmemCarveoutMaxShared));
            OCSB_CUDA(cudaFuncSetAttribute(l10_kernel,
                                           cudaFuncAttributePreferredSharedMemoryCarveout,
                                           (int)cudaSharedmemCarveoutMaxShared));
        }
        int lo_pri = 0, hi_pri = 0;
        OCSB_CUDA(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
        OCSB_CUDA(cudaStreamCreateWithPriority(&ctx->s_split, cudaStreamNonBlocking, lo_pri));
        for (auto &e : ctx->ev_split)
            OCSB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    static const int split_reserve = [] {
        // CTA slots the sweep leaves free.  Measured (profiles/r2c_split_lab.txt):
        // the pyramid needs the whole machine for ~0.16 ms rather than a few slots
        // for longer, so leaving slots free only slows the sweep down; with 0 the
        // pyramid's CTAs move in as the sweep's CTAs retire (its tail)
        const char *e = getenv("OCSB_PASS_SPLIT_RESERVE");
        return e ? atoi(e) : 0;
    }();

    for (int attempt = 0; attempt < 6; ++attempt) {
        ScratchGuard pools(ctx);
        pools.share(sg);        // carve on from where the call-level pieces end
        OCSB_CUDA(cudaMemsetAsync(ctrs, 0, zero_words * sizeof(*ctrs), ctx->stream));
        int *sticky = d.sticky ? d.ext_overflow : nullptr;
        TileArgs T;
        memset(&T, 0, sizeof(T));
        double2 *slots0 = nullptr, *pool0 = nullptr;
        T.wins = d_wins; T.n_wins = d.n_wins;
        T.in = d.in; T.out = d.out;
        T.in_stride = d.in_stride; T.out_stride = d.out_stride;
        T.nf = d.nf; T.nl = d.nl;
        T.lo = ocsb_is_none(d.ghmin) ? -finf : (float)d.ghmin;
        T.hi = ocsb_is_none(d.ghmax) ? finf : (float)d.ghmax;
        T.dem = (const float *)d.dem; T.dem_stride = d.dem_stride;
        T.dem_rows = d.dem_rows; T.dem_cols = d.dem_cols;
        T.use_tma = use_tma;
        T.overflow = overflow;
        T.sticky = sticky;
        for (int l = 0; l < d.nl; ++l) {
            const PassLim &pl = d.lims[l];
            LimiterDev<float> &p = T.lim[l];
            p.hmin = ocsb_is_none(pl.hmin) ? -finf : (float)pl.hmin;
            p.hmax = ocsb_is_none(pl.hmax) ? finf : (float)pl.hmax;
            p.ghmin = T.lo; p.ghmax = T.hi;
            p.lb = ocsb_is_none(pl.lb) ? -finf : (float)pl.lb;
            p.ub = ocsb_is_none(pl.ub) ? finf : (float)pl.ub;
            p.d1_0 = make_cdiv<float>(d.dx);
            p.d2_0 = make_cdiv<float>(2.0 * d.dx);
            p.d1_1 = make_cdiv<float>(d.dy);
            p.d2_1 = make_cdiv<float>(2.0 * d.dy);
            p.fuse_clamp = pl.fuse_clamp;
            T.lim_band[l] = (!ocsb_is_none(pl.lb) || !ocsb_is_none(pl.ub)) ? 1 : 0;
        }
        int *active = nullptr, *n_active = reinterpret_cast<int *>(ctrs + nctr) + 8;
        if (split) OCSB_TRY(pools.get(&active, (size_t)n0));

        for (int f = 0; f < d.nf; ++f) {
            const PassFeat &pf = d.feats[f];
            OCSB_REQUIRE(((uintptr_t)pf.verts & 15) == 0,
                         "ocsb_hfun_feature: verts_xy must be 16-byte aligned");
            const int nv = (int)pf.nv;
            const bool bounded = pf.r < inf;
            const double r2 = bounded ? pf.r * pf.r : inf;
            const double r2lim = bounded ? (pf.r > 0 ? r2 * (1.0 + kDomEps) : -1.0) : inf;
            VGrid G = grids[f];
            const int ncell = (G.ncx + 2) * (G.ncy + 2);
            int *counts = counts_all + cnt_off[f], *cursor = nullptr;
            OCSB_TRY(pools.get(&cursor, (size_t)ncell));
            OCSB_TRY(pools.get(&G.start, (size_t)ncell + 1));
            OCSB_TRY(pools.get(&G.sv, (size_t)nv));
            OCSB_TRY(pools.get(&G.sid, (size_t)nv));
            const double2 *verts = reinterpret_cast<const double2 *>(pf.verts);
            if (ncell <= kVgSmallCells && nv <= kVgSmallVerts) {
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                vg_build_small_kernel<<<1, 1024, 0, ctx->stream>>>(G, verts, nv, far_flag);
                OCSB_LAUNCHED(ctx);
            } else if (ncell <= kVgFusedCells && (nv + 255) / 256 <= 2 * ctx->num_sms &&
                       ctx->keep_ok) {      // (not while passes run side by side on worker streams)
                // histogram, scan and scatter in ONE launch (grid-wide counter barrier)
                int *bar = reinterpret_cast<int *>(ctrs + nctr) + 10 + f;
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                vg_scan_scatter_kernel<true><<<(nv + 255) / 256, 256, 0, ctx->stream>>>(
                    G, verts, nv, counts, counts + ncell, far_flag, bar, overflow, sticky);
                OCSB_LAUNCHED(ctx);
            } else {
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                vg_hist_kernel<<<(nv + 255) / 256, 256, 0, ctx->stream>>>(G, verts, nv, counts, far_flag);
                OCSB_LAUNCHED(ctx);
                if (ncell <= kVgFusedCells) {
                    OCSB_TIME(ctx, KC_FEATURE_REFINE);
                    vg_scan_scatter_kernel<false><<<(nv + 255) / 256, 256, 0, ctx->stream>>>(
                        G, verts, nv, counts, counts + ncell, far_flag, nullptr, overflow,
                        sticky);
                    OCSB_LAUNCHED(ctx);
                } else {
                    OCSB_TIME(ctx, KC_FEATURE_REFINE);
                    vg_scan_kernel<<<1, 1024, 0, ctx->stream>>>(counts, ncell, G.start, cursor);
                    OCSB_LAUNCHED(ctx);
                    OCSB_TIME(ctx, KC_FEATURE_REFINE);
                    vg_scatter_kernel<<<(nv + 255) / 256, 256, 0, ctx->stream>>>(G, verts, nv,
                                                                               cursor);
                    OCSB_LAUNCHED(ctx);
                }
            }
            OCSB_CHECK_LAUNCH();
            if (split) {
                // fork: the limiter sweep goes to the second stream, after the
                // memset (its tile counter, the flags), the window table and the
                // vertex-grid launches (their 1024-thread CTAs would not fit an
                // SM beside the sweep's resident CTAs and would wait for its end)
                OCSB_CUDA(cudaEventRecord(ctx->ev_split[0], ctx->stream));
                OCSB_CUDA(cudaStreamWaitEvent(ctx->s_split, ctx->ev_split[0], 0));
                P2Args P;
                memset(&P, 0, sizeof(P));
                P.wins = d_wins; P.n_wins = d.n_wins; P.n_tiles = (int)n0;
                P.in = d.in; P.out = d.out;
                P.in_stride = d.in_stride; P.out_stride = d.out_stride;
                P.lim = T.lim[0];
                P.lo = T.lo; P.hi = T.hi;
                P.counter = reinterpret_cast<int *>(ctrs + nctr) + 6;
                P.overflow = overflow;
                P.sticky = sticky;
                cudaStream_t main_stream = ctx->stream;
                ctx->stream = ctx->s_split;        // launch + timing events on that stream
                OCSB_TIME(ctx, KC_PASS_LIMITER);
                const int32_t st = launch_pass2(ctx, tmap, P, d.in != nullptr, false, true,
                                                T.lim_band[0] != 0, split_reserve);
                OCSB_LAUNCHED(ctx);
                ctx->stream = main_stream;
                OCSB_TRY(st);
                OCSB_CHECK_LAUNCH();
                OCSB_CUDA(cudaEventRecord(ctx->ev_split[1], ctx->s_split));
            }

            BuildArgs B;
            B.wins = d_wins; B.n_wins = d.n_wins;
            B.verts = verts;
            B.grid = G;
            unsigned long long cap2 = 4ull * nv + 48ull * (unsigned long long)n2 + (1ull << 16);
            unsigned long long cap1 = 4ull * nv + 48ull * (unsigned long long)n1 + (1ull << 16);
            cap2 <<= (2 * attempt); cap1 <<= (2 * attempt);
            if (async_mode) { cap2 <<= 1; cap1 <<= 1; }
            cap2 = std::min<unsigned long long>(cap2, 0x7fffffffull);
            cap1 = std::min<unsigned long long>(cap1, 0x7fffffffull);
            OCSB_TRY(pools.get(&B.l2.off, (size_t)n2));
            OCSB_TRY(pools.get(&B.l2.cnt, (size_t)n2));
            OCSB_TRY(pools.get(&B.l2.ids, (size_t)cap2));
            OCSB_TRY(pools.get(&B.l1.off, (size_t)n1));
            OCSB_TRY(pools.get(&B.l1.cnt, (size_t)n1));
            OCSB_TRY(pools.get(&B.l1.ids, (size_t)cap1));
            OCSB_TRY(pools.get(&B.l1.xy, (size_t)cap1));
            OCSB_TRY(pools.get(&B.l2.xy, (size_t)cap2));
            B.ctr2 = ctrs + 2 * f; B.ctr1 = ctrs + 2 * f + 1;
            B.cap2 = cap2; B.cap1 = cap1;
            B.overflow = overflow;
            B.sticky = sticky;
            B.r = pf.r; B.r2lim = r2lim;
            OCSB_TIME(ctx, KC_FEATURE_REFINE);
            root_kernel<<<(unsigned)n2, kRootThreads, 0, ctx->stream>>>(B);
            OCSB_LAUNCHED(ctx);
            if (!v2) {
                // (the persistent kernel's slots come from l10_kernel below, which
                //  filters the 1024-px lists itself)
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                l1_kernel<<<(unsigned)n1, kL1Threads, 0, ctx->stream>>>(B);
                OCSB_LAUNCHED(ctx);
            }
            OCSB_CHECK_LAUNCH();

            FeatDev &F = T.feat[f];
            F.verts = verts;
            F.off1 = B.l1.off; F.cnt1 = B.l1.cnt; F.ids1 = B.l1.ids;
            if (v2) {
                // 64-px level: one slot per tile (+ overflow pool) for the persistent kernel
                // overflow lists of 64-px tiles + the 256-px lists too long for
                // shared memory (dense vertex sets)
                unsigned long long cap0 = 8ull * nv + 8ull * (unsigned long long)n0 +
                                          48ull * (unsigned long long)n1 + (1ull << 17);
                cap0 <<= (2 * attempt);
                if (async_mode) cap0 <<= 1;
                cap0 = std::min<unsigned long long>(cap0, 0x7fffffffull);
                OCSB_TRY(pools.get(&slots0, (size_t)n0 * kSlot));
                OCSB_TRY(pools.get(&pool0, (size_t)cap0));
                L0Args L;
                L.wins = d_wins; L.n_wins = d.n_wins; L.n_tiles = (int)n0;
                L.cnt2 = B.l2.cnt; L.off2 = B.l2.off; L.xy2 = B.l2.xy;
                L.r2lim = r2lim;
                L.slots = slots0; L.pool = pool0;
                L.ctr = ctrs + 2 * kMaxFeat; L.cap = cap0;
                L.overflow = overflow;
                L.sticky = sticky;
                L.active = active; L.n_active = n_active;
                OCSB_TIME(ctx, KC_FEATURE_REFINE);
                L.n_parents = (int)n1;
                l10_kernel<<<(unsigned)n1, kL10Threads, 0, ctx->stream>>>(L);
                OCSB_LAUNCHED(ctx);
                OCSB_CHECK_LAUNCH();
            }
            F.rt = pf.rate_target; F.target = pf.target;
            F.r2 = (bounded && pf.r <= 0) ? -1.0 : r2;
            F.r2lim = r2lim;
            {
                volatile double m = pf.rate_target * pf.r;   // no FMA contraction
                volatile double s = m + pf.target;
                F.v_far_f = bounded ? (float)s : finf;
            }
            F.fuse = pf.fuse_clamp;
            // the fast size law needs v64 = rt*d + target to be a normal float32
            // with |error| <= 2^-39 v64 (k_feature_eval.cuh)
            const bool fast_ok = pf.rate_target >= 0.0 && pf.rate_target <= 1e15 &&
                                 pf.target >= 1e-30 && pf.target <= 1e30 && coords_ok;
            {
                const double lo = 1e-30;
                uint64_t b; memcpy(&b, &lo, 8);
                F.hi_lo = fast_ok ? (uint32_t)(b >> 32) + 1 : 0xffffffffu;
            }
        }
        if (split) {
            // join: the feature op reads what the sweep wrote
            OCSB_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_split[1], 0));
            P2Args P;
            memset(&P, 0, sizeof(P));
            P.wins = d_wins; P.n_wins = d.n_wins; P.n_tiles = (int)n0;
            P.in = d.out; P.out = d.out;
            P.in_stride = d.out_stride; P.out_stride = d.out_stride;
            P.feat = T.feat[0]; P.slots = slots0; P.pool0 = pool0;
            P.lo = T.lo; P.hi = T.hi;
            P.counter = reinterpret_cast<int *>(ctrs + nctr) + 4;
            P.overflow = overflow;
            P.sticky = sticky;
            P.active = active; P.n_active = n_active;
            OCSB_TIME(ctx, KC_PASS_FEATURE);
            OCSB_TRY(launch_pass2(ctx, tmap, P, true, true, false, false));
        } else if (v2) {
            OCSB_TIME(ctx, KC_PASS_TILE);
            P2Args P;
            memset(&P, 0, sizeof(P));
            P.wins = d_wins; P.n_wins = d.n_wins; P.n_tiles = (int)n0;
            P.in = d.in; P.out = d.out;
            P.in_stride = d.in_stride; P.out_stride = d.out_stride;
            if (d.nf) { P.feat = T.feat[0]; P.slots = slots0; P.pool0 = pool0; }
            if (d.nl) P.lim = T.lim[0];
            P.lo = T.lo; P.hi = T.hi;
            P.counter = reinterpret_cast<int *>(ctrs + nctr) + 4;
            P.overflow = overflow;
            P.sticky = sticky;
            const bool band = d.nl && T.lim_band[0];
            OCSB_TRY(launch_pass2(ctx, tmap, P, d.in != nullptr, d.nf > 0, d.nl > 0, band));
        } else {
            OCSB_TIME(ctx, KC_PASS_TILE);
            const dim3 tgrid = d.n_wins == 1
                                   ? dim3((unsigned)wins[0].n0x, (unsigned)wins[0].n0y)
                                   : dim3((unsigned)n0);
            if (d.in) tile_kernel<true><<<tgrid, kPassThreads, 0, ctx->stream>>>(tmap, T);
            else tile_kernel<false><<<tgrid, kPassThreads, 0, ctx->stream>>>(tmap, T);
        }
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
#ifdef OCSB_TIMELINE
        if (getenv("OCSB_TIMELINE_PRINT")) {
            static unsigned long long tl[8][2];
            static int calls = 0;
            cudaStreamSynchronize(ctx->stream);
            cudaMemcpyFromSymbol(tl, g_timeline, sizeof(tl));
            if (++calls % 8 == 0) {
                unsigned long long t0 = ~0ull;
                for (int i = 0; i < 8; ++i)
                    if (tl[i][1]) t0 = std::min(t0, tl[i][0]);
                fprintf(stderr, "PASS_STATS timeline(us) vg_hist,vg_scatter,root,l10,A,B:");
                for (int i = 0; i < 6; ++i)
                    fprintf(stderr, " [%.1f %.1f]", tl[i][1] ? (tl[i][0] - t0) * 1e-3 : -1.0,
                            tl[i][1] ? (tl[i][1] - t0) * 1e-3 : -1.0);
                fprintf(stderr, "\n");
            }
            for (int i = 0; i < 8; ++i) { tl[i][0] = ~0ull; tl[i][1] = 0; }
            cudaMemcpyToSymbol(g_timeline, tl, sizeof(tl));
        }
#endif
#ifdef OCSB_PASS_STATS
        if (getenv("OCSB_PASS_STATS_PRINT")) {
            static unsigned long long hs[160][32];
            cudaStreamSynchronize(ctx->stream);
            cudaMemcpyFromSymbol(hs, g_pass_stats, sizeof(hs));
            fprintf(stderr, "PASS_STATS");
            for (int i = 0; i < 32; ++i) {
                unsigned long long t = 0;
                for (int m = 0; m < 160; ++m) t += hs[m][i];
                fprintf(stderr, " %llu", t);
            }
            fprintf(stderr, "\n");
            memset(hs, 0, sizeof(hs));
            cudaMemcpyToSymbol(g_pass_stats, hs, sizeof(hs));
        }
#endif

        if (d.nf == 0) return OCSB_OK;
        if (async_mode) {
            // the caller looks at its own flag later (and redoes the band); a
            // sticky status word was written by the kernels themselves
            if (!d.sticky)
                OCSB_CUDA(cudaMemcpyAsync(d.ext_overflow, overflow, sizeof(int),
                                          cudaMemcpyDeviceToDevice, ctx->stream));
            return OCSB_OK;
        }
        int *flag_h = reinterpret_cast<int *>(ctx->pinned);
        int flag_local = 0;
        if (!flag_h) flag_h = &flag_local;
        OCSB_CUDA(cudaMemcpyAsync(flag_h, overflow, sizeof(int), cudaMemcpyDeviceToHost,
                                  ctx->stream));
        OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
        if (*flag_h == 0) return OCSB_OK;
    }
    ocsb_set_error("ocsb_hfun_pass: candidate pools overflowed after 6 attempts "
                   "(degenerate vertex set?)");
    return OCSB_EOVERFLOW;
}
