// Nearest-vertex query for ARBITRARY query points: the second cKDTree site of
// the reference, Constraint._apply_rate (ocsmesh/features/constraint.py:
// 104-105; the raster-pixel site is K1 in k_feature.cu).
//
// Grid-bucketed search: the vertices are counting-sorted into a uniform grid
// (~2 per cell) over their bounding box; one thread per query walks square
// rings of cells outwards from its own cell, keeping the exact f64 minimum
// d2 = fl(fl(dx*dx) + fl(dy*dy)) (FMA-free, what scipy computes) and the
// lowest vertex id among exact ties, and stops once every unvisited cell is
// provably farther than the current best.
#include "ocsb_common.cuh"

#include <algorithm>
#include <cmath>
#include <limits>
#include <mutex>

namespace {

struct NGrid {
    double x0, y0, cell, inv;
    int gx, gy;
};

__device__ __forceinline__ int cell_of(double v, double o, double inv, int n) {
    double t = floor((v - o) * inv);
    if (!(t > 0.0)) return 0;           // also NaN
    if (t > (double)(n - 1)) return n - 1;
    return (int)t;
}

__global__ void __launch_bounds__(256)
bbox_kernel(const double2 *__restrict__ v, int64_t n, double *__restrict__ bb) {
    // bb: [xmin, ymin, xmax, ymax] as ordered-integer keys, atomics per warp
    double x0 = INFINITY, y0 = INFINITY, x1 = -INFINITY, y1 = -INFINITY;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const double2 p = v[i];
        x0 = fmin(x0, p.x); x1 = fmax(x1, p.x);
        y0 = fmin(y0, p.y); y1 = fmax(y1, p.y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        x0 = fmin(x0, __shfl_xor_sync(0xffffffffu, x0, o));
        y0 = fmin(y0, __shfl_xor_sync(0xffffffffu, y0, o));
        x1 = fmax(x1, __shfl_xor_sync(0xffffffffu, x1, o));
        y1 = fmax(y1, __shfl_xor_sync(0xffffffffu, y1, o));
    }
    if ((threadIdx.x & 31) == 0) {
        auto key = [](double d) {   // monotone map double -> signed 64-bit
            long long b = __double_as_longlong(d);
            return b >= 0 ? b : (long long)(0x8000000000000000ull - (unsigned long long)b);
        };
        long long *k = reinterpret_cast<long long *>(bb);
        atomicMin(k + 0, key(x0)); atomicMin(k + 1, key(y0));
        atomicMax(k + 2, key(x1)); atomicMax(k + 3, key(y1));
    }
}

__global__ void __launch_bounds__(256)
cell_hist_kernel(const double2 *__restrict__ v, int64_t n, NGrid G,
                 int *__restrict__ cnt) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const double2 p = v[i];
        const int c = cell_of(p.y, G.y0, G.inv, G.gy) * G.gx + cell_of(p.x, G.x0, G.inv, G.gx);
        atomicAdd(cnt + c, 1);
    }
}

// exclusive scan by one CTA (ncell <= 2^22); start[ncell] = n
__global__ void __launch_bounds__(1024)
cell_scan_kernel(const int *__restrict__ cnt, int *__restrict__ start,
                 int *__restrict__ cursor, int ncell) {
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < ncell; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < ncell ? cnt[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_warp[wid] = x;
        __syncthreads();
        if (wid == 0) {
            int t = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t += y;
            }
            s_warp[lane] = t;
        }
        __syncthreads();
        const int carry = s_carry;
        const int ex = carry + (wid ? s_warp[wid - 1] : 0) + (x - v);
        if (i < ncell) { start[i] = ex; cursor[i] = ex; }
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = ex + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) start[ncell] = s_carry;
}

__global__ void __launch_bounds__(256)
cell_scatter_kernel(const double2 *__restrict__ v, int64_t n, NGrid G,
                    int *__restrict__ cursor, int *__restrict__ ids) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const double2 p = v[i];
        const int c = cell_of(p.y, G.y0, G.inv, G.gy) * G.gx + cell_of(p.x, G.x0, G.inv, G.gx);
        ids[atomicAdd(cursor + c, 1)] = (int)i;
    }
}

__global__ void __launch_bounds__(128)
query_kernel(const double2 *__restrict__ verts, int64_t nv, NGrid G,
             const int *__restrict__ start, const int *__restrict__ ids,
             const double2 *__restrict__ q, int64_t nq, double r2,
             double *__restrict__ dist, int64_t *__restrict__ idx) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq) return;
    const double2 p = q[t];
    const int cx = cell_of(p.x, G.x0, G.inv, G.gx), cy = cell_of(p.y, G.y0, G.inv, G.gy);
    double best = INFINITY;
    int bid = 0x7fffffff;
    auto visit = [&](int ix, int iy) {
        const int c = iy * G.gx + ix;
        const int e = start[c + 1];
        for (int s = start[c]; s < e; ++s) {
            const int id = ids[s];
            const double2 v = verts[id];
            const double dx = __dsub_rn(p.x, v.x), dy = __dsub_rn(p.y, v.y);
            const double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
            if (d2 < best || (d2 == best && id < bid)) { best = d2; bid = id; }
        }
    };
    const int kmax = max(max(cx, G.gx - 1 - cx), max(cy, G.gy - 1 - cy));
    const double slack = 1e-11 * (fabs(G.x0) + fabs(G.y0) + fabs(p.x) + fabs(p.y) +
                                  G.cell * (double)(G.gx + G.gy));
    for (int k = 0; k <= kmax; ++k) {
        const int xa = cx - k, xb = cx + k, ya = cy - k, yb = cy + k;
        if (k == 0) {
            visit(cx, cy);
        } else {
            for (int ix = max(xa, 0); ix <= min(xb, G.gx - 1); ++ix) {
                if (ya >= 0) visit(ix, ya);
                if (yb < G.gy) visit(ix, yb);
            }
            for (int iy = max(ya + 1, 0); iy <= min(yb - 1, G.gy - 1); ++iy) {
                if (xa >= 0) visit(xa, iy);
                if (xb < G.gx) visit(xb, iy);
            }
        }
        // every vertex in an unvisited cell lies outside the block of cells
        // [xa, xb] x [ya, yb]; sides of the block that coincide with the grid
        // border have nothing beyond them (all vertices are inside the grid)
        double m = INFINITY;
        if (xa > 0) m = fmin(m, p.x - (G.x0 + xa * G.cell));
        if (xb < G.gx - 1) m = fmin(m, (G.x0 + (xb + 1) * G.cell) - p.x);
        if (ya > 0) m = fmin(m, p.y - (G.y0 + ya * G.cell));
        if (yb < G.gy - 1) m = fmin(m, (G.y0 + (yb + 1) * G.cell) - p.y);
        // strict, with slack for the rounding of cell indices and edges: an
        // unvisited vertex at exactly `best` (a tie) must still be seen
        m -= slack;
        if (m > 0.0 && m * m > best) break;
        if (m > 0.0 && m * m > r2) break;
    }
    const bool hit = bid != 0x7fffffff && best < r2;
    if (dist) dist[t] = hit ? __dsqrt_rn(best) : INFINITY;
    if (idx) idx[t] = hit ? (int64_t)bid : nv;
}

}  // namespace

int32_t k_nearest_vertex(ocsb_ctx *ctx, const double *verts, int64_t nv,
                         const double *q, int64_t nq, double r, double *dist,
                         int64_t *idx) {
    OCSB_REQUIRE((((uintptr_t)verts | (uintptr_t)q) & 15) == 0,
                 "ocsb_nearest_vertex: coordinate arrays must be 16-byte aligned");
    const double inf = std::numeric_limits<double>::infinity();
    const double r2 = r < inf ? (r > 0 ? r * r : -1.0) : inf;
    const double2 *v2 = reinterpret_cast<const double2 *>(verts);
    ScratchGuard sg(ctx);
    double *bb = nullptr;
    OCSB_TRY(sg.get(&bb, 4));
    {
        // ordered-key identities of +inf / -inf
        const long long kmax = 0x7ff0000000000000LL, kmin = -0x7ff0000000000000LL;
        long long init[4] = {kmax, kmax, kmin, kmin};
        long long *bounce = ctx->pinned ? (long long *)ctx->pinned : nullptr;
        if (bounce) memcpy(bounce, init, sizeof(init));
        OCSB_CUDA(cudaMemcpyAsync(bb, bounce ? bounce : init, sizeof(init),
                                  cudaMemcpyHostToDevice, ctx->stream));
        if (!bounce) OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    const int sweep = ctx->num_sms * 8;
    OCSB_TIME(ctx, KC_NEAREST);
    bbox_kernel<<<(int)std::min<int64_t>(sweep, (nv + 255) / 256), 256, 0, ctx->stream>>>(v2, nv, bb);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    long long keys[4];
    OCSB_CUDA(cudaMemcpyAsync(keys, bb, sizeof(keys), cudaMemcpyDeviceToHost, ctx->stream));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    double box[4];
    for (int i = 0; i < 4; ++i) {
        long long b = keys[i] >= 0 ? keys[i]
                                   : (long long)(0x8000000000000000ull - (unsigned long long)keys[i]);
        memcpy(&box[i], &b, 8);
    }
    OCSB_REQUIRE(std::isfinite(box[0]) && std::isfinite(box[1]) &&
                     std::isfinite(box[2]) && std::isfinite(box[3]),
                 "ocsb_nearest_vertex: non-finite vertex coordinates");
    NGrid G;
    const double wx = box[2] - box[0], wy = box[3] - box[1];
    double ncell_want = std::max<double>(1.0, (double)nv / 2.0);
    if (ncell_want > 4194304.0) ncell_want = 4194304.0;
    double cell;
    if (wx > 0 && wy > 0) cell = sqrt(wx * wy / ncell_want);
    else cell = std::max(wx, wy) / ncell_want;
    if (!(cell > 0)) cell = 1.0;
    // keep each axis at <= 2048 cells (thin clouds would otherwise explode)
    cell = std::max(cell, std::max(wx, wy) / 2048.0);
    G.x0 = box[0]; G.y0 = box[1]; G.cell = cell; G.inv = 1.0 / cell;
    G.gx = (int)std::min(2048.0, floor(wx / cell) + 1.0);
    G.gy = (int)std::min(2048.0, floor(wy / cell) + 1.0);
    const int ncell = G.gx * G.gy;
    int *cnt = nullptr, *start = nullptr, *cursor = nullptr, *ids = nullptr;
    OCSB_TRY(sg.get(&cnt, (size_t)ncell));
    OCSB_TRY(sg.get(&start, (size_t)ncell + 1));
    OCSB_TRY(sg.get(&cursor, (size_t)ncell));
    OCSB_TRY(sg.get(&ids, (size_t)nv));
    OCSB_CUDA(cudaMemsetAsync(cnt, 0, (size_t)ncell * sizeof(int), ctx->stream));
    const int vgrid = (int)std::min<int64_t>(sweep, (nv + 255) / 256);
    OCSB_TIME(ctx, KC_NEAREST);
    cell_hist_kernel<<<vgrid, 256, 0, ctx->stream>>>(v2, nv, G, cnt);
    OCSB_LAUNCHED(ctx);
    OCSB_TIME(ctx, KC_NEAREST);
    cell_scan_kernel<<<1, 1024, 0, ctx->stream>>>(cnt, start, cursor, ncell);
    OCSB_LAUNCHED(ctx);
    OCSB_TIME(ctx, KC_NEAREST);
    cell_scatter_kernel<<<vgrid, 256, 0, ctx->stream>>>(v2, nv, G, cursor, ids);
    OCSB_LAUNCHED(ctx);
    OCSB_TIME(ctx, KC_NEAREST);
    query_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, ctx->stream>>>(
        v2, nv, G, start, ids, reinterpret_cast<const double2 *>(q), nq, r2, dist, idx);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

// ---------------------------------------------------------------------------
// ocsb_nearest_vertex_static: the same query with NO allocation of any kind.
// No counterpart in the reference; added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
// loops, harnesses that forbid it).
//
// The vertex grid (bounding-box keys, cell counts, cell starts, cursors, sorted
// ids) lives in ONE file-scope device array of fixed size; every GPU that loads
// the library holds these ~4 MiB until the process exits.  All calls share the
// array and it keeps the last call's contents: what must start at zero (the
// bounding-box keys, the cell counts) is cleared by every call, and two calls
// that overlap on the GPU (different streams, different contexts on one device)
// race on it -- serialise them.
// ---------------------------------------------------------------------------
namespace {

constexpr int64_t kStaticMaxVerts = OCSB_STATIC_MAX_VERTS;     // vertices per call
constexpr int kStaticMaxCells = 262144;         // grid cells per call
// layout in ints: [bb: 4 x int64, padded to 256 B][cnt][start (+1, padded)][cursor][ids]
constexpr size_t kStOffCnt = 64;
constexpr size_t kStOffStart = kStOffCnt + (size_t)kStaticMaxCells;
constexpr size_t kStOffCursor = kStOffStart + (size_t)kStaticMaxCells + 64;
constexpr size_t kStOffIds = kStOffCursor + (size_t)kStaticMaxCells;
constexpr size_t kStaticInts = kStOffIds + (size_t)kStaticMaxVerts;

static __device__ __align__(256) int g_nn_static[kStaticInts];

// In host code `g_nn_static` is only a stand-in for the symbol, not a device
// address: the address is asked for once per device, with that device current.
constexpr int kStaticMaxDev = 64;
int *g_nn_static_ptr[kStaticMaxDev] = {};
std::mutex g_nn_static_mu;

int32_t nn_static_scratch(int device, int **out) {
    OCSB_REQUIRE(device >= 0 && device < kStaticMaxDev,
                 "ocsb_nearest_vertex_static: device %d out of range", device);
    std::lock_guard<std::mutex> lk(g_nn_static_mu);
    if (!g_nn_static_ptr[device]) {
        void *p = nullptr;
        OCSB_CUDA(cudaGetSymbolAddress(&p, g_nn_static));
        OCSB_REQUIRE(p != nullptr, "ocsb_nearest_vertex_static: no address for the scratch array");
        g_nn_static_ptr[device] = static_cast<int *>(p);
    }
    *out = g_nn_static_ptr[device];
    return OCSB_OK;
}

// bounding-box keys back to the identities of min / max (device code may name
// the array directly)
__global__ void nn_static_init_kernel() {
    long long *k = reinterpret_cast<long long *>(g_nn_static);
    const long long kmax = 0x7ff0000000000000LL, kmin = -0x7ff0000000000000LL;
    if (threadIdx.x < 4) k[threadIdx.x] = threadIdx.x < 2 ? kmax : kmin;
}

}  // namespace

extern "C" int32_t ocsb_nearest_vertex_static(ocsb_ctx *ctx, const double *verts_xy,
                                              int64_t nv, const double *query_xy,
                                              int64_t nq, double r, double *dist_out,
                                              int64_t *idx_out) {
    const char *who = "ocsb_nearest_vertex_static";
    OCSB_REQUIRE(ctx != nullptr, "%s: ctx is NULL", who);
    OCSB_CUDA(cudaSetDevice(ctx->device));
    OCSB_REQUIRE(verts_xy && nv > 0, "%s: empty vertex set", who);
    OCSB_REQUIRE(nv <= kStaticMaxVerts,
                 "%s: %lld vertices exceed the fixed scratch array (at most %lld); "
                 "use ocsb_nearest_vertex",
                 who, (long long)nv, (long long)kStaticMaxVerts);
    OCSB_REQUIRE(nq >= 0, "%s: negative nq", who);
    if (nq == 0) return OCSB_OK;
    OCSB_REQUIRE(query_xy && (dist_out || idx_out), "%s: NULL buffers", who);
    // nothing is staged here: staging would allocate
    OCSB_REQUIRE(ocsb_is_device_ptr(verts_xy) && ocsb_is_device_ptr(query_xy) &&
                     (!dist_out || ocsb_is_device_ptr(dist_out)) &&
                     (!idx_out || ocsb_is_device_ptr(idx_out)),
                 "%s: inputs and outputs must be device memory (this entry point "
                 "allocates no staging buffers)", who);
    OCSB_REQUIRE((((uintptr_t)verts_xy | (uintptr_t)query_xy) & 15) == 0,
                 "%s: coordinate arrays must be 16-byte aligned", who);
    int *scr = nullptr;
    OCSB_TRY(nn_static_scratch(ctx->device, &scr));
    double *bb = reinterpret_cast<double *>(scr);
    int *cnt = scr + kStOffCnt, *start = scr + kStOffStart;
    int *cursor = scr + kStOffCursor, *ids = scr + kStOffIds;

    const double inf = std::numeric_limits<double>::infinity();
    const double rr = ocsb_is_none(r) ? inf : r;
    const double r2 = rr < inf ? (rr > 0 ? rr * rr : -1.0) : inf;
    const double2 *v2 = reinterpret_cast<const double2 *>(verts_xy);
    const int sweep = ctx->num_sms * 8;
    const int vgrid = (int)std::min<int64_t>(sweep, (nv + 255) / 256);

    nn_static_init_kernel<<<1, 32, 0, ctx->stream>>>();
    OCSB_LAUNCHED(ctx);
    bbox_kernel<<<vgrid, 256, 0, ctx->stream>>>(v2, nv, bb);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    long long keys[4];
    OCSB_CUDA(cudaMemcpyAsync(keys, bb, sizeof(keys), cudaMemcpyDeviceToHost, ctx->stream));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    double box[4];
    for (int i = 0; i < 4; ++i) {
        long long b = keys[i] >= 0 ? keys[i]
                                   : (long long)(0x8000000000000000ull - (unsigned long long)keys[i]);
        memcpy(&box[i], &b, 8);
    }
    OCSB_REQUIRE(std::isfinite(box[0]) && std::isfinite(box[1]) &&
                     std::isfinite(box[2]) && std::isfinite(box[3]),
                 "%s: non-finite vertex coordinates", who);
    // ~2 vertices per cell, at most 2048 cells per axis, and never more cells
    // than the array holds (the grid only orders the search: results do not
    // depend on it)
    NGrid G;
    const double wx = box[2] - box[0], wy = box[3] - box[1];
    const double ncell_want = std::max<double>(1.0, (double)nv / 2.0);
    double cell;
    if (wx > 0 && wy > 0) cell = sqrt(wx * wy / ncell_want);
    else cell = std::max(wx, wy) / ncell_want;
    if (!(cell > 0)) cell = 1.0;
    cell = std::max(cell, std::max(wx, wy) / 2048.0);
    for (int it = 0;; ++it) {
        G.gx = (int)std::min(2048.0, floor(wx / cell) + 1.0);
        G.gy = (int)std::min(2048.0, floor(wy / cell) + 1.0);
        if ((int64_t)G.gx * G.gy <= kStaticMaxCells) break;
        OCSB_REQUIRE(it < 64, "%s: vertex grid does not fit the scratch array", who);
        cell *= 1.25;
    }
    G.x0 = box[0]; G.y0 = box[1]; G.cell = cell; G.inv = 1.0 / cell;
    const int ncell = G.gx * G.gy;
    OCSB_CUDA(cudaMemsetAsync(cnt, 0, (size_t)ncell * sizeof(int), ctx->stream));
    cell_hist_kernel<<<vgrid, 256, 0, ctx->stream>>>(v2, nv, G, cnt);
    OCSB_LAUNCHED(ctx);
    cell_scan_kernel<<<1, 1024, 0, ctx->stream>>>(cnt, start, cursor, ncell);
    OCSB_LAUNCHED(ctx);
    cell_scatter_kernel<<<vgrid, 256, 0, ctx->stream>>>(v2, nv, G, cursor, ids);
    OCSB_LAUNCHED(ctx);
    query_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, ctx->stream>>>(
        v2, nv, G, start, ids, reinterpret_cast<const double2 *>(query_xy), nq, r2,
        dist_out, idx_out);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    // device buffers: stream-ordered like the other entry points (the one host
    // wait above is for the bounding box, as in ocsb_nearest_vertex)
    return OCSB_OK;
}
