// Contour extraction, the O(pixels) half:
// Raster.get_contour -> _get_raster_contour_windowed   ocsmesh/raster.py:1340-1372, 1426-1470
// The reference traces iso-lines with matplotlib/contourpy (absent here, its
// algorithm is restated as marching squares with linear edge interpolation
// and the cell mean as saddle decider) and merges them with shapely.  This
// kernel sweeps the DEM once per pass (4 B/px): every 2x2 cell is classified,
// its 0/1/2 line segments are compacted in cell order (count / scan / emit)
// as pairs of GLOBAL EDGE IDS plus the two crossing points in window
// coordinates.  Stitching the O(segments) list into polylines stays on the
// host (ocsmesh_b200/features.py), as does the arc-length resampling.
//
//   corner bits : 1 = (j,i)  2 = (j,i+1)  4 = (j+1,i+1)  8 = (j+1,i), z > level
//   edge ids    : horizontal (j,i)-(j,i+1) ->  j*(w-1) + i
//                 vertical   (j,i)-(j+1,i) -> -(j*w + i) - 1
//   crossing    : t = (level - z0) / (z1 - z0);  p = p0 + t*(p1 - p0)   (f64)
#include "ocsb_common.cuh"
#include "k_scan.cuh"

namespace {

constexpr int kCellThreads = 1024;

struct CellSegs {
    int n;              // 0, 1, 2
    int e[2][2];        // per segment: two edge slots (0 top, 1 rig, 2 bot, 3 lef)
};

template <typename T>
__device__ __forceinline__ CellSegs classify(const T *__restrict__ dem, int64_t zs,
                                             int j, int i, double level,
                                             double z4[4]) {
    const T *r0 = dem + (int64_t)j * zs + i;
    const T *r1 = r0 + zs;
    const double z00 = (double)r0[0], z01 = (double)r0[1];
    const double z10 = (double)r1[0], z11 = (double)r1[1];
    z4[0] = z00; z4[1] = z01; z4[2] = z11; z4[3] = z10;
    const int code = (z00 > level ? 1 : 0) | (z01 > level ? 2 : 0) |
                     (z11 > level ? 4 : 0) | (z10 > level ? 8 : 0);
    CellSegs s;
    s.n = 0;
    s.e[0][0] = s.e[0][1] = s.e[1][0] = s.e[1][1] = 0;
    auto seg = [&](int a, int b) { s.e[s.n][0] = a; s.e[s.n][1] = b; ++s.n; };
    enum { TOP = 0, RIG = 1, BOT = 2, LEF = 3 };
    switch (code) {
    case 1: case 14: seg(TOP, LEF); break;
    case 2: case 13: seg(TOP, RIG); break;
    case 4: case 11: seg(RIG, BOT); break;
    case 8: case 7:  seg(LEF, BOT); break;
    case 3: case 12: seg(LEF, RIG); break;
    case 6: case 9:  seg(TOP, BOT); break;
    case 5: case 10: {
        // saddle: the cell mean decides (sum in numpy's left-to-right order)
        const double m = 0.25 * (((z00 + z01) + z10) + z11);
        const bool cen = m > level;
        if ((code == 5) == cen) { seg(TOP, RIG); seg(LEF, BOT); }
        else { seg(TOP, LEF); seg(RIG, BOT); }
        break;
    }
    default: break;
    }
    return s;
}

template <typename T>
__global__ void __launch_bounds__(kCellThreads)
contour_count_kernel(const T *__restrict__ dem, int64_t zs, int w, int jb, int je,
                     double level, int *__restrict__ counts) {
    const int64_t ncell = (int64_t)(je - jb) * (w - 1);
    const int64_t c = (int64_t)blockIdx.x * kCellThreads + threadIdx.x;
    int n = 0;
    if (c < ncell) {
        const int j = jb + (int)(c / (w - 1)), i = (int)(c % (w - 1));
        double z4[4];
        n = classify(dem, zs, j, i, level, z4).n;
    }
    // block sum of n in {0,1,2}
    const int n1 = __syncthreads_count(n >= 1);
    const int n2 = __syncthreads_count(n == 2);
    if (threadIdx.x == 0) counts[blockIdx.x] = n1 + n2;
}

template <typename T>
__global__ void __launch_bounds__(kCellThreads)
contour_emit_kernel(const T *__restrict__ dem, int64_t zs, int w, int jb, int je,
                    double level, LinAxis ax, LinAxis ay,
                    const int *__restrict__ offsets, int64_t *__restrict__ edges,
                    double *__restrict__ pts) {
    __shared__ int s_warp[kCellThreads / 32];
    const int64_t ncell = (int64_t)(je - jb) * (w - 1);
    const int64_t c = (int64_t)blockIdx.x * kCellThreads + threadIdx.x;
    CellSegs s;
    s.n = 0;
    double z4[4] = {0, 0, 0, 0};
    int j = 0, i = 0;
    if (c < ncell) {
        j = jb + (int)(c / (w - 1));
        i = (int)(c % (w - 1));
        s = classify(dem, zs, j, i, level, z4);
    }
    // exclusive scan of s.n within the CTA
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int x = s.n;
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
    if (s.n == 0) return;
    int64_t pos = (int64_t)offsets[blockIdx.x] + (wid ? s_warp[wid - 1] : 0) + (x - s.n);
    const double x0 = lin_at(ax, i), x1 = lin_at(ax, i + 1);
    const double y0 = lin_at(ay, j), y1 = lin_at(ay, j + 1);
    // z4 = (j,i), (j,i+1), (j+1,i+1), (j+1,i)
    auto cross = [&](int slot, int64_t *id, double *px, double *py) {
        double za, zb;
        switch (slot) {
        case 0:  // top: (j,i)-(j,i+1)
            za = z4[0]; zb = z4[1];
            *id = (int64_t)j * (w - 1) + i;
            *px = __dadd_rn(x0, __dmul_rn(__ddiv_rn(level - za, zb - za), x1 - x0));
            *py = y0;
            break;
        case 2:  // bottom: (j+1,i)-(j+1,i+1)
            za = z4[3]; zb = z4[2];
            *id = (int64_t)(j + 1) * (w - 1) + i;
            *px = __dadd_rn(x0, __dmul_rn(__ddiv_rn(level - za, zb - za), x1 - x0));
            *py = y1;
            break;
        case 3:  // left: (j,i)-(j+1,i)
            za = z4[0]; zb = z4[3];
            *id = -((int64_t)j * w + i) - 1;
            *px = x0;
            *py = __dadd_rn(y0, __dmul_rn(__ddiv_rn(level - za, zb - za), y1 - y0));
            break;
        default:  // right: (j,i+1)-(j+1,i+1)
            za = z4[1]; zb = z4[2];
            *id = -((int64_t)j * w + i + 1) - 1;
            *px = x1;
            *py = __dadd_rn(y0, __dmul_rn(__ddiv_rn(level - za, zb - za), y1 - y0));
            break;
        }
    };
    for (int k = 0; k < s.n; ++k, ++pos) {
        int64_t ia, ib;
        double ax_, ay_, bx_, by_;
        cross(s.e[k][0], &ia, &ax_, &ay_);
        cross(s.e[k][1], &ib, &bx_, &by_);
        edges[2 * pos] = ia;
        edges[2 * pos + 1] = ib;
        pts[4 * pos + 0] = ax_; pts[4 * pos + 1] = ay_;
        pts[4 * pos + 2] = bx_; pts[4 * pos + 3] = by_;
    }
}

}  // namespace

// cells of window rows [jb, je) (je <= h - 1).  edges == nullptr: count only.
int32_t k_contour_segments(ocsb_ctx *ctx, const void *dem, int dem_dtype,
                           int64_t zs, int w, int jb, int je, double level,
                           LinAxis ax, LinAxis ay, int64_t cap, int64_t *edges,
                           double *pts, int64_t *n_out) {
    *n_out = 0;
    if (je <= jb || w < 2) return OCSB_OK;
    const int64_t ncell = (int64_t)(je - jb) * (w - 1);
    const int64_t nblk = (ncell + kCellThreads - 1) / kCellThreads;
    ScratchGuard sg(ctx);
    int *counts = nullptr;
    unsigned long long *total = nullptr;
    OCSB_TRY(sg.get(&counts, (size_t)nblk));
    OCSB_TRY(sg.get(&total, 1));
    OCSB_TIME(ctx, KC_CONTOUR);
    if (dem_dtype == OCSB_F32)
        contour_count_kernel<float><<<(unsigned)nblk, kCellThreads, 0, ctx->stream>>>(
            (const float *)dem, zs, w, jb, je, level, counts);
    else
        contour_count_kernel<double><<<(unsigned)nblk, kCellThreads, 0, ctx->stream>>>(
            (const double *)dem, zs, w, jb, je, level, counts);
    OCSB_LAUNCHED(ctx);
    OCSB_TIME(ctx, KC_CONTOUR);
    scan_counts_kernel<<<1, 1024, 0, ctx->stream>>>(counts, nblk, total);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    unsigned long long host_total = 0;
    unsigned long long *bounce =
        ctx->pinned ? (unsigned long long *)ctx->pinned : &host_total;
    OCSB_CUDA(cudaMemcpyAsync(bounce, total, sizeof(*total), cudaMemcpyDeviceToHost,
                              ctx->stream));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    host_total = *bounce;
    OCSB_REQUIRE(host_total < 2147483647ull,
                 "ocsb_contour_segments: %llu segments exceed the limit", host_total);
    *n_out = (int64_t)host_total;
    if (!edges || host_total == 0) return OCSB_OK;
    if ((int64_t)host_total > cap) {
        ocsb_set_error("ocsb_contour_segments: %llu segments, capacity %lld",
                       host_total, (long long)cap);
        return OCSB_EOVERFLOW;
    }
    OCSB_TIME(ctx, KC_CONTOUR);
    if (dem_dtype == OCSB_F32)
        contour_emit_kernel<float><<<(unsigned)nblk, kCellThreads, 0, ctx->stream>>>(
            (const float *)dem, zs, w, jb, je, level, ax, ay, counts, edges, pts);
    else
        contour_emit_kernel<double><<<(unsigned)nblk, kCellThreads, 0, ctx->stream>>>(
            (const double *)dem, zs, w, jb, je, level, ax, ay, counts, edges, pts);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
