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
// A CTA = 256 threads x 4 consecutive cells of ONE cell row (float4 loads of
// the two DEM rows), CTAs in row-major order.
//
//   corner bits : 1 = (j,i)  2 = (j,i+1)  4 = (j+1,i+1)  8 = (j+1,i), z > level
//   edge ids    : horizontal (j,i)-(j,i+1) ->  j*(w-1) + i
//                 vertical   (j,i)-(j+1,i) -> -(j*w + i) - 1
//   crossing    : t = (level - z0) / (z1 - z0);  p = p0 + t*(p1 - p0)   (f64)
#include "ocsb_common.cuh"
#include "k_scan.cuh"

#include <algorithm>

namespace {

constexpr int kCellsPerThread = 4;
constexpr int kCellThreads = 256;
constexpr int kCellsPerCta = kCellsPerThread * kCellThreads;   // 1024 cells of ONE cell row

enum { TOP = 0, RIG = 1, BOT = 2, LEF = 3 };

// segments of one cell, packed: bits 0-1 = count, then (a, b) edge slots of
// segment 0 in bits 2-3 / 4-5 and of segment 1 in bits 6-7 / 8-9
__device__ __forceinline__ unsigned pack1(int a, int b) { return 1u | (a << 2) | (b << 4); }
__device__ __forceinline__ unsigned pack2(int a, int b, int c, int d) {
    return 2u | (a << 2) | (b << 4) | (c << 6) | (d << 8);
}

// corner code of a cell.  `lt` is the threshold in the DEM dtype: for a
// float32 DEM the largest float <= level, so that z > lt  <=>  (double)z > level
// holds exactly and the sweep needs no f32 -> f64 conversions (XU pipe).
template <typename T>
__device__ __forceinline__ int cell_code(T z00, T z01, T z10, T z11, T lt) {
    return (z00 > lt ? 1 : 0) | (z01 > lt ? 2 : 0) | (z11 > lt ? 4 : 0) |
           (z10 > lt ? 8 : 0);
}

__device__ __forceinline__ int code_count(int code) {
    // 0, 15 -> 0; 5, 10 (saddles) -> 2; everything else -> 1
    return (code == 0 || code == 15) ? 0 : ((code == 5 || code == 10) ? 2 : 1);
}

__device__ __forceinline__ unsigned classify(int code, double z00, double z01,
                                             double z10, double z11, double level) {
    switch (code) {
    case 1: case 14: return pack1(TOP, LEF);
    case 2: case 13: return pack1(TOP, RIG);
    case 4: case 11: return pack1(RIG, BOT);
    case 8: case 7:  return pack1(LEF, BOT);
    case 3: case 12: return pack1(LEF, RIG);
    case 6: case 9:  return pack1(TOP, BOT);
    case 5: case 10: {
        // saddle: the cell mean decides (sum in numpy's left-to-right order)
        const double m = 0.25 * (((z00 + z01) + z10) + z11);
        const bool cen = m > level;
        return ((code == 5) == cen) ? pack2(TOP, RIG, LEF, BOT)
                                    : pack2(TOP, LEF, RIG, BOT);
    }
    default: return 0u;
    }
}

// rows j and j + 1, columns i .. i + 4 (the 4 cells of one thread), as f64;
// columns past the raster are filled with the last valid value (their cells
// are never classified)
template <typename T>
__device__ __forceinline__ void load_rows(const T *__restrict__ dem, int64_t zs,
                                          int j, int i, int w, bool vec,
                                          T r0[5], T r1[5]) {
    const T *p0 = dem + (int64_t)j * zs + i;
    const T *p1 = p0 + zs;
    if (vec && i + 4 < w && sizeof(T) == 4) {
        const float4 a = *reinterpret_cast<const float4 *>(p0);
        const float4 b = *reinterpret_cast<const float4 *>(p1);
        r0[0] = (T)a.x; r0[1] = (T)a.y; r0[2] = (T)a.z; r0[3] = (T)a.w; r0[4] = p0[4];
        r1[0] = (T)b.x; r1[1] = (T)b.y; r1[2] = (T)b.z; r1[3] = (T)b.w; r1[4] = p1[4];
    } else {
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const int c = min(i + k, w - 1) - i;
            r0[k] = p0[c];
            r1[k] = p1[c];
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(kCellThreads)
contour_count_kernel(const T *__restrict__ dem, int64_t zs, int w, int jb, int nbx,
                     int nblk, T lt, int vec, int *__restrict__ counts) {
    __shared__ int s_warp[2][kCellThreads / 32];
    int buf = 0;
    // persistent CTAs: chunk b = 1024 cells of one cell row, row-major
    for (int b = blockIdx.x; b < nblk; b += gridDim.x, buf ^= 1) {
        const int jr = b / nbx;
        const int j = jb + jr;
        const int i = (b - jr * nbx) * kCellsPerCta + threadIdx.x * kCellsPerThread;
        int n = 0;
        if (i < w - 1) {
            T r0[5], r1[5];
            load_rows(dem, zs, j, i, w, vec != 0, r0, r1);
#pragma unroll
            for (int k = 0; k < kCellsPerThread; ++k)
                if (i + k < w - 1)
                    n += code_count(cell_code(r0[k], r0[k + 1], r1[k], r1[k + 1], lt));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
        if ((threadIdx.x & 31) == 0) s_warp[buf][threadIdx.x >> 5] = n;
        __syncthreads();   // double-buffered: one barrier per chunk is enough
        if (threadIdx.x == 0) {
            int t = 0;
#pragma unroll
            for (int k = 0; k < kCellThreads / 32; ++k) t += s_warp[buf][k];
            counts[b] = t;
        }
    }
}

// the same for up to kMaxLevels levels in ONE sweep of the DEM (add_contour with
// a level list: the count sweep is the O(pixels) part, the emit sweep skips the
// chunks without segments): counts[l * nblk + b]
constexpr int kMaxLevels = 4;
template <typename T> struct LevelSet { T lt[kMaxLevels]; int n; };

template <typename T>
__global__ void __launch_bounds__(kCellThreads)
contour_count_multi_kernel(const T *__restrict__ dem, int64_t zs, int w, int jb, int nbx,
                           int nblk, LevelSet<T> L, int vec, int *__restrict__ counts) {
    __shared__ int s_warp[2][kMaxLevels][kCellThreads / 32];
    int buf = 0;
    for (int b = blockIdx.x; b < nblk; b += gridDim.x, buf ^= 1) {
        const int jr = b / nbx;
        const int j = jb + jr;
        const int i = (b - jr * nbx) * kCellsPerCta + threadIdx.x * kCellsPerThread;
        int n[kMaxLevels];
#pragma unroll
        for (int l = 0; l < kMaxLevels; ++l) n[l] = 0;
        if (i < w - 1) {
            T r0[5], r1[5];
            load_rows(dem, zs, j, i, w, vec != 0, r0, r1);
#pragma unroll
            for (int l = 0; l < kMaxLevels; ++l)
                if (l < L.n) {
#pragma unroll
                    for (int k = 0; k < kCellsPerThread; ++k)
                        if (i + k < w - 1)
                            n[l] += code_count(
                                cell_code(r0[k], r0[k + 1], r1[k], r1[k + 1], L.lt[l]));
                }
        }
#pragma unroll
        for (int l = 0; l < kMaxLevels; ++l) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) n[l] += __shfl_xor_sync(0xffffffffu, n[l], o);
            if ((threadIdx.x & 31) == 0) s_warp[buf][l][threadIdx.x >> 5] = n[l];
        }
        __syncthreads();   // double-buffered: one barrier per chunk is enough
        if ((int)threadIdx.x < L.n) {
            int t = 0;
#pragma unroll
            for (int k = 0; k < kCellThreads / 32; ++k) t += s_warp[buf][threadIdx.x][k];
            counts[(int64_t)threadIdx.x * nblk + b] = t;
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(kCellThreads)
contour_emit_kernel(const T *__restrict__ dem, int64_t zs, int w, int jb, int nbx,
                    int nblk, double level, T lt, int vec, LinAxis ax, LinAxis ay,
                    const int *__restrict__ offsets,
                    const unsigned long long *__restrict__ total,
                    int64_t *__restrict__ edges, double *__restrict__ pts) {
    __shared__ int s_warp[2][kCellThreads / 32];
    int buf = 0;
    const int ntot = (int)*total;
    for (int b = blockIdx.x; b < nblk; b += gridDim.x) {
    // the count pass already knows which chunks hold segments: the others
    // (the vast majority) cost two integer loads
    const int off_b = offsets[b];
    if ((b + 1 < nblk ? offsets[b + 1] : ntot) == off_b) continue;
    buf ^= 1;
    const int jr = b / nbx;
    const int j = jb + jr;
    const int i = (b - jr * nbx) * kCellsPerCta + threadIdx.x * kCellsPerThread;
    T r0[5], r1[5];
    unsigned pk[kCellsPerThread];
    int n = 0;
#pragma unroll
    for (int k = 0; k < kCellsPerThread; ++k) pk[k] = 0u;
    if (i < w - 1) {
        load_rows(dem, zs, j, i, w, vec != 0, r0, r1);
#pragma unroll
        for (int k = 0; k < kCellsPerThread; ++k)
            if (i + k < w - 1) {
                const int code = cell_code(r0[k], r0[k + 1], r1[k], r1[k + 1], lt);
                if (code != 0 && code != 15)
                    pk[k] = classify(code, (double)r0[k], (double)r0[k + 1],
                                     (double)r1[k], (double)r1[k + 1], level);
                n += (int)(pk[k] & 3u);
            }
    }
    // exclusive scan of n within the CTA
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int x = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[buf][wid] = x;
    __syncthreads();
    int before = x - n;
    for (int k = 0; k < wid; ++k) before += s_warp[buf][k];
    if (n == 0) continue;
    int64_t pos = (int64_t)off_b + before;
    const double y0 = lin_at(ay, j), y1 = lin_at(ay, j + 1);
#pragma unroll
    for (int k = 0; k < kCellsPerThread; ++k) {
        const int ns = (int)(pk[k] & 3u);
        if (ns == 0) continue;
        const int ic = i + k;
        const double x0 = lin_at(ax, ic), x1 = lin_at(ax, ic + 1);
        const double z00 = (double)r0[k], z01 = (double)r0[k + 1];
        const double z10 = (double)r1[k], z11 = (double)r1[k + 1];
        auto cross = [&](int slot, int64_t *id, double *px, double *py) {
            double za, zb;
            switch (slot) {
            case TOP:  // (j,i)-(j,i+1)
                za = z00; zb = z01;
                *id = (int64_t)j * (w - 1) + ic;
                *px = __dadd_rn(x0, __dmul_rn(__ddiv_rn(level - za, zb - za), x1 - x0));
                *py = y0;
                break;
            case BOT:  // (j+1,i)-(j+1,i+1)
                za = z10; zb = z11;
                *id = (int64_t)(j + 1) * (w - 1) + ic;
                *px = __dadd_rn(x0, __dmul_rn(__ddiv_rn(level - za, zb - za), x1 - x0));
                *py = y1;
                break;
            case LEF:  // (j,i)-(j+1,i)
                za = z00; zb = z10;
                *id = -((int64_t)j * w + ic) - 1;
                *px = x0;
                *py = __dadd_rn(y0, __dmul_rn(__ddiv_rn(level - za, zb - za), y1 - y0));
                break;
            default:   // RIG: (j,i+1)-(j+1,i+1)
                za = z01; zb = z11;
                *id = -((int64_t)j * w + ic + 1) - 1;
                *px = x1;
                *py = __dadd_rn(y0, __dmul_rn(__ddiv_rn(level - za, zb - za), y1 - y0));
                break;
            }
        };
        for (int q = 0; q < ns; ++q, ++pos) {
            int64_t ia, ib;
            double ax_, ay_, bx_, by_;
            cross((pk[k] >> (2 + 4 * q)) & 3u, &ia, &ax_, &ay_);
            cross((pk[k] >> (4 + 4 * q)) & 3u, &ib, &bx_, &by_);
            reinterpret_cast<longlong2 *>(edges)[pos] = make_longlong2(ia, ib);
            reinterpret_cast<double2 *>(pts)[2 * pos] = make_double2(ax_, ay_);
            reinterpret_cast<double2 *>(pts)[2 * pos + 1] = make_double2(bx_, by_);
        }
    }
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
    OCSB_REQUIRE((((uintptr_t)edges | (uintptr_t)pts) & 15) == 0,
                 "ocsb_contour_segments: edges / pts must be 16-byte aligned");
    const int nbx = (w - 1 + kCellsPerCta - 1) / kCellsPerCta;
    const int vec = (((uintptr_t)dem & 15) == 0 && zs % 4 == 0) ? 1 : 0;
    const int64_t nblk = (int64_t)nbx * (je - jb);
    OCSB_REQUIRE(nblk < 2147483647ll, "ocsb_contour_segments: window too large");
    const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)ctx->num_sms * 8);
    // largest float <= level: z > ltf  <=>  (double)z > level for every float z
    float ltf = (float)level;
    if ((double)ltf > level) ltf = nextafterf(ltf, -INFINITY);
    ScratchGuard sg(ctx);
    int *counts = nullptr, *sums = nullptr;
    unsigned long long *total = nullptr;
    OCSB_TRY(sg.get(&counts, (size_t)nblk));
    OCSB_TRY(sg.get(&sums, (size_t)((nblk + 1023) / 1024)));
    OCSB_TRY(sg.get(&total, 1));
    OCSB_TIME(ctx, KC_CONTOUR);
    if (dem_dtype == OCSB_F32)
        contour_count_kernel<float><<<grid, kCellThreads, 0, ctx->stream>>>(
            (const float *)dem, zs, w, jb, nbx, (int)nblk, ltf, vec, counts);
    else
        contour_count_kernel<double><<<grid, kCellThreads, 0, ctx->stream>>>(
            (const double *)dem, zs, w, jb, nbx, (int)nblk, level, vec, counts);
    OCSB_LAUNCHED(ctx);
    OCSB_TIME(ctx, KC_CONTOUR);
    scan_counts(counts, nblk, sums, total, ctx->stream);
    OCSB_LAUNCHED(ctx);
    ctx->launches += 2;
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
        contour_emit_kernel<float><<<grid, kCellThreads, 0, ctx->stream>>>(
            (const float *)dem, zs, w, jb, nbx, (int)nblk, level, ltf, vec, ax, ay, counts, total,
            edges, pts);
    else
        contour_emit_kernel<double><<<grid, kCellThreads, 0, ctx->stream>>>(
            (const double *)dem, zs, w, jb, nbx, (int)nblk, level, level, vec, ax, ay, counts, total,
            edges, pts);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}


// ---- level lists: one count sweep for all levels, one emit launch per level ----
int k_contour_max_levels() { return kMaxLevels; }

static float level_threshold_f32(double level) {
    // largest float <= level: z > ltf  <=>  (double)z > level for every float z
    float ltf = (float)level;
    if ((double)ltf > level) ltf = nextafterf(ltf, -INFINITY);
    return ltf;
}

// counts / offsets of `nl` levels (nl <= k_contour_max_levels()): after the call
// offsets[l * nblk + b] is the exclusive scan of level l's chunk counts,
// totals_dev[l] / totals[l] its number of segments.  One host sync.
int32_t k_contour_count_levels(ocsb_ctx *ctx, ScratchGuard &sg, const void *dem, int dem_dtype,
                               int64_t zs, int w, int jb, int je, const double *levels, int nl,
                               int **offsets, unsigned long long **totals_dev,
                               int64_t *totals, int64_t *nblk_out) {
    OCSB_REQUIRE(nl >= 1 && nl <= kMaxLevels, "k_contour_count_levels: bad level count");
    const int nbx = (w - 1 + kCellsPerCta - 1) / kCellsPerCta;
    const int vec = (((uintptr_t)dem & 15) == 0 && zs % 4 == 0) ? 1 : 0;
    const int64_t nblk = (int64_t)nbx * (je - jb);
    OCSB_REQUIRE(nblk * nl < 2147483647ll, "ocsb_contour: window too large");
    *nblk_out = nblk;
    const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)ctx->num_sms * 8);
    int *counts = nullptr, *sums = nullptr;
    unsigned long long *tot = nullptr;
    OCSB_TRY(sg.get(&counts, (size_t)(nblk * nl)));
    OCSB_TRY(sg.get(&sums, (size_t)((nblk + 1023) / 1024)));
    OCSB_TRY(sg.get(&tot, (size_t)kMaxLevels));
    OCSB_TIME(ctx, KC_CONTOUR);
    if (dem_dtype == OCSB_F32) {
        LevelSet<float> L;
        L.n = nl;
        for (int l = 0; l < kMaxLevels; ++l) L.lt[l] = level_threshold_f32(levels[l < nl ? l : 0]);
        contour_count_multi_kernel<float><<<grid, kCellThreads, 0, ctx->stream>>>(
            (const float *)dem, zs, w, jb, nbx, (int)nblk, L, vec, counts);
    } else {
        LevelSet<double> L;
        L.n = nl;
        for (int l = 0; l < kMaxLevels; ++l) L.lt[l] = levels[l < nl ? l : 0];
        contour_count_multi_kernel<double><<<grid, kCellThreads, 0, ctx->stream>>>(
            (const double *)dem, zs, w, jb, nbx, (int)nblk, L, vec, counts);
    }
    OCSB_LAUNCHED(ctx);
    for (int l = 0; l < nl; ++l) {
        OCSB_TIME(ctx, KC_CONTOUR);
        scan_counts(counts + (int64_t)l * nblk, nblk, sums, tot + l, ctx->stream);
        OCSB_LAUNCHED(ctx);
        ctx->launches += 2;
    }
    OCSB_CHECK_LAUNCH();
    unsigned long long host_tot[kMaxLevels] = {};
    unsigned long long *bounce = ctx->pinned ? (unsigned long long *)ctx->pinned : host_tot;
    OCSB_CUDA(cudaMemcpyAsync(bounce, tot, sizeof(*tot) * nl, cudaMemcpyDeviceToHost,
                              ctx->stream));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int l = 0; l < nl; ++l) {
        OCSB_REQUIRE(bounce[l] < 2147483647ull,
                     "ocsb_contour_segments: %llu segments exceed the limit", bounce[l]);
        totals[l] = (int64_t)bounce[l];
    }
    *offsets = counts;
    *totals_dev = tot;
    return OCSB_OK;
}

// emit launch of one level from its scanned chunk offsets (edges / pts hold
// `total` segments)
int32_t k_contour_emit_level(ocsb_ctx *ctx, const void *dem, int dem_dtype, int64_t zs, int w,
                             int jb, int je, double level, LinAxis ax, LinAxis ay,
                             const int *offsets, const unsigned long long *total_dev,
                             int64_t *edges, double *pts) {
    OCSB_REQUIRE((((uintptr_t)edges | (uintptr_t)pts) & 15) == 0,
                 "ocsb_contour_segments: edges / pts must be 16-byte aligned");
    const int nbx = (w - 1 + kCellsPerCta - 1) / kCellsPerCta;
    const int vec = (((uintptr_t)dem & 15) == 0 && zs % 4 == 0) ? 1 : 0;
    const int64_t nblk = (int64_t)nbx * (je - jb);
    const unsigned grid = (unsigned)std::min<int64_t>(nblk, (int64_t)ctx->num_sms * 8);
    OCSB_TIME(ctx, KC_CONTOUR);
    if (dem_dtype == OCSB_F32)
        contour_emit_kernel<float><<<grid, kCellThreads, 0, ctx->stream>>>(
            (const float *)dem, zs, w, jb, nbx, (int)nblk, level, level_threshold_f32(level), vec,
            ax, ay, offsets, total_dev, edges, pts);
    else
        contour_emit_kernel<double><<<grid, kCellThreads, 0, ctx->stream>>>(
            (const double *)dem, zs, w, jb, nbx, (int)nblk, level, level, vec, ax, ay, offsets,
            total_dev, edges, pts);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
