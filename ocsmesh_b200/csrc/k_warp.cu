// Nearest-neighbour warp of one north-up raster onto another in the SAME
// projected CRS: what HfunCollector._create_big_raster
// asks of rasterio.warp.reproject(..., resampling=nearest,
// init_dest_nodata=False)                  ocsmesh/hfun/collector.py:2296-2442
//
// GDAL is not part of the reference tree (rasterio dependency, absent here), so
// the rule is restated from GDAL's published nearest warp kernel
// (alg/gdalwarpkernel.cpp, GWKNearestThread / GWKCheckAndComputeSrcOffsets):
//   destination pixel (r, c) -> its CENTRE (c + 0.5, r + 0.5)
//   -> georeferenced (X, Y) through the destination geotransform
//   -> source pixel coordinates (u, v) through the inverse source geotransform
//   -> source pixel (floor(u + 1e-10), floor(v + 1e-10));
//   outside the source, or a source nodata pixel: the destination pixel is left
//   as it is (init_dest_nodata=False), otherwise it takes float32(source value).
// Rasters later in the caller's list overwrite earlier ones (collector.py:2427).
// Parity with GDAL is UNPINNED (no GDAL in this image); the oracle restates the
// same rule in numpy.
//
// The index maps are separable: a CTA covers 256 destination columns x 32 rows,
// the 32 source rows go to shared memory once, every thread resolves its source
// column once -- the sweep itself is a gather + coalesced store:
// <= 4 B read + 4 B written per covered destination pixel.
#include "ocsb_common.cuh"

namespace {

struct WarpAxis {
    double d0, dstep;   // destination: coordinate of the pixel EDGE 0, pixel size
    double s0, sstep;   // source: the same
    int n_src;
};

constexpr int WARP_ROWS = 32;

// source index of destination index i along one axis, -1 = outside the source
__device__ __forceinline__ int src_index(const WarpAxis &a, int i) {
    const double X = __dadd_rn(a.d0, __dmul_rn((double)i + 0.5, a.dstep));
    const double u = __ddiv_rn(__dsub_rn(X, a.s0), a.sstep);
    const double f = floor(u + 1e-10);
    if (!(f >= 0.0) || f >= (double)a.n_src) return -1;
    return (int)f;
}

template <typename T>
__global__ void __launch_bounds__(256)
warp_nearest_kernel(WarpAxis ax, WarpAxis ay, const T *__restrict__ src, int64_t ss,
                    const uint8_t *__restrict__ mask, int64_t ms,
                    float *__restrict__ dst, int64_t ds, int c0, int c1, int r0, int r1) {
    __shared__ int rows[WARP_ROWS];
    const int rbase = r0 + (int)blockIdx.y * WARP_ROWS;
    if (threadIdx.x < WARP_ROWS) {
        const int r = rbase + (int)threadIdx.x;
        rows[threadIdx.x] = r < r1 ? src_index(ay, r) : -1;
    }
    __syncthreads();
    const int c = c0 + (int)blockIdx.x * 256 + (int)threadIdx.x;
    if (c >= c1) return;
    const int sc = src_index(ax, c);
    if (sc < 0) return;
#pragma unroll 8
    for (int k = 0; k < WARP_ROWS; ++k) {
        const int sr = rows[k];
        if (sr < 0) continue;
        if (mask && mask[(int64_t)sr * ms + sc]) continue;
        dst[(int64_t)(rbase + k) * ds + c] = (float)src[(int64_t)sr * ss + sc];
    }
}

}  // namespace

// Destination index range [lo, hi) whose centres can fall inside the source
// along one axis (one pixel of slack either side; the kernel decides exactly).
static void covered_range(const WarpAxis &a, int n_dst, int *lo, int *hi) {
    const double e0 = (a.s0 - a.d0) / a.dstep;
    const double e1 = (a.s0 + a.sstep * (double)a.n_src - a.d0) / a.dstep;
    double l = floor(e0 < e1 ? e0 : e1) - 2.0, h = ceil(e0 < e1 ? e1 : e0) + 2.0;
    if (l < 0.0) l = 0.0;
    if (h > (double)n_dst) h = (double)n_dst;
    *lo = (int)l;
    *hi = h > l ? (int)h : (int)l;
}

int32_t k_warp_nearest(ocsb_ctx *ctx, const void *src, int src_dtype, int64_t ss, int sw,
                       int sh, const double sgt[6], const uint8_t *mask, int64_t ms,
                       float *dst, int64_t ds, int dw, int dh, const double dgt[6]) {
    // gt = (a, b, c, d, e, f): x = a*col + c, y = e*row + f (b = d = 0)
    WarpAxis ax{dgt[2], dgt[0], sgt[2], sgt[0], sw};
    WarpAxis ay{dgt[5], dgt[4], sgt[5], sgt[4], sh};
    int c0, c1, r0, r1;
    covered_range(ax, dw, &c0, &c1);
    covered_range(ay, dh, &r0, &r1);
    if (c1 <= c0 || r1 <= r0) return OCSB_OK;
    dim3 grid((unsigned)((c1 - c0 + 255) / 256), (unsigned)((r1 - r0 + WARP_ROWS - 1) / WARP_ROWS));
    OCSB_TIME(ctx, KC_SAMPLE);
    if (src_dtype == OCSB_F32)
        warp_nearest_kernel<float><<<grid, 256, 0, ctx->stream>>>(
            ax, ay, (const float *)src, ss, mask, ms, dst, ds, c0, c1, r0, r1);
    else
        warp_nearest_kernel<double><<<grid, 256, 0, ctx->stream>>>(
            ax, ay, (const double *)src, ss, mask, ms, dst, ds, c0, c1, r0, r1);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
