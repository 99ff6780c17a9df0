// Sampling side of HfunRaster.msh_t:
//   source  : coords = get_xy(win)[::step, ::step], values = float64(hfun)
//             [::step, ::step], clamped to [hmin, hmax] in float64
//                                               ocsmesh/hfun/raster.py:345-377
//   resample: utils.interpolate(src, dst, method='nearest') = scipy griddata
//             nearest = cKDTree(src.coords).query(dst.coords) -> src.values[i]
//                                               ocsmesh/utils.py:1122-1140
// The source points form a regular (strided) grid, so the Euclidean nearest
// source point of a node is the per-axis nearest column and row: the kernel
// locates them from the affine guess and settles the choice by the exact f64
// per-axis distances (what the kd-tree compares), one thread per node:
// 16 B coords in + one 4 B texel + 8 B value out = 28 B/node.
#include "ocsb_common.cuh"

namespace {

// index m in [0, n) of the strided axis sample lin_at(a, m*step) nearest to q
__device__ __forceinline__ int nearest_axis(const LinAxis &a, int step, int n,
                                            double q) {
    const double pitch = a.step * (double)step;
    int g = 0;
    if (n > 1 && pitch != 0.0) {
        double t = floor((q - a.start) / pitch);
        if (!(t > 0.0)) t = 0.0;
        if (t > (double)(n - 1)) t = (double)(n - 1);
        g = (int)t;
    }
    int best = -1;
    double bd = INFINITY;
    // the guess is off by at most one sample; look one further for safety
    for (int m = max(0, g - 1); m <= min(n - 1, g + 2); ++m) {
        const double d = q - lin_at(a, m * step);
        const double d2 = __dmul_rn(d, d);
        if (d2 < bd) { bd = d2; best = m; }
    }
    return best;
}

__global__ void __launch_bounds__(256)
sample_nearest_kernel(LinAxis ax, LinAxis ay, const float *__restrict__ hfun,
                      int64_t hs, int step, int nx, int ny, int has_lo, double lo,
                      int has_hi, double hi, const double2 *__restrict__ xy,
                      int64_t K, double *__restrict__ out) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < K;
         t += (int64_t)gridDim.x * blockDim.x) {
        const double2 p = xy[t];
        const int m = nearest_axis(ax, step, nx, p.x);
        const int r = nearest_axis(ay, step, ny, p.y);
        double v = (double)hfun[(int64_t)(r * step) * hs + (int64_t)m * step];
        if (has_lo && v < lo) v = lo;
        if (has_hi && v > hi) v = hi;
        out[t] = v;
    }
}

__global__ void __launch_bounds__(256)
sample_source_kernel(LinAxis ax, LinAxis ay, const float *__restrict__ hfun,
                     int64_t hs, int step, int nx, int ny, int has_lo, double lo,
                     int has_hi, double hi, double2 *__restrict__ coords,
                     double *__restrict__ values) {
    const int64_t total = (int64_t)nx * ny;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(t / nx), m = (int)(t - (int64_t)r * nx);
        if (coords) coords[t] = make_double2(lin_at(ax, m * step), lin_at(ay, r * step));
        if (values) {
            double v = (double)hfun[(int64_t)(r * step) * hs + (int64_t)m * step];
            if (has_lo && v < lo) v = lo;
            if (has_hi && v > hi) v = hi;
            values[t] = v;
        }
    }
}

}  // namespace

int32_t k_sample_nearest(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const float *hfun,
                         int64_t hs, int step, double hmin, double hmax,
                         const double *xy, int64_t K, double *out) {
    const int nx = (ax.n + step - 1) / step, ny = (ay.n + step - 1) / step;
    int64_t blocks = (K + 255) / 256;
    if (blocks > (int64_t)ctx->num_sms * 16) blocks = (int64_t)ctx->num_sms * 16;
    OCSB_TIME(ctx, KC_SAMPLE);
    sample_nearest_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(
        ax, ay, hfun, hs, step, nx, ny, !ocsb_is_none(hmin), hmin,
        !ocsb_is_none(hmax), hmax, reinterpret_cast<const double2 *>(xy), K, out);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_sample_source(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const float *hfun,
                        int64_t hs, int step, double hmin, double hmax,
                        double *coords, double *values) {
    const int nx = (ax.n + step - 1) / step, ny = (ay.n + step - 1) / step;
    int64_t blocks = ((int64_t)nx * ny + 255) / 256;
    if (blocks > (int64_t)ctx->num_sms * 16) blocks = (int64_t)ctx->num_sms * 16;
    OCSB_TIME(ctx, KC_SAMPLE);
    sample_source_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(
        ax, ay, hfun, hs, step, nx, ny, !ocsb_is_none(hmin), hmin,
        !ocsb_is_none(hmax), hmax, reinterpret_cast<double2 *>(coords), values);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
