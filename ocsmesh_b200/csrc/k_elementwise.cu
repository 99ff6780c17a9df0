// Streaming kernels: init fill (A4), global clamp (A10), min-combine (A12),
// window coordinates (A2).  All HBM-bound, < 1 flop/B: 16-byte accesses,
// grids sized in multiples of the SM count, grid-stride loops.
#include "ocsb_common.cuh"

namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
fill_kernel(float *__restrict__ dst, int64_t n, float v) {
    // head (unaligned) + float4 body + tail
    uintptr_t addr = (uintptr_t)dst;
    int64_t head = ((16 - (addr & 15)) & 15) / 4;
    if (head > n) head = n;
    int64_t body4 = (n - head) / 4;
    int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t nth = (int64_t)gridDim.x * blockDim.x;
    if (tid < head) dst[tid] = v;
    float4 *d4 = reinterpret_cast<float4 *>(dst + head);
    const float4 v4 = make_float4(v, v, v, v);
    for (int64_t i = tid; i < body4; i += nth) __stcs(d4 + i, v4);
    int64_t tail0 = head + body4 * 4;
    if (tid < n - tail0) dst[tail0 + tid] = v;
}

// generic 2-D row-window elementwise driver: each thread owns 4 consecutive
// pixels of one row; vector access when the row segment is 16-byte aligned
template <typename F>
__device__ __forceinline__ void rows4(int w, int h, F f) {
    const int w4 = (w + 3) >> 2;
    const int64_t total = (int64_t)w4 * h;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int row = (int)(i / w4);
        int col = (int)(i - (int64_t)row * w4) << 2;
        f(row, col, min(4, w - col));
    }
}

__global__ void __launch_bounds__(kThreads)
clamp_kernel(float *__restrict__ p, int64_t stride, int w, int h, int has_lo,
             float lo, int has_hi, float hi, int vec_ok) {
    rows4(w, h, [&](int row, int col, int n) {
        float *q = p + (int64_t)row * stride + col;
        if (vec_ok && n == 4) {
            float4 v = *reinterpret_cast<const float4 *>(q);
            v.x = clamp_f32(v.x, has_lo, lo, has_hi, hi);
            v.y = clamp_f32(v.y, has_lo, lo, has_hi, hi);
            v.z = clamp_f32(v.z, has_lo, lo, has_hi, hi);
            v.w = clamp_f32(v.w, has_lo, lo, has_hi, hi);
            *reinterpret_cast<float4 *>(q) = v;
        } else {
            for (int k = 0; k < n; ++k)
                q[k] = clamp_f32(q[k], has_lo, lo, has_hi, hi);
        }
    });
}

__global__ void __launch_bounds__(kThreads)
min_kernel(float *__restrict__ d, int64_t ds, const float *__restrict__ s,
           int64_t ss, int w, int h, int vec_ok) {
    rows4(w, h, [&](int row, int col, int n) {
        float *q = d + (int64_t)row * ds + col;
        const float *r = s + (int64_t)row * ss + col;
        if (vec_ok && n == 4) {
            float4 a = *reinterpret_cast<const float4 *>(q);
            float4 b = __ldcs(reinterpret_cast<const float4 *>(r));
            a.x = np_minimum_f(a.x, b.x);
            a.y = np_minimum_f(a.y, b.y);
            a.z = np_minimum_f(a.z, b.z);
            a.w = np_minimum_f(a.w, b.w);
            *reinterpret_cast<float4 *>(q) = a;
        } else {
            for (int k = 0; k < n; ++k) q[k] = np_minimum_f(q[k], r[k]);
        }
    });
}

__global__ void coords_kernel(LinAxis ax, LinAxis ay, double *__restrict__ x,
                              double *__restrict__ y) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ax.n) x[i] = lin_at(ax, i);
    if (i < ay.n) y[i] = lin_at(ay, i);
}

inline int grid_for(const ocsb_ctx *ctx, int64_t work_items) {
    int64_t blocks = (work_items + kThreads - 1) / kThreads;
    int64_t cap = (int64_t)ctx->num_sms * 8;  // 8 resident CTAs of 256 per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

inline int vec_ok2d(const void *p, int64_t stride) {
    return (((uintptr_t)p & 15) == 0) && (stride % 4 == 0);
}

}  // namespace

int32_t k_fill_f32(ocsb_ctx *ctx, float *dst, int64_t n, float v) {
    OCSB_TIME(ctx, KC_FILL);
    fill_kernel<<<grid_for(ctx, n / 4 + 1), kThreads, 0, ctx->stream>>>(dst, n, v);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_clamp(ocsb_ctx *ctx, float *hfun, int64_t stride, int w, int h,
                double hmin, double hmax) {
    int has_lo = !ocsb_is_none(hmin), has_hi = !ocsb_is_none(hmax);
    if (!has_lo && !has_hi) return OCSB_OK;
    int64_t items = (int64_t)((w + 3) / 4) * h;
    OCSB_TIME(ctx, KC_CLAMP);
    clamp_kernel<<<grid_for(ctx, items), kThreads, 0, ctx->stream>>>(
        hfun, stride, w, h, has_lo, (float)hmin, has_hi, (float)hmax,
        vec_ok2d(hfun, stride));
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_min_combine(ocsb_ctx *ctx, float *dst, int64_t dstride,
                      const float *src, int64_t sstride, int w, int h) {
    int64_t items = (int64_t)((w + 3) / 4) * h;
    OCSB_TIME(ctx, KC_MIN_COMBINE);
    min_kernel<<<grid_for(ctx, items), kThreads, 0, ctx->stream>>>(
        dst, dstride, src, sstride, w, h,
        vec_ok2d(dst, dstride) && vec_ok2d(src, sstride));
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_coords(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, double *x, double *y) {
    int n = ax.n > ay.n ? ax.n : ay.n;
    OCSB_TIME(ctx, KC_COORDS);
    coords_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ax, ay, x, y);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}


// ---------------------------------------------------------------------------
// per-step exchange of a row-band run (ocsmesh_b200.distributed.BandExchange):
// ONE launch packs a rank's block [first owned DEM row | last owned DEM row |
// its contour vertices, padded to the block capacity with repeats of the last
// vertex (or a far point when it has none)], ONE launch unpacks the gathered
// blocks into the contiguous vertex array and the two halo rows.
// ---------------------------------------------------------------------------
namespace {

__global__ void band_pack_kernel(const double2 *verts, long long n, long long cap,
                                 const float *row_first, const float *row_last, int w,
                                 long long row_bytes, unsigned char *send, double far) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nrow = 2ll * w;
    if (i < nrow) {
        float *dst = reinterpret_cast<float *>(send);
        dst[i] = i < w ? row_first[i] : row_last[i - w];
    }
    if (i < cap) {
        double2 *dv = reinterpret_cast<double2 *>(send + row_bytes);
        dv[i] = n > 0 ? verts[i < n ? i : n - 1] : make_double2(far, far);
    }
}

__global__ void band_unpack_kernel(const unsigned char *recv, int world, int rank,
                                   long long blk_bytes, long long row_bytes, long long cap,
                                   int w, double2 *verts_out, float *halo_above,
                                   float *halo_below) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)world * cap) {
        const long long r = i / cap, k = i - r * cap;
        verts_out[i] = reinterpret_cast<const double2 *>(recv + r * blk_bytes + row_bytes)[k];
    }
    if (i < w) {
        if (halo_above)     // last owned row of the rank above
            halo_above[i] = reinterpret_cast<const float *>(recv + (rank - 1) * blk_bytes)[w + i];
        if (halo_below)     // first owned row of the rank below
            halo_below[i] = reinterpret_cast<const float *>(recv + (rank + 1) * blk_bytes)[i];
    }
}

}  // namespace

int32_t k_band_pack(ocsb_ctx *ctx, const double *verts, int64_t n, int64_t cap,
                    const float *row_first, const float *row_last, int w, int64_t row_bytes,
                    void *send, double far) {
    const long long items = std::max<long long>(2ll * w, cap);
    if (items <= 0) return OCSB_OK;
    OCSB_TIME(ctx, KC_COORDS);
    band_pack_kernel<<<(unsigned)((items + 255) / 256), 256, 0, ctx->stream>>>(
        reinterpret_cast<const double2 *>(verts), n, cap, row_first, row_last, w, row_bytes,
        reinterpret_cast<unsigned char *>(send), far);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_band_unpack(ocsb_ctx *ctx, const void *recv, int world, int rank, int64_t blk_bytes,
                      int64_t row_bytes, int64_t cap, int w, double *verts_out,
                      float *halo_above, float *halo_below) {
    const long long items = std::max<long long>((long long)world * cap, w);
    if (items <= 0) return OCSB_OK;
    OCSB_TIME(ctx, KC_COORDS);
    band_unpack_kernel<<<(unsigned)((items + 255) / 256), 256, 0, ctx->stream>>>(
        reinterpret_cast<const unsigned char *>(recv), world, rank, blk_bytes, row_bytes, cap, w,
        reinterpret_cast<double2 *>(verts_out), halo_above, halo_below);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
