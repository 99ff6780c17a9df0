// Per-pixel hfun ops next to the contour/limiter path:
//   add_constant_value              ocsmesh/hfun/raster.py:1438-1502
//   TopoConst/TopoFuncConstraint    ocsmesh/features/constraint.py:135-224
//   CourantNumConstraint            ocsmesh/features/constraint.py:227-317
//   Constraint._apply_rate prologue ocsmesh/features/constraint.py:83-103
//     (mask pixels -> kd-tree points; here: the mask's boundary pixels,
//      compacted in row-major order, feed the K1 nearest-vertex machinery)
// All HBM-bound streaming kernels: one thread = 4 consecutive pixels of a row
// (16-byte accesses when the row segment is aligned), grid-stride.
#include "ocsb_common.cuh"
#include "k_scan.cuh"

namespace {

constexpr int kThreads = 256;

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

inline int grid_for(const ocsb_ctx *ctx, int64_t work_items) {
    int64_t blocks = (work_items + kThreads - 1) / kThreads;
    int64_t cap = (int64_t)ctx->num_sms * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

template <typename T> struct Vec4;
template <> struct Vec4<float> { using type = float4; };
template <> struct Vec4<double> { using type = double4; };

// load n <= 4 consecutive elements (vector load when aligned and n == 4)
template <typename T>
__device__ __forceinline__ void load4(const T *p, int n, bool vec, T out[4]) {
    if (vec && n == 4) {
        if (sizeof(T) == 4) {
            const float4 v = __ldcs(reinterpret_cast<const float4 *>(p));
            out[0] = (T)v.x; out[1] = (T)v.y; out[2] = (T)v.z; out[3] = (T)v.w;
        } else {
            const double2 a = __ldcs(reinterpret_cast<const double2 *>(p));
            const double2 b = __ldcs(reinterpret_cast<const double2 *>(p) + 1);
            out[0] = (T)a.x; out[1] = (T)a.y; out[2] = (T)b.x; out[3] = (T)b.y;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) out[k] = k < n ? p[k] : (T)0;
    }
}

__device__ __forceinline__ void store4f(float *p, int n, bool vec,
                                        const float v[4]) {
    if (vec && n == 4) {
        *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
        for (int k = 0; k < n; ++k) p[k] = v[k];
    }
}

// ---- add_constant_value ----------------------------------------------------
// hfun_values[(rast > lb) & (rast < ub)] = value; min(old, .)  (:1492-1498).
// The compares run in the DEM's dtype (NEP 50: the Python-float bounds are
// cast to the array dtype), `value` is rounded to float32 by the assignment.
template <typename T>
__global__ void __launch_bounds__(kThreads)
constant_value_kernel(int w, int h, const T *__restrict__ dem, int64_t zs,
                      const float *__restrict__ old_, int64_t os,
                      float *__restrict__ new_, int64_t ns, float value, T lb,
                      T ub, int fuse, int has_lo, float lo, int has_hi,
                      float hi, int vz, int vo, int vn) {
    rows4(w, h, [&](int row, int col, int n) {
        T z[4];
        float o[4], r[4];
        load4(dem + (int64_t)row * zs + col, n, vz, z);
        if (old_) load4(old_ + (int64_t)row * os + col, n, vo, o);
        else { o[0] = o[1] = o[2] = o[3] = 3.402823466e+38f; }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float cand = (z[k] > lb && z[k] < ub) ? value : o[k];
            float v = np_minimum_f(o[k], cand);
            if (fuse) v = clamp_f32(v, has_lo, lo, has_hi, hi);
            r[k] = v;
        }
        store4f(new_ + (int64_t)row * ns + col, n, vn, r);
    });
}

// ---- TopoConstConstraint / TopoFuncConstraint.apply (mask part) -------------
//   mask = (ref > lb) & (ref < ub) & ~satisfies(old, val);  old[mask] = val
// satisfies = np.greater (MIN) / np.less (MAX).  `val` is either a constant
// (compared in float32, NEP 50) or a per-pixel array V (float32 / float64; the
// compare promotes like numpy does, the assignment rounds to float32).
template <typename T, typename V>
__global__ void __launch_bounds__(kThreads)
topo_mask_kernel(int w, int h, const T *__restrict__ dem, int64_t zs,
                 float *__restrict__ hfun, int64_t hs,
                 const V *__restrict__ vals, int64_t vs, float vconst, T lb,
                 T ub, int is_min, uint8_t *__restrict__ mask, int vz, int vh,
                 int vv, unsigned long long *__restrict__ nmask) {
    unsigned long long local = 0;
    rows4(w, h, [&](int row, int col, int n) {
        T z[4];
        float o[4];
        V v[4];
        load4(dem + (int64_t)row * zs + col, n, vz, z);
        load4((const float *)hfun + (int64_t)row * hs + col, n, vh, o);
        if (vals) load4(vals + (int64_t)row * vs + col, n, vv, v);
        else { v[0] = v[1] = v[2] = v[3] = (V)vconst; }
        bool any = false;
        uint8_t m[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool sat = is_min ? ((V)o[k] > v[k]) : ((V)o[k] < v[k]);
            const bool mk = k < n && z[k] > lb && z[k] < ub && !sat;
            m[k] = mk ? 1 : 0;
            if (mk) { o[k] = (float)v[k]; any = true; ++local; }
        }
        if (any) store4f(hfun + (int64_t)row * hs + col, n, vh, o);
        if (mask) {
            uint8_t *mp = mask + (int64_t)row * w + col;
            if (n == 4 && (w & 3) == 0) {
                *reinterpret_cast<uchar4 *>(mp) = make_uchar4(m[0], m[1], m[2], m[3]);
            } else {
                for (int k = 0; k < n; ++k) mp[k] = m[k];
            }
        }
    });
    if (nmask) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
        if ((threadIdx.x & 31) == 0 && local) atomicAdd(nmask, local);
    }
}

// ---- CourantNumConstraint.apply ---------------------------------------------
// All arithmetic in the DEM dtype T (numpy keeps float32 for a float32 DEM:
// np.zeros_like(depth), Python-float scalars are weak), IEEE sqrt and divide:
//   lin   = depth <= -|nu|                                utils.py:1741
//   u     = lin ? nu*sqrt(g/|depth|) : sqrt(g*nu)         utils.py:1784-1786
//   c     = lin ? u + sqrt(g*|depth|) : 2*u               constraint.py:296-298
//   size  = c*dt/C                                        utils.py:1704
//   C_old = c*dt/old                                      utils.py:1835
//   old[~satisfies(C_old, C)] = size                      constraint.py:309-310
template <typename T>
__device__ __forceinline__ T t_sqrt(T x);
template <> __device__ __forceinline__ float t_sqrt<float>(float x) { return __fsqrt_rn(x); }
template <> __device__ __forceinline__ double t_sqrt<double>(double x) { return __dsqrt_rn(x); }
template <typename T>
__device__ __forceinline__ T t_div(T a, T b);
template <> __device__ __forceinline__ float t_div<float>(float a, float b) { return __fdiv_rn(a, b); }
template <> __device__ __forceinline__ double t_div<double>(double a, double b) { return __ddiv_rn(a, b); }
template <typename T>
__device__ __forceinline__ T t_mul(T a, T b);
template <> __device__ __forceinline__ float t_mul<float>(float a, float b) { return __fmul_rn(a, b); }
template <> __device__ __forceinline__ double t_mul<double>(double a, double b) { return __dmul_rn(a, b); }

template <typename T>
__global__ void __launch_bounds__(kThreads)
courant_kernel(int w, int h, const T *__restrict__ dem, int64_t zs,
               float *__restrict__ hfun, int64_t hs, T g, T nu, T neg_abs_nu,
               T u_land, T dt, T cval, int is_min, int vz, int vh) {
    rows4(w, h, [&](int row, int col, int n) {
        T z[4];
        float o[4];
        load4(dem + (int64_t)row * zs + col, n, vz, z);
        load4((const float *)hfun + (int64_t)row * hs + col, n, vh, o);
        bool any = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const T az = fabs(z[k]);   // np.abs
            const bool lin = z[k] <= neg_abs_nu;
            T c;
            if (lin) {
                const T u = t_mul(nu, t_sqrt(t_div(g, az)));
                c = u + t_sqrt(t_mul(g, az));
            } else {
                c = t_mul((T)2, u_land);
            }
            const T cdt = t_mul(c, dt);
            const T size = t_div(cdt, cval);
            // element_size is the float32 hfun; numpy promotes it to T
            const T c_old = t_div(cdt, (T)o[k]);
            const bool sat = is_min ? (c_old > cval) : (c_old < cval);
            if (k < n && !sat) { o[k] = (float)size; any = true; }
        }
        if (any) store4f(hfun + (int64_t)row * hs + col, n, vh, o);
    });
}

inline int vec_ok(const void *p, int64_t stride, size_t elem) {
    const size_t per16 = 16 / elem;
    return (((uintptr_t)p & 15) == 0) && (stride % (int64_t)(per16 < 4 ? 4 : per16) == 0);
}

}  // namespace

// ---------------------------------------------------------------------------
// Constraint._apply_rate prologue: kd-tree points = coords[mask]
// (features/constraint.py:99-104).  A mask pixel whose in-window 4-neighbours
// are all mask pixels can never be the nearest mask pixel of a non-mask pixel
// (its neighbour one step towards the query is a mask pixel and strictly
// closer), so only the mask's boundary pixels become vertices.  They are
// compacted in row-major order (count / scan / emit), which makes vertex ids
// deterministic; each carries its coordinates (the window's linspace axes,
// bit-exact) and its post-assignment hfun value.
// ---------------------------------------------------------------------------
namespace {

constexpr int kCompThreads = 256;    // x 4 consecutive pixels per thread: a CTA
constexpr int kCompPx = 1024;        // = 1024 consecutive pixels of ONE row,
                                     // CTAs in row-major order (so is the output)

// boundary flags (bit k = pixel col + k) of 4 consecutive pixels of one row
// open_top / open_bot: the rows beyond the first / last row belong to another
// rank's band (unknown here), so mask pixels on that edge count as boundary --
// a superset of the true boundary, which cannot change a nearest-pixel answer
__device__ __forceinline__ unsigned boundary4(const uint8_t *__restrict__ mask, int w,
                                              int h, int row, int col, bool vec,
                                              int open_top, int open_bot) {
    const uint8_t *p = mask + (int64_t)row * w + col;
    unsigned c = 0;                          // bit k: pixel is a mask pixel
    unsigned up = open_top ? 0u : 0xfu, dn = open_bot ? 0u : 0xfu;
    if (vec && col + 4 <= w) {
        const uchar4 m = *reinterpret_cast<const uchar4 *>(p);
        c = (m.x ? 1u : 0u) | (m.y ? 2u : 0u) | (m.z ? 4u : 0u) | (m.w ? 8u : 0u);
        if (!c) return 0u;
        if (row > 0) {
            const uchar4 u = *reinterpret_cast<const uchar4 *>(p - w);
            up = (u.x ? 1u : 0u) | (u.y ? 2u : 0u) | (u.z ? 4u : 0u) | (u.w ? 8u : 0u);
        }
        if (row + 1 < h) {
            const uchar4 d = *reinterpret_cast<const uchar4 *>(p + w);
            dn = (d.x ? 1u : 0u) | (d.y ? 2u : 0u) | (d.z ? 4u : 0u) | (d.w ? 8u : 0u);
        }
    } else {
        up = dn = 0u;
        for (int k = 0; k < 4 && col + k < w; ++k) {
            c |= (p[k] ? 1u : 0u) << k;
            up |= ((row > 0) ? (p[k - w] ? 1u : 0u) : (open_top ? 0u : 1u)) << k;
            dn |= ((row + 1 < h) ? (p[k + w] ? 1u : 0u) : (open_bot ? 0u : 1u)) << k;
        }
        if (!c) return 0u;
    }
    // horizontal neighbours: out-of-window counts as mask (no query pixel there)
    const unsigned lf = (col > 0) ? (p[-1] ? 1u : 0u) : 1u;
    const unsigned rt = (col + 4 < w) ? (p[4] ? 1u : 0u) : 1u;
    const unsigned left = ((c << 1) | lf) & 0xfu;            // bit k: pixel k-1 is mask
    unsigned right = (c >> 1) | (rt << 3);                   // bit k: pixel k+1 is mask
    // pixels past the row end have no right neighbour inside the window
    for (int k = 0; k < 4; ++k)
        if (col + k + 1 >= w) right |= 1u << k;
    return c & ~(left & right & up & dn) & 0xfu;
}

__global__ void __launch_bounds__(kCompThreads)
boundary_count_kernel(const uint8_t *__restrict__ mask, int w, int h, int nbx,
                      int vec, int open_top, int open_bot, int *__restrict__ counts) {
    __shared__ int s_warp[kCompThreads / 32];
    const int row = blockIdx.x / nbx;
    const int col = (blockIdx.x - row * nbx) * kCompPx + threadIdx.x * 4;
    int n = 0;
    if (col < w) n = __popc(boundary4(mask, w, h, row, col, vec != 0, open_top, open_bot));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int k = 0; k < kCompThreads / 32; ++k) t += s_warp[k];
        counts[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(kCompThreads)
boundary_emit_kernel(const uint8_t *__restrict__ mask, int w, int h, int nbx, int vec,
                     int open_top, int open_bot, int row0,
                     const int *__restrict__ offsets, LinAxis ax, LinAxis ay,
                     const float *__restrict__ hfun, int64_t hs,
                     double2 *__restrict__ verts, float *__restrict__ vals) {
    __shared__ int s_warp[kCompThreads / 32];
    const int row = blockIdx.x / nbx;
    const int col = (blockIdx.x - row * nbx) * kCompPx + threadIdx.x * 4;
    unsigned f = 0;
    if (col < w) f = boundary4(mask, w, h, row, col, vec != 0, open_top, open_bot);
    const int n = __popc(f);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int x = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[wid] = x;
    __syncthreads();
    if (!f) return;
    int pos = offsets[blockIdx.x] + (x - n);
    for (int k = 0; k < wid; ++k) pos += s_warp[k];
    const double y = lin_at(ay, row0 + row);     // window row of band row `row`
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if ((f >> k) & 1u) {
            verts[pos] = make_double2(lin_at(ax, col + k), y);
            vals[pos] = hfun[(int64_t)row * hs + col + k];
            ++pos;
        }
}

}  // namespace

// ---------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------
int32_t k_constant_value(ocsb_ctx *ctx, int w, int h, const void *dem,
                         int dem_dtype, int64_t zs, const float *old_,
                         int64_t os, float *new_, int64_t ns, double value,
                         double lb, double ub, double ghmin, double ghmax,
                         int fuse_clamp) {
    const int has_lo = !ocsb_is_none(ghmin), has_hi = !ocsb_is_none(ghmax);
    const int64_t items = (int64_t)((w + 3) / 4) * h;
    const int vo = old_ ? vec_ok(old_, os, 4) : 0, vn = vec_ok(new_, ns, 4);
    OCSB_TIME(ctx, KC_CONSTRAINT);
    if (dem_dtype == OCSB_F32) {
        constant_value_kernel<float><<<grid_for(ctx, items), kThreads, 0, ctx->stream>>>(
            w, h, (const float *)dem, zs, old_, os, new_, ns, (float)value,
            (float)lb, (float)ub, fuse_clamp, has_lo, (float)ghmin, has_hi,
            (float)ghmax, vec_ok(dem, zs, 4), vo, vn);
    } else {
        constant_value_kernel<double><<<grid_for(ctx, items), kThreads, 0, ctx->stream>>>(
            w, h, (const double *)dem, zs, old_, os, new_, ns, (float)value, lb,
            ub, fuse_clamp, has_lo, (float)ghmin, has_hi, (float)ghmax,
            vec_ok(dem, zs, 8), vo, vn);
    }
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_topo_mask(ocsb_ctx *ctx, int w, int h, const void *dem, int dem_dtype,
                    int64_t zs, float *hfun, int64_t hs, const void *vals,
                    int vals_dtype, int64_t vs, double vconst, double lb,
                    double ub, int is_min, uint8_t *mask,
                    unsigned long long *nmask) {
    const int64_t items = (int64_t)((w + 3) / 4) * h;
    const int grid = grid_for(ctx, items);
    const int vh = vec_ok(hfun, hs, 4);
    OCSB_TIME(ctx, KC_CONSTRAINT);
#define OCSB_TOPO(T, V)                                                        \
    topo_mask_kernel<T, V><<<grid, kThreads, 0, ctx->stream>>>(                \
        w, h, (const T *)dem, zs, hfun, hs, (const V *)vals, vs, (float)vconst,\
        (T)lb, (T)ub, is_min, mask, vec_ok(dem, zs, sizeof(T)), vh,            \
        vals ? vec_ok(vals, vs, sizeof(V)) : 0, nmask)
    const bool v64 = vals && vals_dtype == OCSB_F64;
    if (dem_dtype == OCSB_F32) {
        if (v64) OCSB_TOPO(float, double); else OCSB_TOPO(float, float);
    } else {
        if (v64) OCSB_TOPO(double, double); else OCSB_TOPO(double, float);
    }
#undef OCSB_TOPO
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_courant(ocsb_ctx *ctx, int w, int h, const void *dem, int dem_dtype,
                  int64_t zs, float *hfun, int64_t hs, double cval, double dt,
                  double nu, int is_min) {
    const double g = 9.80665;   // scipy.constants.g
    const int64_t items = (int64_t)((w + 3) / 4) * h;
    const int grid = grid_for(ctx, items);
    const int vh = vec_ok(hfun, hs, 4);
    const double u_land = sqrt(g * nu);   // np.sqrt(constants.g * wave_amplitude)
    OCSB_TIME(ctx, KC_CONSTRAINT);
    if (dem_dtype == OCSB_F32) {
        courant_kernel<float><<<grid, kThreads, 0, ctx->stream>>>(
            w, h, (const float *)dem, zs, hfun, hs, (float)g, (float)nu,
            (float)(-fabs(nu)), (float)u_land, (float)dt, (float)cval, is_min,
            vec_ok(dem, zs, 4), vh);
    } else {
        courant_kernel<double><<<grid, kThreads, 0, ctx->stream>>>(
            w, h, (const double *)dem, zs, hfun, hs, g, nu, -fabs(nu), u_land,
            dt, cval, is_min, vec_ok(dem, zs, 8), vh);
    }
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

// boundary pixels of `mask` (h x w, compact; its row 0 is window row `row0`,
// hfun points at that row too) -> verts[*nv], vals[*nv] (stream-ordered scratch
// owned by `sg`); synchronises once for the count
int32_t k_mask_boundary(ocsb_ctx *ctx, ScratchGuard &sg, const uint8_t *mask,
                        int w, int h, LinAxis ax, LinAxis ay, const float *hfun,
                        int64_t hs, double **verts_out, float **vals_out,
                        int64_t *nv_out, int row0, int open_top, int open_bot) {
    const int nbx = (w + kCompPx - 1) / kCompPx;
    const int64_t nblk = (int64_t)nbx * h;
    const int vec = (((uintptr_t)mask & 3) == 0 && (w & 3) == 0) ? 1 : 0;
    OCSB_REQUIRE(nblk < 2147483647ll, "_apply_rate: window too large");
    int *counts = nullptr, *sums = nullptr;
    unsigned long long *total = nullptr;
    OCSB_TRY(sg.get(&counts, (size_t)nblk));
    OCSB_TRY(sg.get(&sums, (size_t)((nblk + 1023) / 1024)));
    OCSB_TRY(sg.get(&total, 1));
    OCSB_TIME(ctx, KC_CONSTRAINT);
    boundary_count_kernel<<<(unsigned)nblk, kCompThreads, 0, ctx->stream>>>(
        mask, w, h, nbx, vec, open_top, open_bot, counts);
    OCSB_LAUNCHED(ctx);
    OCSB_TIME(ctx, KC_CONSTRAINT);
    scan_counts(counts, nblk, sums, total, ctx->stream);
    OCSB_LAUNCHED(ctx);
    ctx->launches += 2;
    OCSB_CHECK_LAUNCH();
    unsigned long long host_total = 0;
    unsigned long long *bounce = ctx->pinned ? (unsigned long long *)ctx->pinned : &host_total;
    OCSB_CUDA(cudaMemcpyAsync(bounce, total, sizeof(*total), cudaMemcpyDeviceToHost, ctx->stream));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    host_total = *bounce;
    OCSB_REQUIRE(host_total < 2147483647ull,
                 "_apply_rate: %llu mask-boundary pixels exceed the vertex limit",
                 host_total);
    *nv_out = (int64_t)host_total;
    *verts_out = nullptr;
    *vals_out = nullptr;
    if (host_total == 0) return OCSB_OK;
    double2 *verts = nullptr;
    float *vals = nullptr;
    OCSB_TRY(sg.get(&verts, (size_t)host_total));
    OCSB_TRY(sg.get(&vals, (size_t)host_total));
    OCSB_TIME(ctx, KC_CONSTRAINT);
    boundary_emit_kernel<<<(unsigned)nblk, kCompThreads, 0, ctx->stream>>>(
        mask, w, h, nbx, vec, open_top, open_bot, row0, counts, ax, ay, hfun, hs, verts,
        vals);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    *verts_out = reinterpret_cast<double *>(verts);
    *vals_out = vals;
    return OCSB_OK;
}
