// K4 -- Mesh.interpolate(method='spline')
// (reference ocsmesh/mesh/mesh.py:2306-2314:
//   RectBivariateSpline(xi, flip(yi), flipud(zi).T, kx=3, ky=3, s=0).ev(x, y)).
//
// FITPACK's regrid with s=0 is the tensor product of two 1-D not-a-knot cubic
// interpolating B-splines (knots t = [g0]*4 + g[2:-2] + [g_end]*4, SURVEY.md
// A.5).  The 1-D collocation matrix B_j(g_i) is pentadiagonal (tridiagonal
// inside; the rows next to each end have 4 entries because g_1 / g_{n-2} are
// not knots) and depends only on the axis, so it is LU-factorised ONCE on the
// host in O(n) (no pivoting is needed: the pivots stay >= 1/3) and every
// raster line is then a forward/back substitution on the device.
//
// The triangular solves are first-order-dominated recurrences whose memory
// decays like (2 - sqrt 3)^k = 0.268^k (the ratio l1 = u1/d of the cubic
// B-spline collocation matrix), so a line need not be walked end to end by one
// thread: it is cut into chunks of R points and every chunk warms its
// recurrence up over the kWarm = 48 points before it from a zero state --
// 0.268^48 = 4e-28 relative, far below the f64 rounding of the walk itself
// (a chunk that starts at the end of the line needs no warm-up and is exact).
// That turns n dependent steps on nx threads (5000 threads for a 5000 x 5000
// tile: 1 % of the machine, latency-bound) into R + 48 steps on nx * n / R
// threads, all loads coalesced across the fast axis:
//
//   fwd A   L y = b    along y (ascending), DEM (any dtype) -> f64   src  -> coef
//   bwd A   U x = y    along y                                       coef -> tmp
//   transpose (32x32 shared tiles), (ny, nx) -> (nx, ny)             tmp  -> coef
//   fwd B / bwd B      along x (ascending)                  coef -> tmp -> coef
//   result in scipy's layout  c[i * ny + j]
//   eval    one thread per node: knot interval from the linspace axes,
//           de Boor (fpbspl) basis, 4x4 coefficient gather (88 B/node)
//
// Algorithmic HBM bytes: prefilter 2 passes x 8 B/px (one-off per raster),
// evaluation 16 + 16*... see DESIGN.md.
#include "ocsb_common.cuh"

#include <cstring>
#include <vector>

namespace {

struct AscAxisS {
    LinAxis a;
    int flip;
};

inline AscAxisS make_asc(const LinAxis &a) {
    AscAxisS g;
    g.a = a;
    g.flip = a.step < 0.0;
    return g;
}

inline double asc_at_host(const AscAxisS &g, int m) {
    return lin_at_host(g.a, g.flip ? g.a.n - 1 - m : m);
}

__device__ __forceinline__ double asc_at(const AscAxisS &g, int m) {
    return lin_at(g.a, g.flip ? g.a.n - 1 - m : m);
}

// knot p of the not-a-knot cubic knot vector over an ascending grid of n
// points: t[p] = g[0] (p < 4), g[n-1] (p >= n), g[p-2] otherwise
struct HostKnots {
    const double *g;
    int n;
    __host__ __device__ double operator()(int p) const {
        return g[p < 4 ? 0 : (p >= n ? n - 1 : p - 2)];
    }
};
struct DevKnots {
    AscAxisS g;
    __device__ double operator()(int p) const {
        const int n = g.a.n;
        return asc_at(g, p < 4 ? 0 : (p >= n ? n - 1 : p - 2));
    }
};

// fpbspl for k = 3: the four non-zero cubic B-splines on [t[l], t[l+1])
template <typename KnotF>
__host__ __device__ __forceinline__ void bspl3(KnotF t, int l, double x,
                                               double h[4]) {
    double hh[3];
    h[0] = 1.0;
    h[1] = h[2] = h[3] = 0.0;
    // (the host pass of this __host__ __device__ function has no `unroll`)
#ifdef __CUDA_ARCH__
#define OCSB_UNROLL _Pragma("unroll")
#else
#define OCSB_UNROLL
#endif
    OCSB_UNROLL
    for (int j = 1; j <= 3; ++j) {
        OCSB_UNROLL
        for (int i = 0; i < j; ++i) hh[i] = h[i];
        h[0] = 0.0;
        OCSB_UNROLL
        for (int i = 0; i < j; ++i) {
            const int li = l + i + 1, lj = li - j;
            const double tli = t(li), tlj = t(lj);
            const double f = hh[i] / (tli - tlj);
            h[i] = h[i] + f * (tli - x);
            h[i + 1] = f * (x - tlj);
        }
    }
#undef OCSB_UNROLL
}

// LU factors of the collocation matrix, 5 doubles per row:
// (l2, l1, 1/d, u1, u2)
int32_t factor_axis(const AscAxisS &g, std::vector<double> *out) {
    const int n = g.a.n;
    std::vector<double> gv(n);
    for (int i = 0; i < n; ++i) gv[i] = asc_at_host(g, i);
    const HostKnots knot{gv.data(), n};
    std::vector<double> band((size_t)n * 5, 0.0);
    for (int i = 0; i < n; ++i) {
        int l = i + 2;
        if (l < 3) l = 3;
        if (l > n - 1) l = n - 1;
        double h[4];
        bspl3(knot, l, gv[i], h);
        for (int q = 0; q < 4; ++q) {
            const int c = l - 3 + q;
            if (c - i >= -2 && c - i <= 2) band[(size_t)i * 5 + (c - i + 2)] = h[q];
            else if (fabs(h[q]) > 1e-12) {
                ocsb_set_error("spline prefilter: collocation matrix is not "
                               "pentadiagonal (non-monotone axis?)");
                return OCSB_EINVAL;
            }
        }
    }
    out->assign((size_t)n * 5, 0.0);
    std::vector<double> d(n), u1(n), u2(n);
    for (int i = 0; i < n; ++i) {
        double *b = &band[(size_t)i * 5];
        double l2 = 0.0, l1 = 0.0;
        if (i >= 2) {
            l2 = b[0] / d[i - 2];
            b[1] -= l2 * u1[i - 2];
            b[2] -= l2 * u2[i - 2];
        }
        if (i >= 1) {
            l1 = b[1] / d[i - 1];
            b[2] -= l1 * u1[i - 1];
            b[3] -= l1 * u2[i - 1];
        }
        d[i] = b[2]; u1[i] = b[3]; u2[i] = b[4];
        if (!(fabs(d[i]) > 1e-8)) {
            ocsb_set_error("spline prefilter: singular collocation matrix "
                           "(pivot %g at row %d)", d[i], i);
            return OCSB_EINVAL;
        }
        double *o = &(*out)[(size_t)i * 5];
        o[0] = l2; o[1] = l1; o[2] = 1.0 / d[i]; o[3] = u1[i]; o[4] = u2[i];
    }
    return OCSB_OK;
}

constexpr int kWarm = 48;   // warm-up points of a chunked recurrence (see header)

// Forward substitution L y = b along the slow axis of an (n, m) array for every
// fast index k, chunk blockIdx.y of R slow indices: y(i, k) -> dst[i*m + k].
// Source element (i, k) is src[(src_flip ? n-1-i : i) * ss + k]  (type T).
template <typename T>
__global__ void __launch_bounds__(128)
solve_fwd_kernel(const T *__restrict__ src, int64_t ss, int src_flip,
                 double *__restrict__ dst, int n, int m, int R,
                 const double *__restrict__ lu) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    const int a = blockIdx.y * R;
    const int b = min(n, a + R);
    const int s = max(0, a - kWarm);
    double y1 = 0.0, y2 = 0.0;  // y[i-1], y[i-2]
    // (issuing 8 or 16 rows of loads ahead of the FMA chain was measured
    // SLOWER: 0.75 / 0.79 vs 0.54 ms per 5000^2 tile, profiles/r1h_spline.txt)
#pragma unroll 4
    for (int i = s; i < b; ++i) {
        const int64_t si = src_flip ? (n - 1 - i) : i;
        const double v = (double)__ldg(src + si * ss + k);
        const double l2 = __ldg(lu + (int64_t)i * 5 + 0);
        const double l1 = __ldg(lu + (int64_t)i * 5 + 1);
        // one dependent FMA per step: the y2 term does not wait for y1
        const double y = fma(-l1, y1, fma(-l2, y2, v));
        if (i >= a) dst[(int64_t)i * m + k] = y;
        y2 = y1; y1 = y;
    }
}

// Back substitution U x = y along the slow axis, same chunking, walking down
// from kWarm points above the chunk: x(i, k) -> dst[i*m + k].
__global__ void __launch_bounds__(128)
solve_bwd_kernel(const double *__restrict__ ysrc, double *__restrict__ dst, int n,
                 int m, int R, const double *__restrict__ lu) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    const int a = blockIdx.y * R;
    const int b = min(n, a + R);
    const int e = min(n, b + kWarm);
    double x1 = 0.0, x2 = 0.0;  // x[i+1], x[i+2]
#pragma unroll 4
    for (int i = e - 1; i >= a; --i) {
        const double invd = __ldg(lu + (int64_t)i * 5 + 2);
        const double u1 = __ldg(lu + (int64_t)i * 5 + 3);
        const double u2 = __ldg(lu + (int64_t)i * 5 + 4);
        const double y = __ldg(ysrc + (int64_t)i * m + k);
        // x = (y - u1 x1 - u2 x2) / d with only one FMA on the x1 chain
        const double x = fma(-(u1 * invd), x1, fma(-u2, x2, y) * invd);
        if (i < b) dst[(int64_t)i * m + k] = x;
        x2 = x1; x1 = x;
    }
}

// chunk length: enough threads to fill the machine, warm-up overhead bounded
inline int chunk_rows(int n, int m, int num_sms) {
    const int64_t want = (int64_t)num_sms * 2048;          // resident threads
    int64_t r = ((int64_t)n * m + want - 1) / want;
    r = (r + 31) / 32 * 32;
    if (r < 32) r = 32;
    if (r > 256) r = 256;
    return (int)r;
}

// out[c * rows + r] = in[r * cols + c]
__global__ void __launch_bounds__(256)
transpose_kernel(const double *__restrict__ in, double *__restrict__ out,
                 int rows, int cols) {
    __shared__ double tile[32][33];
    const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int j = ty; j < 32; j += 8) {
        const int r = by + j, c = bx + tx;
        if (r < rows && c < cols) tile[j][tx] = in[(int64_t)r * cols + c];
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int c = bx + j, r = by + tx;
        if (r < rows && c < cols) out[(int64_t)c * rows + r] = tile[tx][j];
    }
}

// data interval i with g[i] <= v < g[i+1] clipped to [0, n-2]; the analytic
// guess is checked against the exact grid lines and walked when it is off
__device__ __forceinline__ int find_interval(const AscAxisS &g, double v, double lo,
                                             double inv) {
    const int n = g.a.n;
    int i = __double2int_rd((v - lo) * inv);
    i = max(0, min(n - 2, i));
    while (i > 0 && v < asc_at(g, i)) --i;
    while (i < n - 2 && v >= asc_at(g, i + 1)) ++i;
    return i;
}

// the six knots t[l-2 .. l+3] around knot interval l, read once
struct LocalKnots {
    double t[6];
    int l;
    __device__ double operator()(int p) const { return t[p - l + 2]; }
};

__device__ __forceinline__ LocalKnots local_knots(const AscAxisS &g, int l) {
    const DevKnots k{g};
    LocalKnots L;
    L.l = l;
#pragma unroll
    for (int q = 0; q < 6; ++q) L.t[q] = k(l - 2 + q);
    return L;
}

struct SplineRaster {
    AscAxisS gx, gy;
    double xmin, xmax, ymin, ymax;
    double lox, invx, loy, invy;   // asc[0] and 1/|step| per axis
    const double *coef;
};

// RectBivariateSpline.ev of one raster at p (fpbisp order: i over x, j over y)
__device__ __forceinline__ double spline_at(const SplineRaster &R, double2 p) {
    const int nx = R.gx.a.n, ny = R.gy.a.n;
    // knot interval: l = clamp(i + 2, 3, n - 1) (see header comment)
    int lx = find_interval(R.gx, p.x, R.lox, R.invx) + 2;
    lx = max(3, min(nx - 1, lx));
    int ly = find_interval(R.gy, p.y, R.loy, R.invy) + 2;
    ly = max(3, min(ny - 1, ly));
    double hx[4], hy[4];
    bspl3(local_knots(R.gx, lx), lx, p.x, hx);
    bspl3(local_knots(R.gy, ly), ly, p.y, hy);
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double *row = R.coef + (int64_t)(lx - 3 + i) * ny + (ly - 3);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            acc = acc + __ldg(row + j) * hx[i] * hy[j];
    }
    return acc;
}

// One launch for a list of rasters: the reference applies them in list order
// and the last one whose window contains the node wins (mesh.py:372-376; the
// spline branch has no nodata rule), so every node walks the list backwards
// and evaluates exactly one spline -- instead of one launch per raster in
// which 15 of 16 lanes of a warp of randomly ordered nodes sit idle.
__global__ void __launch_bounds__(256)
spline_eval_kernel(const SplineRaster *__restrict__ rasters, int n_rasters,
                   const double2 *__restrict__ xy, int64_t K,
                   uint8_t *__restrict__ touched, double *__restrict__ values) {
    extern __shared__ unsigned char s_raw[];
    SplineRaster *s_r = reinterpret_cast<SplineRaster *>(s_raw);
    for (int i = threadIdx.x; i < n_rasters * (int)(sizeof(SplineRaster) / 8);
         i += blockDim.x)
        reinterpret_cast<uint64_t *>(s_r)[i] =
            reinterpret_cast<const uint64_t *>(rasters)[i];
    __syncthreads();
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < K;
         k += (int64_t)gridDim.x * blockDim.x) {
        const double2 p = __ldcs(xy + k);
        for (int r = n_rasters - 1; r >= 0; --r) {
            const SplineRaster &R = s_r[r];
            if (!(R.xmin <= p.x && R.xmax >= p.x && R.ymin <= p.y && R.ymax >= p.y))
                continue;
            values[k] = spline_at(R, p);
            touched[k] = 1;
            break;
        }
    }
}

}  // namespace

int32_t k_spline_prefilter(ocsb_ctx *ctx, LinAxis ax, LinAxis ay,
                           const void *dem, int dem_dtype, int64_t dem_stride,
                           double *coef) {
    const AscAxisS gx = make_asc(ax), gy = make_asc(ay);
    const int nx = ax.n, ny = ay.n;
    std::vector<double> fx, fy;
    OCSB_TRY(factor_axis(gx, &fx));
    OCSB_TRY(factor_axis(gy, &fy));
    ScratchGuard sg(ctx);
    double *lux = nullptr, *luy = nullptr, *tmp = nullptr;
    OCSB_TRY(sg.get(&lux, fx.size()));
    OCSB_TRY(sg.get(&luy, fy.size()));
    OCSB_TRY(sg.get(&tmp, (size_t)nx * ny));
    // pageable host vectors: the copies complete before cudaMemcpyAsync returns
    OCSB_CUDA(cudaMemcpyAsync(lux, fx.data(), fx.size() * 8,
                              cudaMemcpyHostToDevice, ctx->stream));
    OCSB_CUDA(cudaMemcpyAsync(luy, fy.data(), fy.size() * 8,
                              cudaMemcpyHostToDevice, ctx->stream));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    const dim3 tb(128);
    // along y for every pixel column c: (ny, nx) arrays, element (j, c) at [j*nx + c]
    {
        const int R = chunk_rows(ny, nx, ctx->num_sms);
        const dim3 grid((nx + 127) / 128, (ny + R - 1) / R);
        OCSB_TIME(ctx, KC_SPLINE_PREFILTER);
        if (dem_dtype == OCSB_F32)
            solve_fwd_kernel<float><<<grid, tb, 0, ctx->stream>>>(
                (const float *)dem, dem_stride, gy.flip, coef, ny, nx, R, luy);
        else
            solve_fwd_kernel<double><<<grid, tb, 0, ctx->stream>>>(
                (const double *)dem, dem_stride, gy.flip, coef, ny, nx, R, luy);
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
        OCSB_TIME(ctx, KC_SPLINE_PREFILTER);
        solve_bwd_kernel<<<grid, tb, 0, ctx->stream>>>(coef, tmp, ny, nx, R, luy);
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
    }
    // transpose tmp (ny, nx) -> coef (nx, ny): coef[c*ny + j]
    {
        dim3 grid((nx + 31) / 32, (ny + 31) / 32);
        OCSB_TIME(ctx, KC_SPLINE_PREFILTER);
        transpose_kernel<<<grid, 256, 0, ctx->stream>>>(tmp, coef, ny, nx);
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
    }
    // along x (ascending) for every j; source row = pixel column
    {
        const int R = chunk_rows(nx, ny, ctx->num_sms);
        const dim3 grid((ny + 127) / 128, (nx + R - 1) / R);
        OCSB_TIME(ctx, KC_SPLINE_PREFILTER);
        solve_fwd_kernel<double><<<grid, tb, 0, ctx->stream>>>(
            coef, ny, gx.flip, tmp, nx, ny, R, lux);
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
        OCSB_TIME(ctx, KC_SPLINE_PREFILTER);
        solve_bwd_kernel<<<grid, tb, 0, ctx->stream>>>(tmp, coef, nx, ny, R, lux);
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
    }
    return OCSB_OK;
}

int32_t k_spline_eval_multi(ocsb_ctx *ctx, const SplineEvalRaster *rs, int n,
                            const double *xy, int64_t K, uint8_t *touched,
                            double *values) {
    std::vector<SplineRaster> host(n);
    for (int i = 0; i < n; ++i) {
        SplineRaster &R = host[i];
        memset(&R, 0, sizeof(R));
        R.gx = make_asc(rs[i].ax);
        R.gy = make_asc(rs[i].ay);
        const LinAxis &ax = rs[i].ax, &ay = rs[i].ay;
        const double x_a = lin_at_host(ax, 0), x_b = lin_at_host(ax, ax.n - 1);
        const double y_a = lin_at_host(ay, 0), y_b = lin_at_host(ay, ay.n - 1);
        R.xmin = fmin(x_a, x_b); R.xmax = fmax(x_a, x_b);
        R.ymin = fmin(y_a, y_b); R.ymax = fmax(y_a, y_b);
        R.lox = asc_at_host(R.gx, 0); R.loy = asc_at_host(R.gy, 0);
        R.invx = ax.step != 0.0 ? 1.0 / fabs(ax.step) : 0.0;
        R.invy = ay.step != 0.0 ? 1.0 / fabs(ay.step) : 0.0;
        R.coef = rs[i].coef;
    }
    static_assert(sizeof(SplineRaster) % 8 == 0, "copied as 64-bit words");
    const size_t smem = sizeof(SplineRaster) * n;
    OCSB_REQUIRE(smem <= 40 * 1024, "spline evaluation: too many rasters (%d)", n);
    ScratchGuard sg(ctx);
    SplineRaster *dev = nullptr;
    OCSB_TRY(sg.get(&dev, (size_t)n));
    // pageable source: staged before cudaMemcpyAsync returns
    OCSB_CUDA(cudaMemcpyAsync(dev, host.data(), smem, cudaMemcpyHostToDevice,
                              ctx->stream));
    int64_t blocks = (K + 255) / 256;
    int64_t cap = (int64_t)ctx->num_sms * 16;
    if (blocks > cap) blocks = cap;
    OCSB_TIME(ctx, KC_SPLINE_EVAL);
    spline_eval_kernel<<<(unsigned)blocks, 256, smem, ctx->stream>>>(
        dev, n, reinterpret_cast<const double2 *>(xy), K, touched, values);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_spline_eval(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const double *coef,
                      const double *xy, int64_t K, uint8_t *touched,
                      double *values) {
    SplineEvalRaster r{ax, ay, coef};
    return k_spline_eval_multi(ctx, &r, 1, xy, K, touched, values);
}
