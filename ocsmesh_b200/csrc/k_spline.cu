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
//   pass A  one thread per raster column: solve along y (ascending),
//           DEM (any dtype) -> f64, coalesced across columns
//   transpose (32x32 shared tiles)
//   pass B  one thread per y index: solve along x (ascending), writes the
//           coefficients in scipy's layout  c[i * ny + j]
//   eval    one thread per node: knot interval from the linspace axes,
//           de Boor (fpbspl) basis, 4x4 coefficient gather (88 B/node)
//
// Algorithmic HBM bytes: prefilter 2 passes x 8 B/px (one-off per raster),
// evaluation 16 + 16*... see DESIGN.md.
#include "ocsb_common.cuh"

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
    for (int j = 1; j <= 3; ++j) {
        for (int i = 0; i < j; ++i) hh[i] = h[i];
        h[0] = 0.0;
        for (int i = 0; i < j; ++i) {
            const int li = l + i + 1, lj = li - j;
            const double tli = t(li), tlj = t(lj);
            const double f = hh[i] / (tli - tlj);
            h[i] = h[i] + f * (tli - x);
            h[i + 1] = f * (x - tlj);
        }
    }
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

// Solve along the slow axis of a (n, m) f64 array `dst` (element (i, k) at
// dst[i*m + k]) for every k.  Source element (i, k) is read from
//   src[(src_flip ? n-1-i : i) * ss + k]   (type T).
template <typename T>
__global__ void __launch_bounds__(128)
solve_lines_kernel(const T *__restrict__ src, int64_t ss, int src_flip,
                   double *__restrict__ dst, int n, int m,
                   const double *__restrict__ lu) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    double y1 = 0.0, y2 = 0.0;  // y[i-1], y[i-2]
    for (int i = 0; i < n; ++i) {
        const int64_t si = src_flip ? (n - 1 - i) : i;
        const double b = (double)src[si * ss + k];
        const double l2 = __ldg(lu + (int64_t)i * 5 + 0);
        const double l1 = __ldg(lu + (int64_t)i * 5 + 1);
        const double y = b - l1 * y1 - l2 * y2;
        dst[(int64_t)i * m + k] = y;
        y2 = y1; y1 = y;
    }
    double x1 = 0.0, x2 = 0.0;  // x[i+1], x[i+2]
    for (int i = n - 1; i >= 0; --i) {
        const double invd = __ldg(lu + (int64_t)i * 5 + 2);
        const double u1 = __ldg(lu + (int64_t)i * 5 + 3);
        const double u2 = __ldg(lu + (int64_t)i * 5 + 4);
        const double y = dst[(int64_t)i * m + k];
        const double x = (y - u1 * x1 - u2 * x2) * invd;
        dst[(int64_t)i * m + k] = x;
        x2 = x1; x1 = x;
    }
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

// data interval i with g[i] <= v < g[i+1] clipped to [0, n-2]
__device__ __forceinline__ int find_interval(const AscAxisS &g, double v) {
    const int n = g.a.n;
    const double lo = asc_at(g, 0);
    int i = (int)floor((v - lo) / fabs(g.a.step));
    i = max(0, min(n - 2, i));
    while (i > 0 && v < asc_at(g, i)) --i;
    while (i < n - 2 && v >= asc_at(g, i + 1)) ++i;
    return i;
}

__global__ void __launch_bounds__(256)
spline_eval_kernel(AscAxisS gx, AscAxisS gy, double xmin, double xmax,
                   double ymin, double ymax, const double *__restrict__ coef,
                   const double2 *__restrict__ xy, int64_t K,
                   uint8_t *__restrict__ touched, double *__restrict__ values) {
    const int nx = gx.a.n, ny = gy.a.n;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < K;
         k += (int64_t)gridDim.x * blockDim.x) {
        const double2 p = __ldcs(xy + k);
        if (!(xmin <= p.x && xmax >= p.x && ymin <= p.y && ymax >= p.y))
            continue;
        // knot interval: l = clamp(i + 2, 3, n - 1) (see header comment)
        int lx = find_interval(gx, p.x) + 2;
        lx = max(3, min(nx - 1, lx));
        int ly = find_interval(gy, p.y) + 2;
        ly = max(3, min(ny - 1, ly));
        double hx[4], hy[4];
        bspl3(DevKnots{gx}, lx, p.x, hx);
        bspl3(DevKnots{gy}, ly, p.y, hy);
        // fpbisp accumulation order: i over x, j over y
        double acc = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const double *row = coef + (int64_t)(lx - 3 + i) * ny + (ly - 3);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                acc = acc + __ldg(row + j) * hx[i] * hy[j];
        }
        values[k] = acc;
        touched[k] = 1;
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
    // pass A: along y for every pixel column.  Result in `coef` viewed as
    // (ny, nx): element (j, c) = coef[j*nx + c] (c = pixel column)
    {
        const int blocks = (nx + 127) / 128;
        OCSB_TIME(ctx, KC_SPLINE_PREFILTER);
        if (dem_dtype == OCSB_F32)
            solve_lines_kernel<float><<<blocks, 128, 0, ctx->stream>>>(
                (const float *)dem, dem_stride, gy.flip, coef, ny, nx, luy);
        else
            solve_lines_kernel<double><<<blocks, 128, 0, ctx->stream>>>(
                (const double *)dem, dem_stride, gy.flip, coef, ny, nx, luy);
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
    }
    // transpose (ny, nx) -> tmp (nx, ny): tmp[c*ny + j]
    {
        dim3 grid((nx + 31) / 32, (ny + 31) / 32);
        OCSB_TIME(ctx, KC_SPLINE_PREFILTER);
        transpose_kernel<<<grid, 256, 0, ctx->stream>>>(coef, tmp, ny, nx);
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
    }
    // pass B: along x (ascending) for every j; source row = pixel column
    {
        const int blocks = (ny + 127) / 128;
        OCSB_TIME(ctx, KC_SPLINE_PREFILTER);
        solve_lines_kernel<double><<<blocks, 128, 0, ctx->stream>>>(
            tmp, ny, gx.flip, coef, nx, ny, lux);
        OCSB_LAUNCHED(ctx);
        OCSB_CHECK_LAUNCH();
    }
    return OCSB_OK;
}

int32_t k_spline_eval(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const double *coef,
                      const double *xy, int64_t K, uint8_t *touched,
                      double *values) {
    const AscAxisS gx = make_asc(ax), gy = make_asc(ay);
    double x_a = lin_at_host(ax, 0), x_b = lin_at_host(ax, ax.n - 1);
    double y_a = lin_at_host(ay, 0), y_b = lin_at_host(ay, ay.n - 1);
    int64_t blocks = (K + 255) / 256;
    int64_t cap = (int64_t)ctx->num_sms * 16;
    if (blocks > cap) blocks = cap;
    OCSB_TIME(ctx, KC_SPLINE_EVAL);
    spline_eval_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(
        gx, gy, fmin(x_a, x_b), fmax(x_a, x_b), fmin(y_a, y_b), fmax(y_a, y_b),
        coef, reinterpret_cast<const double2 *>(xy), K, touched, values);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
