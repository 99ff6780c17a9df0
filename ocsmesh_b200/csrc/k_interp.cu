// K3 -- DEM -> mesh-node sampling, linear / nearest
// (reference ocsmesh/mesh/mesh.py:2285-2347; scipy RegularGridInterpolator
//  _rgi.py _find_indices / _evaluate_linear / _evaluate_nearest).
//
// Gather kernel, HBM/sector-bound: 40 B/node algorithmic for linear (16 B
// coords + 4 x 4 B texels + 8 B value), 28 B/node for nearest.  One thread per
// node; all index math in float64 with numpy's linspace rule so that cell
// indices and nearest picks are bit-exact.
//
// ONE launch serves a whole raster list: the reference applies the rasters in
// list order and the last one that covers a node wins
// (ocsmesh/mesh/mesh.py:372-376), so each thread walks the list from the LAST
// raster to the first and stops at the first raster that yields a value.  The
// node coordinates are read once and the value written once, instead of once
// per raster (16 tiles x 10 M nodes: 2.6 GB of coordinate re-reads saved).
//
// Raster selection: the host lays a coarse SEL_G x SEL_G grid over the union
// of the window bounding boxes; each cell holds the bit set of the rasters
// whose box touches the cell grown by one cell on every side (so the cell
// lookup may be computed sloppily), and a node tests only those boxes -- in
// f64, highest index first -- instead of walking the whole list.
//
// Index search: the analytic guess floor((v - lo) * (1/|step|)) is checked
// against the exact linspace values of its two grid lines and walked when it
// is off (rare); nearest needs no division at all: t <= 0.5 is decided
// exactly by  v - g[i] <= 0.5 * (g[i+1] - g[i]).
//
// Grid convention: scipy always works on ASCENDING grids (descending inputs
// are flipped together with the values), so for each axis we build the
// ascending view asc[m] = L[m] (step > 0) or L[n-1-m] (step < 0) of the
// window's linspace L and map m back to the pixel index.
#include "ocsb_common.cuh"

namespace {

#ifndef OCSB_INTERP_MIN_CTAS
#define OCSB_INTERP_MIN_CTAS 4
#endif
constexpr int SEL_G = 64;          // selection grid: SEL_G x SEL_G cells
constexpr int SEL_MAX_RASTERS = 64;  // one bit per raster

struct AscAxis {
    LinAxis a;
    double lo;   // asc[0]
    double inv;  // 1 / |step|
    int flip;    // 1: ascending index m <-> linspace index n-1-m
    int pad;
};

struct RasterDev {
    AscAxis gx, gy;
    double xmin, xmax, ymin, ymax;
    const void *dem;
    int64_t ds;
    const uint8_t *mask;
    int64_t ms;
    int dtype;
    int pad;
};

struct SelGrid {
    double x0, y0, inv_x, inv_y;   // cell = (p - origin) * inv
    const uint64_t *cells;         // SEL_G * SEL_G bit sets, NULL: walk the list
};

__device__ __forceinline__ double asc_at(const AscAxis &g, int m) {
    return lin_at(g.a, g.flip ? g.a.n - 1 - m : m);
}
__device__ __forceinline__ int asc_to_pix(const AscAxis &g, int m) {
    return g.flip ? g.a.n - 1 - m : m;
}

// interval i with asc[i] <= v < asc[i+1], clipped to [0, n-2]
// (== clip(searchsorted(asc, v, 'right') - 1, 0, n-2)); *num = v - asc[i],
// *den = asc[i+1] - asc[i]  (t = num / den)
__device__ __forceinline__ int find_cell(const AscAxis &g, double v, double *num,
                                         double *den) {
    const int n = g.a.n;
    int i = __double2int_rd(__dmul_rn(__dsub_rn(v, g.lo), g.inv));
    i = max(0, min(n - 2, i));
    double gi = asc_at(g, i);
    double gi1 = asc_at(g, i + 1);
    // the guess is exact except within an ulp or so of a grid line
    if (v < gi) {
        while (i > 0 && v < gi) {
            --i;
            gi1 = gi;
            gi = asc_at(g, i);
        }
    } else if (v >= gi1) {
        while (i < n - 2 && v >= gi1) {
            ++i;
            gi = gi1;
            gi1 = asc_at(g, i + 1);
        }
    }
    *num = __dsub_rn(v, gi);
    *den = __dsub_rn(gi1, gi);
    return i;
}

template <typename T>
__device__ __forceinline__ double texel(const RasterDev &R, int64_t row, int64_t col) {
    return (double)__ldg(reinterpret_cast<const T *>(R.dem) + row * R.ds + col);
}

// one raster, one node: true (and *out) if this raster sets the node's value
template <typename T, int METHOD>
__device__ __forceinline__ bool sample(const RasterDev &R, double2 p, double *out) {
    double nx, dx, ny, dy;
    const int i = find_cell(R.gx, p.x, &nx, &dx);
    const int j = find_cell(R.gy, p.y, &ny, &dy);
    if (METHOD == OCSB_INTERP_NEAREST) {
        // t = fl(num / den) <= 0.5  <=>  num <= den / 2: the halving is exact
        // and a quotient above 1/2 exceeds it by more than half an ulp of 1/2
        const int ii = (nx <= __dmul_rn(0.5, dx)) ? i : i + 1;
        const int jj = (ny <= __dmul_rn(0.5, dy)) ? j : j + 1;
        const int64_t col = asc_to_pix(R.gx, ii);
        const int64_t row = asc_to_pix(R.gy, jj);
        if (R.mask && R.mask[row * R.ms + col]) return false;  // picked texel is nodata
        *out = texel<T>(R, row, col);
        return true;
    }
    const int64_t c0 = asc_to_pix(R.gx, i), c1 = asc_to_pix(R.gx, i + 1);
    const int64_t r0 = asc_to_pix(R.gy, j), r1 = asc_to_pix(R.gy, j + 1);
    // the four texels first: their latency hides the two divisions
    const double v00 = texel<T>(R, r0, c0);
    const double v01 = texel<T>(R, r1, c0);
    const double v10 = texel<T>(R, r0, c1);
    const double v11 = texel<T>(R, r1, c1);
    const double tx = __ddiv_rn(nx, dx), ty = __ddiv_rn(ny, dy);
    const double ux = __dsub_rn(1.0, tx), uy = __dsub_rn(1.0, ty);
    if (R.mask) {
        // node is dropped iff the same interpolation of the 0/1 mask is > 0:
        // a masked corner whose two weights are both non-zero
        const uint8_t *m = R.mask;
        const bool bad = (m[r0 * R.ms + c0] && ux != 0.0 && uy != 0.0) ||
                         (m[r1 * R.ms + c0] && ux != 0.0 && ty != 0.0) ||
                         (m[r0 * R.ms + c1] && tx != 0.0 && uy != 0.0) ||
                         (m[r1 * R.ms + c1] && tx != 0.0 && ty != 0.0);
        if (bad) return false;
    }
    // V[i,j] = z[row(j), col(i)];  hypercube order of _evaluate_linear:
    // (i,j) (i,j+1) (i+1,j) (i+1,j+1), weights (1-tx)(1-ty) ... tx*ty,
    // summed left to right starting from 0.0, no FMA contraction
    double acc;
    if (sizeof(T) == 4) {
        // generic python path: term = V * (w0 * w1)
        acc = __dadd_rn(0.0, __dmul_rn(v00, __dmul_rn(ux, uy)));
        acc = __dadd_rn(acc, __dmul_rn(v01, __dmul_rn(ux, ty)));
        acc = __dadd_rn(acc, __dmul_rn(v10, __dmul_rn(tx, uy)));
        acc = __dadd_rn(acc, __dmul_rn(v11, __dmul_rn(tx, ty)));
    } else {
        // float64 values take scipy's cython fast path
        // (evaluate_linear_2d): (V * w0) * w1 accumulated in order
        acc = __dmul_rn(__dmul_rn(v00, ux), uy);
        acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(v01, ux), ty));
        acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(v10, tx), uy));
        acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(v11, tx), ty));
    }
    *out = acc;
    return true;
}

// DT: OCSB_F32 / OCSB_F64 = every raster of the list has that dtype, 2 = mixed
template <int METHOD, int DT>
__device__ __forceinline__ bool try_raster(const RasterDev &R, double2 p, double *v) {
    // inclusive bbox test against min/max of the window coordinates
    // (mesh.py:2290-2297); NaN coordinates fail every compare
    if (!(R.xmin <= p.x && R.xmax >= p.x && R.ymin <= p.y && R.ymax >= p.y))
        return false;
    if (DT == OCSB_F32) return sample<float, METHOD>(R, p, v);
    if (DT == OCSB_F64) return sample<double, METHOD>(R, p, v);
    return (R.dtype == OCSB_F32) ? sample<float, METHOD>(R, p, v)
                                 : sample<double, METHOD>(R, p, v);
}

template <int METHOD, int DT, bool SEL>
__global__ void __launch_bounds__(256, OCSB_INTERP_MIN_CTAS)
interp_kernel(const RasterDev *__restrict__ rasters, int n_rasters, SelGrid sel,
              const double2 *__restrict__ xy, int64_t K,
              uint8_t *__restrict__ touched, double *__restrict__ values,
              int32_t *__restrict__ source, const int32_t *__restrict__ perm) {
    extern __shared__ unsigned char s_raw[];
    RasterDev *s_r = reinterpret_cast<RasterDev *>(s_raw);
    for (int i = threadIdx.x; i < n_rasters * (int)(sizeof(RasterDev) / 8);
         i += blockDim.x)
        reinterpret_cast<uint64_t *>(s_r)[i] =
            reinterpret_cast<const uint64_t *>(rasters)[i];
    __syncthreads();
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < K;
         k += (int64_t)gridDim.x * blockDim.x) {
        const double2 p = __ldcs(xy + k);
        double v;
        int hit = -1;
        if (SEL) {
            // sloppy cell lookup (the cell sets carry a one-cell margin);
            // NaN -> cell 0, whose boxes then reject the node exactly
            const float fx = (float)__dmul_rn(__dsub_rn(p.x, sel.x0), sel.inv_x);
            const float fy = (float)__dmul_rn(__dsub_rn(p.y, sel.y0), sel.inv_y);
            const int cx = max(0, min(SEL_G - 1, __float2int_rd(fx)));
            const int cy = max(0, min(SEL_G - 1, __float2int_rd(fy)));
            uint64_t m = __ldg(sel.cells + cy * SEL_G + cx);
            while (m) {
                const int r = 63 - __clzll((long long)m);
                m ^= 1ull << r;
                if (try_raster<METHOD, DT>(s_r[r], p, &v)) {
                    hit = r;
                    break;
                }
            }
        } else {
            for (int r = n_rasters - 1; r >= 0; --r)
                if (try_raster<METHOD, DT>(s_r[r], p, &v)) {
                    hit = r;
                    break;
                }
        }
        if (hit >= 0) {
            // bucketed input: results go back to the original node index
            const int64_t o = perm ? (int64_t)__ldg(perm + k) : k;
            __stcs(values + o, v);
            touched[o] = 1;
            if (source) source[o] = hit;
        }
    }
}

AscAxis make_asc(const LinAxis &a) {
    AscAxis g;
    g.a = a;
    g.flip = a.step < 0.0;
    g.lo = lin_at_host(a, g.flip ? a.n - 1 : 0);
    g.inv = a.step != 0.0 ? 1.0 / fabs(a.step) : 0.0;
    g.pad = 0;
    return g;
}

template <int METHOD, int DT>
void launch_dt(unsigned blocks, size_t smem, cudaStream_t st, const RasterDev *dev,
               int n, const SelGrid &sel, const double2 *q, int64_t K,
               uint8_t *touched, double *values, int32_t *source,
               const int32_t *perm) {
    if (sel.cells)
        interp_kernel<METHOD, DT, true><<<blocks, 256, smem, st>>>(
            dev, n, sel, q, K, touched, values, source, perm);
    else
        interp_kernel<METHOD, DT, false><<<blocks, 256, smem, st>>>(
            dev, n, sel, q, K, touched, values, source, perm);
}

template <int METHOD>
void launch(int dt, unsigned blocks, size_t smem, cudaStream_t st, const RasterDev *dev,
            int n, const SelGrid &sel, const double2 *q, int64_t K, uint8_t *touched,
            double *values, int32_t *source, const int32_t *perm) {
    if (dt == OCSB_F32)
        launch_dt<METHOD, OCSB_F32>(blocks, smem, st, dev, n, sel, q, K, touched, values,
                                    source, perm);
    else if (dt == OCSB_F64)
        launch_dt<METHOD, OCSB_F64>(blocks, smem, st, dev, n, sel, q, K, touched, values,
                                    source, perm);
    else
        launch_dt<METHOD, 2>(blocks, smem, st, dev, n, sel, q, K, touched, values, source,
                             perm);
}

}  // namespace

int32_t k_interp_multi(ocsb_ctx *ctx, const InterpRaster *rs, int n,
                       const double *xy, int64_t K, int method, uint8_t *touched,
                       double *values, int32_t *source, const int32_t *perm) {
    std::vector<RasterDev> host(n);
    int dt = n > 0 ? rs[0].dem_dtype : OCSB_F32;
    double ux0 = INFINITY, uy0 = INFINITY, ux1 = -INFINITY, uy1 = -INFINITY;
    for (int i = 0; i < n; ++i) {
        RasterDev &R = host[i];
        R.gx = make_asc(rs[i].ax);
        R.gy = make_asc(rs[i].ay);
        // np.min / np.max of the linspace vectors: endpoints (pinned stop)
        const double x_a = lin_at_host(rs[i].ax, 0), x_b = lin_at_host(rs[i].ax, rs[i].ax.n - 1);
        const double y_a = lin_at_host(rs[i].ay, 0), y_b = lin_at_host(rs[i].ay, rs[i].ay.n - 1);
        R.xmin = fmin(x_a, x_b); R.xmax = fmax(x_a, x_b);
        R.ymin = fmin(y_a, y_b); R.ymax = fmax(y_a, y_b);
        R.dem = rs[i].dem; R.ds = rs[i].dem_stride;
        R.mask = rs[i].mask; R.ms = rs[i].mask_stride;
        R.dtype = rs[i].dem_dtype;
        R.pad = 0;
        if (R.dtype != dt) dt = 2;
        ux0 = fmin(ux0, R.xmin); ux1 = fmax(ux1, R.xmax);
        uy0 = fmin(uy0, R.ymin); uy1 = fmax(uy1, R.ymax);
    }
    static_assert(sizeof(RasterDev) % 8 == 0, "RasterDev is copied as 64-bit words");
    ScratchGuard sg(ctx);
    RasterDev *dev = nullptr;
    OCSB_TRY(sg.get(&dev, (size_t)n));
    // pageable source: the copy has been staged when cudaMemcpyAsync returns
    OCSB_CUDA(cudaMemcpyAsync(dev, host.data(), sizeof(RasterDev) * n,
                              cudaMemcpyHostToDevice, ctx->stream));
    // selection grid (only worth it for a list; needs finite, non-empty extent)
    SelGrid sel{0.0, 0.0, 0.0, 0.0, nullptr};
    std::vector<uint64_t> cells;
    if (n >= 3 && n <= SEL_MAX_RASTERS && ux1 > ux0 && uy1 > uy0 &&
        std::isfinite(ux1 - ux0) && std::isfinite(uy1 - uy0)) {
        const double cw = (ux1 - ux0) / SEL_G, ch = (uy1 - uy0) / SEL_G;
        cells.assign((size_t)SEL_G * SEL_G, 0);
        for (int i = 0; i < n; ++i) {
            const RasterDev &R = host[i];
            // cells touched by the box, grown by one cell on every side
            int a = (int)floor((R.xmin - ux0) / cw) - 1, b = (int)floor((R.xmax - ux0) / cw) + 1;
            int c = (int)floor((R.ymin - uy0) / ch) - 1, d = (int)floor((R.ymax - uy0) / ch) + 1;
            a = std::max(a, 0); c = std::max(c, 0);
            b = std::min(b, SEL_G - 1); d = std::min(d, SEL_G - 1);
            for (int cy = c; cy <= d; ++cy)
                for (int cx = a; cx <= b; ++cx)
                    cells[(size_t)cy * SEL_G + cx] |= 1ull << i;
        }
        uint64_t *dcells = nullptr;
        OCSB_TRY(sg.get(&dcells, cells.size()));
        OCSB_CUDA(cudaMemcpyAsync(dcells, cells.data(), cells.size() * sizeof(uint64_t),
                                  cudaMemcpyHostToDevice, ctx->stream));
        sel.x0 = ux0; sel.y0 = uy0;
        sel.inv_x = 1.0 / cw; sel.inv_y = 1.0 / ch;
        sel.cells = dcells;
    }
    int64_t blocks = (K + 255) / 256;
    const int64_t cap = (int64_t)ctx->num_sms * 8;
    if (blocks > cap) blocks = cap;
    const size_t smem = sizeof(RasterDev) * n;
    OCSB_REQUIRE(smem <= 40 * 1024, "ocsb_interp_multi: too many rasters (%d)", n);
    const double2 *q = reinterpret_cast<const double2 *>(xy);
    OCSB_TIME(ctx, KC_INTERP);
    if (method == OCSB_INTERP_NEAREST)
        launch<OCSB_INTERP_NEAREST>(dt, (unsigned)blocks, smem, ctx->stream, dev, n, sel, q,
                                    K, touched, values, source, perm);
    else
        launch<OCSB_INTERP_LINEAR>(dt, (unsigned)blocks, smem, ctx->stream, dev, n, sel, q,
                                   K, touched, values, source, perm);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

int32_t k_interp(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const void *dem,
                 int dem_dtype, int64_t dem_stride, const uint8_t *mask,
                 int64_t mask_stride, const double *xy, int64_t K, int method,
                 uint8_t *touched, double *values) {
    InterpRaster r{ax, ay, dem, dem_dtype, dem_stride, mask, mask_stride};
    return k_interp_multi(ctx, &r, 1, xy, K, method, touched, values, nullptr,
                          nullptr);
}
