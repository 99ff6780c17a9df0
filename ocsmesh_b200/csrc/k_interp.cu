// K3 -- DEM -> mesh-node sampling, linear / nearest
// (reference ocsmesh/mesh/mesh.py:2285-2347; scipy RegularGridInterpolator
//  _rgi.py _find_indices / _evaluate_linear / _evaluate_nearest).
//
// Gather kernel, HBM/sector-bound: 40 B/node algorithmic for linear (16 B
// coords + 4 x 4 B texels + 8 B value), 28 B/node for nearest.  One thread per
// node; all index math in float64 with numpy's linspace rule so that cell
// indices and nearest picks are bit-exact.
//
// Grid convention: scipy always works on ASCENDING grids (descending inputs
// are flipped together with the values), so for each axis we build the
// ascending view asc[m] = L[m] (step > 0) or L[n-1-m] (step < 0) of the
// window's linspace L and map m back to the pixel index.
#include "ocsb_common.cuh"

namespace {

struct AscAxis {
    LinAxis a;
    int flip;  // 1: ascending index m <-> linspace index n-1-m
};

__device__ __forceinline__ double asc_at(const AscAxis &g, int m) {
    return lin_at(g.a, g.flip ? g.a.n - 1 - m : m);
}
__device__ __forceinline__ int asc_to_pix(const AscAxis &g, int m) {
    return g.flip ? g.a.n - 1 - m : m;
}

// interval i with asc[i] <= v < asc[i+1], clipped to [0, n-2]
// (== clip(searchsorted(asc, v, 'right') - 1, 0, n-2)), plus
// t = (v - asc[i]) / (asc[i+1] - asc[i])
__device__ __forceinline__ void find_cell(const AscAxis &g, double v, int *i_out,
                                          double *t_out) {
    const int n = g.a.n;
    const double lo = asc_at(g, 0);
    const double astep = fabs(g.a.step);
    int i = (int)floor((v - lo) / astep);
    i = max(0, min(n - 2, i));
    // correct the analytic guess against the exact grid values
    while (i > 0 && v < asc_at(g, i)) --i;
    while (i < n - 2 && v >= asc_at(g, i + 1)) ++i;
    const double gi = asc_at(g, i);
    const double gi1 = asc_at(g, i + 1);
    *i_out = i;
    *t_out = __ddiv_rn(__dsub_rn(v, gi), __dsub_rn(gi1, gi));
}

template <typename T, int METHOD>
__global__ void __launch_bounds__(256)
interp_kernel(AscAxis gx, AscAxis gy, double xmin, double xmax, double ymin,
              double ymax, const T *__restrict__ dem, int64_t ds,
              const uint8_t *__restrict__ mask, int64_t ms,
              const double2 *__restrict__ xy, int64_t K,
              uint8_t *__restrict__ touched, double *__restrict__ values) {
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < K;
         k += (int64_t)gridDim.x * blockDim.x) {
        const double2 p = __ldcs(xy + k);
        // inclusive bbox test against min/max of the window coordinates
        // (mesh.py:2290-2297); NaN coordinates fail every compare
        if (!(xmin <= p.x && xmax >= p.x && ymin <= p.y && ymax >= p.y))
            continue;
        int i, j;
        double tx, ty;
        find_cell(gx, p.x, &i, &tx);
        find_cell(gy, p.y, &j, &ty);
        if (METHOD == OCSB_INTERP_NEAREST) {
            const int ii = (tx <= 0.5) ? i : i + 1;
            const int jj = (ty <= 0.5) ? j : j + 1;
            const int64_t col = asc_to_pix(gx, ii);
            const int64_t row = asc_to_pix(gy, jj);
            if (mask && mask[row * ms + col]) continue;  // picked texel is nodata
            values[k] = (double)__ldg(dem + row * ds + col);
            touched[k] = 1;
        } else {
            const int64_t c0 = asc_to_pix(gx, i), c1 = asc_to_pix(gx, i + 1);
            const int64_t r0 = asc_to_pix(gy, j), r1 = asc_to_pix(gy, j + 1);
            const double ux = __dsub_rn(1.0, tx), uy = __dsub_rn(1.0, ty);
            if (mask) {
                // node is dropped iff the same interpolation of the 0/1 mask
                // is > 0: a masked corner whose two weights are both non-zero
                bool bad = (mask[r0 * ms + c0] && ux != 0.0 && uy != 0.0) ||
                           (mask[r1 * ms + c0] && ux != 0.0 && ty != 0.0) ||
                           (mask[r0 * ms + c1] && tx != 0.0 && uy != 0.0) ||
                           (mask[r1 * ms + c1] && tx != 0.0 && ty != 0.0);
                if (bad) continue;
            }
            // V[i,j] = z[row(j), col(i)];  hypercube order of _evaluate_linear:
            // (i,j) (i,j+1) (i+1,j) (i+1,j+1), weights (1-tx)(1-ty) ... tx*ty,
            // summed left to right starting from 0.0, no FMA contraction
            const double v00 = (double)__ldg(dem + r0 * ds + c0);
            const double v01 = (double)__ldg(dem + r1 * ds + c0);
            const double v10 = (double)__ldg(dem + r0 * ds + c1);
            const double v11 = (double)__ldg(dem + r1 * ds + c1);
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
            values[k] = acc;
            touched[k] = 1;
        }
    }
}

AscAxis make_asc(const LinAxis &a) {
    AscAxis g;
    g.a = a;
    g.flip = a.step < 0.0;
    return g;
}

}  // namespace

int32_t k_interp(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const void *dem,
                 int dem_dtype, int64_t dem_stride, const uint8_t *mask,
                 int64_t mask_stride, const double *xy, int64_t K, int method,
                 uint8_t *touched, double *values) {
    AscAxis gx = make_asc(ax), gy = make_asc(ay);
    // np.min / np.max of the linspace vectors: endpoints (pinned stop)
    double x_a = lin_at_host(ax, 0), x_b = lin_at_host(ax, ax.n - 1);
    double y_a = lin_at_host(ay, 0), y_b = lin_at_host(ay, ay.n - 1);
    double xmin = fmin(x_a, x_b), xmax = fmax(x_a, x_b);
    double ymin = fmin(y_a, y_b), ymax = fmax(y_a, y_b);
    int64_t blocks = (K + 255) / 256;
    int64_t cap = (int64_t)ctx->num_sms * 16;
    if (blocks > cap) blocks = cap;
    const double2 *q = reinterpret_cast<const double2 *>(xy);
#define LAUNCH(T, M)                                                           \
    interp_kernel<T, M><<<(unsigned)blocks, 256, 0, ctx->stream>>>(            \
        gx, gy, xmin, xmax, ymin, ymax, (const T *)dem, dem_stride, mask,      \
        mask_stride, q, K, touched, values)
    OCSB_TIME(ctx, KC_INTERP);
    if (dem_dtype == OCSB_F32) {
        if (method == OCSB_INTERP_NEAREST) LAUNCH(float, OCSB_INTERP_NEAREST);
        else LAUNCH(float, OCSB_INTERP_LINEAR);
    } else {
        if (method == OCSB_INTERP_NEAREST) LAUNCH(double, OCSB_INTERP_NEAREST);
        else LAUNCH(double, OCSB_INTERP_LINEAR);
    }
#undef LAUNCH
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
