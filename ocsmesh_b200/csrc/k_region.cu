// K8 -- polygon regions on the raster grid
// (reference: RegionConstraint.apply, ocsmesh/features/constraint.py:357-395:
//  geopandas sjoin(points, region, predicate='within') on the pixel
//  coordinates; HfunRaster.add_patch, ocsmesh/hfun/raster.py:890-933:
//  rasterio.mask.raster_geometry_mask(all_touched=True, invert=True)).
//
// shapely / GDAL are not importable here, so both predicates are RESTATED and
// their parity with the reference is UNPINNED:
//   * points  : even-odd crossing number of the +x ray from the pixel's window
//               coordinate (np.linspace, like get_xy) over ALL ring edges
//               (holes and the polygons of a multipolygon need no special
//               case); half-open in y, float64;
//   * touched : the pixel's area (edge-to-edge rectangle of the window
//               transform) is touched when its centre is inside (same count)
//               or any ring edge intersects the closed rectangle (Liang-Barsky
//               clip, float64) -- the geometric meaning of all_touched=True.
//
// One CTA = a 32 x 8 pixel tile.  Edges are culled per tile (y-range of the
// tile, nothing entirely left of it) into shared memory in chunks of 256, then
// every pixel walks the short list.  Output: float32 1 / 0 per pixel, so the
// result can stand in for the DEM of a TopoConstConstraint (a region
// constraint is that constraint on a 0 / 1 raster).
#include "ocsb_common.cuh"

#include <algorithm>

namespace {

constexpr int kTW = 32, kTH = 8, kThreads = kTW * kTH;

struct RegionArgs {
    LinAxis ax, ay;             // window coordinates (points mode)
    double x0, ytop, pw, ph;    // pixel rectangles (touched mode), north-up
    int rb, re, w;
    int touched;
    const double4 *edges;
    long long n_edges;
    float *out;                 // (virtual) window row 0
    long long stride;
    double bx0, by0, bx1, by1;  // bounding box of the edges
};

__device__ __forceinline__ bool seg_hits_rect(double4 e, double xa, double ya, double xb,
                                              double yb) {
    // Liang-Barsky: clip the segment against the closed rectangle
    const double dx = e.z - e.x, dy = e.w - e.y;
    double t0 = 0.0, t1 = 1.0;
    const double p[4] = {-dx, dx, -dy, dy};
    const double q[4] = {e.x - xa, xb - e.x, e.y - ya, yb - e.y};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (p[i] == 0.0) {
            if (q[i] < 0.0) return false;
        } else {
            const double r = q[i] / p[i];
            if (p[i] < 0.0) { if (r > t1) return false; if (r > t0) t0 = r; }
            else            { if (r < t0) return false; if (r < t1) t1 = r; }
        }
    }
    return t0 <= t1;
}

__global__ void __launch_bounds__(kThreads) region_kernel(RegionArgs A) {
    __shared__ double4 s_e[kThreads];
    __shared__ int s_n;
    const int tx = threadIdx.x % kTW, ty = threadIdx.x / kTW;
    const int c = blockIdx.x * kTW + tx, r = A.rb + blockIdx.y * kTH + ty;
    const bool live = c < A.w && r < A.re;
    // this pixel's point and rectangle
    double px = 0.0, py = 0.0, xa = 0.0, xb = 0.0, ya = 0.0, yb = 0.0;
    if (A.touched) {
        const int cc = min(c, A.w - 1), rr = min(r, A.re - 1);
        xa = A.x0 + cc * A.pw; xb = A.x0 + (cc + 1) * A.pw;
        yb = A.ytop - rr * A.ph; ya = A.ytop - (rr + 1) * A.ph;
        px = 0.5 * (xa + xb); py = 0.5 * (ya + yb);
    } else {
        px = lin_at(A.ax, min(c, A.w - 1));
        py = lin_at(A.ay, min(r, A.re - 1));
        xa = xb = px; ya = yb = py;
    }
    // tile extent (all threads compute their own; reduce through shared memory
    // is not needed: the tile is an axis-aligned block of the same formulas)
    const int c0 = blockIdx.x * kTW, c1 = min(A.w, c0 + kTW) - 1;
    const int r0 = A.rb + blockIdx.y * kTH, r1 = min(A.re, r0 + kTH) - 1;
    double tx0, tx1, ty0, ty1;
    if (A.touched) {
        tx0 = A.x0 + c0 * A.pw; tx1 = A.x0 + (c1 + 1) * A.pw;
        ty1 = A.ytop - r0 * A.ph; ty0 = A.ytop - (r1 + 1) * A.ph;
    } else {
        const double xa_ = lin_at(A.ax, c0), xb_ = lin_at(A.ax, c1);
        const double ya_ = lin_at(A.ay, r0), yb_ = lin_at(A.ay, r1);
        tx0 = fmin(xa_, xb_); tx1 = fmax(xa_, xb_);
        ty0 = fmin(ya_, yb_); ty1 = fmax(ya_, yb_);
    }
    int cnt = 0;
    bool hit = false;
    // a tile outside the edges' bounding box is outside every ring
    const bool tile_out = tx1 < A.bx0 || tx0 > A.bx1 || ty1 < A.by0 || ty0 > A.by1;
    if (!tile_out) {
        for (long long base = 0; base < A.n_edges; base += kThreads) {
            if (threadIdx.x == 0) s_n = 0;
            __syncthreads();
            const long long i = base + threadIdx.x;
            if (i < A.n_edges) {
                const double4 e = A.edges[i];
                const double ey0 = fmin(e.y, e.w), ey1 = fmax(e.y, e.w);
                // crossing a +x ray of the tile, or touching one of its pixels
                if (ey1 >= ty0 && ey0 <= ty1 && fmax(e.x, e.z) >= tx0)
                    s_e[atomicAdd(&s_n, 1)] = e;
            }
            __syncthreads();
            const int n = s_n;
            for (int k = 0; k < n; ++k) {
                const double4 e = s_e[k];
                if ((e.y > py) != (e.w > py)) {
                    const double xint = e.x + (py - e.y) * (e.z - e.x) / (e.w - e.y);
                    if (px < xint) ++cnt;
                }
                if (A.touched && !hit) hit = seg_hits_rect(e, xa, ya, xb, yb);
            }
            __syncthreads();
        }
    }
    if (live) A.out[(long long)r * A.stride + c] = ((cnt & 1) || hit) ? 1.0f : 0.0f;
}

}  // namespace

int32_t k_region_mask(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, double x0, double ytop, double pw,
                      double ph, int rb, int re, int w, int touched, const double *edges,
                      int64_t n_edges, const double bbox[4], float *out, int64_t stride) {
    if (re <= rb || w <= 0) return OCSB_OK;
    RegionArgs A;
    A.ax = ax; A.ay = ay;
    A.x0 = x0; A.ytop = ytop; A.pw = pw; A.ph = ph;
    A.rb = rb; A.re = re; A.w = w; A.touched = touched;
    A.edges = reinterpret_cast<const double4 *>(edges);
    A.n_edges = n_edges;
    A.out = out; A.stride = stride;
    A.bx0 = bbox[0]; A.by0 = bbox[1]; A.bx1 = bbox[2]; A.by1 = bbox[3];
    const dim3 grid((unsigned)((w + kTW - 1) / kTW), (unsigned)((re - rb + kTH - 1) / kTH));
    OCSB_REQUIRE(grid.y <= 65535u * 64u, "ocsb_region_mask: band too tall");
    OCSB_TIME(ctx, KC_CONSTRAINT);
    if (grid.y <= 65535u) {
        region_kernel<<<grid, kThreads, 0, ctx->stream>>>(A);
        OCSB_LAUNCHED(ctx);
    } else {
        // more rows than one grid dimension holds: slabs of 65535 tile rows
        for (int a = rb; a < re; a += 65535 * kTH) {
            RegionArgs B = A;
            B.rb = a; B.re = std::min(re, a + 65535 * kTH);
            const dim3 g(grid.x, (unsigned)((B.re - B.rb + kTH - 1) / kTH));
            region_kernel<<<g, kThreads, 0, ctx->stream>>>(B);
        }
        OCSB_LAUNCHED(ctx);
    }
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
