// Geometry helpers shared by the K1 translation units (k_feature.cu, k_pass.cu);
// included INSIDE each unit's anonymous namespace.
//
// Candidate test (exact, conservative; see k_feature.cu header): a vertex v can
// be dropped from a rect's list when mindist(v, rect) > r, or when some kept
// vertex u is at least as close as v at all four rect corners.
#pragma once

constexpr int kLeafW = 32, kLeafH = 8;
constexpr int kTile0 = 64;            // finest list level (pixels)
constexpr int kEvalThreads = 256;
constexpr int kCandCap = 1024;        // per-CTA shared staging of a tile list
constexpr double kDomEps = 1e-12;     // relative safety margin of the tests
constexpr int kMaxLevels = 12;
constexpr int kMaxSplit = 64;

struct Geom {
    LinAxis ax, ay;
    int w, h;     // extent of the device tile (band) in pixels
    int row0;     // first window row of the band: tile row r is window row row0 + r
};

struct Rect {
    double xlo, xhi, ylo, yhi;
};

__device__ __forceinline__ Rect tile_rect(const Geom &g, int c0, int c1, int r0,
                                          int r1) {
    double xa = lin_at(g.ax, c0), xb = lin_at(g.ax, c1);
    double ya = lin_at(g.ay, g.row0 + r0), yb = lin_at(g.ay, g.row0 + r1);
    Rect q;
    q.xlo = fmin(xa, xb); q.xhi = fmax(xa, xb);
    q.ylo = fmin(ya, yb); q.yhi = fmax(ya, yb);
    return q;
}

// squared distances of v to the four corners (00 = (xlo,ylo), 01 = (xlo,yhi),
// 10 = (xhi,ylo), 11 = (xhi,yhi)) and to the rect itself
struct CornerD2 {
    double d00, d01, d10, d11;
};

__device__ __forceinline__ CornerD2 corner_d2(const Rect &q, double2 v) {
    double a = v.x - q.xlo, b = v.x - q.xhi, c = v.y - q.ylo, d = v.y - q.yhi;
    a *= a; b *= b; c *= c; d *= d;
    CornerD2 o;
    o.d00 = a + c; o.d01 = a + d; o.d10 = b + c; o.d11 = b + d;
    return o;
}

__device__ __forceinline__ double rect_d2(const Rect &q, double2 v) {
    double mx = fmax(fmax(q.xlo - v.x, v.x - q.xhi), 0.0);
    double my = fmax(fmax(q.ylo - v.y, v.y - q.yhi), 0.0);
    return mx * mx + my * my;
}

__device__ __forceinline__ double center_d2(const Rect &q, double2 v) {
    double cx = 0.5 * (q.xlo + q.xhi), cy = 0.5 * (q.ylo + q.yhi);
    double dx = v.x - cx, dy = v.y - cy;
    return dx * dx + dy * dy;
}

struct Best {
    double d2;
    int id;
};
__device__ __forceinline__ void best_upd(Best &b, double d2, int id) {
    if (d2 < b.d2 || (d2 == b.d2 && id < b.id)) { b.d2 = d2; b.id = id; }
}
__device__ __forceinline__ Best best_warp_min(Best b) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double d = __shfl_xor_sync(0xffffffffu, b.d2, o);
        int i = __shfl_xor_sync(0xffffffffu, b.id, o);
        best_upd(b, d, i);
    }
    return b;
}

// dominator table: squared corner distances (already inflated by 1+eps) of up
// to 5 dominators
struct Dom {
    double c[5][4];
};

__device__ __forceinline__ bool dominated(const Dom &D, const CornerD2 &v) {
    bool dom = false;
#pragma unroll
    for (int j = 0; j < 5; ++j)
        dom = dom || (v.d00 > D.c[j][0] && v.d01 > D.c[j][1] &&
                      v.d10 > D.c[j][2] && v.d11 > D.c[j][3]);
    return dom;
}

