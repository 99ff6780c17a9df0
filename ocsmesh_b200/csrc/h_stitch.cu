// Host side of the contour path: join the segments
// that ocsb_contour_segments emits into polylines.  The reference gets
// ordered paths from matplotlib and merges them with shapely.ops.linemerge
// (ocsmesh/raster.py:1426-1470, 1340-1372); here two segments are joined where
// they share a cell-edge id.  O(n log n) on the host, no device work: the
// segment list is ~1e5 entries for a 20000^2 DEM.
//
// The result is defined by these rules (ocsmesh_b200/features.py:_stitch is
// the same algorithm in numpy and the test oracle of this function):
//   * nodes = distinct edge ids, numbered in increasing id order
//   * segments are ordered by (min id, max id); a node's first neighbour is
//     the first segment that lists it as its smaller end, ...
//   * open lines first, each walked from its lower-numbered end node; then
//     closed rings from their lowest node, closed by repeating the start
#include "ocsb_common.cuh"

#include <algorithm>
#include <numeric>

extern "C" int32_t ocsb_stitch_segments(const int64_t *edges, const double *pts,
                                        int64_t n, double *xy_out, int64_t cap_xy,
                                        int64_t *offsets_out, int64_t cap_lines,
                                        int64_t *n_xy, int64_t *n_lines) {
    OCSB_REQUIRE(n >= 0 && n_xy && n_lines, "ocsb_stitch_segments: bad arguments");
    *n_xy = 0;
    *n_lines = 0;
    if (n == 0) return OCSB_OK;
    OCSB_REQUIRE(edges && pts && xy_out && offsets_out,
                 "ocsb_stitch_segments: NULL buffers");
    OCSB_REQUIRE(n < (int64_t)1 << 30, "ocsb_stitch_segments: too many segments");
    // ---- nodes: sorted distinct edge ids, coordinates of their first use ------
    std::vector<int64_t> ids(edges, edges + 2 * n);
    std::sort(ids.begin(), ids.end());
    ids.erase(std::unique(ids.begin(), ids.end()), ids.end());
    const int nn = (int)ids.size();
    auto node_of = [&](int64_t id) {
        return (int)(std::lower_bound(ids.begin(), ids.end(), id) - ids.begin());
    };
    std::vector<double> px(nn), py(nn);
    std::vector<char> have(nn, 0);
    std::vector<int> lo(n), hi(n);
    for (int64_t s = 0; s < n; ++s) {
        const int a = node_of(edges[2 * s]), b = node_of(edges[2 * s + 1]);
        if (!have[a]) { have[a] = 1; px[a] = pts[4 * s]; py[a] = pts[4 * s + 1]; }
        if (!have[b]) { have[b] = 1; px[b] = pts[4 * s + 2]; py[b] = pts[4 * s + 3]; }
        lo[s] = std::min(a, b);     // node numbers are monotone in the ids
        hi[s] = std::max(a, b);
    }
    // ---- canonical segment order ---------------------------------------------------
    std::vector<int> order(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int u, int v) {
        return lo[u] != lo[v] ? lo[u] < lo[v] : hi[u] < hi[v];
    });
    // ---- neighbour table: first use -> slot 0, every later use -> slot 1 -----------
    std::vector<int> n0(nn, -1), n1(nn, -1);
    std::vector<char> used(nn, 0);
    for (int64_t k = 0; k < n; ++k) {        // uses as the smaller end ...
        const int s = order[k], v = lo[s];
        if (!used[v]) { used[v] = 1; n0[v] = hi[s]; } else n1[v] = hi[s];
    }
    for (int64_t k = 0; k < n; ++k) {        // ... then as the larger end
        const int s = order[k], v = hi[s];
        if (!used[v]) { used[v] = 1; n0[v] = lo[s]; } else n1[v] = lo[s];
    }
    // ---- walks -----------------------------------------------------------------------
    std::vector<char> visited(nn, 0);
    int64_t nxy = 0, nl = 0;
    std::vector<int> path;
    bool overflow = false;
    auto walk = [&](int start) {
        path.clear();
        path.push_back(start);
        visited[start] = 1;
        int cur = start;
        for (;;) {
            const int a = n0[cur], b = n1[cur];
            int nxt;
            if (a >= 0 && !visited[a]) nxt = a;
            else if (b >= 0 && !visited[b]) nxt = b;
            else {
                if (path.size() > 2 && (start == a || start == b)) path.push_back(start);
                break;
            }
            path.push_back(nxt);
            visited[nxt] = 1;
            cur = nxt;
        }
        if (path.size() < 2) return;
        if (nl + 1 > cap_lines || nxy + (int64_t)path.size() > cap_xy) {
            overflow = true;
            return;
        }
        offsets_out[nl] = nxy;
        for (int v : path) {
            xy_out[2 * nxy] = px[v];
            xy_out[2 * nxy + 1] = py[v];
            ++nxy;
        }
        ++nl;
    };
    auto degree = [&](int v) { return (n0[v] >= 0 ? 1 : 0) + (n1[v] >= 0 ? 1 : 0); };
    for (int v = 0; v < nn && !overflow; ++v)
        if (degree(v) == 1 && !visited[v]) walk(v);
    for (int v = 0; v < nn && !overflow; ++v)
        if (degree(v) == 2 && !visited[v]) walk(v);
    if (overflow) {
        ocsb_set_error("ocsb_stitch_segments: output capacity too small "
                       "(2 n vertices and n lines always suffice)");
        return OCSB_EOVERFLOW;
    }
    offsets_out[nl] = nxy;
    *n_xy = nxy;
    *n_lines = nl;
    return OCSB_OK;
}

// ---------------------------------------------------------------------------
// Feature prologue of HfunRaster.add_feature (ocsmesh/hfun/raster.py:1219-1225):
//   utils.repartition_features  ocsmesh/utils.py:247-277  chunks of <= max_verts
//                               vertices sharing their end vertex
//   utils.transform_linestring  ocsmesh/utils.py:280-312  a point every
//                               target_size metres along each chunk (+ ends)
// restated operation by operation in the order numpy evaluates them
// (sequential cumsum, Python's float floor division, no FMA contraction), so
// the vertex array is bit-identical to ocsmesh_b200/features.py:feature_points
// and to the goldens of tests/golden (tests/test_host_cpu.py).
// ---------------------------------------------------------------------------
namespace {

// Python's float.__floordiv__ (Objects/floatobject.c:float_floor_div)
double py_floordiv(double x, double y) {
    double mod = fmod(x, y);
    double div = (x - mod) / y;
    if (mod != 0.0 && ((y < 0) != (mod < 0))) div -= 1.0;
    if (div == 0.0) return copysign(0.0, x / y);
    double fl = floor(div);
    if (div - fl > 0.5) fl += 1.0;
    return fl;
}

// one chunk c[0..m): appends to out (if non-null) and returns the point count
int64_t resample_chunk(const double *c, int64_t m, double target,
                       std::vector<double> &len, std::vector<double> &cum,
                       double *out) {
    auto put = [&](int64_t k, double x, double y) {
        if (out) { out[2 * k] = x; out[2 * k + 1] = y; }
    };
    if (m < 2) {
        for (int64_t k = 0; k < m; ++k) put(k, c[2 * k], c[2 * k + 1]);
        return m;
    }
    len.resize(m - 1);
    cum.resize(m - 1);
    double acc = 0.0;
    for (int64_t k = 0; k + 1 < m; ++k) {
        const double dx = c[2 * k + 2] - c[2 * k], dy = c[2 * k + 3] - c[2 * k + 1];
        volatile double sx = dx * dx, sy = dy * dy;     // no contraction
        const double l = sqrt(sx + sy);
        len[k] = l;
        acc = (k == 0) ? l : acc + l;
        cum[k] = acc;
    }
    const double total = cum[m - 2];
    const double q = py_floordiv(total, target);
    const int64_t nst = (q > 0.0 && q < 9e15) ? (int64_t)q : 0;
    if (nst == 0) {                                     // shorter than target
        for (int64_t k = 0; k < m; ++k) put(k, c[2 * k], c[2 * k + 1]);
        return m;
    }
    int64_t n = 0;
    put(n++, c[0], c[1]);
    // stations: cumsum(ones(nst) * target), then `total` if the last one is short
    double d = 0.0;
    int64_t seg = 0;
    auto station = [&](double dist) {
        // np.searchsorted(cum, dist, 'left'), clipped to the last segment
        while (seg < m - 2 && cum[seg] < dist) ++seg;
        const double ratio = (cum[seg] - dist) / len[seg];
        const double om = 1.0 - ratio;
        volatile double ax = c[2 * seg] * ratio, bx = c[2 * seg + 2] * om;
        volatile double ay = c[2 * seg + 1] * ratio, by = c[2 * seg + 3] * om;
        put(n++, ax + bx, ay + by);
    };
    for (int64_t k = 0; k < nst; ++k) {
        d = (k == 0) ? target : d + target;
        station(d);
    }
    if (d < total) station(total);
    return n;
}

}  // namespace

extern "C" int32_t ocsb_feature_points(const double *xy, const int64_t *offsets,
                                       int64_t n_lines, double target_size,
                                       int32_t max_verts, double *out, int64_t cap,
                                       int64_t *n_out) {
    OCSB_REQUIRE(n_lines >= 0 && n_out, "ocsb_feature_points: bad arguments");
    OCSB_REQUIRE(target_size > 0 && target_size == target_size,
                 "Argument target_size must be greater than zero.");
    OCSB_REQUIRE(max_verts >= 2, "ocsb_feature_points: max_verts must be >= 2");
    *n_out = 0;
    if (n_lines == 0) return OCSB_OK;
    OCSB_REQUIRE(xy && offsets, "ocsb_feature_points: NULL buffers");
    std::vector<double> len, cum;
    int64_t n = 0;
    for (int pass = 0; pass < 2; ++pass) {
        // pass 0 counts, pass 1 writes (skipped when out is NULL or too small)
        n = 0;
        for (int64_t l = 0; l < n_lines; ++l) {
            const double *c = xy + 2 * offsets[l];
            const int64_t m = offsets[l + 1] - offsets[l];
            OCSB_REQUIRE(m >= 0, "ocsb_feature_points: offsets must not decrease");
            // repartition: chunks of max_verts vertices, consecutive chunks
            // share one vertex (utils.py:263-274)
            const int64_t step = max_verts - 1;
            int64_t first = 0;
            do {
                int64_t last = (m <= max_verts) ? m - 1 : std::min(first + step, m - 1);
                const int64_t cm = last - first + 1;
                n += resample_chunk(c + 2 * first, cm, target_size, len, cum,
                                    pass ? out + 2 * n : nullptr);
                first = last;
            } while (first < m - 1);
        }
        if (pass == 0) {
            *n_out = n;
            if (!out) return OCSB_OK;
            if (n > cap) {
                ocsb_set_error("ocsb_feature_points: %lld points, capacity %lld",
                               (long long)n, (long long)cap);
                return OCSB_EOVERFLOW;
            }
        }
    }
    return OCSB_OK;
}
