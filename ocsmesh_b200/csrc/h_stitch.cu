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
