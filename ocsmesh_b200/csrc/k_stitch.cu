// K5' -- contour polylines and the feature prologue on the device
// (reference: Raster.get_contour -> _get_raster_contour_windowed + shapely
//  linemerge, ocsmesh/raster.py:1340-1372, 1426-1470; repartition_features /
//  transform_linestring, ocsmesh/utils.py:247-312, called at
//  ocsmesh/hfun/raster.py:1219-1225).
//
// k_contour.cu leaves (edge-id pair, two crossing points) per line segment.
// Here the segments become polylines without a sequential walk:
//
//   1. nodes     : the 2n segment ends sorted by edge id (radix sort); equal
//                  ids = one node (a crossing point shared by the two cells
//                  of a grid edge); node index = rank of its edge id;
//   2. adjacency : a node has one neighbour (a line end on the window border)
//                  or two;
//   3. ranking   : every undirected link is two half-edges u->v; pointer
//                  jumping along "previous half-edge" gives each half-edge the
//                  head of its directed chain and its distance from it
//                  (Wyllie list ranking, ceil(log2) rounds); closed rings are
//                  found by the same jumps (no head is ever reached), cut at
//                  their smallest node and ranked again;
//   4. selection : of the two directed chains of a line the one that starts at
//                  the smaller end node is kept (a ring: from its smallest node
//                  towards its first neighbour) -- the order and orientation
//                  of features._stitch / h_stitch.cu, so the polylines are
//                  identical vertex for vertex; lines are ordered open lines
//                  first (by start node), then rings (by smallest node);
//   5. resample  : one thread per chunk of <= max_verts vertices walks its
//                  chunk sequentially (the running-sum order of np.cumsum, so
//                  the stations are bit-identical to features.feature_points).
#include "ocsb_common.cuh"

#include <cub/cub.cuh>

#include <algorithm>

namespace {

constexpr int kT = 256;
inline unsigned nblk(long long n) { return (unsigned)((n + kT - 1) / kT); }

__global__ void st_keys_kernel(const long long *edges, long long n2, long long *key, int *val) {
    const long long s = (long long)blockIdx.x * kT + threadIdx.x;
    if (s >= n2) return;
    key[s] = edges[s];
    val[s] = (int)s;
}

__global__ void st_heads_kernel(const long long *key, long long n2, int *head) {
    const long long j = (long long)blockIdx.x * kT + threadIdx.x;
    if (j >= n2) return;
    head[j] = (j == 0 || key[j] != key[j - 1]) ? 1 : 0;
}

// per run of equal keys: the node, its point, its (up to two) slots
__global__ void st_nodes_kernel(const long long *key, const int *val, const int *nid,
                                long long n2, const double2 *pts, double2 *node_xy,
                                int2 *node_slots, int *node_of_slot) {
    const long long j = (long long)blockIdx.x * kT + threadIdx.x;
    if (j >= n2) return;
    if (j > 0 && key[j] == key[j - 1]) return;
    const int u = nid[j] - 1;
    const int s1 = val[j];
    int s2 = -1;
    if (j + 1 < n2 && key[j + 1] == key[j]) s2 = val[j + 1];
    node_xy[u] = pts[s1];                 // first occurrence in segment order
    node_slots[u] = make_int2(s1, s2);
    node_of_slot[s1] = u;
    if (s2 >= 0) node_of_slot[s2] = u;
    // a third segment end on one grid edge cannot come out of the sweep
    for (long long k = j + 2; k < n2 && key[k] == key[j]; ++k) node_of_slot[val[k]] = u;
}

// neighbours in the order of features._stitch: ends where the node is the
// smaller end of its segment first, then by the other end
__global__ void st_nbrs_kernel(const int2 *node_slots, const int *node_of_slot, int nn,
                               int2 *nbr) {
    const int u = blockIdx.x * kT + threadIdx.x;
    if (u >= nn) return;
    const int2 s = node_slots[u];
    int o1 = node_of_slot[s.x ^ 1], o2 = s.y >= 0 ? node_of_slot[s.y ^ 1] : -1;
    if (o2 >= 0) {
        const bool l1 = o1 > u, l2 = o2 > u;      // u is the smaller end of that segment
        const bool swap = (l1 != l2) ? l2 : (o2 < o1);
        if (swap) { const int t = o1; o1 = o2; o2 = t; }
    }
    nbr[u] = make_int2(o1, o2);
}

__device__ __forceinline__ int nbr_k(const int2 n, int k) { return k ? n.y : n.x; }

// half-edge h = 2u + k : u -> nbr[u][k].  prev(h) = (t -> u), t the other
// neighbour of u; -1: h starts its chain; -2: no such half-edge.
// cut != nullptr: ring half-edges leaving their ring's smallest node start it.
__global__ void st_half_init_kernel(const int2 *nbr, int nn, const int *ring_min,
                                    int *p, int *d, int *aux, int mode) {
    const int h = blockIdx.x * kT + threadIdx.x;
    if (h >= 2 * nn) return;
    const int u = h >> 1, k = h & 1;
    const int2 nu = nbr[u];
    const int v = nbr_k(nu, k);
    if (v < 0) { p[h] = -2; d[h] = 0; aux[h] = mode ? h : u; return; }
    const int t = nbr_k(nu, 1 - k);
    int prev = -1;
    if (t >= 0 && !(mode && ring_min[h] == u)) {
        const int2 nt = nbr[t];
        prev = 2 * t + (nt.x == u ? 0 : 1);
    }
    p[h] = prev;
    d[h] = prev >= 0 ? 1 : 0;
    aux[h] = mode ? h : u;      // mode 0: running min of source nodes; 1: head
}

// one round of pointer jumping (double buffered)
__global__ void st_jump_kernel(const int *p_in, const int *d_in, const int *a_in, int n,
                               int *p_out, int *d_out, int *a_out, int mode) {
    const int h = blockIdx.x * kT + threadIdx.x;
    if (h >= n) return;
    const int p = p_in[h];
    int d = d_in[h], a = a_in[h], q = p;
    if (p >= 0) {
        d += d_in[p];
        a = mode ? a_in[p] : min(a, a_in[p]);
        q = p_in[p];
    }
    p_out[h] = q; d_out[h] = d; a_out[h] = a;
}

// after the first ranking: ring_min[h] = smallest node of h's ring, -1 for lines
__global__ void st_ring_kernel(const int *p, const int *mn, int n, int *ring_min) {
    const int h = blockIdx.x * kT + threadIdx.x;
    if (h >= n) return;
    ring_min[h] = p[h] >= 0 ? mn[h] : -1;
}

// which chain of a line is kept, its length, its head flag
__global__ void st_choose_kernel(const int2 *nbr, const int *p, const int *d, const int *head,
                                 const int *ring_min, int nn, int *chosen, int *len,
                                 int *is_head, long long *head_key) {
    const int h = blockIdx.x * kT + threadIdx.x;
    if (h >= 2 * nn) return;
    chosen[h] = 0; is_head[h] = 0;
    if (p[h] == -2) return;
    const int u = h >> 1, k = h & 1;
    const int H = head[h];
    bool keep;
    if (ring_min[h] >= 0) {
        keep = (H & 1) == 0;                // from the smallest node to its first neighbour
    } else {
        const int v = nbr_k(nbr[u], k);
        const int2 nv = nbr[v];
        const int twin = 2 * v + (nv.x == u ? 0 : 1);
        keep = (H >> 1) < (head[twin] >> 1);        // start at the smaller end node
    }
    if (!keep) return;
    chosen[h] = 1;
    atomicMax(&len[H], d[h] + 1);
    if (h == H) {
        is_head[h] = 1;
        head_key[h] = ((long long)(ring_min[h] >= 0 ? 1 : 0) << 40) | (long long)(H >> 1);
    }
}

__global__ void st_gather_heads_kernel(const int *is_head, const int *pos, const long long *head_key,
                                       int n, long long *keys, int *heads) {
    const int h = blockIdx.x * kT + threadIdx.x;
    if (h >= n || !is_head[h]) return;
    const int i = pos[h] - 1;
    keys[i] = head_key[h];
    heads[i] = h;
}

__global__ void st_line_len_kernel(const int *heads_sorted, const int *len, int nl, int *npts,
                                   int *line_of_head) {
    const int i = blockIdx.x * kT + threadIdx.x;
    if (i >= nl) return;
    const int H = heads_sorted[i];
    npts[i] = len[H] + 1;
    line_of_head[H] = i;
}

__global__ void st_scatter_kernel(const int2 *nbr, const int *chosen, const int *d,
                                  const int *head, const int *line_of_head, const long long *off,
                                  const double2 *node_xy, int nn, double2 *out) {
    const int h = blockIdx.x * kT + threadIdx.x;
    if (h >= 2 * nn || !chosen[h]) return;
    const int u = h >> 1, k = h & 1;
    const int H = head[h];
    const long long base = off[line_of_head[H]];
    out[base + d[h] + 1] = node_xy[nbr_k(nbr[u], k)];
    if (h == H) out[base] = node_xy[u];
}

// ---- resampling ----------------------------------------------------------------
__device__ double py_floordiv_dev(double x, double y) {
    // Python's float.__floordiv__ (what `total // target_size` evaluates)
    const double mod = fmod(x, y);
    double div = __ddiv_rn(__dsub_rn(x, mod), y);
    if (mod != 0.0 && ((y < 0) != (mod < 0))) div = __dsub_rn(div, 1.0);
    if (div == 0.0) return copysign(0.0, __ddiv_rn(x, y));
    double fl = floor(div);
    if (__dsub_rn(div, fl) > 0.5) fl = __dadd_rn(fl, 1.0);
    return fl;
}

__global__ void rs_nchunks_kernel(const long long *off, int nl, int max_verts, int *nch) {
    const int i = blockIdx.x * kT + threadIdx.x;
    if (i >= nl) return;
    const long long m = off[i + 1] - off[i];
    const long long step = max_verts - 1;
    nch[i] = m <= max_verts ? 1 : (int)((m - 1 + step - 1) / step);
}

// one chunk: c[0..m), sequential like h_stitch.cu:resample_chunk; out == nullptr counts
__device__ long long rs_chunk(const double2 *c, long long m, double target, double2 *out) {
    if (m < 2) {
        if (out) for (long long k = 0; k < m; ++k) out[k] = c[k];
        return m;
    }
    auto seg_len = [&](long long k) {
        const double dx = __dsub_rn(c[k + 1].x, c[k].x), dy = __dsub_rn(c[k + 1].y, c[k].y);
        return __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    };
    double total = 0.0;
    for (long long k = 0; k + 1 < m; ++k) {
        const double l = seg_len(k);
        total = (k == 0) ? l : __dadd_rn(total, l);
    }
    const double q = py_floordiv_dev(total, target);
    const long long nst = (q > 0.0 && q < 9e15) ? (long long)q : 0;
    if (nst == 0) {
        if (out) for (long long k = 0; k < m; ++k) out[k] = c[k];
        return m;
    }
    long long n = 0;
    if (out) out[n] = c[0];
    ++n;
    long long seg = 0;
    double len = seg_len(0), cum = len;
    auto station = [&](double dist) {
        // np.searchsorted(cum_len, dist, 'left'), clipped to the last segment
        while (seg < m - 2 && cum < dist) {
            ++seg;
            len = seg_len(seg);
            cum = __dadd_rn(cum, len);
        }
        if (out) {
            const double ratio = __ddiv_rn(__dsub_rn(cum, dist), len);
            const double om = __dsub_rn(1.0, ratio);
            out[n] = make_double2(
                __dadd_rn(__dmul_rn(c[seg].x, ratio), __dmul_rn(c[seg + 1].x, om)),
                __dadd_rn(__dmul_rn(c[seg].y, ratio), __dmul_rn(c[seg + 1].y, om)));
        }
        ++n;
    };
    double d = 0.0;
    for (long long k = 0; k < nst; ++k) {
        d = (k == 0) ? target : __dadd_rn(d, target);
        station(d);
    }
    if (d < total) station(total);
    return n;
}

// chunk g of the line list: its vertex range
__device__ __forceinline__ void rs_locate(const long long *off, const long long *chunk_base,
                                          int nl, int max_verts, long long g, long long *first,
                                          long long *cm) {
    int lo = 0, hi = nl - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (chunk_base[mid] <= g) lo = mid; else hi = mid - 1;
    }
    const long long m = off[lo + 1] - off[lo];
    const long long c = g - chunk_base[lo];
    const long long step = max_verts - 1;
    const long long f = m <= max_verts ? 0 : c * step;
    const long long l = m <= max_verts ? m - 1 : min(f + step, m - 1);
    *first = off[lo] + f;
    *cm = l - f + 1;
}

// One WARP per chunk.  Segment lengths are independent (lanes, coalesced
// loads); the running sums that decide the stations are taken by lane 0 in
// np.cumsum order from shared memory (bit-identical to the sequential walk,
// without its chain of dependent global loads); the stations themselves are
// independent again (binary search in the running sums).  Chunks longer than
// kRsSeg segments or with more than kRsSt stations take the sequential walk.
constexpr int kRsSeg = 256, kRsSt = 256, kRsWarps = 4;

struct RsScr {
    double len[kRsSeg], cum[kRsSeg], st[kRsSt];
};

template <bool WRITE>
__global__ void __launch_bounds__(kRsWarps * 32)
rs_kernel(const double2 *xy, const long long *off, const long long *chunk_base, int nl,
          long long nchunks, int max_verts, double target, long long *cnt,
          const long long *out_off, double2 *out) {
    __shared__ RsScr s_scr[kRsWarps];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long long g = (long long)blockIdx.x * kRsWarps + wid;
    if (g >= nchunks) return;
    long long first, m;
    rs_locate(off, chunk_base, nl, max_verts, g, &first, &m);
    const double2 *c = xy + first;
    double2 *o = WRITE ? out + out_off[g] : nullptr;
    RsScr &S = s_scr[wid];
    if (m < 2 || m - 1 > kRsSeg) {
        if (lane == 0) {
            const long long n = rs_chunk(c, m, target, o);
            if (!WRITE) cnt[g] = n;
        }
        return;
    }
    const int ns = (int)(m - 1);
    for (int k = lane; k < ns; k += 32) {
        const double dx = __dsub_rn(c[k + 1].x, c[k].x), dy = __dsub_rn(c[k + 1].y, c[k].y);
        S.len[k] = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    }
    __syncwarp();
    long long nst = 0;
    int extra = 0;
    if (lane == 0) {
        double acc = 0.0;
        for (int k = 0; k < ns; ++k) {
            acc = (k == 0) ? S.len[0] : __dadd_rn(acc, S.len[k]);
            S.cum[k] = acc;
        }
        const double total = acc;
        const double q = py_floordiv_dev(total, target);
        nst = (q > 0.0 && q < 9e15) ? (long long)q : 0;
        if (nst > 0 && nst <= kRsSt) {
            // stations: cumsum(ones(nst) * target), then `total` if the last is short
            double d = 0.0;
            for (int k = 0; k < (int)nst; ++k) {
                d = (k == 0) ? target : __dadd_rn(d, target);
                S.st[k] = d;
            }
            extra = d < total ? 1 : 0;
        }
    }
    nst = __shfl_sync(0xffffffffu, nst, 0);
    extra = __shfl_sync(0xffffffffu, extra, 0);
    if (nst > kRsSt) {                      // very fine target: sequential walk
        if (lane == 0) {
            const long long n = rs_chunk(c, m, target, o);
            if (!WRITE) cnt[g] = n;
        }
        return;
    }
    if (!WRITE) {
        if (lane == 0) cnt[g] = nst == 0 ? m : 1 + nst + extra;
        return;
    }
    __syncwarp();
    if (nst == 0) {                          // shorter than target: unchanged
        for (long long k = lane; k < m; k += 32) o[k] = c[k];
        return;
    }
    if (lane == 0) o[0] = c[0];
    const int nstations = (int)nst + extra;
    const double total = S.cum[ns - 1];
    for (int k = lane; k < nstations; k += 32) {
        const double dist = k < (int)nst ? S.st[k] : total;
        // np.searchsorted(cum_len, dist, 'left'), clipped to the last segment
        int lo = 0, hi = ns - 1;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (S.cum[mid] < dist) lo = mid + 1; else hi = mid;
        }
        const int seg = lo;
        const double ratio = __ddiv_rn(__dsub_rn(S.cum[seg], dist), S.len[seg]);
        const double om = __dsub_rn(1.0, ratio);
        o[1 + k] = make_double2(
            __dadd_rn(__dmul_rn(c[seg].x, ratio), __dmul_rn(c[seg + 1].x, om)),
            __dadd_rn(__dmul_rn(c[seg].y, ratio), __dmul_rn(c[seg + 1].y, om)));
    }
}

__global__ void widen_kernel(const int *a, int n, long long *b) {
    const int i = blockIdx.x * kT + threadIdx.x;
    if (i < n) b[i] = a[i];
}

// device -> host scalar through the pinned bounce buffer (one sync)
template <typename T> int32_t read_back(ocsb_ctx *ctx, const T *dev, T *host) {
    T local;
    T *bounce = ctx->pinned ? reinterpret_cast<T *>(ctx->pinned) : &local;
    OCSB_CUDA(cudaMemcpyAsync(bounce, dev, sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    *host = *bounce;
    return OCSB_OK;
}

template <typename In, typename Out>
int32_t excl_sum(ocsb_ctx *ctx, ScratchGuard &sg, const In *in, Out *out, long long n) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (int)n, ctx->stream);
    char *tmp = nullptr;
    OCSB_TRY(sg.get(&tmp, bytes + 16));
    OCSB_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int)n, ctx->stream));
    return OCSB_OK;
}

template <typename In, typename Out>
int32_t incl_sum(ocsb_ctx *ctx, ScratchGuard &sg, const In *in, Out *out, long long n) {
    size_t bytes = 0;
    cub::DeviceScan::InclusiveSum(nullptr, bytes, in, out, (int)n, ctx->stream);
    char *tmp = nullptr;
    OCSB_TRY(sg.get(&tmp, bytes + 16));
    OCSB_CUDA(cub::DeviceScan::InclusiveSum(tmp, bytes, in, out, (int)n, ctx->stream));
    return OCSB_OK;
}

template <typename K, typename V>
int32_t sort_pairs(ocsb_ctx *ctx, ScratchGuard &sg, const K *kin, K *kout, const V *vin, V *vout,
                   long long n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, kin, kout, vin, vout, (int)n, 0,
                                    (int)sizeof(K) * 8, ctx->stream);
    char *tmp = nullptr;
    OCSB_TRY(sg.get(&tmp, bytes + 16));
    OCSB_CUDA(cub::DeviceRadixSort::SortPairs(tmp, bytes, kin, kout, vin, vout, (int)n, 0,
                                              (int)sizeof(K) * 8, ctx->stream));
    return OCSB_OK;
}

}  // namespace

// Polylines of n segments (device arrays of k_contour_segments): *xy (n_xy
// points) and *off (n_lines + 1 offsets) are carved from `sg`.
int32_t k_stitch_segments(ocsb_ctx *ctx, ScratchGuard &sg, const int64_t *edges,
                          const double *pts, int64_t n, double **xy_out, int64_t **off_out,
                          int64_t *n_xy, int64_t *n_lines) {
    *n_xy = 0; *n_lines = 0; *xy_out = nullptr; *off_out = nullptr;
    if (n <= 0) return OCSB_OK;
    OCSB_REQUIRE(n < (1ll << 29), "stitch: too many segments");
    const long long n2 = 2 * n;
    cudaStream_t st = ctx->stream;
    long long *key = nullptr, *skey = nullptr;
    int *val = nullptr, *sval = nullptr, *headf = nullptr, *nid = nullptr, *node_of_slot = nullptr;
    OCSB_TRY(sg.get(&key, (size_t)n2)); OCSB_TRY(sg.get(&skey, (size_t)n2));
    OCSB_TRY(sg.get(&val, (size_t)n2)); OCSB_TRY(sg.get(&sval, (size_t)n2));
    OCSB_TRY(sg.get(&headf, (size_t)n2)); OCSB_TRY(sg.get(&nid, (size_t)n2));
    OCSB_TRY(sg.get(&node_of_slot, (size_t)n2));
    OCSB_TIME(ctx, KC_CONTOUR);
    st_keys_kernel<<<nblk(n2), kT, 0, st>>>((const long long *)edges, n2, key, val);
    OCSB_LAUNCHED(ctx);
    OCSB_TRY(sort_pairs(ctx, sg, key, skey, val, sval, n2));
    st_heads_kernel<<<nblk(n2), kT, 0, st>>>(skey, n2, headf);
    OCSB_TRY(incl_sum(ctx, sg, headf, nid, n2));
    int nn = 0;
    OCSB_TRY(read_back(ctx, nid + n2 - 1, &nn));
    double2 *node_xy = nullptr;
    int2 *node_slots = nullptr, *nbr = nullptr;
    OCSB_TRY(sg.get(&node_xy, (size_t)nn)); OCSB_TRY(sg.get(&node_slots, (size_t)nn));
    OCSB_TRY(sg.get(&nbr, (size_t)nn));
    st_nodes_kernel<<<nblk(n2), kT, 0, st>>>(skey, sval, nid, n2, (const double2 *)pts, node_xy,
                                             node_slots, node_of_slot);
    st_nbrs_kernel<<<nblk(nn), kT, 0, st>>>(node_slots, node_of_slot, nn, nbr);
    // ---- list ranking over the 2 nn half-edges --------------------------------------------
    const int nh = 2 * nn;
    int *buf[6];
    for (auto &b : buf) OCSB_TRY(sg.get(&b, (size_t)nh));
    int *ring_min = nullptr;
    OCSB_TRY(sg.get(&ring_min, (size_t)nh));
    int rounds = 1;
    while ((1ll << rounds) < nh) ++rounds;
    ++rounds;
    int *p = buf[0], *d = buf[1], *a = buf[2], *p2 = buf[3], *d2 = buf[4], *a2 = buf[5];
    for (int mode = 0; mode < 2; ++mode) {
        OCSB_TIME(ctx, KC_CONTOUR);
        st_half_init_kernel<<<nblk(nh), kT, 0, st>>>(nbr, nn, ring_min, p, d, a, mode);
        OCSB_LAUNCHED(ctx);
        for (int r = 0; r < rounds; ++r) {
            st_jump_kernel<<<nblk(nh), kT, 0, st>>>(p, d, a, nh, p2, d2, a2, mode);
            std::swap(p, p2); std::swap(d, d2); std::swap(a, a2);
        }
        if (mode == 0) st_ring_kernel<<<nblk(nh), kT, 0, st>>>(p, a, nh, ring_min);
    }
    // now: d = distance from the chain head, a = head half-edge, p = -1 / -2
    // (p == -2 marks absent half-edges: recompute from nbr in the consumers)
    int *chosen = p2, *len = d2, *is_head = a2;      // reuse the spare buffers
    long long *head_key = nullptr;
    OCSB_TRY(sg.get(&head_key, (size_t)nh));
    OCSB_CUDA(cudaMemsetAsync(len, 0, sizeof(int) * nh, st));
    // absent half-edges: p stayed -2 through the jumps
    st_choose_kernel<<<nblk(nh), kT, 0, st>>>(nbr, p, d, a, ring_min, nn, chosen, len, is_head,
                                              head_key);
    int *pos = nullptr;
    OCSB_TRY(sg.get(&pos, (size_t)nh));
    OCSB_TRY(incl_sum(ctx, sg, is_head, pos, nh));
    int nl = 0;
    OCSB_TRY(read_back(ctx, pos + nh - 1, &nl));
    if (nl == 0) return OCSB_OK;
    long long *lkeys = nullptr, *lkeys_s = nullptr, *off = nullptr;
    int *heads = nullptr, *heads_s = nullptr, *npts = nullptr, *line_of_head = nullptr;
    OCSB_TRY(sg.get(&lkeys, (size_t)nl)); OCSB_TRY(sg.get(&lkeys_s, (size_t)nl));
    OCSB_TRY(sg.get(&heads, (size_t)nl)); OCSB_TRY(sg.get(&heads_s, (size_t)nl));
    OCSB_TRY(sg.get(&npts, (size_t)nl + 1)); OCSB_TRY(sg.get(&line_of_head, (size_t)nh));
    OCSB_TRY(sg.get(&off, (size_t)nl + 1));
    st_gather_heads_kernel<<<nblk(nh), kT, 0, st>>>(is_head, pos, head_key, nh, lkeys, heads);
    OCSB_TRY(sort_pairs(ctx, sg, lkeys, lkeys_s, heads, heads_s, nl));
    OCSB_CUDA(cudaMemsetAsync(npts + nl, 0, sizeof(int), st));
    st_line_len_kernel<<<nblk(nl), kT, 0, st>>>(heads_s, len, nl, npts, line_of_head);
    OCSB_TRY(excl_sum(ctx, sg, npts, off, (long long)nl + 1));
    long long total = 0;
    OCSB_TRY(read_back(ctx, off + nl, &total));
    double2 *xy = nullptr;
    OCSB_TRY(sg.get(&xy, (size_t)total));
    OCSB_TIME(ctx, KC_CONTOUR);
    st_scatter_kernel<<<nblk(nh), kT, 0, st>>>(nbr, chosen, d, a, line_of_head, off, node_xy, nn,
                                               xy);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    *xy_out = reinterpret_cast<double *>(xy);
    *off_out = reinterpret_cast<int64_t *>(off);
    *n_xy = total;
    *n_lines = nl;
    return OCSB_OK;
}

// The vertex array of hfun/raster.py:1230 from device polylines: every line cut
// into chunks of max_verts vertices (sharing their end vertex), every chunk
// resampled at target_size.  *out is carved from `sg`.
int32_t k_resample_lines(ocsb_ctx *ctx, ScratchGuard &sg, const double *xy, const int64_t *off,
                         int64_t n_lines, double target, int max_verts, double **out,
                         int64_t *n_out) {
    *out = nullptr; *n_out = 0;
    if (n_lines <= 0) return OCSB_OK;
    OCSB_REQUIRE(n_lines < (1ll << 30), "resample: too many lines");
    cudaStream_t st = ctx->stream;
    const int nl = (int)n_lines;
    int *nch = nullptr;
    long long *nch64 = nullptr, *chunk_base = nullptr;
    OCSB_TRY(sg.get(&nch, (size_t)nl + 1));
    OCSB_TRY(sg.get(&nch64, (size_t)nl + 1));
    OCSB_TRY(sg.get(&chunk_base, (size_t)nl + 1));
    OCSB_CUDA(cudaMemsetAsync(nch + nl, 0, sizeof(int), st));
    rs_nchunks_kernel<<<nblk(nl), kT, 0, st>>>((const long long *)off, nl, max_verts, nch);
    widen_kernel<<<nblk(nl + 1), kT, 0, st>>>(nch, nl + 1, nch64);
    OCSB_TRY(excl_sum(ctx, sg, nch64, chunk_base, (long long)nl + 1));
    long long nchunks = 0;
    OCSB_TRY(read_back(ctx, chunk_base + nl, &nchunks));
    OCSB_REQUIRE(nchunks < (1ll << 30), "resample: too many chunks");
    long long *cnt = nullptr, *out_off = nullptr;
    OCSB_TRY(sg.get(&cnt, (size_t)nchunks + 1));
    OCSB_TRY(sg.get(&out_off, (size_t)nchunks + 1));
    OCSB_CUDA(cudaMemsetAsync(cnt + nchunks, 0, sizeof(long long), st));
    OCSB_TIME(ctx, KC_CONTOUR);
    rs_kernel<false><<<(unsigned)((nchunks + kRsWarps - 1) / kRsWarps), kRsWarps * 32, 0, st>>>(
        (const double2 *)xy, (const long long *)off, chunk_base, nl, nchunks, max_verts, target,
        cnt, nullptr, nullptr);
    OCSB_LAUNCHED(ctx);
    OCSB_TRY(excl_sum(ctx, sg, cnt, out_off, nchunks + 1));
    long long total = 0;
    OCSB_TRY(read_back(ctx, out_off + nchunks, &total));
    double2 *o = nullptr;
    OCSB_TRY(sg.get(&o, (size_t)std::max<long long>(total, 1)));
    OCSB_TIME(ctx, KC_CONTOUR);
    rs_kernel<true><<<(unsigned)((nchunks + kRsWarps - 1) / kRsWarps), kRsWarps * 32, 0, st>>>(
        (const double2 *)xy, (const long long *)off, chunk_base, nl, nchunks, max_verts, target,
        nullptr, out_off, o);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    *out = reinterpret_cast<double *>(o);
    *n_out = total;
    return OCSB_OK;
}
