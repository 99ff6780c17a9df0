// K-pass, persistent form -- included INSIDE k_pass.cu's anonymous namespace.
//
// Same arithmetic as tile_kernel (one feature op and / or one limiter op, fused
// clamp), organised as a persistent producer / consumer pipeline in which the
// eight warps of a CTA never meet at a CTA barrier:
//
//   * grid = resident CTAs only; 64x64-px tiles are handed out in chunks of up to 4
//     by one atomic counter (a tile with candidates costs ~2x a tile without;
//     one atomic per chunk: its latency is the producer's only stall);
//   * thread 0 is the producer.  Two tiles ahead it issues ONE mbarrier-tracked
//     transaction per stage (3-stage ring): the DEM box
//     (cp.async.bulk.tensor.2d), the tile's candidate slot (512 B: header + up
//     to 31 vertices, built by l0_kernel from the 256-px lists) and the window
//     record (cp.async.bulk), so neither the DEM latency (light tiles have
//     nothing else to hide it behind) nor dependent list loads are waited for;
//   * consumers: a warp owns a 32x16 rect of the tile -- own coordinates, own
//     filter of the slot to the rect, own pixels -- and hands the stage back
//     through an "empty" mbarrier (8 arrivals).  Interior tiles run a variant
//     without any per-pixel bound or border test.
//
// Reference: the window loops of ocsmesh/hfun/raster.py:1212-1260, 1393-1436.

constexpr int kStages = 3;
constexpr int kDemStage = 5024;      // floats per stage: 66 x 76 = 5016, padded to 128 B
constexpr int kSlot = 32;            // double2 per tile slot: header + 31 candidates
constexpr int kSlotCap = kSlot - 1;
constexpr int kSurv2 = 32;           // survivors of a warp's rect kept in shared memory
constexpr int kP2Threads = 256;
constexpr int kP2Warps = kP2Threads / 32;
#ifndef OCSB_PASS2_CHUNK
#define OCSB_PASS2_CHUNK 4
#endif
constexpr int kChunkMax = OCSB_PASS2_CHUNK; // consecutive tiles per fetch of the tile counter
                                            // (fewer when a CTA only gets a few tiles)

// slot header (16 bytes): candidates of the tile; more than kSlotCap live in the
// overflow pool at element `off`
struct SlotHead {
    int cnt, pad;
    long long off;
};

struct P2Info {
    int t, tx, ty, pad;
};

struct P2Args {
    const PWin *wins;
    int n_wins, n_tiles;
    const float *in;
    float *out;
    long long in_stride, out_stride;
    FeatDev feat;
    const double2 *slots;   // n_tiles x kSlot
    const double2 *pool0;   // overflow lists of the 64-px level
    LimiterDev<float> lim;
    float lo, hi;
    int *counter;           // zeroed before the launch
    int chunk;              // tiles per fetch of the counter
    // LIST form (the feature half of a split pass): the units of work are the
    // entries of `active` (tiles with candidates, any order), *n_active of them
    const int *active;
    const int *n_active;
    const int *overflow;
    const int *sticky;      // status word of a pass sequence (or nullptr)
};

struct W2Scr {
    DomTable dom;           // 160 B
    double2 sv[kSurv2];     // survivors of the warp's rect
    double x[32], y[16];    // coordinates of the rect's columns / rows
};

struct __align__(128) P2Smem {
    float dem[kStages][kDemStage];
    double2 slot[kStages][kSlot];
    PWin win[kStages];
    P2Info info[kStages];
    W2Scr w[kP2Warps];
    unsigned long long full[kStages], empty[kStages];
    // producer state: the chunk [q_next, q_end) being handed out, the decode of
    // q_next (q_tx < 0: not decoded) and a cache of its window (tile -> window
    // is monotone)
    int q_next, q_end, q_tx, q_ty, ended;
    int n_work;             // tiles (or list entries) of this launch
    int cur_wi, cur_base0, cur_n0x, cur_next;
    int cur_demc, cur_demr;
};

__device__ __forceinline__ void p2_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    }
}

__device__ __forceinline__ void p2_bulk(void *dst_smem, const void *src, uint32_t bytes,
                                        uint32_t bar) {
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(dst_smem);
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1], %2, [%3];"
        ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// producer: the next tile -> stage (info, window record, slot, DEM box); past
// the last tile it publishes the end marker instead
template <bool HAS_FEAT, bool HAS_LIM, bool LIST>
__device__ __forceinline__ void p2_produce(const CUtensorMap *tmap, const P2Args &A,
                                           P2Smem &S, int stage) {
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&S.full[stage]);
    int u = S.q_next, tx = S.q_tx, ty = S.q_ty;
    if (u == S.q_end) {
        u = atomicAdd(A.counter, A.chunk) + (int)gridDim.x * A.chunk;
        S.q_end = u + A.chunk;
        tx = -1;
    }
    if (u >= S.n_work) {
        S.ended = 1;
        S.info[stage].t = -1;
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
        return;
    }
    int t = u;
    if (LIST) { t = A.active[u]; tx = -1; }     // list entries are in no order
    if (t >= S.cur_next || (LIST && t < S.cur_base0)) {
        int wi = S.cur_wi;
        if (LIST) wi = find_win(A.wins, A.n_wins, t, 0);
        else while (wi + 1 < A.n_wins && t >= A.wins[wi + 1].base0) ++wi;
        const PWin &G = A.wins[wi];
        S.cur_wi = wi; S.cur_base0 = G.base0; S.cur_n0x = G.n0x;
        S.cur_next = wi + 1 < A.n_wins ? A.wins[wi + 1].base0 : 0x7fffffff;
        S.cur_demc = G.dem_c; S.cur_demr = G.dem_r + G.rb;
        tx = -1;
    }
    const int n0x = S.cur_n0x;
    if (tx < 0) {
        const int lt = t - S.cur_base0;
        ty = lt / n0x;
        tx = lt - ty * n0x;
    }
    P2Info I;
    I.t = t; I.tx = tx; I.ty = ty; I.pad = 0;
    S.info[stage] = I;
    uint32_t bytes = (uint32_t)sizeof(PWin);
    if (HAS_FEAT) bytes += (uint32_t)(kSlot * 16);
    if (HAS_LIM) bytes += (uint32_t)(kDemRows * kDemPitch * 4);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 ::"r"(bar), "r"(bytes) : "memory");
    p2_bulk(&S.win[stage], &A.wins[S.cur_wi], (uint32_t)sizeof(PWin), bar);
    if (HAS_FEAT) p2_bulk(&S.slot[stage][0], A.slots + (size_t)t * kSlot, kSlot * 16, bar);
    if (HAS_LIM) {
        // the box starts on a 16-byte boundary of the DEM rows (TMA requirement):
        // up to 3 columns further left, the consumers shift by dem_c & 3
        const int cx = (S.cur_demc + tx * kT0 - 4) & ~3, cy = S.cur_demr + ty * kT0 - 1;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&S.dem[stage][0]);
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
            " [%0], [%1, {%2, %3}], [%4];"
            ::"r"(dst), "l"(tmap), "r"(cx), "r"(cy), "r"(bar) : "memory");
    }
    // the tile after this one (a window change re-decodes it)
    S.q_next = u + 1;
    if (++tx == n0x) { tx = 0; ++ty; }
    S.q_tx = tx; S.q_ty = ty;
}

// ---- the pixels of one lane: 4 columns x 4 rows, two rows at a time -----------
// FULL: the tile lies strictly inside its window (no border pixel, everything
// written, 16-byte accesses aligned): no per-pixel test survives.
template <bool HAS_OLD, bool HAS_FEAT, bool HAS_LIM, bool BAND, bool FULL>
__device__ __forceinline__ void p2_pixels(const P2Args &A, const float *sdem,
                                          const PWin &W, int c0t, int r0t, int lcol,
                                          int lrow, int ntile, int ns, bool refilter,
                                          const double2 *list, bool far, int lane,
                                          W2Scr &WS, int sc0, int sr0, int sc1, int sr1) {
    const FeatDev &F = A.feat;
    const int bw = W.ax.n, rb = W.rb;
    const int gcol = c0t + lcol, grow = r0t + lrow;
    const int wc1 = FULL ? sc1 : min(sc1, W.keep_w - 1);
    const int wr1 = FULL ? sr1 : min(sr1, W.keep_h - 1 - rb);
    const bool vec_r = FULL || (W.vec && gcol + 3 <= sc1);
    const bool vec_w = FULL || (W.vec && gcol + 3 <= wc1);
    float *const out0 = A.out + W.hoff + (long long)(rb + grow) * A.out_stride + gcol;
    const float *const in0 =
        HAS_OLD ? A.in + W.hoff + (long long)(rb + grow) * A.in_stride + gcol : nullptr;
    // the warp's own coordinates: column sc0 + i -> WS.x[i], row sr0 + j -> WS.y[j]
    const double *sx = WS.x + (lane & 7) * 4, *sy = WS.y + (lane >> 3) * 4;

#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
        const int hr = grow + half * 2;
        // lanes outside a partial rect leave early -- unless the chunk-wise path
        // with its warp-wide votes follows (their loads / stores are masked)
        if (!FULL && !refilter && (hr > sr1 || gcol > sc1)) break;
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = 3.402823466e+38f;
        auto load_old = [&]() {
            if (!HAS_OLD) return;
            const float *src = in0 + (long long)(half * 2) * A.in_stride;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (!FULL && hr + j > sr1) continue;
                const float *orow = src + j * A.in_stride;
                if (vec_r) {
                    const float4 v4 = __ldcs(reinterpret_cast<const float4 *>(orow));
                    o[j * 4 + 0] = v4.x; o[j * 4 + 1] = v4.y;
                    o[j * 4 + 2] = v4.z; o[j * 4 + 3] = v4.w;
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (gcol + k <= sc1) o[j * 4 + k] = orow[k];
                }
            }
        };
        // the feature-only form (second half of a split pass) is bound by the
        // latency of these loads when they follow the distance evaluation
        // (long-scoreboard stalls, profiles/r2h_ncu_full_summary.txt): issue them first
        constexpr bool EARLY_OLD = HAS_OLD && HAS_FEAT && !HAS_LIM;
        if (EARLY_OLD) load_old();
        float vf[8];
        if (HAS_FEAT) {
            if (ntile == 0) {
#pragma unroll
                for (int k = 0; k < 8; ++k) vf[k] = F.v_far_f;
            } else {
                double px[4];
                {
                    const double2 a = *reinterpret_cast<const double2 *>(&sx[0]);
                    const double2 c = *reinterpret_cast<const double2 *>(&sx[2]);
                    px[0] = a.x; px[1] = a.y; px[2] = c.x; px[3] = c.y;
                }
                const double2 py = *reinterpret_cast<const double2 *>(&sy[half * 2]);
                double best[8];
                auto eval_one = [&](const double2 sv, bool first) {
                    double dx2[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const double d = __dsub_rn(px[k], sv.x);
                        dx2[k] = __dmul_rn(d, d);
                    }
                    const double e0 = __dsub_rn(py.x, sv.y), e1 = __dsub_rn(py.y, sv.y);
                    const double dy2[2] = {__dmul_rn(e0, e0), __dmul_rn(e1, e1)};
#pragma unroll
                    for (int j = 0; j < 2; ++j)
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const double d2 = __dadd_rn(dx2[k], dy2[j]);
                            best[j * 4 + k] = (first || d2 < best[j * 4 + k])
                                                  ? d2 : best[j * 4 + k];
                        }
                };
                if (!refilter) {
                    if (ns > 0) eval_one(WS.sv[0], true);
                    else {
#pragma unroll
                        for (int k = 0; k < 8; ++k) best[k] = kBig;
                    }
                    for (int s = 1; s < ns; ++s) eval_one(WS.sv[s], false);
                } else {
                    // more than kSurv2 survivors: filter and evaluate chunk-wise
#pragma unroll
                    for (int k = 0; k < 8; ++k) best[k] = kBig;
                    Rect q;
                    {
                        const double xa = WS.x[0], xb = WS.x[sc1 - sc0];
                        const double ya = WS.y[0], yb = WS.y[sr1 - sr0];
                        q.xlo = xa < xb ? xa : xb; q.xhi = xa < xb ? xb : xa;
                        q.ylo = ya < yb ? ya : yb; q.yhi = ya < yb ? yb : ya;
                    }
                    for (int b0 = 0; b0 < ntile; b0 += 32) {
                        const int i = b0 + lane;
                        const bool valid = i < ntile;
                        double2 v = make_double2(0.0, 0.0);
                        if (valid) v = list[i];
                        const bool keep = valid && !(rect_d2(q, v) > F.r2lim) &&
                                          !dominated_tab(WS.dom, corner_d2(q, v));
                        const unsigned m = __ballot_sync(0xffffffffu, keep);
                        const int n = __popc(m);
                        if (keep) WS.sv[__popc(m & ((1u << lane) - 1u))] = v;
                        __syncwarp();
                        for (int s = 0; s < n; ++s) eval_one(WS.sv[s], false);
                        __syncwarp();
                    }
                }
                size_law8(best, F, far ? 0xffffffffu : F.hi_lo, vf);
            }
        }
        // ---- pre-op values -------------------------------------------------------------
        if (!EARLY_OLD) load_old();
        if (HAS_FEAT) {
            // clamp(vf); min(old, .); A10 clamp: fused (monotone, idempotent)
#pragma unroll
            for (int k = 0; k < 8; ++k)
                o[k] = clamp_nan(min3_nan(o[k], vf[k], F.v_far_f), A.lo, A.hi);
        }
        // ---- limiter op ------------------------------------------------------------------
        if (HAS_LIM) {
            const int H = W.ay.n;
            const int shift = FULL ? 0 : (W.dem_c & 3);
            const float *sd = sdem + (lrow + half * 2) * kDemPitch + 4 + lcol + shift;
            float4 rowv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) rowv[j] = dem_row4(sd + j * kDemPitch, shift);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (!FULL && hr + j > sr1) continue;
                const int wr = rb + hr + j;        // window row
                const float4 up = rowv[j], zc = rowv[j + 1], dn = rowv[j + 2];
                const float west = sd[(j + 1) * kDemPitch - 1];
                const float east = sd[(j + 1) * kDemPitch + 4];
                const bool interior = FULL || (wr >= 1 && wr <= H - 2 && gcol >= 1 &&
                                               gcol + 4 <= bw - 1);
                float h[4];
                bool done = false;
                if (interior) done = limiter_raw4(up, zc, dn, west, east, A.lim, h);
                const float zz[4] = {zc.x, zc.y, zc.z, zc.w};
                if (!done) {
                    // generic IEEE path: window borders, zero / subnormal / huge
                    // gradients, NaN (bit-identical, just slower)
                    const bool top = (wr == 0), bot = (wr == H - 1);
                    const float uu[4] = {up.x, up.y, up.z, up.w};
                    const float dd[4] = {dn.x, dn.y, dn.z, dn.w};
                    const LimiterDev<float> &p = A.lim;
                    const ConstDiv<float> d0 = (top || bot) ? p.d1_0 : p.d2_0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int c = gcol + k;
                        const float u = top ? zz[k] : uu[k];
                        const float d = bot ? zz[k] : dd[k];
                        const float g0 = cdiv(__fsub_rn(d, u), d0);
                        const float zl0 = (k == 0) ? west : zz[k > 0 ? k - 1 : 0];
                        const float zr0 = (k == 3) ? east : zz[k < 3 ? k + 1 : 3];
                        const bool lft = (c == 0), rgt = (c >= bw - 1);
                        const float zl = lft ? zz[k] : zl0;
                        const float zr = rgt ? zz[k] : zr0;
                        const ConstDiv<float> d1 = (lft || rgt) ? p.d1_1 : p.d2_1;
                        const float g1 = cdiv(__fsub_rn(zr, zl), d1);
                        const float dh = __fsqrt_rn(
                            __fadd_rn(__fmul_rn(g0, g0), __fmul_rn(g1, g1)));
                        const float qv = __fdiv_rn(zz[k], dh);
                        const float hv = (float)fabs(__dmul_rn(1.0 / 3.0, (double)qv));
                        h[k] = (dh == 0.0f) ? (float)(1.0 / 3.0) : hv;
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    o[j * 4 + k] = limiter_tail<BAND, true, float>(h[k], zz[k], o[j * 4 + k],
                                                                   A.lim);
            }
        }
        // ---- store -------------------------------------------------------------------------
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            if (!FULL && (hr + j > wr1 || gcol > wc1)) continue;
            float *nrow = out0 + (long long)(half * 2 + j) * A.out_stride;
            if (vec_w) {
                __stcs(reinterpret_cast<float4 *>(nrow),
                       make_float4(o[j * 4 + 0], o[j * 4 + 1], o[j * 4 + 2], o[j * 4 + 3]));
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (gcol + k <= wc1) nrow[k] = o[j * 4 + k];
            }
        }
    }
}

#ifndef OCSB_PASS2_MIN_CTAS
#define OCSB_PASS2_MIN_CTAS 3
#endif

template <bool HAS_OLD, bool HAS_FEAT, bool HAS_LIM, bool BAND, bool LIST>
__global__ void __launch_bounds__(kP2Threads, OCSB_PASS2_MIN_CTAS)
pass2_kernel(const __grid_constant__ CUtensorMap tmap, const P2Args A) {
    extern __shared__ __align__(128) unsigned char p2_raw[];
    P2Smem &S = *reinterpret_cast<P2Smem *>(p2_raw);
    TL_SCOPE(LIST ? 5 : 4);

    const int2 flags = *reinterpret_cast<const int2 *>(A.overflow);   // {overflow, far}
    if (flags.x || (A.sticky && *A.sticky)) return;
    const bool far = flags.y != 0;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int lcol = (wid & 1) * 32 + (lane & 7) * 4;      // tile-local
    const int lrow = (wid >> 1) * 16 + (lane >> 3) * 4;
    W2Scr &WS = S.w[wid];
    const FeatDev &F = A.feat;

    // ---- producer prologue: stages of iterations 0 and 1 --------------------------------
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            const uint32_t fb = (uint32_t)__cvta_generic_to_shared(&S.full[s]);
            const uint32_t eb = (uint32_t)__cvta_generic_to_shared(&S.empty[s]);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(fb));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(eb), "r"(kP2Warps));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        S.cur_wi = 0; S.cur_next = -1;       // forces the first window lookup
        S.q_next = (int)blockIdx.x * A.chunk;  // the first chunk is static
        S.q_end = S.q_next + A.chunk;
        S.q_tx = -1; S.q_ty = 0; S.ended = 0;
        S.cur_base0 = 0;
        S.n_work = LIST ? min(*A.n_active, A.n_tiles) : A.n_tiles;
        p2_produce<HAS_FEAT, HAS_LIM, LIST>(&tmap, A, S, 0);
        if (!S.ended) p2_produce<HAS_FEAT, HAS_LIM, LIST>(&tmap, A, S, 1);
    }
    __syncthreads();         // barrier words initialised before anyone polls them

    for (int it = 0;; ++it) {
        const int s = it % kStages;
        const uint32_t use = (uint32_t)(it / kStages);
        // ---- producer: stage of iteration it + 2 ------------------------------------------------
        if (threadIdx.x == 0 && !S.ended) {
            const int s2 = (it + 2) % kStages;
            if (it >= 1) {
                // its previous use was iteration it - 1: wait until all warps left it
                const uint32_t eb = (uint32_t)__cvta_generic_to_shared(&S.empty[s2]);
                p2_wait(eb, (uint32_t)((it - 1) / kStages) & 1u);
            }
            p2_produce<HAS_FEAT, HAS_LIM, LIST>(&tmap, A, S, s2);
        }
        __syncwarp();
        // ---- consumer: wait for the stage ----------------------------------------------------
        p2_wait((uint32_t)__cvta_generic_to_shared(&S.full[s]), use & 1u);
        const P2Info I = S.info[s];
        if (I.t < 0) break;
        const PWin &W = S.win[s];
        const int c0t = I.tx * kT0, r0t = I.ty * kT0;
        const int bw = W.ax.n, rb = W.rb, bh = W.re - rb;
        const int sc0 = c0t + (wid & 1) * 32, sr0 = r0t + (wid >> 1) * 16;
        const bool sub_valid = sc0 < bw && sr0 < bh;
        if (sub_valid) {
            const int sc1 = min(bw, sc0 + 32) - 1, sr1 = min(bh, sr0 + 16) - 1;
            int ntile = 0, ns = 0;
            bool refilter = false;
            const double2 *list = &S.slot[s][1];
            if (HAS_FEAT) {
                const SlotHead hd = *reinterpret_cast<const SlotHead *>(&S.slot[s][0]);
                ntile = hd.cnt;
                if (ntile > kSlotCap) list = A.pool0 + hd.off;
            }
            if (HAS_FEAT && ntile > 0) {
                // the rect's coordinates (the linspace of the window, bit-exact)
                WS.x[lane] = lin_at(W.ax, min(sc0 + lane, bw - 1));
                if (lane < 16) WS.y[lane] = lin_at(W.ay, rb + min(sr0 + lane, bh - 1));
                __syncwarp();
                Rect q;
                {
                    const double xa = WS.x[0], xb = WS.x[sc1 - sc0];
                    const double ya = WS.y[0], yb = WS.y[sr1 - sr0];
                    q.xlo = xa < xb ? xa : xb; q.xhi = xa < xb ? xb : xa;
                    q.ylo = ya < yb ? ya : yb; q.yhi = ya < yb ? yb : ya;
                }
                unsigned key[5];
#pragma unroll
                for (int j = 0; j < 5; ++j) key[j] = 0xffffffffu;
                for (int b0 = 0; b0 < ntile; b0 += 32) {
                    const int i = b0 + lane;
                    const bool valid = i < ntile;
                    double2 v = make_double2(0.0, 0.0);
                    if (valid) v = list[i];
                    dom_step(q, valid, v, key, WS.dom, lane);
                }
                __syncwarp();
                for (int b0 = 0; b0 < ntile; b0 += 32) {
                    const int i = b0 + lane;
                    const bool valid = i < ntile;
                    double2 v = make_double2(0.0, 0.0);
                    if (valid) v = list[i];
                    const bool keep = valid && !(rect_d2(q, v) > F.r2lim) &&
                                      !dominated_tab(WS.dom, corner_d2(q, v));
                    const unsigned m = __ballot_sync(0xffffffffu, keep);
                    const int n = __popc(m);
                    if (ns + n > kSurv2) { refilter = true; break; }
                    if (keep) WS.sv[ns + __popc(m & ((1u << lane) - 1u))] = v;
                    ns += n;
                }
                __syncwarp();
            }
            // strictly inside the window, fully written, aligned
            const bool full = W.vec && (W.dem_c & 3) == 0 && c0t >= 1 && c0t + kT0 <= min(bw - 1, W.keep_w) &&
                              rb + r0t >= 1 && r0t + kT0 <= bh &&
                              rb + r0t + kT0 <= min(W.ay.n - 1, W.keep_h);
            if (full)
                p2_pixels<HAS_OLD, HAS_FEAT, HAS_LIM, BAND, true>(
                    A, &S.dem[s][0], W, c0t, r0t, lcol, lrow, ntile, ns, refilter, list, far,
                    lane, WS, sc0, sr0, sc1, sr1);
            else
                p2_pixels<HAS_OLD, HAS_FEAT, HAS_LIM, BAND, false>(
                    A, &S.dem[s][0], W, c0t, r0t, lcol, lrow, ntile, ns, refilter, list, far,
                    lane, WS, sc0, sr0, sc1, sr1);
        }
        // ---- hand the stage back -------------------------------------------------------------------
        __syncwarp();
        if (lane == 0) {
            const uint32_t eb = (uint32_t)__cvta_generic_to_shared(&S.empty[s]);
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(eb) : "memory");
        }
    }
}

// ---------------------------------------------------------------------------
// levels 1 + 0 in one launch: one CTA per 256-px tile filters its 1024-px parent
// list down to the tile (the work of l1_kernel) INTO SHARED MEMORY and goes
// straight on to the slots of its (up to 16) 64-px children, a warp per two
// children; a tile without candidates only writes empty headers.  The 256-px
// lists never go to global memory and one launch of the chain is gone.
// ---------------------------------------------------------------------------
struct L0Args {
    const PWin *wins;
    int n_wins, n_tiles, n_parents;
    const int *cnt2, *off2;         // 1024-px lists (coordinates)
    const double2 *xy2;
    double r2lim;
    double2 *slots, *pool;
    unsigned long long *ctr;
    unsigned long long cap;
    int *overflow;
    int *sticky;
    // split pass: the tiles that end with candidates (nullptr: not wanted)
    int *active, *n_active;
};

#ifndef OCSB_L10_THREADS
#define OCSB_L10_THREADS 256
#endif
constexpr int kL10Threads = OCSB_L10_THREADS;
constexpr int kL0Stage = 1024;       // tile candidates held in shared memory

__global__ void __launch_bounds__(kL10Threads) l10_kernel(L0Args A) {
    constexpr int NW = kL10Threads / 32;
    __shared__ unsigned s_key[NW][5];
    __shared__ DomTable s_tab[NW];
    __shared__ DomTable s_fin;
    __shared__ unsigned s_key0;
    __shared__ __align__(16) double2 s_xy[kL0Stage];
    __shared__ int s_cnt;
    __shared__ unsigned long long s_base;
    __shared__ int s_nact, s_act[16];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int t1 = blockIdx.x;
    TL_SCOPE(3);
    if (t1 >= A.n_parents || *A.overflow) return;
    const int wi = A.n_wins == 1 ? 0 : find_win(A.wins, A.n_wins, t1, 1);
    const PWin &W = A.wins[wi];
    const int lt1 = t1 - W.base1;
    const int ty1 = lt1 / W.n1x, tx1 = lt1 - ty1 * W.n1x;
    const int pt = W.base2 + (ty1 >> 2) * W.n2x + (tx1 >> 2);
    const int pcnt2 = A.cnt2[pt], poff2 = A.off2[pt];
    const int ncx = min(4, W.n0x - 4 * tx1), ncy = min(4, W.n0y - 4 * ty1);
    SlotHead hd;
    hd.cnt = 0; hd.pad = 0; hd.off = 0;
    auto empty_headers = [&]() {
        if ((int)threadIdx.x < 16) {
            const int cx = threadIdx.x & 3, cy = threadIdx.x >> 2;
            if (cx < ncx && cy < ncy) {
                const int t = W.base0 + (4 * ty1 + cy) * W.n0x + 4 * tx1 + cx;
                *reinterpret_cast<SlotHead *>(A.slots + (size_t)t * kSlot) = hd;
            }
        }
    };
    if (pcnt2 == 0) { empty_headers(); return; }
    const int bw = W.ax.n, bh = W.re - W.rb;
    // ---- level 1: the 1024-px list -> this 256-px tile ----------------------------
    const double2 *src2 = A.xy2 + poff2;
    {
        const int c0 = tx1 * kT1, r0 = ty1 * kT1;
        const Rect q = win_rect(W, c0, min(bw, c0 + kT1) - 1, r0, min(bh, r0 + kT1) - 1);
        unsigned key[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) key[j] = 0xffffffffu;
        {
            DomAcc acc;
            dom_acc_init(acc);
            for (int i = threadIdx.x; i < pcnt2; i += kL10Threads)
                dom_acc_step(q, src2[i], i, acc);
            dom_acc_finish(q, acc, key, s_tab[wid], lane, [&](int i) { return src2[i]; });
        }
        __syncwarp();
        if (threadIdx.x == 0) { s_cnt = 0; s_nact = 0; }
        dom_combine<NW>(key, s_key, s_tab, &s_fin, &s_key0);
        for (int b0 = wid * 32; b0 < pcnt2; b0 += kL10Threads) {
            const int i = b0 + lane;
            const bool valid = i < pcnt2;
            double2 v = make_double2(0.0, 0.0);
            if (valid) v = src2[i];
            const bool keep = valid && !(rect_d2(q, v) > A.r2lim) &&
                              !dominated_tab(s_fin, corner_d2(q, v));
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            int base = 0;
            if (lane == 0 && m) base = atomicAdd(&s_cnt, __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            const int pos = base + __popc(m & ((1u << lane) - 1u));
            if (keep && pos < kL0Stage) s_xy[pos] = v;
        }
        __syncthreads();
    }
    const int pcnt = s_cnt;
    if (pcnt == 0) { empty_headers(); return; }
    // more survivors than the staging holds (dense vertex sets): the tile's list
    // goes through the overflow pool instead of shared memory
    const bool staged = pcnt <= kL0Stage;
    if (!staged) {
        if (threadIdx.x == 0) {
            unsigned long long base = atomicAdd(A.ctr, (unsigned long long)pcnt);
            if (base + pcnt > A.cap) { flag_overflow(A.overflow, A.sticky); base = ~0ull; }
            s_base = base;
            s_cnt = 0;
        }
        __syncthreads();
        if (s_base == ~0ull) return;
        const int c0 = tx1 * kT1, r0 = ty1 * kT1;
        const Rect q = win_rect(W, c0, min(bw, c0 + kT1) - 1, r0, min(bh, r0 + kT1) - 1);
        for (int b0 = wid * 32; b0 < pcnt2; b0 += kL10Threads) {
            const int i = b0 + lane;
            const bool valid = i < pcnt2;
            double2 v = make_double2(0.0, 0.0);
            if (valid) v = src2[i];
            const bool keep = valid && !(rect_d2(q, v) > A.r2lim) &&
                              !dominated_tab(s_fin, corner_d2(q, v));
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            int base = 0;
            if (lane == 0 && m) base = atomicAdd(&s_cnt, __popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (keep) A.pool[s_base + base + __popc(m & ((1u << lane) - 1u))] = v;
        }
        __syncthreads();
    }
    const double2 *src = staged ? s_xy : A.pool + s_base;
    const int nsrc = pcnt;
    DomTable &tab = s_tab[wid];
    __syncthreads();        // s_tab[] is reused per warp below
    for (int child = wid; child < 16; child += NW) {
        const int cx = child & 3, cy = child >> 2;
        if (cx >= ncx || cy >= ncy) continue;           // warp-uniform
        const int tx = 4 * tx1 + cx, ty = 4 * ty1 + cy;
        const int t = W.base0 + ty * W.n0x + tx;
        double2 *slot = A.slots + (size_t)t * kSlot;
        const int c0 = tx * kT0, r0 = ty * kT0;
        const Rect q = win_rect(W, c0, min(bw, c0 + kT0) - 1, r0, min(bh, r0 + kT0) - 1);
        unsigned key[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) key[j] = 0xffffffffu;
        {
            DomAcc acc;
            dom_acc_init(acc);
            for (int i = lane; i < nsrc; i += 32) dom_acc_step(q, src[i], i, acc);
            dom_acc_finish(q, acc, key, tab, lane, [&](int i) { return src[i]; });
        }
        __syncwarp();
        int total = 0;
        for (int b0 = 0; b0 < nsrc; b0 += 32) {
            const int i = b0 + lane;
            const bool valid = i < nsrc;
            double2 v = make_double2(0.0, 0.0);
            if (valid) v = src[i];
            const bool keep = valid && !(rect_d2(q, v) > A.r2lim) &&
                              !dominated_tab(tab, corner_d2(q, v));
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            const int pos = total + __popc(m & ((1u << lane) - 1u));
            if (keep && pos < kSlotCap) slot[1 + pos] = v;
            total += __popc(m);
        }
        hd.cnt = total; hd.off = 0;
        if (total > kSlotCap) {
            // the whole list goes to the overflow pool (contiguous)
            unsigned long long base = 0;
            if (lane == 0) {
                base = atomicAdd(A.ctr, (unsigned long long)total);
                if (base + total > A.cap) { flag_overflow(A.overflow, A.sticky); base = ~0ull; }
            }
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base == ~0ull) continue;
            int w = 0;
            for (int b0 = 0; b0 < nsrc; b0 += 32) {
                const int i = b0 + lane;
                const bool valid = i < nsrc;
                double2 v = make_double2(0.0, 0.0);
                if (valid) v = src[i];
                const bool keep = valid && !(rect_d2(q, v) > A.r2lim) &&
                                  !dominated_tab(tab, corner_d2(q, v));
                const unsigned m = __ballot_sync(0xffffffffu, keep);
                if (keep) A.pool[base + w + __popc(m & ((1u << lane) - 1u))] = v;
                w += __popc(m);
            }
            hd.off = (long long)base;
        }
        __syncwarp();
        if (lane == 0) {
            *reinterpret_cast<SlotHead *>(slot) = hd;
            if (A.active && total > 0) s_act[atomicAdd(&s_nact, 1)] = t;
        }
    }
    if (A.active) {
        __syncthreads();
        if (threadIdx.x == 0 && s_nact > 0) {
            const int base = atomicAdd(A.n_active, s_nact);
            for (int i = 0; i < s_nact; ++i) A.active[base + i] = s_act[i];
        }
    }
}
