// K-pass, persistent form -- included INSIDE k_pass.cu's anonymous namespace.
//
// Same arithmetic as tile_kernel (one feature op and / or one limiter op, fused
// clamp), organised as a persistent producer / consumer pipeline:
//
//   * grid = resident CTAs only; tiles are handed out by one atomic counter
//     (heavy tiles -- those with candidates -- cost ~2x a light one);
//   * thread 0 is the producer: two tiles ahead it decodes the tile and looks
//     up its parent list, one tile ahead it issues ONE mbarrier-tracked
//     transaction per stage = the DEM box (cp.async.bulk.tensor.2d) plus the
//     tile's parent candidate list (cp.async.bulk of cnt x double2).  So neither
//     the DEM latency (light tiles used to expose all of it: nothing else to
//     do) nor the dependent list loads (heavy tiles) are waited for;
//   * one CTA barrier per light tile; interior tiles run a variant without any
//     per-pixel bound or border test.
//
// Reference: the window loops of ocsmesh/hfun/raster.py:1212-1260, 1393-1436.

constexpr int kDemStage = 5024;      // floats per stage: 66 x 76 = 5016, padded to 128 B
constexpr int kCand = 256;           // prefetched parent list (vertices)
constexpr int kP2Threads = 256;

struct P2Info {
    int t, wi, tx, ty, cnt, off, pad0, pad1;
};

struct P2Args {
    const PWin *wins;
    int n_wins, n_tiles;
    const float *in;
    float *out;
    long long in_stride, out_stride;
    FeatDev feat;
    const double2 *xy1;     // level-1 lists as coordinates (same offsets as ids1)
    LimiterDev<float> lim;
    float lo, hi;
    int *counter;           // zeroed before the launch
    const int *overflow;
};

struct __align__(128) P2Smem {
    float dem[2][kDemStage];
    double2 cand[2][kCand];
    double2 v[kListCap];
    WarpScr w[kP2Threads / 32];
    double x[2][kT0], y[2][kT0];
    PWin win[2];
    P2Info info[2];
    DomTable fin;
    unsigned key[kP2Threads / 32][5];
    int n;
    int pend;
    unsigned long long bar[2];
};

__device__ __forceinline__ void p2_decode(const PWin *wins, int n_wins, int t, int &wi,
                                          int &tx, int &ty) {
    wi = n_wins == 1 ? 0 : find_win(wins, n_wins, t, 0);
    const int lt = t - wins[wi].base0, n0x = wins[wi].n0x;
    ty = lt / n0x;
    tx = lt - ty * n0x;
}

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

// the loads of one stage (producer thread): DEM box + parent candidate list
template <bool HAS_FEAT, bool HAS_LIM>
__device__ __forceinline__ void p2_issue(const CUtensorMap *tmap, const P2Args &A,
                                         P2Smem &S, int stage) {
    const P2Info I = S.info[stage];
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&S.bar[stage]);
    const bool list = HAS_FEAT && I.cnt > 0 && I.cnt <= kCand;
    uint32_t bytes = 0;
    if (HAS_LIM) bytes += (uint32_t)(kDemRows * kDemPitch * 4);
    if (list) bytes += (uint32_t)I.cnt * 16u;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 ::"r"(bar), "r"(bytes) : "memory");
    if (HAS_LIM) {
        const PWin &G = A.wins[I.wi];
        const int cx = G.dem_c + I.tx * kT0 - 4, cy = G.dem_r + G.rb + I.ty * kT0 - 1;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&S.dem[stage][0]);
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
            " [%0], [%1, {%2, %3}], [%4];"
            ::"r"(dst), "l"(tmap), "r"(cx), "r"(cy), "r"(bar) : "memory");
    }
    if (list) {
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&S.cand[stage][0]);
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes"
            " [%0], [%1], %2, [%3];"
            ::"r"(dst), "l"(A.xy1 + I.off), "r"((uint32_t)I.cnt * 16u), "r"(bar) : "memory");
    }
}

// parent list of a tile: {cnt, off} (0, 0 without a feature op)
template <bool HAS_FEAT>
__device__ __forceinline__ int2 p2_parent(const P2Args &A, int wi, int tx, int ty) {
    if (!HAS_FEAT) return make_int2(0, 0);
    const PWin &G = A.wins[wi];
    const int t1 = G.base1 + (ty >> 2) * G.n1x + (tx >> 2);
    return make_int2(A.feat.cnt1[t1], A.feat.off1[t1]);
}

// ---- the pixels of one lane: 4 columns x 4 rows, two rows at a time -----------
// FULL: the tile lies strictly inside its window (no border pixel, everything
// written, 16-byte accesses aligned): no per-pixel test survives.
template <bool HAS_OLD, bool HAS_FEAT, bool HAS_LIM, bool BAND, bool FULL>
__device__ __forceinline__ void p2_pixels(const P2Args &A, P2Smem &S, const PWin &W,
                                          int b, int c0t, int r0t, int lcol, int lrow,
                                          int ntile, int ns, bool refilter, bool in_smem,
                                          const double2 *glist, bool far, int lane,
                                          WarpScr &WS, int sc0, int sr0, int sc1, int sr1) {
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
    const double *sx = S.x[b], *sy = S.y[b];

#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
        const int hr = grow + half * 2;
        if (!FULL && (hr > sr1 || gcol > sc1)) break;
        float vf[8];
        if (HAS_FEAT) {
            if (ntile == 0) {
#pragma unroll
                for (int k = 0; k < 8; ++k) vf[k] = F.v_far_f;
            } else {
                double px[4];
                {
                    const double2 a = *reinterpret_cast<const double2 *>(&sx[lcol]);
                    const double2 c = *reinterpret_cast<const double2 *>(&sx[lcol + 2]);
                    px[0] = a.x; px[1] = a.y; px[2] = c.x; px[3] = c.y;
                }
                const double2 py = *reinterpret_cast<const double2 *>(&sy[lrow + half * 2]);
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
                    // more than kSurv survivors: filter and evaluate chunk-wise
#pragma unroll
                    for (int k = 0; k < 8; ++k) best[k] = kBig;
                    Rect q;
                    {
                        const double xa = sx[sc0 - c0t], xb = sx[sc1 - c0t];
                        const double ya = sy[sr0 - r0t], yb = sy[sr1 - r0t];
                        q.xlo = xa < xb ? xa : xb; q.xhi = xa < xb ? xb : xa;
                        q.ylo = ya < yb ? ya : yb; q.yhi = ya < yb ? yb : ya;
                    }
                    for (int b0 = 0; b0 < ntile; b0 += 32) {
                        const int i = b0 + lane;
                        const bool valid = i < ntile;
                        double2 v = make_double2(0.0, 0.0);
                        if (valid) v = in_smem ? S.v[i] : glist[i];
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
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = 3.402823466e+38f;
        if (HAS_OLD) {
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
        }
        if (HAS_FEAT) {
            // clamp(vf); min(old, .); A10 clamp: fused (monotone, idempotent)
#pragma unroll
            for (int k = 0; k < 8; ++k)
                o[k] = clamp_nan(min3_nan(o[k], vf[k], F.v_far_f), A.lo, A.hi);
        }
        // ---- limiter op ------------------------------------------------------------------
        if (HAS_LIM) {
            const int H = W.ay.n;
            const float *sd = &S.dem[b][0] + (lrow + half * 2) * kDemPitch + 4 + lcol;
            float4 rowv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                rowv[j] = *reinterpret_cast<const float4 *>(sd + j * kDemPitch);
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

template <bool HAS_OLD, bool HAS_FEAT, bool HAS_LIM, bool BAND>
__global__ void __launch_bounds__(kP2Threads, OCSB_PASS2_MIN_CTAS)
pass2_kernel(const __grid_constant__ CUtensorMap tmap, const P2Args A) {
    extern __shared__ __align__(128) unsigned char p2_raw[];
    P2Smem &S = *reinterpret_cast<P2Smem *>(p2_raw);

    const int2 flags = *reinterpret_cast<const int2 *>(A.overflow);   // {overflow, far}
    if (flags.x) return;
    const bool far = flags.y != 0;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int lcol = (wid & 1) * 32 + (lane & 7) * 4;      // tile-local
    const int lrow = (wid >> 1) * 16 + (lane >> 3) * 4;
    WarpScr &WS = S.w[wid];
    const FeatDev &F = A.feat;
    const int grid = gridDim.x;

    // ---- producer prologue: tiles of iterations 0 and 1, loads of iteration 0 ---------
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&S.bar[s]);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            P2Info I;
            I.t = (int)blockIdx.x + s * grid;
            I.wi = I.tx = I.ty = I.cnt = I.off = I.pad0 = I.pad1 = 0;
            if (I.t < A.n_tiles) {
                p2_decode(A.wins, A.n_wins, I.t, I.wi, I.tx, I.ty);
                const int2 pl = p2_parent<HAS_FEAT>(A, I.wi, I.tx, I.ty);
                I.cnt = pl.x; I.off = pl.y;
            } else {
                I.t = -1;
            }
            S.info[s] = I;
        }
        S.pend = atomicAdd(A.counter, 1) + 2 * grid;
        p2_issue<HAS_FEAT, HAS_LIM>(&tmap, A, S, 0);
    }
    __syncthreads();

    for (int it = 0;; ++it) {
        const int b = it & 1;
        const P2Info I = S.info[b];
        if (I.t < 0) break;
        const int c0t = I.tx * kT0, r0t = I.ty * kT0;
        // ---- window record + coordinates of the tile's columns / rows -----------------
        {
            const PWin &G = A.wins[I.wi];
            if (threadIdx.x < kT0) {
                S.x[b][threadIdx.x] = lin_at(G.ax, min(c0t + (int)threadIdx.x, G.ax.n - 1));
            } else if (threadIdx.x < 2 * kT0) {
                const int j = threadIdx.x - kT0;
                S.y[b][j] = lin_at(G.ay, G.rb + min(r0t + j, G.re - G.rb - 1));
            } else if (threadIdx.x < 2 * kT0 + (int)(sizeof(PWin) / 4)) {
                const int i = threadIdx.x - 2 * kT0;
                reinterpret_cast<int *>(&S.win[b])[i] = reinterpret_cast<const int *>(&G)[i];
            }
        }
        __syncthreads();                                   // (1)
        // ---- producer: loads of iteration it+1, tile of iteration it+2 ---------------
        if (threadIdx.x == 0) {
            if (S.info[b ^ 1].t >= 0) p2_issue<HAS_FEAT, HAS_LIM>(&tmap, A, S, b ^ 1);
            P2Info N;
            N.t = S.pend;
            N.wi = N.tx = N.ty = N.cnt = N.off = N.pad0 = N.pad1 = 0;
            if (N.t < A.n_tiles) {
                p2_decode(A.wins, A.n_wins, N.t, N.wi, N.tx, N.ty);
                const int2 pl = p2_parent<HAS_FEAT>(A, N.wi, N.tx, N.ty);
                S.pend = atomicAdd(A.counter, 1) + 2 * grid;
                N.cnt = pl.x; N.off = pl.y;
            } else {
                N.t = -1;
            }
            // slot b: every thread has read I (they all passed barrier 1)
            S.info[b] = N;
        }
        const PWin &W = S.win[b];
        const int bw = W.ax.n, rb = W.rb, bh = W.re - rb;
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&S.bar[b]);
        const uint32_t parity = (uint32_t)(it >> 1) & 1u;
        bool waited = false;

        // ---- feature op: parent list -> tile list ------------------------------------------
        int ntile = 0;
        bool in_smem = true;
        const double2 *glist = nullptr;
        if (HAS_FEAT && I.cnt > 0) {
            const int pcnt = I.cnt;
            const bool staged = pcnt <= kCand;
            const double2 *gsrc = A.xy1 + I.off;
            if (staged) { p2_wait(bar, parity); waited = true; }
            auto pcand = [&](int i) -> double2 { return staged ? S.cand[b][i] : gsrc[i]; };
            Rect qT;
            {
                const int tc1 = min(bw, c0t + kT0) - 1 - c0t;
                const int tr1 = min(bh, r0t + kT0) - 1 - r0t;
                const double xa = S.x[b][0], xb = S.x[b][tc1];
                const double ya = S.y[b][0], yb = S.y[b][tr1];
                qT.xlo = xa < xb ? xa : xb; qT.xhi = xa < xb ? xb : xa;
                qT.ylo = ya < yb ? ya : yb; qT.yhi = ya < yb ? yb : ya;
            }
            unsigned key[5];
#pragma unroll
            for (int j = 0; j < 5; ++j) key[j] = 0xffffffffu;
            for (int b0 = wid * 32; b0 < pcnt; b0 += kP2Threads) {
                const int i = b0 + lane;
                const bool valid = i < pcnt;
                double2 v = make_double2(0.0, 0.0);
                if (valid) v = pcand(i);
                dom_step(qT, valid, v, key, WS.dom, lane);
            }
            __syncwarp();
            if (threadIdx.x == 0) S.n = 0;
            if (lane < 5) {
                unsigned k = 0;
#pragma unroll
                for (int j = 0; j < 5; ++j) if (lane == j) k = key[j];
                S.key[wid][lane] = k;
            }
            __syncthreads();                               // (2)
            if (threadIdx.x < 5) {
                unsigned best = 0xffffffffu;
                int bwp = 0;
                for (int w = 0; w < kP2Threads / 32; ++w)
                    if (S.key[w][threadIdx.x] < best) { best = S.key[w][threadIdx.x]; bwp = w; }
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    S.fin.c[threadIdx.x][c] = S.w[bwp].dom.c[threadIdx.x][c];
            }
            __syncthreads();                               // (3)
            for (int b0 = wid * 32; b0 < pcnt; b0 += kP2Threads) {
                const int i = b0 + lane;
                const bool valid = i < pcnt;
                double2 v = make_double2(0.0, 0.0);
                if (valid) v = pcand(i);
                const bool keep = valid && !(rect_d2(qT, v) > F.r2lim) &&
                                  !dominated_tab(S.fin, corner_d2(qT, v));
                const unsigned m = __ballot_sync(0xffffffffu, keep);
                int base = 0;
                if (lane == 0 && m) base = atomicAdd(&S.n, __popc(m));
                base = __shfl_sync(0xffffffffu, base, 0);
                const int pos = base + __popc(m & ((1u << lane) - 1u));
                if (keep && pos < kListCap) S.v[pos] = v;
            }
            __syncthreads();                               // (4)
            ntile = S.n;
            if (ntile > kListCap) {
                // the parent list itself (its coordinates are in global memory)
                in_smem = false; ntile = pcnt; glist = gsrc;
            }
        }

        // ---- this warp's 32x16 rect -------------------------------------------------------
        const int sc0 = c0t + (wid & 1) * 32, sr0 = r0t + (wid >> 1) * 16;
        const bool sub_valid = sc0 < bw && sr0 < bh;
        const int sc1 = min(bw, sc0 + 32) - 1, sr1 = min(bh, sr0 + 16) - 1;
        if (sub_valid) {
            int ns = 0;
            bool refilter = false;
            if (HAS_FEAT && ntile > 0) {
                Rect q;
                {
                    const double xa = S.x[b][sc0 - c0t], xb = S.x[b][sc1 - c0t];
                    const double ya = S.y[b][sr0 - r0t], yb = S.y[b][sr1 - r0t];
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
                    if (valid) v = in_smem ? S.v[i] : glist[i];
                    dom_step(q, valid, v, key, WS.dom, lane);
                }
                __syncwarp();
                for (int b0 = 0; b0 < ntile; b0 += 32) {
                    const int i = b0 + lane;
                    const bool valid = i < ntile;
                    double2 v = make_double2(0.0, 0.0);
                    if (valid) v = in_smem ? S.v[i] : glist[i];
                    const bool keep = valid && !(rect_d2(q, v) > F.r2lim) &&
                                      !dominated_tab(WS.dom, corner_d2(q, v));
                    const unsigned m = __ballot_sync(0xffffffffu, keep);
                    const int n = __popc(m);
                    if (ns + n > kSurv) { refilter = true; break; }
                    if (keep) WS.sv[ns + __popc(m & ((1u << lane) - 1u))] = v;
                    ns += n;
                }
                __syncwarp();
            }
            if (HAS_LIM && !waited) { p2_wait(bar, parity); waited = true; }
            // strictly inside the window, fully written, aligned
            const bool full = W.vec && c0t >= 1 && c0t + kT0 <= min(bw - 1, W.keep_w) &&
                              rb + r0t >= 1 && r0t + kT0 <= bh &&
                              rb + r0t + kT0 <= min(W.ay.n - 1, W.keep_h);
            if (full)
                p2_pixels<HAS_OLD, HAS_FEAT, HAS_LIM, BAND, true>(
                    A, S, W, b, c0t, r0t, lcol, lrow, ntile, ns, refilter, in_smem, glist,
                    far, lane, WS, sc0, sr0, sc1, sr1);
            else
                p2_pixels<HAS_OLD, HAS_FEAT, HAS_LIM, BAND, false>(
                    A, S, W, b, c0t, r0t, lcol, lrow, ntile, ns, refilter, in_smem, glist,
                    far, lane, WS, sc0, sr0, sc1, sr1);
        }
        // every thread observes the stage's phase (a thread that skipped the wait
        // could otherwise run a whole phase ahead of the barrier)
        if (!waited) p2_wait(bar, parity);
    }
}
