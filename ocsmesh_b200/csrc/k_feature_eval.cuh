// K1 evaluation kernel (included by k_feature.cu inside its anonymous
// namespace): fused leaf filter + exact nearest-vertex distance + size law.
//
// One CTA per 64x64 tile, one warp per 32x16-pixel rect = two 32x8 leaves, one
// lane per 4x2 pixels of a leaf.  Per rect:
//   pass A  lane = candidate: the 5 dominators (nearest to the leaf centre and
//           to its 4 corners) by redux.sync min; their corner distances go to
//           a per-warp shared table (they are warp-uniform, so keeping them
//           in registers only cost occupancy)
//   pass B  lane = candidate: radius + dominance filter, survivors compacted
//           into per-warp shared memory (ballot + popc), then every lane
//           evaluates its 8 pixels against each survivor read as one LDS.128
//           broadcast
//   Both passes run once per 32x16 rect (two stacked leaves): the filter is
//   warp-level work that does not shrink with the rect, the survivor list
//   barely grows, so sharing it between the two leaves halves its cost per
//   pixel.  More than kSurvCap survivors -> the leaves re-filter chunk-wise.
//   epilogue  float32(rate*target*sqrt(d2) + target), clamps, min(old, .)
//
// Exactness of the epilogue.  The reference value is
//   v = float32( fl(fl(rt * RN(sqrt(d2))) + target) )        (all f64)
// RN(sqrt) in f64 costs ~28 instructions/pixel, which dominated the first
// version of this kernel.  The fast path computes d = sqrt(d2) to 2^-39.4
// relative (MUFU.RSQ64H seed, measured 2^-20.0 on B200 by tools/ubench, + one
// f64 Newton step), forms v64 the same way and accepts it only if the 29
// mantissa bits that the f32 rounding drops are not within 2^15 ulp64 of the
// round-to-nearest tie -- then float32(v64) is provably the reference's
// float32; otherwise (1 pixel in 8 k) it falls back to __dsqrt_rn.  The
// domain of the fast path (1e-30 < d2 < min(r^2, 1e30), decided on the high
// word) and the tie test are integer ops on the two halves of the doubles:
// on sm_100 the XU pipe (F2F, MUFU) issues one warp instruction per 8 cycles
// and the ALU pipe one per 2, so conversions and compare+select pairs are
// what the epilogue has to avoid.  Requesting dist_out / idx_out selects the
// exact path.
#pragma once

struct EvalArgs {
    Geom g;
    LevelDev lvl;           // the 64-px level
    const double2 *verts;
    const float *old_;
    int64_t os;
    float *new_;
    int64_t ns;
    double *dist;
    int64_t *idx;
    int64_t nv;
    double rate_target, target, r, r2, r2lim;
    double v_far;           // fl(fl(rate_target * r) + target): beyond the radius
    float lo, hi;           // f32-rounded [hmin, hmax], -/+inf when None
    int fuse_clamp;
    int fast_ok;            // parameters in the range the error bound covers
    unsigned hi_lo, hi_span;  // fast domain on the high word of d2
    float v_far_f;          // float32(v_far)
    float old_const;        // pre-op value everywhere when old_ == nullptr
    int vec_old, vec_new;
    const int *overflow;
    // Constraint._apply_rate epilogue (exact path only, see FeatureArgs)
    const float *rate_vals;
    const uint8_t *rate_mask;
    double rate, rate_sign;
    int rate_is_min;
};

constexpr int kSubH = 16;   // filter rect: kLeafW x kSubH px = two stacked leaves
constexpr int kSurvCap = 64;

struct WarpScratch {
    double dom[5][4];       // dominator corner distances (inflated by 1+eps)
    double2 sv[kSurvCap];   // compacted survivors of the warp's filter rect
    int sid[kSurvCap];
};

__device__ __forceinline__ float min_nan_f32(float a, float b) {
    float r;   // np.minimum: NaN-propagating
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

__device__ __forceinline__ bool dominated_smem(const WarpScratch &W,
                                               const CornerD2 &v) {
    bool dom = false;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const double2 a = *reinterpret_cast<const double2 *>(&W.dom[j][0]);
        const double2 b = *reinterpret_cast<const double2 *>(&W.dom[j][2]);
        dom = dom || (v.d00 > a.x && v.d01 > a.y && v.d10 > b.x && v.d11 > b.y);
    }
    return dom;
}

__device__ __forceinline__ float max_nan_f32(float a, float b) {
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

__device__ __forceinline__ double rsqrt_approx_f64(double x) {
    double r;   // MUFU.RSQ64H: ~20-bit seed from the high word, no conversions
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    return r;
}

// float32 of the size law for squared distance d2 (see header comment)
__device__ __forceinline__ float size_law_f32(double d2, const EvalArgs &A) {
    // fast path, evaluated unconditionally (garbage outside its domain)
    const double y = rsqrt_approx_f64(d2);
    const double g = d2 * y;
    const double e = fma(-g, g, d2);
    const double d = fma(e * y, 0.5, g);
    const double v64 = __dadd_rn(__dmul_rn(A.rate_target, d), A.target);
    // domain: hi_lo <= hi(d2) < hi_lo + hi_span  (1e-30 < d2 < min(r2, 1e30))
    const bool in_dom = (unsigned)(__double2hiint(d2) - (int)A.hi_lo) < A.hi_span;
    // the 29 dropped bits, shifted to the top of a word: tie at 2^31
    const unsigned t = ((unsigned)__double2loint(v64) << 3) - (0x80000000u - (1u << 18));
    const bool near_tie = t < (1u << 19);
    float vf = (float)v64;
    if (!in_dom || near_tie) {
        if (d2 < A.r2) {
            const double dd = __dsqrt_rn(d2);
            vf = (float)__dadd_rn(__dmul_rn(A.rate_target, dd), A.target);
        } else {
            vf = A.v_far_f;
        }
    }
    return vf;
}

template <bool WANT_OUT, bool HAS_OLD>
__global__ void __launch_bounds__(kEvalThreads, WANT_OUT ? 3 : 4) eval_kernel(EvalArgs A) {
    __shared__ double2 s_v[kCandCap];
    __shared__ int s_id[kCandCap];
    __shared__ WarpScratch s_w[kEvalThreads / 32];
    if (*A.overflow) return;
    const int tx = blockIdx.x, ty = blockIdx.y;
    const int t = ty * A.lvl.ntx + tx;
    const int cnt = A.lvl.cnt[t];
    const int off = A.lvl.off[t];
    const bool in_smem = cnt <= kCandCap;
    if (in_smem) {
        for (int i = threadIdx.x; i < cnt; i += kEvalThreads) {
            int id = A.lvl.ids[off + i];
            s_id[i] = id;
            s_v[i] = A.verts[id];
        }
        __syncthreads();
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const Geom &g = A.g;
    WarpScratch &W = s_w[wid];

    // candidate i of the tile list (lane = candidate)
    auto load_cand = [&](int i, double2 &v, int &id) {
        if (in_smem) { v = s_v[i]; id = s_id[i]; }
        else { id = A.lvl.ids[off + i]; v = A.verts[id]; }
    };

    for (int sub = wid; sub < 2 * (kTile0 / kSubH); sub += kEvalThreads / 32) {
        const int c0 = tx * kTile0 + (sub & 1) * kLeafW;
        const int R0 = ty * kTile0 + (sub >> 1) * kSubH;
        if (c0 >= g.w || R0 >= g.h) continue;
        const int c1 = min(g.w, c0 + kLeafW) - 1;
        const int R1 = min(g.h, R0 + kSubH) - 1;
        const int gcol = c0 + (lane & 7) * 4;
        const bool full = gcol + 3 <= c1;
        const Rect q = tile_rect(g, c0, c1, R0, R1);

        // ---- pass A: dominators (nearest to centre and corners) ------------
        int ns = 0;
        bool refilter = false;
        if (cnt > 0) {
            unsigned domkey[5];
#pragma unroll
            for (int j = 0; j < 5; ++j) domkey[j] = 0xffffffffu;
            for (int b0 = 0; b0 < cnt; b0 += 32) {
                const int i = b0 + lane;
                const bool valid = i < cnt;
                double2 v = make_double2(0.0, 0.0);
                int id = 0;
                if (valid) load_cand(i, v, id);
                const CornerD2 c = corner_d2(q, v);
                const double dd[5] = {center_d2(q, v), c.d00, c.d01, c.d10, c.d11};
#pragma unroll
                for (int j = 0; j < 5; ++j) {
                    const unsigned key = valid ? __float_as_uint((float)dd[j]) : 0xffffffffu;
                    const unsigned m = __reduce_min_sync(0xffffffffu, key);
                    if (m < domkey[j]) {   // warp-uniform
                        domkey[j] = m;
                        const int src = __ffs(__ballot_sync(0xffffffffu, key == m)) - 1;
                        if (lane == src) {
                            const double mg = 1.0 + kDomEps;
                            W.dom[j][0] = c.d00 * mg; W.dom[j][1] = c.d01 * mg;
                            W.dom[j][2] = c.d10 * mg; W.dom[j][3] = c.d11 * mg;
                        }
                    }
                }
            }
            __syncwarp();
        }

        // ---- pass B: radius + dominance filter of one 32-candidate chunk;
        // survivors are appended at W.sv[base ...]; returns their number (they
        // are only stored while they fit)
        auto filter_chunk = [&](int b0, int base) -> int {
            const int i = b0 + lane;
            const bool valid = i < cnt;
            double2 v = make_double2(0.0, 0.0);
            int id = 0;
            if (valid) load_cand(i, v, id);
            const bool keep = valid && !(rect_d2(q, v) > A.r2lim) &&
                              !dominated_smem(W, corner_d2(q, v));
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            const int n = __popc(mask);
            if (keep && base + n <= kSurvCap) {
                const int pos = base + __popc(mask & lt_mask);
                W.sv[pos] = v;
                if (WANT_OUT) W.sid[pos] = id;
            }
            return n;
        };
        if (cnt > 0) {
            for (int b0 = 0; b0 < cnt; b0 += 32) {
                const int n = filter_chunk(b0, ns);
                if (ns + n > kSurvCap) { refilter = true; break; }
                ns += n;
            }
            __syncwarp();
        }

        // this lane's pixel columns
        double px[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) px[k] = lin_at(g.ax, min(gcol + k, g.w - 1));

        for (int half = 0; half < kSubH / kLeafH; ++half) {
            const int r0 = R0 + half * kLeafH;
            if (r0 > R1) break;
            const int r1 = min(R1, r0 + kLeafH - 1);
            // this lane's pixels: 4 consecutive columns x 2 rows (row, row + 4)
            const int grow[2] = {r0 + (lane >> 3), r0 + (lane >> 3) + 4};

            // old hfun values: issued first, consumed in the epilogue
            float o[2][4];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
#pragma unroll
                for (int k = 0; k < 4; ++k) o[j][k] = A.old_const;
                if (!HAS_OLD) continue;
                if (grow[j] > r1 || gcol > c1) continue;
                const float *orow = A.old_ + (int64_t)(g.row0 + grow[j]) * A.os;
                if (A.vec_old && full) {
                    const float4 v4 = __ldcs(reinterpret_cast<const float4 *>(orow + gcol));
                    o[j][0] = v4.x; o[j][1] = v4.y; o[j][2] = v4.z; o[j][3] = v4.w;
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (gcol + k <= c1) o[j][k] = orow[gcol + k];
                }
            }

            double best[8];
            int bid[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) { best[k] = INFINITY; bid[k] = 0x7fffffff; }

            if (cnt > 0) {
                double py[2];
#pragma unroll
                for (int j = 0; j < 2; ++j)
                    py[j] = lin_at(g.ay, g.row0 + min(grow[j], g.h - 1));

                // every lane evaluates its 8 pixels against survivors
                // W.sv[0 .. n), each read as one LDS.128 broadcast
                auto eval_survivors = [&](int n) {
                    for (int s = 0; s < n; ++s) {
                        const double2 sv = W.sv[s];
                        int vid = 0;
                        if (WANT_OUT) vid = W.sid[s];
                        double dx2[4], dy2[2];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const double d = __dsub_rn(px[k], sv.x);
                            dx2[k] = __dmul_rn(d, d);
                        }
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const double d = __dsub_rn(py[j], sv.y);
                            dy2[j] = __dmul_rn(d, d);
                        }
#pragma unroll
                        for (int j = 0; j < 2; ++j)
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const double d2 = __dadd_rn(dx2[k], dy2[j]);
                                if (WANT_OUT) {
                                    if (d2 < best[j * 4 + k] ||
                                        (d2 == best[j * 4 + k] && vid < bid[j * 4 + k])) {
                                        best[j * 4 + k] = d2;
                                        bid[j * 4 + k] = vid;
                                    }
                                } else {
                                    // d2 is never NaN (finite coordinates): a
                                    // compare + select, not the 7-op IEEE fmin
                                    best[j * 4 + k] =
                                        (d2 < best[j * 4 + k]) ? d2 : best[j * 4 + k];
                                }
                            }
                    }
                };
                if (!refilter) {
                    eval_survivors(ns);
                } else {
                    // more than kSurvCap survivors: filter and evaluate chunk
                    // by chunk (the leaves share the rect's filter criteria)
                    for (int b0 = 0; b0 < cnt; b0 += 32) {
                        const int n = filter_chunk(b0, 0);
                        __syncwarp();
                        eval_survivors(n);
                        __syncwarp();
                    }
                }
            }

            // ---- epilogue --------------------------------------------------
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int row = grow[j];
                if (row > r1 || gcol > c1) continue;
                float *nrow = A.new_ + (int64_t)(g.row0 + row) * A.ns;
                float out[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const double d2 = best[j * 4 + k];
                    float vf;
                    if (WANT_OUT) {
                        const bool hit = d2 < A.r2;
                        const double d = hit ? __dsqrt_rn(d2) : A.r;
                        if (A.rate_vals) {
                            // temp = values[mask][neighbors] * (1 + d*rate*sign);
                            // non-mask pixels that do not satisfy it take it
                            // (features/constraint.py:106-117); no clamp, no min
                            float res = o[j][k];
                            if (gcol + k <= c1 &&
                                !A.rate_mask[(int64_t)(g.row0 + row) * g.w + gcol + k]) {
                                const double f = __dadd_rn(
                                    1.0, __dmul_rn(__dmul_rn(d, A.rate), A.rate_sign));
                                const double t =
                                    __dmul_rn((double)A.rate_vals[bid[j * 4 + k]], f);
                                const bool sat = A.rate_is_min ? ((double)res > t)
                                                               : ((double)res < t);
                                if (!sat) res = (float)t;
                            }
                            out[k] = res;
                            continue;
                        }
                        // values = expansion_rate*target_size*distances + target_size
                        vf = (float)__dadd_rn(__dmul_rn(A.rate_target, d), A.target);
                        if (gcol + k <= c1) {
                            const int64_t lin = (int64_t)(g.row0 + row) * g.w + gcol + k;
                            if (A.dist) A.dist[lin] = d;
                            if (A.idx) A.idx[lin] = hit ? (int64_t)bid[j * 4 + k] : A.nv;
                        }
                    } else {
                        vf = size_law_f32(d2, A);
                    }
                    // reference: clamp(vf); min(old, .); [A10: clamp again].  The
                    // clamp is monotone and idempotent, so with A10 fused
                    // clamp(min(old, clamp(vf))) == clamp(min(old, vf)); max.NaN /
                    // min.NaN keep NaN like the reference's compares do.
                    float res;
                    if (A.fuse_clamp) {
                        res = min_nan_f32(o[j][k], vf);
                        res = min_nan_f32(max_nan_f32(res, A.lo), A.hi);
                    } else {
                        vf = min_nan_f32(max_nan_f32(vf, A.lo), A.hi);
                        res = min_nan_f32(o[j][k], vf);
                    }
                    out[k] = res;
                }
                if (A.vec_new && full) {
                    __stcs(reinterpret_cast<float4 *>(nrow + gcol),
                           make_float4(out[0], out[1], out[2], out[3]));
                } else {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (gcol + k <= c1) nrow[gcol + k] = out[k];
                }
            }
        }
        __syncwarp();   // W is reused by the warp's next rect
    }
}
