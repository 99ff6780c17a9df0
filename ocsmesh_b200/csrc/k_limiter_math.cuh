// Per-pixel arithmetic of the flow limiter (reference
// ocsmesh/hfun/raster.py:1396-1435), shared by k_limiter.cu and k_pass.cu;
// included INSIDE each unit's anonymous namespace.  See k_limiter.cu for the
// derivation of every sequence.
#pragma once

template <typename T> struct Arith;
template <> struct Arith<float> {
    __device__ static float sub(float a, float b) { return __fsub_rn(a, b); }
    __device__ static float add(float a, float b) { return __fadd_rn(a, b); }
    __device__ static float mul(float a, float b) { return __fmul_rn(a, b); }
    __device__ static float div(float a, float b) { return __fdiv_rn(a, b); }
    __device__ static float sqrt(float a) { return __fsqrt_rn(a); }
    __device__ static float fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
    __device__ static bool normal_range(float q) {
        const float m = fabsf(q);
        return m >= 1e-30f && m <= 1e30f;
    }
};
template <> struct Arith<double> {
    __device__ static double sub(double a, double b) { return __dsub_rn(a, b); }
    __device__ static double add(double a, double b) { return __dadd_rn(a, b); }
    __device__ static double mul(double a, double b) { return __dmul_rn(a, b); }
    __device__ static double div(double a, double b) { return __ddiv_rn(a, b); }
    __device__ static double sqrt(double a) { return __dsqrt_rn(a); }
    __device__ static double fma(double a, double b, double c) { return __fma_rn(a, b, c); }
    __device__ static bool normal_range(double q) {
        const double m = fabs(q);
        return m >= 1e-280 && m <= 1e280;
    }
};

// division by a loop-invariant constant, bit-identical to IEEE a / c
template <typename T> struct ConstDiv {
    T c, r;   // r = RN(1 / c), computed on the host in T arithmetic
};

template <typename T>
__device__ __forceinline__ T cdiv(T a, const ConstDiv<T> &d) {
    using A = Arith<T>;
    const T q0 = A::mul(a, d.r);
    if (A::normal_range(q0)) {
        const T rem = A::fma(-q0, d.c, a);     // exact
        return A::fma(rem, d.r, q0);           // Markstein: == RN(a / c)
    }
    return A::div(a, d.c);  // zero, subnormal, huge, inf, nan
}

__device__ __forceinline__ float min_nan_f32(float a, float b) {
    float r;   // np.minimum: NaN-propagating
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

template <typename T> struct LimiterDev {
    // float32-rounded bounds, +/-inf when None
    float hmin, hmax, ghmin, ghmax;
    T lb, ub;                      // elevation band in the DEM dtype, +/-inf
    ConstDiv<T> d1_0, d2_0;        // axis 0 (rows): dx, 2 dx
    ConstDiv<T> d1_1, d2_1;        // axis 1 (cols): dy, 2 dy
    int fuse_clamp;
};


__device__ __forceinline__ float max_nan_f32(float a, float b) {
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float mufu_rsq(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float mufu_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// v < lo -> lo; v > hi -> hi with float32 compares; NaN stays NaN
__device__ __forceinline__ float clamp_nan(float v, float lo, float hi) {
    return min_nan_f32(max_nan_f32(v, lo), hi);
}

// Tail of the per-pixel law, shared by both paths: local clamp, elevation
// band restore, global clamp, min with old, optional A10 re-clamp.  With FUSE
// clampG(min(old, clampG(x))) == clampG(min(old, x)) (clamp is monotone and
// idempotent, NaN propagates through min.NaN / max.NaN like the compares).
template <bool HAS_BAND, bool FUSE, typename T>
__device__ __forceinline__ float limiter_tail(float h, T z, float old,
                                              const LimiterDev<T> &p) {
    h = clamp_nan(h, p.hmin, p.hmax);
    bool ob = false;
    if (HAS_BAND) ob = (z > p.ub) || (z < p.lb);
    if (FUSE) {
        float v = min_nan_f32(old, h);
        if (HAS_BAND) v = ob ? old : v;
        return clamp_nan(v, p.ghmin, p.ghmax);
    }
    if (HAS_BAND) h = ob ? old : h;
    h = clamp_nan(h, p.ghmin, p.ghmax);
    return min_nan_f32(old, h);
}

template <bool HAS_BAND, bool FUSE, typename T>
__device__ __forceinline__ float limiter_px(T z, T g0, T g1, float old,
                                            const LimiterDev<T> &p) {
    using A = Arith<T>;
    const T dh = A::sqrt(A::add(A::mul(g0, g0), A::mul(g1, g1)));
    // h = |(1/3) * (double)(z / dh)| rounded to f32; masked where dh == 0: the
    // data under the mask is 1/3
    const T q = A::div(z, dh);
    float h = (float)fabs(__dmul_rn(1.0 / 3.0, (double)q));
    h = (dh == T(0)) ? (float)(1.0 / 3.0) : h;
    return limiter_tail<HAS_BAND, FUSE, T>(h, z, old, p);
}

template <typename T> ConstDiv<T> make_cdiv(double c) {
    ConstDiv<T> d;
    d.c = (T)c;                 // numpy divides by the scalar rounded to T
    volatile T one = (T)1, cc = d.c;
    d.r = one / cc;             // correctly rounded reciprocal in T
    return d;
}

