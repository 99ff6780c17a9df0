// Internal shared declarations of libocsb200 (not part of the C ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <vector>
#include <string>
#include <functional>

#include "ocsb200.h"

enum {
    KC_FILL = OCSB_KC_FILL, KC_CLAMP = OCSB_KC_CLAMP,
    KC_MIN_COMBINE = OCSB_KC_MIN_COMBINE, KC_COORDS = OCSB_KC_COORDS,
    KC_LIMITER = OCSB_KC_LIMITER, KC_FEATURE_REFINE = OCSB_KC_FEATURE_REFINE,
    KC_FEATURE_EVAL = OCSB_KC_FEATURE_EVAL, KC_INTERP = OCSB_KC_INTERP,
    KC_SPLINE_PREFILTER = OCSB_KC_SPLINE_PREFILTER,
    KC_SPLINE_EVAL = OCSB_KC_SPLINE_EVAL, KC_NEAREST = OCSB_KC_NEAREST,
    KC_BUCKET = OCSB_KC_BUCKET, KC_CONSTRAINT = OCSB_KC_CONSTRAINT,
    KC_CONTOUR = OCSB_KC_CONTOUR, KC_SAMPLE = OCSB_KC_SAMPLE,
    KC_PASS_TILE = OCSB_KC_PASS_TILE, KC_PASS_LIMITER = OCSB_KC_PASS_LIMITER,
    KC_PASS_FEATURE = OCSB_KC_PASS_FEATURE
};

// ---------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------
void ocsb_set_error(const char *fmt, ...);

#define OCSB_CUDA(call)                                                        \
    do {                                                                       \
        cudaError_t e__ = (call);                                              \
        if (e__ != cudaSuccess) {                                              \
            ocsb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call,       \
                           cudaGetErrorString(e__));                           \
            return e__ == cudaErrorMemoryAllocation ? OCSB_ENOMEM : OCSB_ECUDA;\
        }                                                                      \
    } while (0)

#define OCSB_REQUIRE(cond, ...)                                                \
    do {                                                                       \
        if (!(cond)) {                                                         \
            ocsb_set_error(__VA_ARGS__);                                       \
            return OCSB_EINVAL;                                                \
        }                                                                      \
    } while (0)

#define OCSB_TRY(expr)                                                         \
    do {                                                                       \
        int32_t s__ = (expr);                                                  \
        if (s__ != OCSB_OK) return s__;                                        \
    } while (0)

// ---------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------
// band buffer sets of ocsb_hfun_pass: H2D(k+1..k+2) | ops(k) | D2H(k-1)
#define OCSB_PASS_SETS 3

struct KTimeRec {
    int kclass;
    cudaEvent_t a, b;
};
#ifndef OCSB_WIN_STREAMS
#define OCSB_WIN_STREAMS 8
#endif
struct ocsb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int num_sms = 148;
    int64_t launches = 0;
    // per-kernel-class CUDA-event timing (off unless ocsb_ctx_set_timing)
    bool timing = false;
    bool time_open = false;
    std::vector<KTimeRec> time_recs;
    std::vector<cudaEvent_t> event_pool;
    // pinned bounce buffer for small D2H reads (flags, counters)
    void *pinned = nullptr;
    size_t pinned_bytes = 0;
    // pinned staging ring for small host -> device uploads that must not wait
    // for the stream (window tables of the asynchronous pass)
    struct StageSlot {
        void *host = nullptr;
        size_t bytes = 0;
        cudaEvent_t ev = nullptr;
        bool used = false;
    } stage_ring[32];
    int stage_next = 0;
    // copy streams + events of ocsb_hfun_pass (created on first use)
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
    cudaEvent_t ev_in[OCSB_PASS_SETS] = {}, ev_comp[OCSB_PASS_SETS] = {},
                ev_out[OCSB_PASS_SETS] = {}, ev_alloc = nullptr;
    // worker streams of ocsb_hfun_op_windows (created on first use)
    cudaStream_t s_win[OCSB_WIN_STREAMS] = {};
    cudaEvent_t ev_win[OCSB_WIN_STREAMS] = {};
    // scratch arena kept between passes (k_pass_tiled): a band step of a multi-GPU
    // run is bound by HOST time, and ~40 stream-ordered malloc / free calls per
    // pass were most of it.  Only ever used by work ordered on `stream`
    void *keep_arena = nullptr;
    size_t keep_arena_bytes = 0;
    bool keep_ok = true;        // false while `stream` is swapped to a worker stream
    // the window table last uploaded to the head of the kept arena (a step that
    // repeats the previous table skips the upload)
    std::vector<unsigned char> wins_cache;
    // second stream of the split pass (limiter sweep beside the candidate pyramid)
    cudaStream_t s_split = nullptr;
    cudaEvent_t ev_split[2] = {nullptr, nullptr};
    // one-time cudaFuncSetAttribute calls already made on this device
    unsigned func_attr_done = 0;
    // results kept on the device until the caller takes them (ocsb_take_result):
    // slot 0 = points / polyline vertices, slot 1 = polyline offsets
    void *kept[2] = {nullptr, nullptr};
    size_t kept_bytes[2] = {0, 0};
    // spline coefficient cache (one raster window)
    const void *spl_key = nullptr;
    double *spl_coef = nullptr;
    size_t spl_bytes = 0;
};

// host -> device upload through the pinned ring: returns without waiting for
// earlier work of the stream (a pageable cudaMemcpyAsync would)
int32_t ocsb_stage_upload(ocsb_ctx *ctx, void *dst_dev, const void *src, size_t bytes);

// stream-ordered scratch allocation (cudaMallocAsync pool, reused across calls)
int32_t ocsb_alloc(ocsb_ctx *ctx, void **p, size_t bytes);
void ocsb_free(ocsb_ctx *ctx, void *p);

struct ScratchGuard {
    ocsb_ctx *ctx;
    std::vector<void *> ptrs;
    char *arena = nullptr;      // reserve(): one block that get() carves up
    size_t arena_left = 0;
    explicit ScratchGuard(ocsb_ctx *c) : ctx(c) {}
    ~ScratchGuard() {
        for (void *p : ptrs) ocsb_free(ctx, p);
    }
    // One allocation for the next get() calls (256-byte aligned pieces): the
    // per-window launches of a many-window op are bound by host call count.
    // Requests that do not fit fall back to their own allocation.
    int32_t reserve(size_t bytes) {
        void *p = nullptr;
        int32_t s = ocsb_alloc(ctx, &p, bytes);
        if (s != OCSB_OK) return s;
        ptrs.push_back(p);
        arena = reinterpret_cast<char *>(p);
        arena_left = bytes;
        return OCSB_OK;
    }
    // The same from the context's kept arena (grown when too small, never freed
    // by the guard).  Larger requests than kKeepMax take a plain reserve().
    static constexpr size_t kKeepMax = (size_t)384 << 20;
    int32_t reserve_kept(size_t bytes) {
        if (bytes > kKeepMax || !ctx->keep_ok) return reserve(bytes);
        if (bytes > ctx->keep_arena_bytes) {
            if (ctx->keep_arena) ocsb_free(ctx, ctx->keep_arena);
            ctx->keep_arena = nullptr;
            ctx->keep_arena_bytes = 0;
            ctx->wins_cache.clear();
            void *p = nullptr;
            const size_t want = bytes + bytes / 4;
            int32_t s = ocsb_alloc(ctx, &p, want);
            if (s != OCSB_OK) return s;
            ctx->keep_arena = p;
            ctx->keep_arena_bytes = want;
        }
        arena = reinterpret_cast<char *>(ctx->keep_arena);
        arena_left = ctx->keep_arena_bytes;
        return OCSB_OK;
    }
    // carve from where another guard stands, without owning anything
    void share(const ScratchGuard &o) { arena = o.arena; arena_left = o.arena_left; }
    static size_t piece(size_t bytes) { return (bytes + 255) & ~(size_t)255; }
    template <typename T> int32_t get(T **out, size_t count) {
        const size_t need = piece(count * sizeof(T) ? count * sizeof(T) : 16);
        if (need <= arena_left) {
            *out = reinterpret_cast<T *>(arena);
            arena += need;
            arena_left -= need;
            return OCSB_OK;
        }
        void *p = nullptr;
        int32_t s = ocsb_alloc(ctx, &p, count * sizeof(T));
        if (s != OCSB_OK) return s;
        ptrs.push_back(p);
        *out = reinterpret_cast<T *>(p);
        return OCSB_OK;
    }
};

static inline bool ocsb_is_none(double v) { return v != v; }

// ---------------------------------------------------------------------------
// np.linspace(start, stop, n)[k], bit-exact (numpy/_core/function_base.py):
//   step = (stop - start) / (n - 1);  y[k] = fl(fl(k * step) + start);
//   y[n-1] = stop (n > 1);  n == 1 -> [start]
// ---------------------------------------------------------------------------
struct LinAxis {
    double start, stop, step;
    int n;
};

static inline LinAxis make_axis(double start, double stop, int n) {
    LinAxis a;
    a.start = start;
    a.stop = stop;
    a.n = n;
    a.step = n > 1 ? (stop - start) / (double)(n - 1) : 0.0;
    return a;
}

#ifdef __CUDACC__
__device__ __forceinline__ double lin_at(const LinAxis &a, int k) {
    // separate multiply and add: numpy does y = arange*step; y += start
    // (double)k for 0 <= k < 2^32 without the XU-pipe I2F (one issue slot per
    // 8 cycles on sm_100): 2^52 + k is exact, so is the subtraction
    const double kd = __dsub_rn(__hiloint2double(0x43300000, k), 4503599627370496.0);
    double v = __dadd_rn(__dmul_rn(kd, a.step), a.start);
    return (k == a.n - 1 && a.n > 1) ? a.stop : v;
}
#endif

#ifdef __CUDACC__
// np.minimum: NaN-propagating
__device__ __forceinline__ float np_minimum_f(float a, float b) {
    return (a != a || b != b) ? __int_as_float(0x7fc00000) : fminf(a, b);
}
__device__ __forceinline__ double np_minimum_d(double a, double b) {
    return (a != a || b != b) ? __longlong_as_double(0x7ff8000000000000LL)
                              : fmin(a, b);
}
// values[values < lo] = lo; values[values > hi] = hi  (float32 compares)
__device__ __forceinline__ float clamp_f32(float v, int has_lo, float lo,
                                           int has_hi, float hi) {
    if (has_lo && v < lo) v = lo;
    if (has_hi && v > hi) v = hi;
    return v;
}
#endif

static inline double lin_at_host(const LinAxis &a, int k) {
    volatile double m = (double)k * a.step;
    double v = m + a.start;
    return (k == a.n - 1 && a.n > 1) ? a.stop : v;
}

// ---------------------------------------------------------------------------
// host/device buffer staging (the C ABI accepts either kind of pointer)
// ---------------------------------------------------------------------------
bool ocsb_is_device_ptr(const void *p);

enum StageDir { STAGE_IN = 1, STAGE_OUT = 2, STAGE_INOUT = 3 };

// One 2-D (rows x width, row_stride) buffer of `elem` bytes per element.
// If `user` is a device pointer, dev == user.  Otherwise a compact device
// copy (stride = width) is made (STAGE_IN) and written back (STAGE_OUT).
struct Staged {
    ocsb_ctx *ctx = nullptr;
    void *user = nullptr;
    void *dev = nullptr;     // (virtual) device address of row 0
    void *alloc = nullptr;   // the staging allocation (holds rows row0..)
    int64_t row0 = 0;
    int64_t rows = 0, width = 0, user_stride = 0, dev_stride = 0;
    size_t elem = 0;
    int dir = 0;
    bool staged = false;
    // rows [row0_, row0_ + rows_) of the user buffer are (to be) valid
    int32_t init(ocsb_ctx *c, const void *p, int64_t rows_, int64_t width_,
                 int64_t stride_, size_t elem_, int dir_, int64_t row0_ = 0);
    int32_t finish();  // enqueue copy-back (if staged and OUT)
    ~Staged();
};

// kernel launch accounting: OCSB_TIME(ctx, class) right before a launch,
// OCSB_LAUNCHED(ctx) right after it
void ocsb_time_begin(ocsb_ctx *ctx, int kclass);
void ocsb_time_end(ocsb_ctx *ctx);
#define OCSB_TIME(ctx, kc)                                                     \
    do {                                                                       \
        if ((ctx)->timing) ocsb_time_begin((ctx), (kc));                       \
    } while (0)
#define OCSB_LAUNCHED(ctx)                                                     \
    do {                                                                       \
        (ctx)->launches++;                                                     \
        if ((ctx)->time_open) ocsb_time_end((ctx));                            \
    } while (0)
#define OCSB_CHECK_LAUNCH()                                                    \
    do {                                                                       \
        cudaError_t e__ = cudaGetLastError();                                  \
        if (e__ != cudaSuccess) {                                              \
            ocsb_set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__,   \
                           cudaGetErrorString(e__));                           \
            return OCSB_ECUDA;                                                 \
        }                                                                      \
    } while (0)

// ---------------------------------------------------------------------------
// kernel launchers (device pointers only), one per translation unit
// ---------------------------------------------------------------------------
int32_t k_fill_f32(ocsb_ctx *ctx, float *dst, int64_t n, float v);
int32_t k_clamp(ocsb_ctx *ctx, float *hfun, int64_t stride, int w, int h,
                double hmin, double hmax);
int32_t k_min_combine(ocsb_ctx *ctx, float *dst, int64_t dstride,
                      const float *src, int64_t sstride, int w, int h);
int32_t k_coords(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, double *x, double *y);

struct LimiterArgs {
    double dx, dy;
    double hmin, hmax, lb, ub, ghmin, ghmax;  // NaN = None
    int fuse_clamp;
};
int32_t k_flow_limiter(ocsb_ctx *ctx, int w, int h, int row_begin, int row_end,
                       const void *dem,
                       int dem_dtype, int64_t dem_stride, const float *old_,
                       int64_t old_stride, float *new_, int64_t new_stride,
                       const LimiterArgs &a);

struct FeatureArgs {
    LinAxis ax, ay;         // pixel coordinates of the window
    int row_begin, row_end; // device tile: rows [row_begin, row_end)
    double rate_target;     // expansion_rate * target_size (f64 product)
    double target;
    double r;               // search radius, +inf if unbounded
    float hmin_f, hmax_f;   // f32-rounded global bounds
    int has_hmin, has_hmax;
    int fuse_clamp;
    // non-null: asynchronous mode.  One attempt with generous pools, no
    // host read-back; the pool-overflow flag is left in *ext_overflow (device
    // int, zeroed by the caller), and the evaluation kernel writes nothing if
    // it is set -- the caller checks later and redoes the band synchronously.
    int *ext_overflow = nullptr;
    // Constraint._apply_rate epilogue instead of the size law (exact path):
    //   t = vals[idx] * (1 + d*rate*sign); new = satisfies(old, t) ? old : t
    // for pixels outside `rate_mask` (features/constraint.py:104-117).
    // Overlap hook (resident op lists, ocsb_hfun_pass): the candidate-pyramid
    // refinement runs on `refine_stream` while `*between` (the caller's other,
    // vertex-independent ops of the list) is enqueued on the context's stream;
    // the evaluation kernel joins both.  Called exactly once.
    cudaStream_t refine_stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    std::function<int32_t()> *between = nullptr;
    const float *rate_vals = nullptr;    // per-vertex values, nv floats
    const uint8_t *rate_mask = nullptr;  // window row 0, stride = width
    double rate = 0.0, rate_sign = 0.0;
    int rate_is_min = 0;
};
int32_t k_hfun_feature(ocsb_ctx *ctx, const FeatureArgs &a,
                       const double *verts, int64_t nv, const float *old_,
                       int64_t old_stride, float *new_, int64_t new_stride,
                       double *dist_out, int64_t *idx_out);
// k_pass.cu: the op list of a pass as ONE tiled sweep over a table of semantic
// windows (feature ops first, then limiter ops: with one fuse_clamp setting the
// ops commute bit for bit)
struct PassWin {
    LinAxis ax, ay;         // np.linspace axes of the window
    int rb, re;             // window rows [rb, re) processed by this call
    int keep_w, keep_h;     // window pixels (r < keep_h, c < keep_w) are written
    int64_t hoff;           // element offset of window pixel (0, 0) in in / out
    int dem_c, dem_r;       // DEM block column / row of window pixel (0, 0)
};
struct PassFeat {
    const double *verts;
    int64_t nv;
    double rate_target, target, r;
    int fuse_clamp;
};
struct PassLim {
    double hmin, hmax, lb, ub;   // NaN = None
    int fuse_clamp;
};
struct PassDesc {
    const PassWin *wins = nullptr;
    int n_wins = 0;
    const void *dem = nullptr;   // element (0, 0) of the stored DEM block
    int dem_dtype = OCSB_F32;
    int64_t dem_stride = 0;
    int dem_rows = 0, dem_cols = 0;
    const float *in = nullptr;   // nullptr: initial state (finfo(float32).max)
    int64_t in_stride = 0;
    float *out = nullptr;
    int64_t out_stride = 0;
    double dx = 0, dy = 0, ghmin = 0, ghmax = 0;
    const PassFeat *feats = nullptr;
    int nf = 0;
    const PassLim *lims = nullptr;
    int nl = 0;
    int *ext_overflow = nullptr; // non-null: asynchronous (see FeatureArgs)
    // asynchronous AND sticky: the pass is a no-op on the device when
    // *ext_overflow is already non-zero on entry (ocsb_hfun_pass_windows)
    bool sticky = false;
};
bool k_pass_supported(const PassDesc &d);
int32_t k_pass_tiled(ocsb_ctx *ctx, const PassDesc &d);

int32_t k_nearest_vertex(ocsb_ctx *ctx, const double *verts, int64_t nv,
                         const double *q, int64_t nq, double r, double *dist,
                         int64_t *idx);

// k_sample.cu
int32_t k_sample_nearest(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const float *hfun,
                         int64_t hs, int step, double hmin, double hmax,
                         const double *xy, int64_t K, double *out);
int32_t k_sample_source(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const float *hfun,
                        int64_t hs, int step, double hmin, double hmax,
                        double *coords, double *values);

// k_warp.cu: nearest-neighbour warp between north-up rasters of one CRS
int32_t k_warp_nearest(ocsb_ctx *ctx, const void *src, int src_dtype, int64_t ss, int sw,
                       int sh, const double sgt[6], const uint8_t *mask, int64_t ms,
                       float *dst, int64_t ds, int dw, int dh, const double dgt[6]);

// k_contour.cu
int32_t k_contour_segments(ocsb_ctx *ctx, const void *dem, int dem_dtype,
                           int64_t zs, int w, int jb, int je, double level,
                           LinAxis ax, LinAxis ay, int64_t cap, int64_t *edges,
                           double *pts, int64_t *n_out);

// k_contour.cu, level lists: ONE count sweep for all levels, an emit launch per level
int k_contour_max_levels();
int32_t k_contour_count_levels(ocsb_ctx *ctx, ScratchGuard &sg, const void *dem, int dem_dtype,
                               int64_t zs, int w, int jb, int je, const double *levels, int nl,
                               int **offsets, unsigned long long **totals_dev,
                               int64_t *totals, int64_t *nblk_out);
int32_t k_contour_emit_level(ocsb_ctx *ctx, const void *dem, int dem_dtype, int64_t zs, int w,
                             int jb, int je, double level, LinAxis ax, LinAxis ay,
                             const int *offsets, const unsigned long long *total_dev,
                             int64_t *edges, double *pts);

// k_stitch.cu (results are carved from the caller's ScratchGuard)
int32_t k_stitch_segments(ocsb_ctx *ctx, ScratchGuard &sg, const int64_t *edges,
                          const double *pts, int64_t n, double **xy_out, int64_t **off_out,
                          int64_t *n_xy, int64_t *n_lines);
int32_t k_resample_lines(ocsb_ctx *ctx, ScratchGuard &sg, const double *xy, const int64_t *off,
                         int64_t n_lines, double target, int max_verts, double **out,
                         int64_t *n_out);

// k_elementwise.cu: band exchange blocks
int32_t k_band_pack(ocsb_ctx *ctx, const double *verts, int64_t n, int64_t cap,
                    const float *row_first, const float *row_last, int w, int64_t row_bytes,
                    void *send, double far);
int32_t k_band_unpack(ocsb_ctx *ctx, const void *recv, int world, int rank, int64_t blk_bytes,
                      int64_t row_bytes, int64_t cap, int w, double *verts_out,
                      float *halo_above, float *halo_below);

// k_region.cu
int32_t k_region_mask(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, double x0, double ytop, double pw,
                      double ph, int rb, int re, int w, int touched, const double *edges,
                      int64_t n_edges, const double bbox[4], float *out, int64_t stride);

// k_constraint.cu
int32_t k_constant_value(ocsb_ctx *ctx, int w, int h, const void *dem,
                         int dem_dtype, int64_t zs, const float *old_,
                         int64_t os, float *new_, int64_t ns, double value,
                         double lb, double ub, double ghmin, double ghmax,
                         int fuse_clamp);
int32_t k_topo_mask(ocsb_ctx *ctx, int w, int h, const void *dem, int dem_dtype,
                    int64_t zs, float *hfun, int64_t hs, const void *vals,
                    int vals_dtype, int64_t vs, double vconst, double lb,
                    double ub, int is_min, uint8_t *mask,
                    unsigned long long *nmask);
int32_t k_courant(ocsb_ctx *ctx, int w, int h, const void *dem, int dem_dtype,
                  int64_t zs, float *hfun, int64_t hs, double cval, double dt,
                  double nu, int is_min);
int32_t k_mask_boundary(ocsb_ctx *ctx, ScratchGuard &sg, const uint8_t *mask,
                        int w, int h, LinAxis ax, LinAxis ay, const float *hfun,
                        int64_t hs, double **verts_out, float **vals_out,
                        int64_t *nv_out, int row0 = 0, int open_top = 0,
                        int open_bot = 0);

int32_t k_interp(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const void *dem,
                 int dem_dtype, int64_t dem_stride, const uint8_t *mask,
                 int64_t mask_stride, const double *xy, int64_t K, int method,
                 uint8_t *touched, double *values);
struct InterpRaster {
    LinAxis ax, ay;
    const void *dem;
    int dem_dtype;
    int64_t dem_stride;
    const uint8_t *mask;
    int64_t mask_stride;
};
// rasters in list order, last valid hit wins; source (optional) receives the
// index of the raster that set each node
int32_t k_interp_multi(ocsb_ctx *ctx, const InterpRaster *rs, int n,
                       const double *xy, int64_t K, int method, uint8_t *touched,
                       double *values, int32_t *source, const int32_t *perm);
// counting sort of nodes by grid cell: perm[i] = original index of sorted node i
int32_t k_bucket_nodes(ocsb_ctx *ctx, const double *xy, int64_t K,
                       const double bbox[4], double cell, int32_t *perm,
                       double *xy_sorted);
int32_t k_spline_prefilter(ocsb_ctx *ctx, LinAxis ax, LinAxis ay,
                           const void *dem, int dem_dtype, int64_t dem_stride,
                           double *coef);
struct SplineEvalRaster {
    LinAxis ax, ay;
    const double *coef;   // (nx, ny) coefficients of k_spline_prefilter
};
// rasters in list order, the last one whose window contains a node wins
int32_t k_spline_eval_multi(ocsb_ctx *ctx, const SplineEvalRaster *rs, int n,
                            const double *xy, int64_t K, uint8_t *touched,
                            double *values);
int32_t k_spline_eval(ocsb_ctx *ctx, LinAxis ax, LinAxis ay, const double *coef,
                      const double *xy, int64_t K, uint8_t *touched,
                      double *values);
