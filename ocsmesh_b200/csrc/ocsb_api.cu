// C ABI entry points of libocsb200.so: argument validation, host<->device
// staging, context management.  Kernels live in k_*.cu.
#include "ocsb_common.cuh"

#include <stdarg.h>
#include <limits>

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------
static thread_local char g_err[1024] = "";

void ocsb_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *ocsb_last_error(void) { return g_err; }
extern "C" int32_t ocsb_abi_version(void) { return OCSB_ABI_VERSION; }

// ---------------------------------------------------------------------------
// scratch memory: stream-ordered allocation from the device's default pool
// ---------------------------------------------------------------------------
int32_t ocsb_alloc(ocsb_ctx *ctx, void **p, size_t bytes) {
    if (bytes == 0) bytes = 16;
    cudaError_t e = cudaMallocAsync(p, bytes, ctx->stream);
    if (e != cudaSuccess) {
        ocsb_set_error("cudaMallocAsync(%zu bytes) -> %s", bytes,
                       cudaGetErrorString(e));
        (void)cudaGetLastError();
        return OCSB_ENOMEM;
    }
    return OCSB_OK;
}

void ocsb_free(ocsb_ctx *ctx, void *p) {
    if (p) cudaFreeAsync(p, ctx->stream);
}

bool ocsb_is_device_ptr(const void *p) {
    if (!p) return false;
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, p);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

int32_t Staged::init(ocsb_ctx *c, const void *p, int64_t rows_, int64_t width_,
                     int64_t stride_, size_t elem_, int dir_, int64_t row0_) {
    ctx = c;
    row0 = row0_;
    user = const_cast<void *>(p);
    rows = rows_;
    width = width_;
    user_stride = stride_;
    elem = elem_;
    dir = dir_;
    if (!p) {
        dev = nullptr;
        return OCSB_OK;
    }
    // `p` is the (virtual) address of window row 0; only rows >= row0 exist
    // (a rank's band buffer starts at its first held row), so the memory kind
    // is asked of the first row that does
    if (ocsb_is_device_ptr((const char *)p + (size_t)row0_ * (size_t)stride_ * elem_)) {
        dev = user;
        dev_stride = user_stride;
        staged = false;
        return OCSB_OK;
    }
    staged = true;
    dev_stride = width;
    OCSB_TRY(ocsb_alloc(ctx, &alloc, (size_t)rows * (size_t)width * elem));
    // virtual address of row 0: only rows >= row0 are ever dereferenced
    dev = (char *)alloc - (size_t)row0 * (size_t)width * elem;
    if (dir & STAGE_IN) {
        OCSB_CUDA(cudaMemcpy2DAsync(alloc, (size_t)width * elem,
                                    (const char *)user +
                                        (size_t)row0 * (size_t)user_stride * elem,
                                    (size_t)user_stride * elem,
                                    (size_t)width * elem, (size_t)rows,
                                    cudaMemcpyHostToDevice, ctx->stream));
    }
    return OCSB_OK;
}

int32_t Staged::finish() {
    if (staged && alloc && (dir & STAGE_OUT)) {
        OCSB_CUDA(cudaMemcpy2DAsync((char *)user +
                                        (size_t)row0 * (size_t)user_stride * elem,
                                    (size_t)user_stride * elem, alloc,
                                    (size_t)width * elem, (size_t)width * elem,
                                    (size_t)rows, cudaMemcpyDeviceToHost,
                                    ctx->stream));
    }
    return OCSB_OK;
}

Staged::~Staged() {
    if (staged && alloc) ocsb_free(ctx, alloc);
}

// after the work of one API call was enqueued: if anything was staged the
// call must not return before the copy-back has landed in the caller's memory
static int32_t finish_call(ocsb_ctx *ctx, std::initializer_list<Staged *> bufs) {
    bool any = false;
    for (Staged *s : bufs) {
        OCSB_TRY(s->finish());
        any = any || s->staged;
    }
    if (any) OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    return OCSB_OK;
}

// ---------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_ctx_create(int32_t device, void *stream,
                                   ocsb_ctx **out) {
    OCSB_REQUIRE(out != nullptr, "ocsb_ctx_create: out is NULL");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        (void)cudaGetLastError();
        ocsb_set_error("ocsb_ctx_create: no CUDA device available (%s); "
                       "libocsb200 has no CPU fallback",
                       e != cudaSuccess ? cudaGetErrorString(e) : "0 devices");
        return OCSB_ECUDA;
    }
    OCSB_REQUIRE(device >= 0 && device < ndev,
                 "ocsb_ctx_create: device %d out of range [0,%d)", device, ndev);
    OCSB_CUDA(cudaSetDevice(device));
    ocsb_ctx *c = new ocsb_ctx();
    c->device = device;
    if (stream) {
        c->stream = (cudaStream_t)stream;
        c->own_stream = false;
    } else {
        cudaError_t e2 = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
        if (e2 != cudaSuccess) {
            ocsb_set_error("cudaStreamCreate -> %s", cudaGetErrorString(e2));
            delete c;
            return OCSB_ECUDA;
        }
        c->own_stream = true;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess)
        c->num_sms = prop.multiProcessorCount;
    // keep freed scratch in the pool instead of returning it to the driver
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        uint64_t thr = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    c->pinned_bytes = 4096;
    if (cudaMallocHost(&c->pinned, c->pinned_bytes) != cudaSuccess) {
        (void)cudaGetLastError();
        c->pinned = nullptr;
    }
    *out = c;
    return OCSB_OK;
}

extern "C" int32_t ocsb_ctx_destroy(ocsb_ctx *ctx) {
    if (!ctx) return OCSB_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->spl_coef) cudaFree(ctx->spl_coef);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    for (auto &k : ctx->kept)
        if (k) ocsb_free(ctx, k);
    for (auto &sl : ctx->stage_ring) {
        if (sl.host) cudaFreeHost(sl.host);
        if (sl.ev) cudaEventDestroy(sl.ev);
    }
    for (auto &r : ctx->time_recs) {
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    for (auto e : ctx->event_pool) cudaEventDestroy(e);
    for (int i = 0; i < OCSB_WIN_STREAMS; ++i) {
        if (ctx->s_win[i]) cudaStreamDestroy(ctx->s_win[i]);
        if (ctx->ev_win[i]) cudaEventDestroy(ctx->ev_win[i]);
    }
    if (ctx->keep_arena) cudaFree(ctx->keep_arena);
    if (ctx->s_split) cudaStreamDestroy(ctx->s_split);
    for (auto &e : ctx->ev_split)
        if (e) cudaEventDestroy(e);
    if (ctx->s_h2d) cudaStreamDestroy(ctx->s_h2d);
    if (ctx->s_d2h) cudaStreamDestroy(ctx->s_d2h);
    for (int i = 0; i < OCSB_PASS_SETS; ++i) {
        if (ctx->ev_in[i]) cudaEventDestroy(ctx->ev_in[i]);
        if (ctx->ev_comp[i]) cudaEventDestroy(ctx->ev_comp[i]);
        if (ctx->ev_out[i]) cudaEventDestroy(ctx->ev_out[i]);
    }
    if (ctx->ev_alloc) cudaEventDestroy(ctx->ev_alloc);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return OCSB_OK;
}

int32_t ocsb_stage_upload(ocsb_ctx *ctx, void *dst_dev, const void *src, size_t bytes) {
    if (bytes == 0) return OCSB_OK;
    ocsb_ctx::StageSlot &sl = ctx->stage_ring[ctx->stage_next];
    ctx->stage_next = (ctx->stage_next + 1) % 32;
    if (sl.used) OCSB_CUDA(cudaEventSynchronize(sl.ev));   // its last upload has left
    if (!sl.ev) OCSB_CUDA(cudaEventCreateWithFlags(&sl.ev, cudaEventDisableTiming));
    if (sl.bytes < bytes) {
        if (sl.host) cudaFreeHost(sl.host);
        sl.host = nullptr;
        sl.bytes = 0;
        const size_t want = bytes < 4096 ? 4096 : bytes;
        OCSB_CUDA(cudaMallocHost(&sl.host, want));
        sl.bytes = want;
    }
    memcpy(sl.host, src, bytes);
    OCSB_CUDA(cudaMemcpyAsync(dst_dev, sl.host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    OCSB_CUDA(cudaEventRecord(sl.ev, ctx->stream));
    sl.used = true;
    return OCSB_OK;
}

extern "C" int32_t ocsb_ctx_set_stream(ocsb_ctx *ctx, void *stream) {
    OCSB_REQUIRE(ctx, "ctx is NULL");
    if ((cudaStream_t)stream != ctx->stream && ctx->keep_arena) {
        // the kept scratch arena is only ordered on the old stream
        cudaStreamSynchronize(ctx->stream);
        cudaFree(ctx->keep_arena);
        ctx->keep_arena = nullptr;
        ctx->keep_arena_bytes = 0;
        ctx->wins_cache.clear();
    }
    if (ctx->own_stream) {
        cudaStreamSynchronize(ctx->stream);
        cudaStreamDestroy(ctx->stream);
        ctx->own_stream = false;
    }
    ctx->stream = (cudaStream_t)stream;
    return OCSB_OK;
}

extern "C" int32_t ocsb_ctx_sync(ocsb_ctx *ctx) {
    OCSB_REQUIRE(ctx, "ctx is NULL");
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    return OCSB_OK;
}

extern "C" int64_t ocsb_ctx_launch_count(const ocsb_ctx *ctx) {
    return ctx ? ctx->launches : 0;
}

// ---------------------------------------------------------------------------
// per-kernel-class timing
// ---------------------------------------------------------------------------
static cudaEvent_t take_event(ocsb_ctx *ctx) {
    if (!ctx->event_pool.empty()) {
        cudaEvent_t e = ctx->event_pool.back();
        ctx->event_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

void ocsb_time_begin(ocsb_ctx *ctx, int kclass) {
    KTimeRec r;
    r.kclass = kclass;
    r.a = take_event(ctx);
    r.b = take_event(ctx);
    cudaEventRecord(r.a, ctx->stream);
    ctx->time_recs.push_back(r);
    ctx->time_open = true;
}

void ocsb_time_end(ocsb_ctx *ctx) {
    cudaEventRecord(ctx->time_recs.back().b, ctx->stream);
    ctx->time_open = false;
}

extern "C" int32_t ocsb_ctx_set_timing(ocsb_ctx *ctx, int32_t enabled) {
    OCSB_REQUIRE(ctx, "ctx is NULL");
    ctx->timing = enabled != 0;
    return OCSB_OK;
}

extern "C" int32_t ocsb_ctx_reset_timing(ocsb_ctx *ctx) {
    OCSB_REQUIRE(ctx, "ctx is NULL");
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    for (auto &r : ctx->time_recs) {
        ctx->event_pool.push_back(r.a);
        ctx->event_pool.push_back(r.b);
    }
    ctx->time_recs.clear();
    ctx->time_open = false;
    return OCSB_OK;
}

extern "C" int32_t ocsb_ctx_get_timing(ocsb_ctx *ctx, int32_t kclass,
                                       double *total_ms, int64_t *launches) {
    OCSB_REQUIRE(ctx, "ctx is NULL");
    OCSB_REQUIRE(kclass >= 0 && kclass < OCSB_KC_COUNT, "bad kernel class %d",
                 kclass);
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    double sum = 0.0;
    int64_t n = 0;
    for (auto &r : ctx->time_recs) {
        if (r.kclass != kclass) continue;
        float ms = 0.f;
        OCSB_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
        sum += ms;
        ++n;
    }
    if (total_ms) *total_ms = sum;
    if (launches) *launches = n;
    return OCSB_OK;
}

extern "C" const char *ocsb_kernel_class_name(int32_t kclass) {
    static const char *names[OCSB_KC_COUNT] = {
        "fill", "clamp", "min_combine", "coords", "flow_limiter",
        "feature_refine", "feature_eval", "interp", "spline_prefilter",
        "spline_eval", "nearest_vertex", "bucket_nodes", "constraint", "contour", "sample",
        "pass_tile", "pass_limiter", "pass_feature"};
    return (kclass >= 0 && kclass < OCSB_KC_COUNT) ? names[kclass] : "?";
}

extern "C" int32_t ocsb_ctx_trim(ocsb_ctx *ctx) {
    OCSB_REQUIRE(ctx, "ctx is NULL");
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->spl_coef) {
        cudaFree(ctx->spl_coef);
        ctx->spl_coef = nullptr;
        ctx->spl_key = nullptr;
        ctx->spl_bytes = 0;
    }
    if (ctx->keep_arena) {
        cudaFree(ctx->keep_arena);
        ctx->keep_arena = nullptr;
        ctx->keep_arena_bytes = 0;
        ctx->wins_cache.clear();
    }
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, ctx->device) == cudaSuccess)
        cudaMemPoolTrimTo(pool, 0);
    return OCSB_OK;
}

// ---------------------------------------------------------------------------
// A1 windows (host only)
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_windows(int32_t width, int32_t height,
                                int32_t chunk_size, int32_t overlap,
                                int32_t *out4, int64_t cap, int64_t *n) {
    OCSB_REQUIRE(n != nullptr, "ocsb_windows: n is NULL");
    OCSB_REQUIRE(width > 0 && height > 0, "ocsb_windows: empty raster %dx%d",
                 width, height);
    OCSB_REQUIRE(chunk_size >= 0, "Argument chunk_size must be >= 0.");
    if (chunk_size == 0) {
        *n = 1;
        if (!out4 || cap < 1) return out4 ? OCSB_EOVERFLOW : OCSB_OK;
        out4[0] = 0; out4[1] = 0; out4[2] = width; out4[3] = height;
        return OCSB_OK;
    }
    int64_t nh = ((int64_t)height + chunk_size - 1) / chunk_size;
    int64_t nw = ((int64_t)width + chunk_size - 1) / chunk_size;
    *n = nh * nw;
    if (!out4) return OCSB_OK;
    if (cap < *n) {
        ocsb_set_error("ocsb_windows: %lld windows do not fit cap %lld",
                       (long long)*n, (long long)cap);
        return OCSB_EOVERFLOW;
    }
    int64_t k = 0;
    for (int64_t i = 0; i < nh; ++i) {
        for (int64_t j = 0; j < nw; ++j) {
            int64_t off_h = i * chunk_size, off_w = j * chunk_size;
            int64_t h = (int64_t)chunk_size + overlap;
            if (off_h + h > height) h = h - (off_h + h) % height;
            int64_t w = (int64_t)chunk_size + overlap;
            if (off_w + w > width) w = w - (off_w + w) % width;
            out4[4 * k + 0] = (int32_t)off_w;
            out4[4 * k + 1] = (int32_t)off_h;
            out4[4 * k + 2] = (int32_t)w;
            out4[4 * k + 3] = (int32_t)h;
            ++k;
        }
    }
    return OCSB_OK;
}

static int32_t check_win(const ocsb_win *win, const char *who) {
    OCSB_REQUIRE(win != nullptr, "%s: win is NULL", who);
    OCSB_REQUIRE(win->width > 0 && win->height > 0, "%s: empty window %dx%d",
                 who, win->width, win->height);
    OCSB_REQUIRE(isfinite(win->x0) && isfinite(win->x1) && isfinite(win->y0) &&
                     isfinite(win->y1),
                 "%s: non-finite window bounds", who);
    return OCSB_OK;
}

// the device tile [rb, re) of a window
static int32_t win_band(const ocsb_win *win, const char *who, int *rb, int *re) {
    *rb = win->row_begin;
    *re = win->row_end == 0 ? win->height : win->row_end;
    OCSB_REQUIRE(*rb >= 0 && *rb < *re && *re <= win->height,
                 "%s: bad row band [%d, %d) of a window of height %d", who,
                 win->row_begin, win->row_end, win->height);
    return OCSB_OK;
}

static void win_axes(const ocsb_win *win, LinAxis *ax, LinAxis *ay) {
    *ax = make_axis(win->x0, win->x1, win->width);   // get_x: linspace(x0, x1, w)
    *ay = make_axis(win->y1, win->y0, win->height);  // get_y: linspace(y1, y0, h)
}

#define CTX_ENTER(who)                                                         \
    OCSB_REQUIRE(ctx != nullptr, "%s: ctx is NULL", who);                      \
    OCSB_CUDA(cudaSetDevice(ctx->device))

// ---------------------------------------------------------------------------
// A2 coords
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_window_coords(ocsb_ctx *ctx, const ocsb_win *win,
                                      double *x_out, double *y_out) {
    CTX_ENTER("ocsb_window_coords");
    OCSB_TRY(check_win(win, "ocsb_window_coords"));
    OCSB_REQUIRE(x_out && y_out, "ocsb_window_coords: NULL output");
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    Staged sx, sy;
    OCSB_TRY(sx.init(ctx, x_out, 1, win->width, win->width, 8, STAGE_OUT));
    OCSB_TRY(sy.init(ctx, y_out, 1, win->height, win->height, 8, STAGE_OUT));
    OCSB_TRY(k_coords(ctx, ax, ay, (double *)sx.dev, (double *)sy.dev));
    return finish_call(ctx, {&sx, &sy});
}

// ---------------------------------------------------------------------------
// A4 fill / A10 clamp / A12 min-combine
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_fill_f32(ocsb_ctx *ctx, float *dst, int64_t n,
                                 float value) {
    CTX_ENTER("ocsb_fill_f32");
    OCSB_REQUIRE(n >= 0, "ocsb_fill_f32: negative n");
    if (n == 0) return OCSB_OK;
    OCSB_REQUIRE(dst, "ocsb_fill_f32: dst is NULL");
    Staged s;
    OCSB_TRY(s.init(ctx, dst, 1, n, n, 4, STAGE_OUT));
    OCSB_TRY(k_fill_f32(ctx, (float *)s.dev, n, value));
    return finish_call(ctx, {&s});
}

extern "C" int32_t ocsb_hfun_clamp(ocsb_ctx *ctx, float *hfun,
                                   int64_t row_stride, int32_t width,
                                   int32_t height, double hmin, double hmax) {
    CTX_ENTER("ocsb_hfun_clamp");
    OCSB_REQUIRE(hfun && width > 0 && height > 0 && row_stride >= width,
                 "ocsb_hfun_clamp: bad buffer (%p, %dx%d, stride %lld)",
                 (void *)hfun, width, height, (long long)row_stride);
    Staged s;
    OCSB_TRY(s.init(ctx, hfun, height, width, row_stride, 4, STAGE_INOUT));
    OCSB_TRY(k_clamp(ctx, (float *)s.dev, s.dev_stride, width, height, hmin,
                     hmax));
    return finish_call(ctx, {&s});
}

extern "C" int32_t ocsb_min_combine(ocsb_ctx *ctx, float *dst,
                                    int64_t dst_stride, const float *src,
                                    int64_t src_stride, int32_t width,
                                    int32_t height) {
    CTX_ENTER("ocsb_min_combine");
    OCSB_REQUIRE(dst && src && width > 0 && height > 0 &&
                     dst_stride >= width && src_stride >= width,
                 "ocsb_min_combine: bad buffers");
    Staged d, s;
    OCSB_TRY(d.init(ctx, dst, height, width, dst_stride, 4, STAGE_INOUT));
    OCSB_TRY(s.init(ctx, src, height, width, src_stride, 4, STAGE_IN));
    OCSB_TRY(k_min_combine(ctx, (float *)d.dev, d.dev_stride,
                           (const float *)s.dev, s.dev_stride, width, height));
    return finish_call(ctx, {&d, &s});
}

// ---------------------------------------------------------------------------
// A6 + A7 feature distance sizing
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_hfun_feature(
    ocsb_ctx *ctx, const ocsb_win *win, const double *verts_xy, int64_t nv,
    const float *hfun_old, int64_t old_stride, float *hfun_new,
    int64_t new_stride, double expansion_rate, double target_size, double hmin,
    double hmax, int32_t fuse_clamp, double *dist_out, int64_t *idx_out) {
    CTX_ENTER("ocsb_hfun_feature");
    OCSB_TRY(check_win(win, "ocsb_hfun_feature"));
    OCSB_REQUIRE(verts_xy && nv > 0,
                 "ocsb_hfun_feature: empty vertex set (cKDTree needs a "
                 "non-empty (N,2) array)");
    OCSB_REQUIRE(nv < (int64_t)2147483647, "ocsb_hfun_feature: nv too large");
    OCSB_REQUIRE(hfun_new && new_stride >= win->width &&
                     (!hfun_old || old_stride >= win->width),
                 "ocsb_hfun_feature: bad hfun buffers");
    OCSB_REQUIRE(!ocsb_is_none(target_size) && target_size > 0,
                 "Argument target_size must be greater than zero.");
    OCSB_REQUIRE(!ocsb_is_none(expansion_rate),
                 "ocsb_hfun_feature: expansion_rate is NaN");

    FeatureArgs a;
    win_axes(win, &a.ax, &a.ay);
    a.rate_target = expansion_rate * target_size;
    a.target = target_size;
    a.has_hmin = !ocsb_is_none(hmin);
    a.has_hmax = !ocsb_is_none(hmax);
    a.hmin_f = a.has_hmin ? (float)hmin : 0.f;
    a.hmax_f = a.has_hmax ? (float)hmax : 0.f;
    a.fuse_clamp = fuse_clamp;
    if (a.has_hmax && hmax != 0.0) {  // `if self.hmax:` hfun/raster.py:1240
        OCSB_REQUIRE(a.rate_target != 0.0,
                     "ocsb_hfun_feature: float division by zero "
                     "(expansion_rate * target_size == 0 with hmax set)");
        a.r = (hmax - target_size) / a.rate_target;
    } else {
        a.r = std::numeric_limits<double>::infinity();
    }

    const int w = win->width;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_hfun_feature", &rb, &re));
    a.row_begin = rb;
    a.row_end = re;
    const int h = re - rb;
    Staged sv, so, sn, sd, si;
    OCSB_TRY(sv.init(ctx, verts_xy, 1, nv * 2, nv * 2, 8, STAGE_IN));
    const bool alias = (const void *)hfun_old == (const void *)hfun_new &&
                       old_stride == new_stride;
    if (alias) {
        OCSB_TRY(sn.init(ctx, hfun_new, h, w, new_stride, 4, STAGE_INOUT, rb));
    } else {
        // hfun_old == NULL: the pre-op raster is the A4 initial state
        if (hfun_old)
            OCSB_TRY(so.init(ctx, hfun_old, h, w, old_stride, 4, STAGE_IN, rb));
        OCSB_TRY(sn.init(ctx, hfun_new, h, w, new_stride, 4, STAGE_OUT, rb));
    }
    OCSB_TRY(sd.init(ctx, dist_out, h, w, w, 8, STAGE_OUT, rb));
    OCSB_TRY(si.init(ctx, idx_out, h, w, w, 8, STAGE_OUT, rb));
    const float *oldp = alias ? (const float *)sn.dev : (const float *)so.dev;
    int64_t olds = alias ? sn.dev_stride : so.dev_stride;
    if (!dist_out && !idx_out) {
        // sizes only: the tiled sweep (vertex-grid pyramid + persistent pass
        // kernel, k_pass.cu) -- same arithmetic, bit-identical
        static const bool tiled_on = [] {
            const char *e = getenv("OCSB_PASS_TILED");
            return !(e && atoi(e) == 0);
        }();
        PassWin pw;
        pw.ax = a.ax; pw.ay = a.ay;
        pw.rb = rb; pw.re = re;
        pw.keep_w = w; pw.keep_h = win->height;
        pw.hoff = 0; pw.dem_c = 0; pw.dem_r = 0;
        PassFeat pf;
        pf.verts = (const double *)sv.dev; pf.nv = nv;
        pf.rate_target = a.rate_target; pf.target = a.target; pf.r = a.r;
        pf.fuse_clamp = fuse_clamp;
        PassDesc d;
        d.wins = &pw; d.n_wins = 1;
        d.in = oldp; d.in_stride = olds;
        d.out = (float *)sn.dev; d.out_stride = sn.dev_stride;
        d.ghmin = hmin; d.ghmax = hmax;
        d.feats = &pf; d.nf = 1;
        if (tiled_on && k_pass_supported(d)) {
            OCSB_TRY(k_pass_tiled(ctx, d));
            return finish_call(ctx, {&sv, &so, &sn, &sd, &si});
        }
    }
    OCSB_TRY(k_hfun_feature(ctx, a, (const double *)sv.dev, nv, oldp, olds,
                            (float *)sn.dev, sn.dev_stride, (double *)sd.dev,
                            (int64_t *)si.dev));
    return finish_call(ctx, {&sv, &so, &sn, &sd, &si});
}

extern "C" int32_t ocsb_nearest_vertex(ocsb_ctx *ctx, const double *verts_xy,
                                       int64_t nv, const double *query_xy,
                                       int64_t nq, double r, double *dist_out,
                                       int64_t *idx_out) {
    CTX_ENTER("ocsb_nearest_vertex");
    OCSB_REQUIRE(verts_xy && nv > 0, "ocsb_nearest_vertex: empty vertex set");
    OCSB_REQUIRE(nv < (int64_t)2147483647, "ocsb_nearest_vertex: nv too large");
    OCSB_REQUIRE(nq >= 0, "ocsb_nearest_vertex: negative nq");
    if (nq == 0) return OCSB_OK;
    OCSB_REQUIRE(query_xy && (dist_out || idx_out),
                 "ocsb_nearest_vertex: NULL buffers");
    double rr = ocsb_is_none(r) ? std::numeric_limits<double>::infinity() : r;
    Staged sv, sq, sd, si;
    OCSB_TRY(sv.init(ctx, verts_xy, 1, nv * 2, nv * 2, 8, STAGE_IN));
    OCSB_TRY(sq.init(ctx, query_xy, 1, nq * 2, nq * 2, 8, STAGE_IN));
    OCSB_TRY(sd.init(ctx, dist_out, 1, nq, nq, 8, STAGE_OUT));
    OCSB_TRY(si.init(ctx, idx_out, 1, nq, nq, 8, STAGE_OUT));
    OCSB_TRY(k_nearest_vertex(ctx, (const double *)sv.dev, nv,
                              (const double *)sq.dev, nq, rr,
                              (double *)sd.dev, (int64_t *)si.dev));
    return finish_call(ctx, {&sv, &sq, &sd, &si});
}

// ---------------------------------------------------------------------------
// A8 flow limiter
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_hfun_flow_limiter(
    ocsb_ctx *ctx, const ocsb_win *win, const void *dem, int32_t dem_dtype,
    int64_t dem_stride, const float *hfun_old, int64_t old_stride,
    float *hfun_new, int64_t new_stride, double dx, double dy, double hmin,
    double hmax, double lower_bound, double upper_bound, double ghmin,
    double ghmax, int32_t fuse_clamp) {
    CTX_ENTER("ocsb_hfun_flow_limiter");
    OCSB_TRY(check_win(win, "ocsb_hfun_flow_limiter"));
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "ocsb_hfun_flow_limiter: unsupported DEM dtype %d", dem_dtype);
    OCSB_REQUIRE(win->width >= 2 && win->height >= 2,
                 "Shape of array too small to calculate a numerical gradient, "
                 "at least (edge_order + 1) elements are required.");
    OCSB_REQUIRE(dem && hfun_new && dem_stride >= win->width &&
                     (!hfun_old || old_stride >= win->width) &&
                     new_stride >= win->width,
                 "ocsb_hfun_flow_limiter: bad buffers");
    OCSB_REQUIRE(!ocsb_is_none(dx) && !ocsb_is_none(dy),
                 "ocsb_hfun_flow_limiter: dx/dy is NaN");
    const int w = win->width, h = win->height;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_hfun_flow_limiter", &rb, &re));
    const int hb = re - rb;
    // DEM rows needed: the band plus one halo row on each side
    const int zb = rb > 0 ? rb - 1 : 0, ze = re < h ? re + 1 : h;
    const size_t es = dem_dtype == OCSB_F32 ? 4 : 8;
    Staged sz, so, sn;
    OCSB_TRY(sz.init(ctx, dem, ze - zb, w, dem_stride, es, STAGE_IN, zb));
    const bool alias = (const void *)hfun_old == (const void *)hfun_new &&
                       old_stride == new_stride;
    if (alias) {
        OCSB_TRY(sn.init(ctx, hfun_new, hb, w, new_stride, 4, STAGE_INOUT, rb));
    } else {
        if (hfun_old)
            OCSB_TRY(so.init(ctx, hfun_old, hb, w, old_stride, 4, STAGE_IN, rb));
        OCSB_TRY(sn.init(ctx, hfun_new, hb, w, new_stride, 4, STAGE_OUT, rb));
    }
    LimiterArgs a{dx, dy, hmin, hmax, lower_bound, upper_bound, ghmin, ghmax,
                  fuse_clamp};
    const float *oldp = alias ? (const float *)sn.dev : (const float *)so.dev;
    int64_t olds = alias ? sn.dev_stride : so.dev_stride;
    OCSB_TRY(k_flow_limiter(ctx, w, h, rb, re, sz.dev, dem_dtype, sz.dev_stride, oldp,
                            olds, (float *)sn.dev, sn.dev_stride, a));
    return finish_call(ctx, {&sz, &so, &sn});
}

// ---------------------------------------------------------------------------
// next-2: add_constant_value, Topo / Courant constraints (+ _apply_rate)
// ---------------------------------------------------------------------------
namespace {
inline char *row_ptr(void *base, int64_t row, int64_t stride, size_t elem) {
    return base ? (char *)base + (size_t)row * (size_t)stride * elem : nullptr;
}
}  // namespace

extern "C" int32_t ocsb_hfun_constant_value(
    ocsb_ctx *ctx, const ocsb_win *win, const void *dem, int32_t dem_dtype,
    int64_t dem_stride, const float *hfun_old, int64_t old_stride,
    float *hfun_new, int64_t new_stride, double value, double lower_bound,
    double upper_bound, double ghmin, double ghmax, int32_t fuse_clamp) {
    CTX_ENTER("ocsb_hfun_constant_value");
    OCSB_TRY(check_win(win, "ocsb_hfun_constant_value"));
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "ocsb_hfun_constant_value: unsupported DEM dtype %d", dem_dtype);
    OCSB_REQUIRE(dem && hfun_new && dem_stride >= win->width &&
                     (!hfun_old || old_stride >= win->width) &&
                     new_stride >= win->width,
                 "ocsb_hfun_constant_value: bad buffers");
    OCSB_REQUIRE(!ocsb_is_none(value) && !ocsb_is_none(lower_bound) &&
                     !ocsb_is_none(upper_bound),
                 "ocsb_hfun_constant_value: NaN argument");
    const int w = win->width;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_hfun_constant_value", &rb, &re));
    const int hb = re - rb;
    const size_t es = dem_dtype == OCSB_F32 ? 4 : 8;
    Staged sz, so, sn;
    OCSB_TRY(sz.init(ctx, dem, hb, w, dem_stride, es, STAGE_IN, rb));
    const bool alias = (const void *)hfun_old == (const void *)hfun_new &&
                       old_stride == new_stride;
    if (alias) {
        OCSB_TRY(sn.init(ctx, hfun_new, hb, w, new_stride, 4, STAGE_INOUT, rb));
    } else {
        if (hfun_old)
            OCSB_TRY(so.init(ctx, hfun_old, hb, w, old_stride, 4, STAGE_IN, rb));
        OCSB_TRY(sn.init(ctx, hfun_new, hb, w, new_stride, 4, STAGE_OUT, rb));
    }
    void *oldb = alias ? sn.dev : so.dev;
    const int64_t olds = alias ? sn.dev_stride : so.dev_stride;
    OCSB_TRY(k_constant_value(
        ctx, w, hb, row_ptr(sz.dev, rb, sz.dev_stride, es), dem_dtype,
        sz.dev_stride, (const float *)row_ptr(oldb, rb, olds, 4), olds,
        (float *)row_ptr(sn.dev, rb, sn.dev_stride, 4), sn.dev_stride, value,
        lower_bound, upper_bound, ghmin, ghmax, fuse_clamp));
    return finish_call(ctx, {&sz, &so, &sn});
}

extern "C" int32_t ocsb_hfun_topo_constraint(
    ocsb_ctx *ctx, const ocsb_win *win, const void *dem, int32_t dem_dtype,
    int64_t dem_stride, float *hfun, int64_t hfun_stride, const void *values,
    int32_t values_dtype, int64_t values_stride, double value,
    double lower_bound, double upper_bound, int32_t value_type, double rate) {
    CTX_ENTER("ocsb_hfun_topo_constraint");
    OCSB_TRY(check_win(win, "ocsb_hfun_topo_constraint"));
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "ocsb_hfun_topo_constraint: unsupported DEM dtype %d", dem_dtype);
    OCSB_REQUIRE(value_type == OCSB_VT_MIN || value_type == OCSB_VT_MAX,
                 "Invalid value type for constraint!");
    OCSB_REQUIRE(dem && hfun && dem_stride >= win->width &&
                     hfun_stride >= win->width,
                 "ocsb_hfun_topo_constraint: bad buffers");
    OCSB_REQUIRE(!values || ((values_dtype == OCSB_F32 || values_dtype == OCSB_F64) &&
                             values_stride >= win->width),
                 "ocsb_hfun_topo_constraint: bad values array");
    OCSB_REQUIRE(values || !ocsb_is_none(value),
                 "ocsb_hfun_topo_constraint: value is NaN");
    OCSB_REQUIRE(!ocsb_is_none(lower_bound) && !ocsb_is_none(upper_bound),
                 "ocsb_hfun_topo_constraint: NaN bound");
    const int w = win->width, h = win->height;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_hfun_topo_constraint", &rb, &re));
    const int hb = re - rb;
    const bool with_rate = !ocsb_is_none(rate);
    // _apply_rate looks for the nearest mask pixel of the WHOLE window
    OCSB_REQUIRE(!with_rate || (rb == 0 && re == h),
                 "ocsb_hfun_topo_constraint: a rate needs the whole window on "
                 "one device (row band [%d, %d) of %d rows given)", rb, re, h);
    const size_t es = dem_dtype == OCSB_F32 ? 4 : 8;
    const size_t vs = values_dtype == OCSB_F64 ? 8 : 4;
    Staged sz, sh, sv;
    OCSB_TRY(sz.init(ctx, dem, hb, w, dem_stride, es, STAGE_IN, rb));
    OCSB_TRY(sh.init(ctx, hfun, hb, w, hfun_stride, 4, STAGE_INOUT, rb));
    if (values)
        OCSB_TRY(sv.init(ctx, values, hb, w, values_stride, vs, STAGE_IN, rb));
    ScratchGuard sg(ctx);
    uint8_t *mask = nullptr;
    if (with_rate) OCSB_TRY(sg.get(&mask, (size_t)w * hb));
    const int is_min = value_type == OCSB_VT_MIN;
    float *hb_ptr = (float *)row_ptr(sh.dev, rb, sh.dev_stride, 4);
    OCSB_TRY(k_topo_mask(ctx, w, hb, row_ptr(sz.dev, rb, sz.dev_stride, es),
                         dem_dtype, sz.dev_stride, hb_ptr, sh.dev_stride,
                         row_ptr(sv.dev, rb, sv.dev_stride, vs), values_dtype,
                         sv.dev_stride, value, lower_bound, upper_bound, is_min,
                         mask, nullptr));
    if (with_rate) {
        LinAxis ax, ay;
        win_axes(win, &ax, &ay);
        double *verts = nullptr;
        float *vals = nullptr;
        int64_t nvb = 0;
        OCSB_TRY(k_mask_boundary(ctx, sg, mask, w, h, ax, ay, hb_ptr,
                                 sh.dev_stride, &verts, &vals, &nvb));
        // `if not np.any(mask): return values` (constraint.py:87-88); a mask
        // covering the whole window has no boundary and no pixel to relax
        if (nvb > 0) {
            FeatureArgs fa;
            fa.ax = ax; fa.ay = ay;
            fa.rate_target = 0.0; fa.target = 1.0;
            fa.has_hmin = fa.has_hmax = 0;
            fa.hmin_f = fa.hmax_f = 0.f;
            fa.fuse_clamp = 0;
            fa.r = std::numeric_limits<double>::infinity();
            fa.row_begin = 0; fa.row_end = h;
            fa.rate_vals = vals;
            fa.rate_mask = mask;
            fa.rate = rate;
            fa.rate_sign = is_min ? -1.0 : 1.0;
            fa.rate_is_min = is_min;
            OCSB_TRY(k_hfun_feature(ctx, fa, verts, nvb, hb_ptr, sh.dev_stride,
                                    hb_ptr, sh.dev_stride, nullptr, nullptr));
        }
    }
    return finish_call(ctx, {&sz, &sh, &sv});
}

// ---- the same constraint with a rate on a ROW BAND of the window (one rank of
// a multi-GPU run): the nearest mask pixel of a band pixel may live in another
// rank's band, so the call is split around an all-gather of the ranks'
// mask-boundary pixels (ocsmesh_b200/distributed.py:all_gather_vertices) -----
extern "C" int32_t ocsb_hfun_topo_mask_band(
    ocsb_ctx *ctx, const ocsb_win *win, const void *dem, int32_t dem_dtype,
    int64_t dem_stride, float *hfun, int64_t hfun_stride, const void *values,
    int32_t values_dtype, int64_t values_stride, double value,
    double lower_bound, double upper_bound, int32_t value_type,
    uint8_t *mask_out, double *verts_out, float *vals_out, int64_t cap,
    int64_t *n_out) {
    CTX_ENTER("ocsb_hfun_topo_mask_band");
    OCSB_TRY(check_win(win, "ocsb_hfun_topo_mask_band"));
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "ocsb_hfun_topo_mask_band: unsupported DEM dtype %d", dem_dtype);
    OCSB_REQUIRE(value_type == OCSB_VT_MIN || value_type == OCSB_VT_MAX,
                 "Invalid value type for constraint!");
    OCSB_REQUIRE(dem && hfun && mask_out && n_out && dem_stride >= win->width &&
                     hfun_stride >= win->width && cap >= 0,
                 "ocsb_hfun_topo_mask_band: bad buffers");
    OCSB_REQUIRE(!values || ((values_dtype == OCSB_F32 || values_dtype == OCSB_F64) &&
                             values_stride >= win->width),
                 "ocsb_hfun_topo_mask_band: bad values array");
    OCSB_REQUIRE(values || !ocsb_is_none(value),
                 "ocsb_hfun_topo_mask_band: value is NaN");
    OCSB_REQUIRE(!ocsb_is_none(lower_bound) && !ocsb_is_none(upper_bound),
                 "ocsb_hfun_topo_mask_band: NaN bound");
    OCSB_REQUIRE(ocsb_is_device_ptr(mask_out),
                 "ocsb_hfun_topo_mask_band: mask_out must be device memory");
    const int w = win->width, h = win->height;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_hfun_topo_mask_band", &rb, &re));
    const int hb = re - rb;
    const size_t es = dem_dtype == OCSB_F32 ? 4 : 8;
    const size_t vs = values_dtype == OCSB_F64 ? 8 : 4;
    Staged sz, sh, sv;
    OCSB_TRY(sz.init(ctx, dem, hb, w, dem_stride, es, STAGE_IN, rb));
    OCSB_TRY(sh.init(ctx, hfun, hb, w, hfun_stride, 4, STAGE_INOUT, rb));
    if (values)
        OCSB_TRY(sv.init(ctx, values, hb, w, values_stride, vs, STAGE_IN, rb));
    const int is_min = value_type == OCSB_VT_MIN;
    float *hb_ptr = (float *)row_ptr(sh.dev, rb, sh.dev_stride, 4);
    // idempotent: pixels set to `val` fail satisfies(val, val) again, so the
    // count call and the emit call produce the same mask
    OCSB_TRY(k_topo_mask(ctx, w, hb, row_ptr(sz.dev, rb, sz.dev_stride, es),
                         dem_dtype, sz.dev_stride, hb_ptr, sh.dev_stride,
                         row_ptr(sv.dev, rb, sv.dev_stride, vs), values_dtype,
                         sv.dev_stride, value, lower_bound, upper_bound, is_min,
                         mask_out, nullptr));
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    ScratchGuard sg(ctx);
    double *verts = nullptr;
    float *vals = nullptr;
    int64_t nvb = 0;
    OCSB_TRY(k_mask_boundary(ctx, sg, mask_out, w, hb, ax, ay, hb_ptr, sh.dev_stride,
                             &verts, &vals, &nvb, rb, rb > 0, re < h));
    *n_out = nvb;
    if (verts_out && vals_out && nvb > 0) {
        if (nvb > cap) {
            OCSB_TRY(finish_call(ctx, {&sz, &sh, &sv}));
            ocsb_set_error("ocsb_hfun_topo_mask_band: %lld boundary pixels, "
                           "capacity %lld", (long long)nvb, (long long)cap);
            return OCSB_EOVERFLOW;
        }
        const cudaMemcpyKind kv = ocsb_is_device_ptr(verts_out)
                                      ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
        const cudaMemcpyKind kl = ocsb_is_device_ptr(vals_out)
                                      ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
        OCSB_CUDA(cudaMemcpyAsync(verts_out, verts, (size_t)nvb * 16, kv, ctx->stream));
        OCSB_CUDA(cudaMemcpyAsync(vals_out, vals, (size_t)nvb * 4, kl, ctx->stream));
        OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return finish_call(ctx, {&sz, &sh, &sv});
}

extern "C" int32_t ocsb_hfun_apply_rate(
    ocsb_ctx *ctx, const ocsb_win *win, float *hfun, int64_t hfun_stride,
    const uint8_t *mask, const double *verts_xy, const float *vert_values,
    int64_t nv, double rate, int32_t value_type) {
    CTX_ENTER("ocsb_hfun_apply_rate");
    OCSB_TRY(check_win(win, "ocsb_hfun_apply_rate"));
    OCSB_REQUIRE(value_type == OCSB_VT_MIN || value_type == OCSB_VT_MAX,
                 "Invalid value type for constraint!");
    OCSB_REQUIRE(nv >= 0 && nv < (int64_t)2147483647, "ocsb_hfun_apply_rate: bad nv");
    // `if not np.any(mask): return values` (constraint.py:87-88)
    if (nv == 0) return OCSB_OK;
    OCSB_REQUIRE(hfun && mask && verts_xy && vert_values &&
                     hfun_stride >= win->width && !ocsb_is_none(rate),
                 "ocsb_hfun_apply_rate: bad arguments");
    OCSB_REQUIRE(ocsb_is_device_ptr(mask) && ocsb_is_device_ptr(hfun),
                 "ocsb_hfun_apply_rate: hfun and mask must be device memory");
    const int w = win->width;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_hfun_apply_rate", &rb, &re));
    Staged sv, sl;
    OCSB_TRY(sv.init(ctx, verts_xy, 1, nv * 2, nv * 2, 8, STAGE_IN));
    OCSB_TRY(sl.init(ctx, vert_values, 1, nv, nv, 4, STAGE_IN));
    const int is_min = value_type == OCSB_VT_MIN;
    FeatureArgs fa;
    win_axes(win, &fa.ax, &fa.ay);
    fa.rate_target = 0.0; fa.target = 1.0;
    fa.has_hmin = fa.has_hmax = 0;
    fa.hmin_f = fa.hmax_f = 0.f;
    fa.fuse_clamp = 0;
    fa.r = std::numeric_limits<double>::infinity();
    fa.row_begin = rb; fa.row_end = re;
    fa.rate_vals = (const float *)sl.dev;
    // both indexed by WINDOW row: hand over the (virtual) row-0 addresses
    fa.rate_mask = mask - (int64_t)rb * w;
    fa.rate = rate;
    fa.rate_sign = is_min ? -1.0 : 1.0;
    fa.rate_is_min = is_min;
    // hfun is the caller's (virtual) window-row-0 pointer already
    OCSB_TRY(k_hfun_feature(ctx, fa, (const double *)sv.dev, nv, hfun, hfun_stride,
                            hfun, hfun_stride, nullptr, nullptr));
    return finish_call(ctx, {&sv, &sl});
}

extern "C" int32_t ocsb_hfun_courant_constraint(
    ocsb_ctx *ctx, const ocsb_win *win, const void *dem, int32_t dem_dtype,
    int64_t dem_stride, float *hfun, int64_t hfun_stride, double courant,
    double timestep, double wave_amplitude, int32_t value_type) {
    CTX_ENTER("ocsb_hfun_courant_constraint");
    OCSB_TRY(check_win(win, "ocsb_hfun_courant_constraint"));
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "ocsb_hfun_courant_constraint: unsupported DEM dtype %d", dem_dtype);
    OCSB_REQUIRE(value_type == OCSB_VT_MIN || value_type == OCSB_VT_MAX,
                 "Invalid value type for constraint!");
    OCSB_REQUIRE(dem && hfun && dem_stride >= win->width &&
                     hfun_stride >= win->width,
                 "ocsb_hfun_courant_constraint: bad buffers");
    OCSB_REQUIRE(!ocsb_is_none(courant) && !ocsb_is_none(timestep) &&
                     !ocsb_is_none(wave_amplitude),
                 "ocsb_hfun_courant_constraint: NaN argument");
    const int w = win->width;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_hfun_courant_constraint", &rb, &re));
    const int hb = re - rb;
    const size_t es = dem_dtype == OCSB_F32 ? 4 : 8;
    Staged sz, sh;
    OCSB_TRY(sz.init(ctx, dem, hb, w, dem_stride, es, STAGE_IN, rb));
    OCSB_TRY(sh.init(ctx, hfun, hb, w, hfun_stride, 4, STAGE_INOUT, rb));
    OCSB_TRY(k_courant(ctx, w, hb, row_ptr(sz.dev, rb, sz.dev_stride, es),
                       dem_dtype, sz.dev_stride,
                       (float *)row_ptr(sh.dev, rb, sh.dev_stride, 4),
                       sh.dev_stride, courant, timestep, wave_amplitude,
                       value_type == OCSB_VT_MIN));
    return finish_call(ctx, {&sz, &sh});
}

// ---------------------------------------------------------------------------
// next-1: contour segments
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_contour_segments(ocsb_ctx *ctx, const ocsb_win *win,
                                         const void *dem, int32_t dem_dtype,
                                         int64_t dem_stride, double level,
                                         int64_t cap, int64_t *edges,
                                         double *pts, int64_t *n) {
    CTX_ENTER("ocsb_contour_segments");
    OCSB_TRY(check_win(win, "ocsb_contour_segments"));
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "ocsb_contour_segments: unsupported DEM dtype %d", dem_dtype);
    OCSB_REQUIRE(dem && n && dem_stride >= win->width && cap >= 0,
                 "ocsb_contour_segments: bad buffers");
    OCSB_REQUIRE(!ocsb_is_none(level), "ocsb_contour_segments: level is NaN");
    const int w = win->width, h = win->height;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_contour_segments", &rb, &re));
    *n = 0;
    const int je = re < h - 1 ? re : h - 1;   // cell rows [rb, je)
    if (w < 2 || je <= rb) return OCSB_OK;
    const bool emit = edges && pts && cap > 0;
    const size_t es = dem_dtype == OCSB_F32 ? 4 : 8;
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    Staged sz, se, sp;
    OCSB_TRY(sz.init(ctx, dem, je + 1 - rb, w, dem_stride, es, STAGE_IN, rb));
    if (emit) {
        OCSB_TRY(se.init(ctx, edges, 1, cap * 2, cap * 2, 8, STAGE_OUT));
        OCSB_TRY(sp.init(ctx, pts, 1, cap * 4, cap * 4, 8, STAGE_OUT));
    }
    OCSB_TRY(k_contour_segments(ctx, sz.dev, dem_dtype, sz.dev_stride, w, rb, je,
                                level, ax, ay, emit ? cap : 0,
                                emit ? (int64_t *)se.dev : nullptr,
                                emit ? (double *)sp.dev : nullptr, n));
    // copy back only the segments that exist, not the whole capacity
    se.width = *n * 2;
    sp.width = *n * 4;
    return finish_call(ctx, {&sz, &se, &sp});
}

// ---------------------------------------------------------------------------
// multi-GPU: the blocks of the per-step band exchange (device memory only)
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_band_pack(ocsb_ctx *ctx, const double *verts_xy, int64_t n, int64_t cap,
                                  const float *row_first, const float *row_last, int32_t width,
                                  void *send_block, double far_coord) {
    CTX_ENTER("ocsb_band_pack");
    OCSB_REQUIRE(n >= 0 && cap >= n && cap >= 1 && width >= 1 && send_block && row_first &&
                     row_last && (n == 0 || verts_xy),
                 "ocsb_band_pack: bad arguments");
    OCSB_REQUIRE(ocsb_is_device_ptr(send_block) && ocsb_is_device_ptr(row_first),
                 "ocsb_band_pack: device memory only");
    const int64_t row_bytes = ((2ll * width * 4) + 15) / 16 * 16;
    return k_band_pack(ctx, verts_xy, n, cap, row_first, row_last, width, row_bytes, send_block,
                       far_coord);
}

extern "C" int32_t ocsb_band_unpack(ocsb_ctx *ctx, const void *recv_blocks, int32_t world,
                                    int32_t rank, int64_t cap, int32_t width, double *verts_out,
                                    float *halo_above, float *halo_below) {
    CTX_ENTER("ocsb_band_unpack");
    OCSB_REQUIRE(recv_blocks && world >= 1 && rank >= 0 && rank < world && cap >= 1 &&
                     width >= 1 && verts_out,
                 "ocsb_band_unpack: bad arguments");
    OCSB_REQUIRE((!halo_above || rank > 0) && (!halo_below || rank < world - 1),
                 "ocsb_band_unpack: no neighbour on that side");
    const int64_t row_bytes = ((2ll * width * 4) + 15) / 16 * 16;
    return k_band_unpack(ctx, recv_blocks, world, rank, row_bytes + cap * 16, row_bytes, cap,
                         width, verts_out, halo_above, halo_below);
}

// ---------------------------------------------------------------------------
// next-2: polygon regions (RegionConstraint / add_patch masks)
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_region_mask(ocsb_ctx *ctx, const ocsb_win *win, const double *edges,
                                    int64_t n_edges, int32_t all_touched, float *mask_out,
                                    int64_t mask_stride) {
    CTX_ENTER("ocsb_region_mask");
    OCSB_TRY(check_win(win, "ocsb_region_mask"));
    OCSB_REQUIRE(n_edges >= 0 && (n_edges == 0 || edges), "ocsb_region_mask: bad edge list");
    OCSB_REQUIRE(mask_out && mask_stride >= win->width, "ocsb_region_mask: bad mask buffer");
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_region_mask", &rb, &re));
    const int w = win->width, h = win->height;
    Staged se, sm;
    OCSB_TRY(se.init(ctx, edges, 1, n_edges * 4, n_edges * 4, 8, STAGE_IN));
    OCSB_TRY(sm.init(ctx, mask_out, re - rb, w, mask_stride, 4, STAGE_OUT, rb));
    // bounding box of the rings (host copy of the edge list: small)
    std::vector<double> eh((size_t)n_edges * 4);
    if (n_edges)
        OCSB_CUDA(cudaMemcpyAsync(eh.data(), se.dev, eh.size() * 8, cudaMemcpyDeviceToHost,
                                  ctx->stream));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    double bbox[4] = {INFINITY, INFINITY, -INFINITY, -INFINITY};
    for (int64_t i = 0; i < n_edges; ++i) {
        for (int k = 0; k < 2; ++k) {
            const double x = eh[4 * i + 2 * k], y = eh[4 * i + 2 * k + 1];
            OCSB_REQUIRE(std::isfinite(x) && std::isfinite(y),
                         "ocsb_region_mask: non-finite ring coordinate");
            bbox[0] = fmin(bbox[0], x); bbox[2] = fmax(bbox[2], x);
            bbox[1] = fmin(bbox[1], y); bbox[3] = fmax(bbox[3], y);
        }
    }
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    const double ylo = fmin(win->y0, win->y1), yhi = fmax(win->y0, win->y1);
    OCSB_TRY(k_region_mask(ctx, ax, ay, win->x0, yhi, (win->x1 - win->x0) / w,
                           (yhi - ylo) / h, rb, re, w, all_touched ? 1 : 0,
                           (const double *)se.dev, n_edges, bbox, (float *)sm.dev,
                           sm.dev_stride));
    return finish_call(ctx, {&se, &sm});
}

// ---------------------------------------------------------------------------
// next-1, device side: polylines and the resampled vertex array
// ---------------------------------------------------------------------------
namespace {

int32_t keep_result(ocsb_ctx *ctx, int slot, const void *dev, size_t bytes) {
    // (stream-ordered pool: cudaMalloc / cudaFree would synchronise the device)
    if (ctx->kept[slot]) { ocsb_free(ctx, ctx->kept[slot]); ctx->kept[slot] = nullptr; }
    ctx->kept_bytes[slot] = bytes;
    if (bytes == 0) return OCSB_OK;
    OCSB_TRY(ocsb_alloc(ctx, &ctx->kept[slot], bytes));
    OCSB_CUDA(cudaMemcpyAsync(ctx->kept[slot], dev, bytes, cudaMemcpyDeviceToDevice,
                              ctx->stream));
    return OCSB_OK;
}

// segments of one level on the device (carved from sg): one count + emit call
// when the capacity guess holds, a second call with the exact size otherwise
int32_t level_segments(ocsb_ctx *ctx, ScratchGuard &sg, const ocsb_win *win, const void *dem,
                       int dem_dtype, int64_t dem_stride, double level, int64_t **edges,
                       double **pts, int64_t *n) {
    const int w = win->width, h = win->height;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_contour", &rb, &re));
    *n = 0; *edges = nullptr; *pts = nullptr;
    const int je = re < h - 1 ? re : h - 1;   // cell rows [rb, je)
    if (w < 2 || je <= rb) return OCSB_OK;
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    int64_t cap = 64ll * (w + (je - rb)) + 4096;
    for (int attempt = 0; attempt < 2; ++attempt) {
        OCSB_TRY(sg.get(edges, (size_t)cap * 2));
        OCSB_TRY(sg.get(pts, (size_t)cap * 4));
        const int32_t rc = k_contour_segments(ctx, dem, dem_dtype, dem_stride, w, rb, je, level,
                                              ax, ay, cap, *edges, *pts, n);
        if (rc == OCSB_EOVERFLOW && *n > cap && attempt == 0) { cap = *n; continue; }
        return rc;
    }
    return OCSB_OK;
}

int32_t check_contour_args(const ocsb_win *win, const void *dem, int32_t dem_dtype,
                           int64_t dem_stride, const char *who) {
    OCSB_TRY(check_win(win, who));
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "%s: unsupported DEM dtype %d", who, dem_dtype);
    OCSB_REQUIRE(dem && dem_stride >= win->width, "%s: bad buffers", who);
    OCSB_REQUIRE(ocsb_is_device_ptr(dem), "%s: the DEM must be device memory", who);
    return OCSB_OK;
}

}  // namespace

extern "C" int32_t ocsb_contour_lines(ocsb_ctx *ctx, const ocsb_win *win, const void *dem,
                                      int32_t dem_dtype, int64_t dem_stride, double level,
                                      int64_t *n_xy, int64_t *n_lines) {
    CTX_ENTER("ocsb_contour_lines");
    OCSB_TRY(check_contour_args(win, dem, dem_dtype, dem_stride, "ocsb_contour_lines"));
    OCSB_REQUIRE(n_xy && n_lines && !ocsb_is_none(level), "ocsb_contour_lines: bad arguments");
    ScratchGuard sg(ctx);
    int64_t *edges = nullptr, n = 0, *off = nullptr;
    double *pts = nullptr, *xy = nullptr;
    OCSB_TRY(level_segments(ctx, sg, win, dem, dem_dtype, dem_stride, level, &edges, &pts, &n));
    OCSB_TRY(k_stitch_segments(ctx, sg, edges, pts, n, &xy, &off, n_xy, n_lines));
    OCSB_TRY(keep_result(ctx, 0, xy, (size_t)*n_xy * 16));
    OCSB_TRY(keep_result(ctx, 1, off, *n_lines ? (size_t)(*n_lines + 1) * 8 : 0));
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    return OCSB_OK;
}

extern "C" int32_t ocsb_contour_points(ocsb_ctx *ctx, const ocsb_win *win, const void *dem,
                                       int32_t dem_dtype, int64_t dem_stride,
                                       const double *levels, int32_t n_levels,
                                       double target_size, int32_t max_verts,
                                       int64_t *n_points) {
    CTX_ENTER("ocsb_contour_points");
    OCSB_TRY(check_contour_args(win, dem, dem_dtype, dem_stride, "ocsb_contour_points"));
    OCSB_REQUIRE(levels && n_levels > 0 && n_points, "ocsb_contour_points: bad arguments");
    OCSB_REQUIRE(target_size > 0 && target_size == target_size,
                 "Argument target_size must be greater than zero.");
    OCSB_REQUIRE(max_verts >= 2, "ocsb_contour_points: max_verts must be >= 2");
    ScratchGuard sg(ctx);
    std::vector<double *> parts;
    std::vector<int64_t> counts;
    int64_t total = 0;
    for (int l = 0; l < n_levels; ++l)
        OCSB_REQUIRE(!ocsb_is_none(levels[l]), "ocsb_contour_points: level is NaN");
    // a level list: ONE count sweep of the DEM for all levels (the O(pixels)
    // part), then an emit launch per level that only visits chunks with segments
    const int w_ = win->width, h_ = win->height;
    int rb_ = 0, re_ = 0;
    OCSB_TRY(win_band(win, "ocsb_contour", &rb_, &re_));
    const int je_ = re_ < h_ - 1 ? re_ : h_ - 1;
    const bool multi = n_levels >= 2 && n_levels <= k_contour_max_levels() && w_ >= 2 &&
                       je_ > rb_;
    int *lvl_offsets = nullptr;
    unsigned long long *lvl_totals_dev = nullptr;
    int64_t lvl_totals[8] = {}, lvl_nblk = 0;
    LinAxis cax, cay;
    win_axes(win, &cax, &cay);
    if (multi)
        OCSB_TRY(k_contour_count_levels(ctx, sg, dem, dem_dtype, dem_stride, w_, rb_, je_, levels,
                                        n_levels, &lvl_offsets, &lvl_totals_dev, lvl_totals,
                                        &lvl_nblk));
    for (int l = 0; l < n_levels; ++l) {
        int64_t *edges = nullptr, n = 0, *off = nullptr, n_xy = 0, n_lines = 0, np = 0;
        double *pts = nullptr, *xy = nullptr, *rs = nullptr;
        if (multi) {
            n = lvl_totals[l];
            if (n > 0) {
                OCSB_TRY(sg.get(&edges, (size_t)n * 2));
                OCSB_TRY(sg.get(&pts, (size_t)n * 4));
                OCSB_TRY(k_contour_emit_level(ctx, dem, dem_dtype, dem_stride, w_, rb_, je_,
                                              levels[l], cax, cay, lvl_offsets + l * lvl_nblk,
                                              lvl_totals_dev + l, edges, pts));
            }
        } else
        OCSB_TRY(level_segments(ctx, sg, win, dem, dem_dtype, dem_stride, levels[l], &edges,
                                &pts, &n));
        OCSB_TRY(k_stitch_segments(ctx, sg, edges, pts, n, &xy, &off, &n_xy, &n_lines));
        OCSB_TRY(k_resample_lines(ctx, sg, xy, off, n_lines, target_size, max_verts, &rs, &np));
        parts.push_back(rs);
        counts.push_back(np);
        total += np;
    }
    *n_points = total;
    if (ctx->kept[0]) { ocsb_free(ctx, ctx->kept[0]); ctx->kept[0] = nullptr; }
    ctx->kept_bytes[0] = (size_t)total * 16;
    if (total) {
        OCSB_TRY(ocsb_alloc(ctx, &ctx->kept[0], (size_t)total * 16));
        size_t at = 0;
        for (size_t i = 0; i < parts.size(); ++i) {
            if (counts[i])
                OCSB_CUDA(cudaMemcpyAsync((char *)ctx->kept[0] + at, parts[i],
                                          (size_t)counts[i] * 16, cudaMemcpyDeviceToDevice,
                                          ctx->stream));
            at += (size_t)counts[i] * 16;
        }
    }
    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    return OCSB_OK;
}

extern "C" int32_t ocsb_take_result(ocsb_ctx *ctx, int32_t slot, void *dst, int64_t dst_bytes) {
    CTX_ENTER("ocsb_take_result");
    OCSB_REQUIRE(slot == 0 || slot == 1, "ocsb_take_result: slot must be 0 or 1");
    OCSB_REQUIRE(dst_bytes >= 0 && (size_t)dst_bytes >= ctx->kept_bytes[slot],
                 "ocsb_take_result: the kept result has %lld bytes, dst holds %lld",
                 (long long)ctx->kept_bytes[slot], (long long)dst_bytes);
    if (ctx->kept_bytes[slot]) {
        OCSB_REQUIRE(dst, "ocsb_take_result: NULL destination");
        OCSB_CUDA(cudaMemcpyAsync(dst, ctx->kept[slot], ctx->kept_bytes[slot],
                                  cudaMemcpyDefault, ctx->stream));
        OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
        ocsb_free(ctx, ctx->kept[slot]);
        ctx->kept[slot] = nullptr;
        ctx->kept_bytes[slot] = 0;
    }
    return OCSB_OK;
}

// ---------------------------------------------------------------------------
// next-3: msh_t sampling side
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_hfun_sample_source(ocsb_ctx *ctx, const ocsb_win *win,
                                           const float *hfun, int64_t hfun_stride,
                                           int32_t stride, double hmin, double hmax,
                                           double *coords_out, double *values_out) {
    CTX_ENTER("ocsb_hfun_sample_source");
    OCSB_TRY(check_win(win, "ocsb_hfun_sample_source"));
    OCSB_REQUIRE(stride >= 1, "ocsb_hfun_sample_source: stride must be >= 1");
    OCSB_REQUIRE(hfun && hfun_stride >= win->width && (coords_out || values_out),
                 "ocsb_hfun_sample_source: bad buffers");
    const int w = win->width, h = win->height;
    const int64_t nx = (w + stride - 1) / stride, ny = (h + stride - 1) / stride;
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    Staged sh, sc, sv;
    OCSB_TRY(sh.init(ctx, hfun, h, w, hfun_stride, 4, STAGE_IN));
    OCSB_TRY(sc.init(ctx, coords_out, 1, nx * ny * 2, nx * ny * 2, 8, STAGE_OUT));
    OCSB_TRY(sv.init(ctx, values_out, 1, nx * ny, nx * ny, 8, STAGE_OUT));
    OCSB_REQUIRE(((uintptr_t)sc.dev & 15) == 0,
                 "ocsb_hfun_sample_source: coords_out must be 16-byte aligned");
    OCSB_TRY(k_sample_source(ctx, ax, ay, (const float *)sh.dev, sh.dev_stride,
                             stride, hmin, hmax, (double *)sc.dev, (double *)sv.dev));
    return finish_call(ctx, {&sh, &sc, &sv});
}

extern "C" int32_t ocsb_hfun_sample_nearest(ocsb_ctx *ctx, const ocsb_win *win,
                                            const float *hfun, int64_t hfun_stride,
                                            int32_t stride, double hmin, double hmax,
                                            const double *xy, int64_t K,
                                            double *values_out) {
    CTX_ENTER("ocsb_hfun_sample_nearest");
    OCSB_TRY(check_win(win, "ocsb_hfun_sample_nearest"));
    OCSB_REQUIRE(stride >= 1, "ocsb_hfun_sample_nearest: stride must be >= 1");
    OCSB_REQUIRE(K >= 0, "ocsb_hfun_sample_nearest: negative K");
    if (K == 0) return OCSB_OK;
    OCSB_REQUIRE(hfun && hfun_stride >= win->width && xy && values_out,
                 "ocsb_hfun_sample_nearest: bad buffers");
    const int w = win->width, h = win->height;
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    Staged sh, sx, sv;
    OCSB_TRY(sh.init(ctx, hfun, h, w, hfun_stride, 4, STAGE_IN));
    OCSB_TRY(sx.init(ctx, xy, 1, K * 2, K * 2, 8, STAGE_IN));
    OCSB_TRY(sv.init(ctx, values_out, 1, K, K, 8, STAGE_OUT));
    OCSB_REQUIRE(((uintptr_t)sx.dev & 15) == 0,
                 "ocsb_hfun_sample_nearest: xy must be 16-byte aligned");
    OCSB_TRY(k_sample_nearest(ctx, ax, ay, (const float *)sh.dev, sh.dev_stride,
                              stride, hmin, hmax, (const double *)sx.dev, K,
                              (double *)sv.dev));
    return finish_call(ctx, {&sh, &sx, &sv});
}

// ---------------------------------------------------------------------------
// next-4: nearest warp onto the fast collector's big raster
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_resample_nearest(
    ocsb_ctx *ctx, const void *src, int32_t src_dtype, int64_t src_stride, int32_t src_width,
    int32_t src_height, const double *src_transform, const uint8_t *src_mask,
    int64_t mask_stride, float *dst, int64_t dst_stride, int32_t dst_width,
    int32_t dst_height, const double *dst_transform) {
    CTX_ENTER("ocsb_resample_nearest");
    OCSB_REQUIRE(src_dtype == OCSB_F32 || src_dtype == OCSB_F64,
                 "ocsb_resample_nearest: unsupported source dtype %d", src_dtype);
    OCSB_REQUIRE(src && dst && src_transform && dst_transform && src_width > 0 &&
                     src_height > 0 && dst_width > 0 && dst_height > 0 &&
                     src_stride >= src_width && dst_stride >= dst_width &&
                     (!src_mask || mask_stride >= src_width),
                 "ocsb_resample_nearest: bad buffers");
    for (const double *t : {src_transform, dst_transform}) {
        OCSB_REQUIRE(t[1] == 0.0 && t[3] == 0.0,
                     "ocsb_resample_nearest: rotated / sheared rasters are not supported");
        OCSB_REQUIRE(t[0] != 0.0 && t[4] != 0.0 && std::isfinite(t[0]) &&
                         std::isfinite(t[4]) && std::isfinite(t[2]) && std::isfinite(t[5]),
                     "ocsb_resample_nearest: degenerate transform");
    }
    const int es = src_dtype == OCSB_F32 ? 4 : 8;
    Staged ss, sm, sd;
    OCSB_TRY(ss.init(ctx, src, src_height, src_width, src_stride, es, STAGE_IN));
    OCSB_TRY(sm.init(ctx, src_mask, src_height, src_width, mask_stride, 1, STAGE_IN));
    OCSB_TRY(sd.init(ctx, dst, dst_height, dst_width, dst_stride, 4, STAGE_INOUT));
    OCSB_TRY(k_warp_nearest(ctx, ss.dev, src_dtype, ss.dev_stride, src_width, src_height,
                            src_transform, (const uint8_t *)sm.dev, sm.dev_stride,
                            (float *)sd.dev, sd.dev_stride, dst_width, dst_height,
                            dst_transform));
    return finish_call(ctx, {&ss, &sm, &sd});
}

// ---------------------------------------------------------------------------
// A14 interpolate
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_spline_prefilter(ocsb_ctx *ctx, const ocsb_win *win,
                                         const void *dem, int32_t dem_dtype,
                                         int64_t dem_stride, double *coef) {
    CTX_ENTER("ocsb_spline_prefilter");
    OCSB_TRY(check_win(win, "ocsb_spline_prefilter"));
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "ocsb_spline_prefilter: unsupported DEM dtype %d", dem_dtype);
    OCSB_REQUIRE(win->width >= 4 && win->height >= 4,
                 "ocsb_spline_prefilter: cubic spline needs >= 4 points per "
                 "axis (m > k must hold)");
    OCSB_REQUIRE(dem && coef && dem_stride >= win->width,
                 "ocsb_spline_prefilter: bad buffers");
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    const int w = win->width, h = win->height;
    const size_t es = dem_dtype == OCSB_F32 ? 4 : 8;
    Staged sz, sc;
    OCSB_TRY(sz.init(ctx, dem, h, w, dem_stride, es, STAGE_IN));
    OCSB_TRY(sc.init(ctx, coef, h, w, w, 8, STAGE_OUT));
    OCSB_TRY(k_spline_prefilter(ctx, ax, ay, sz.dev, dem_dtype, sz.dev_stride,
                                (double *)sc.dev));
    return finish_call(ctx, {&sz, &sc});
}

extern "C" int32_t ocsb_interp(ocsb_ctx *ctx, const ocsb_win *win,
                               const void *dem, int32_t dem_dtype,
                               int64_t dem_stride, const uint8_t *nodata_mask,
                               int64_t mask_stride, const double *xy, int64_t K,
                               int32_t method, uint8_t *touched,
                               double *values) {
    CTX_ENTER("ocsb_interp");
    OCSB_TRY(check_win(win, "ocsb_interp"));
    OCSB_REQUIRE(method == OCSB_INTERP_LINEAR || method == OCSB_INTERP_NEAREST ||
                     method == OCSB_INTERP_SPLINE,
                 "Invalid value method specified <%d>!", method);
    OCSB_REQUIRE(dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64,
                 "ocsb_interp: unsupported DEM dtype %d", dem_dtype);
    OCSB_REQUIRE(K >= 0, "ocsb_interp: negative K");
    if (K == 0) return OCSB_OK;
    OCSB_REQUIRE(dem && xy && touched && values && dem_stride >= win->width,
                 "ocsb_interp: bad buffers");
    OCSB_REQUIRE(!nodata_mask || mask_stride >= win->width,
                 "ocsb_interp: bad mask stride");
    const int w = win->width, h = win->height;
    if (method == OCSB_INTERP_SPLINE) {
        OCSB_REQUIRE(w >= 4 && h >= 4,
                     "ocsb_interp: cubic spline needs >= 4 points per axis");
    } else {
        OCSB_REQUIRE(w >= 2 && h >= 2,
                     "ocsb_interp: linear/nearest need >= 2 points per axis");
    }
    OCSB_REQUIRE(win->x0 != win->x1 && win->y0 != win->y1,
                 "ocsb_interp: degenerate window bounds");
    LinAxis ax, ay;
    win_axes(win, &ax, &ay);
    const size_t es = dem_dtype == OCSB_F32 ? 4 : 8;
    Staged sz, sm, sxy, st, sval;
    OCSB_TRY(sz.init(ctx, dem, h, w, dem_stride, es, STAGE_IN));
    OCSB_TRY(sm.init(ctx, nodata_mask, h, w, mask_stride, 1, STAGE_IN));
    OCSB_TRY(sxy.init(ctx, xy, 1, K * 2, K * 2, 8, STAGE_IN));
    OCSB_TRY(st.init(ctx, touched, 1, K, K, 1, STAGE_INOUT));
    OCSB_TRY(sval.init(ctx, values, 1, K, K, 8, STAGE_INOUT));
    if (method == OCSB_INTERP_SPLINE) {
        // coefficients are cached per (dem pointer, shape) so a list of node
        // batches against one raster pays the prefilter once
        size_t need = (size_t)w * h * sizeof(double);
        bool hit = ctx->spl_key == dem && ctx->spl_bytes == need && !sz.staged;
        if (!hit) {
            if (ctx->spl_bytes < need || !ctx->spl_coef) {
                if (ctx->spl_coef) {
                    OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
                    cudaFree(ctx->spl_coef);
                    ctx->spl_coef = nullptr;
                }
                OCSB_CUDA(cudaMalloc((void **)&ctx->spl_coef, need));
            }
            ctx->spl_bytes = need;
            ctx->spl_key = nullptr;  // never trust the cache across calls:
                                     // the caller may have rewritten the DEM
            OCSB_TRY(k_spline_prefilter(ctx, ax, ay, sz.dev, dem_dtype,
                                        sz.dev_stride, ctx->spl_coef));
        }
        OCSB_TRY(k_spline_eval(ctx, ax, ay, ctx->spl_coef,
                               (const double *)sxy.dev, K, (uint8_t *)st.dev,
                               (double *)sval.dev));
    } else {
        OCSB_TRY(k_interp(ctx, ax, ay, sz.dev, dem_dtype, sz.dev_stride,
                          (const uint8_t *)sm.dev, sm.dev_stride,
                          (const double *)sxy.dev, K, method,
                          (uint8_t *)st.dev, (double *)sval.dev));
    }
    return finish_call(ctx, {&sz, &sm, &sxy, &st, &sval});
}

extern "C" int32_t ocsb_interp_multi(ocsb_ctx *ctx,
                                     const ocsb_interp_raster *rasters,
                                     int32_t n_rasters, const double *xy,
                                     int64_t K, int32_t method,
                                     uint8_t *touched, double *values,
                                     int32_t *source, const int32_t *perm) {
    CTX_ENTER("ocsb_interp_multi");
    OCSB_REQUIRE(method == OCSB_INTERP_LINEAR || method == OCSB_INTERP_NEAREST ||
                     method == OCSB_INTERP_SPLINE,
                 "Invalid value method specified <%d>!", method);
    OCSB_REQUIRE(n_rasters >= 0 && K >= 0, "ocsb_interp_multi: negative count");
    if (K == 0 || n_rasters == 0) return OCSB_OK;
    OCSB_REQUIRE(rasters && xy && touched && values,
                 "ocsb_interp_multi: NULL buffers");
    if (method == OCSB_INTERP_SPLINE) {
        // groups of consecutive rasters whose coefficient arrays (8 B/px) fit
        // a budget: prefilter each member, then ONE evaluation launch per
        // group in which every node takes the last member containing it;
        // later groups overwrite earlier ones, so list order still decides
        OCSB_REQUIRE(source == nullptr && perm == nullptr,
                     "ocsb_interp_multi: source / perm are not available for "
                     "spline");
        for (int32_t i = 0; i < n_rasters; ++i) {
            const ocsb_interp_raster &r = rasters[i];
            OCSB_TRY(check_win(&r.win, "ocsb_interp_multi"));
            OCSB_REQUIRE(r.dem_dtype == OCSB_F32 || r.dem_dtype == OCSB_F64,
                         "ocsb_interp_multi: unsupported DEM dtype %d", r.dem_dtype);
            OCSB_REQUIRE(r.win.width >= 4 && r.win.height >= 4,
                         "ocsb_interp: cubic spline needs >= 4 points per axis");
            OCSB_REQUIRE(r.dem && r.dem_stride >= r.win.width,
                         "ocsb_interp_multi: bad raster %d", i);
            OCSB_REQUIRE(r.win.x0 != r.win.x1 && r.win.y0 != r.win.y1,
                         "ocsb_interp_multi: degenerate window bounds");
        }
        Staged sxy, st, sval;
        OCSB_TRY(sxy.init(ctx, xy, 1, K * 2, K * 2, 8, STAGE_IN));
        OCSB_TRY(st.init(ctx, touched, 1, K, K, 1, STAGE_INOUT));
        OCSB_TRY(sval.init(ctx, values, 1, K, K, 8, STAGE_INOUT));
        const size_t budget = (size_t)16 << 30;   // coefficient bytes per group
        const int32_t max_group = 64;
        int32_t g0 = 0;
        while (g0 < n_rasters) {
            size_t bytes = 0;
            int32_t g1 = g0;
            while (g1 < n_rasters && g1 - g0 < max_group) {
                const size_t need = rasters[g1].spline_coef ? 0 :
                                    (size_t)rasters[g1].win.width *
                                    rasters[g1].win.height * sizeof(double);
                if (g1 > g0 && bytes + need > budget) break;
                bytes += need;
                ++g1;
            }
            {
                ScratchGuard sg(ctx);
                std::vector<Staged> sz(g1 - g0);
                std::vector<SplineEvalRaster> rs(g1 - g0);
                for (int32_t i = g0; i < g1; ++i) {
                    const ocsb_interp_raster &r = rasters[i];
                    const int w = r.win.width, h = r.win.height;
                    const size_t es = r.dem_dtype == OCSB_F32 ? 4 : 8;
                    SplineEvalRaster &e = rs[i - g0];
                    win_axes(&r.win, &e.ax, &e.ay);
                    if (r.spline_coef) {
                        // the caller kept this window's coefficients
                        OCSB_REQUIRE(ocsb_is_device_ptr(r.spline_coef),
                                     "ocsb_interp_multi: spline_coef of raster %d "
                                     "must be device memory", i);
                        e.coef = const_cast<double *>(r.spline_coef);
                        continue;
                    }
                    Staged &z = sz[i - g0];
                    OCSB_TRY(z.init(ctx, r.dem, h, w, r.dem_stride, es, STAGE_IN));
                    double *coef = nullptr;
                    OCSB_TRY(sg.get(&coef, (size_t)w * h));
                    OCSB_TRY(k_spline_prefilter(ctx, e.ax, e.ay, z.dev, r.dem_dtype,
                                                z.dev_stride, coef));
                    e.coef = coef;
                }
                OCSB_TRY(k_spline_eval_multi(ctx, rs.data(), g1 - g0,
                                             (const double *)sxy.dev, K,
                                             (uint8_t *)st.dev, (double *)sval.dev));
                // the coefficient arrays are released in stream order
            }
            g0 = g1;
        }
        OCSB_TRY(st.finish());
        OCSB_TRY(sval.finish());
        if (sxy.staged || st.staged || sval.staged)
            OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return OCSB_OK;
    }
    std::vector<Staged> sz(n_rasters), sm(n_rasters);
    std::vector<InterpRaster> rs(n_rasters);
    for (int32_t i = 0; i < n_rasters; ++i) {
        const ocsb_interp_raster &r = rasters[i];
        OCSB_TRY(check_win(&r.win, "ocsb_interp_multi"));
        OCSB_REQUIRE(r.dem_dtype == OCSB_F32 || r.dem_dtype == OCSB_F64,
                     "ocsb_interp_multi: unsupported DEM dtype %d", r.dem_dtype);
        const int w = r.win.width, h = r.win.height;
        OCSB_REQUIRE(w >= 2 && h >= 2 && r.dem && r.dem_stride >= w,
                     "ocsb_interp_multi: bad raster %d", i);
        OCSB_REQUIRE(!r.nodata_mask || r.mask_stride >= w,
                     "ocsb_interp_multi: bad mask stride of raster %d", i);
        OCSB_REQUIRE(r.win.x0 != r.win.x1 && r.win.y0 != r.win.y1,
                     "ocsb_interp_multi: degenerate window bounds");
        const size_t es = r.dem_dtype == OCSB_F32 ? 4 : 8;
        OCSB_TRY(sz[i].init(ctx, r.dem, h, w, r.dem_stride, es, STAGE_IN));
        OCSB_TRY(sm[i].init(ctx, r.nodata_mask, h, w, r.mask_stride, 1, STAGE_IN));
        win_axes(&r.win, &rs[i].ax, &rs[i].ay);
        rs[i].dem = sz[i].dev;
        rs[i].dem_dtype = r.dem_dtype;
        rs[i].dem_stride = sz[i].dev_stride;
        rs[i].mask = (const uint8_t *)sm[i].dev;
        rs[i].mask_stride = sm[i].dev_stride;
    }
    Staged sxy, st, sval, ssrc, sperm;
    OCSB_TRY(sperm.init(ctx, perm, 1, K, K, 4, STAGE_IN));
    OCSB_TRY(sxy.init(ctx, xy, 1, K * 2, K * 2, 8, STAGE_IN));
    OCSB_TRY(st.init(ctx, touched, 1, K, K, 1, STAGE_INOUT));
    OCSB_TRY(sval.init(ctx, values, 1, K, K, 8, STAGE_INOUT));
    OCSB_TRY(ssrc.init(ctx, source, 1, K, K, 4, STAGE_INOUT));
    OCSB_TRY(k_interp_multi(ctx, rs.data(), n_rasters, (const double *)sxy.dev, K,
                            method, (uint8_t *)st.dev, (double *)sval.dev,
                            (int32_t *)ssrc.dev, (const int32_t *)sperm.dev));
    bool any = sperm.staged || sxy.staged || st.staged || sval.staged || ssrc.staged;
    for (auto &s : sz) any = any || s.staged;
    for (auto &s : sm) any = any || s.staged;
    OCSB_TRY(st.finish());
    OCSB_TRY(sval.finish());
    OCSB_TRY(ssrc.finish());
    if (any) OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
    return OCSB_OK;
}

extern "C" int32_t ocsb_bucket_nodes(ocsb_ctx *ctx, const double *xy, int64_t K,
                                     const double *bbox4, double cell_size,
                                     int32_t *perm_out, double *xy_sorted_out) {
    CTX_ENTER("ocsb_bucket_nodes");
    OCSB_REQUIRE(K >= 0 && K < (int64_t)2147483647, "ocsb_bucket_nodes: bad K");
    if (K == 0) return OCSB_OK;
    OCSB_REQUIRE(xy && bbox4 && perm_out && xy_sorted_out,
                 "ocsb_bucket_nodes: NULL buffers");
    OCSB_REQUIRE(cell_size > 0 && isfinite(cell_size) && isfinite(bbox4[0]) &&
                     isfinite(bbox4[1]) && isfinite(bbox4[2]) && isfinite(bbox4[3]),
                 "ocsb_bucket_nodes: bad grid");
    Staged sxy, sp, ss;
    OCSB_TRY(sxy.init(ctx, xy, 1, K * 2, K * 2, 8, STAGE_IN));
    OCSB_TRY(sp.init(ctx, perm_out, 1, K, K, 4, STAGE_OUT));
    OCSB_TRY(ss.init(ctx, xy_sorted_out, 1, K * 2, K * 2, 8, STAGE_OUT));
    OCSB_TRY(k_bucket_nodes(ctx, (const double *)sxy.dev, K, bbox4, cell_size,
                            (int32_t *)sp.dev, (double *)ss.dev));
    return finish_call(ctx, {&sxy, &sp, &ss});
}

// ---------------------------------------------------------------------------
// A4 + A6..A10 as one streamed pass (see ocsb200.h)
// ---------------------------------------------------------------------------
namespace {

// one pixel buffer of the pass: used in place when it is device memory,
// otherwise streamed through two device band buffers
struct PassBuf {
    const char *user = nullptr;   // window row 0 (host or device)
    int64_t stride = 0;           // elements
    size_t elem = 4;
    bool host = false;
    char *band[OCSB_PASS_SETS] = {};      // compact (stride = w) band scratch
};

int32_t pass_streams(ocsb_ctx *ctx) {
    if (ctx->s_h2d) return OCSB_OK;
    OCSB_CUDA(cudaStreamCreateWithFlags(&ctx->s_h2d, cudaStreamNonBlocking));
    OCSB_CUDA(cudaStreamCreateWithFlags(&ctx->s_d2h, cudaStreamNonBlocking));
    for (int i = 0; i < OCSB_PASS_SETS; ++i) {
        OCSB_CUDA(cudaEventCreateWithFlags(&ctx->ev_in[i], cudaEventDisableTiming));
        OCSB_CUDA(cudaEventCreateWithFlags(&ctx->ev_comp[i], cudaEventDisableTiming));
        OCSB_CUDA(cudaEventCreateWithFlags(&ctx->ev_out[i], cudaEventDisableTiming));
    }
    OCSB_CUDA(cudaEventCreateWithFlags(&ctx->ev_alloc, cudaEventDisableTiming));
    return OCSB_OK;
}

// the op list on rows [a, b) of the window; pointers are (virtual) addresses
// of window row 0
int32_t pass_ops(ocsb_ctx *ctx, const ocsb_win *win, int a, int b,
                 const void *dem, int dem_dtype, int64_t dem_stride,
                 const float *in, int64_t in_stride, float *out,
                 int64_t out_stride, double dx, double dy, double ghmin,
                 double ghmax, const ocsb_hfun_op *ops, int n_ops,
                 const std::vector<const double *> &verts,
                 int *flags /* device, n_ops ints; nullptr = synchronous */) {
    // ---- one tiled sweep for the whole list (k_pass.cu) when it qualifies -----
    {
        static const bool tiled_on = [] {
            const char *e = getenv("OCSB_PASS_TILED");
            return !(e && atoi(e) == 0);
        }();
        const int w = win->width, h = win->height;
        PassWin pw;
        win_axes(win, &pw.ax, &pw.ay);
        pw.rb = a; pw.re = b;
        pw.keep_w = w; pw.keep_h = h;
        pw.hoff = 0;
        const int za = a > 0 ? a - 1 : 0, zb = b < h ? b + 1 : h;
        pw.dem_c = 0; pw.dem_r = -za;
        std::vector<PassFeat> pf;
        std::vector<PassLim> pl;
        for (int i = 0; i < n_ops; ++i) {
            const ocsb_hfun_op &op = ops[i];
            if (op.kind == OCSB_OP_FEATURE) {
                PassFeat f;
                f.verts = verts[i]; f.nv = op.nv;
                f.rate_target = op.expansion_rate * op.target_size;
                f.target = op.target_size;
                f.r = (!ocsb_is_none(ghmax) && ghmax != 0.0)
                          ? (ghmax - op.target_size) / f.rate_target
                          : std::numeric_limits<double>::infinity();
                f.fuse_clamp = op.fuse_clamp;
                pf.push_back(f);
            } else {
                pl.push_back(PassLim{op.hmin, op.hmax, op.lower_bound, op.upper_bound,
                                     op.fuse_clamp});
            }
        }
        PassDesc d;
        d.wins = &pw; d.n_wins = 1;
        if (!pl.empty()) {
            d.dem = (const char *)dem + (size_t)za * (size_t)dem_stride *
                                            (dem_dtype == OCSB_F64 ? 8 : 4);
            d.dem_dtype = dem_dtype; d.dem_stride = dem_stride;
            d.dem_rows = zb - za; d.dem_cols = w;
        }
        d.in = in; d.in_stride = in_stride;
        d.out = out; d.out_stride = out_stride;
        d.dx = dx; d.dy = dy; d.ghmin = ghmin; d.ghmax = ghmax;
        d.feats = pf.data(); d.nf = (int)pf.size();
        d.lims = pl.data(); d.nl = (int)pl.size();
        d.ext_overflow = flags;
        if (tiled_on && k_pass_supported(d)) return k_pass_tiled(ctx, d);
    }
    const float *oldp = in;
    int64_t olds = in_stride;
    // Resident lists that start with a feature op: every op is a monotone
    // clamp of min(old, f(pixel)), so with one fuse_clamp setting the order of
    // the ops does not change a single bit.  The vertex-independent ops
    // (limiters) CAN then run first, on the context's stream, while the first
    // feature op's candidate pyramid is refined on a second stream.
    int n_other = 0;
    bool same_fuse = true;
    for (int i = 0; i < n_ops; ++i) {
        if (ops[i].kind != OCSB_OP_FEATURE) ++n_other;
        same_fuse = same_fuse && ops[i].fuse_clamp == ops[0].fuse_clamp;
    }
    // Measured on B200 (profiles/r1f_overlap.txt): NOT a win -- the limiter
    // already fills the machine, so the refinement kernels are starved (0.27 ->
    // 1.07 ms) and the feature op loses its "pre-op raster = initial state"
    // shortcut (evaluation 0.94 -> 1.21 ms): 2.19 -> 2.37 ms per 400 Mpx.  Off
    // unless OCSB_PASS_OVERLAP=1.
    const char *ov = getenv("OCSB_PASS_OVERLAP");
    const bool overlap_on = ov && atoi(ov) != 0;
    const bool overlap = overlap_on && flags == nullptr && n_ops >= 2 &&
                         ops[0].kind == OCSB_OP_FEATURE && n_other > 0 &&
                         same_fuse && pass_streams(ctx) == OCSB_OK;
    auto run_limiter = [&](const ocsb_hfun_op &op) -> int32_t {
        LimiterArgs la{dx, dy, op.hmin, op.hmax, op.lower_bound,
                       op.upper_bound, ghmin, ghmax, op.fuse_clamp};
        OCSB_TRY(k_flow_limiter(ctx, win->width, win->height, a, b, dem,
                                dem_dtype, dem_stride, oldp, olds, out,
                                out_stride, la));
        oldp = out;
        olds = out_stride;
        return OCSB_OK;
    };
    std::function<int32_t()> between = [&]() -> int32_t {
        for (int k = 0; k < n_ops; ++k)
            if (ops[k].kind != OCSB_OP_FEATURE) OCSB_TRY(run_limiter(ops[k]));
        return OCSB_OK;
    };
    for (int i = 0; i < n_ops; ++i) {
        const ocsb_hfun_op &op = ops[i];
        if (overlap && op.kind != OCSB_OP_FEATURE) continue;   // ran in `between`
        if (op.kind == OCSB_OP_FEATURE) {
            FeatureArgs fa;
            win_axes(win, &fa.ax, &fa.ay);
            fa.rate_target = op.expansion_rate * op.target_size;
            fa.target = op.target_size;
            fa.has_hmin = !ocsb_is_none(ghmin);
            fa.has_hmax = !ocsb_is_none(ghmax);
            fa.hmin_f = fa.has_hmin ? (float)ghmin : 0.f;
            fa.hmax_f = fa.has_hmax ? (float)ghmax : 0.f;
            fa.fuse_clamp = op.fuse_clamp;
            fa.r = (fa.has_hmax && ghmax != 0.0)
                       ? (ghmax - op.target_size) / fa.rate_target
                       : std::numeric_limits<double>::infinity();
            fa.row_begin = a;
            fa.row_end = b;
            fa.ext_overflow = flags ? flags + i : nullptr;
            if (overlap && i == 0) {
                fa.refine_stream = ctx->s_h2d;
                fa.ev_fork = ctx->ev_alloc;
                fa.ev_join = ctx->ev_comp[0];
                fa.between = &between;
                // the limiters write `out` first: this op's pre-op raster is `out`
                OCSB_TRY(k_hfun_feature(ctx, fa, verts[i], op.nv, out, out_stride,
                                        out, out_stride, nullptr, nullptr));
                oldp = out;
                olds = out_stride;
                continue;
            }
            OCSB_TRY(k_hfun_feature(ctx, fa, verts[i], op.nv, oldp, olds, out,
                                    out_stride, nullptr, nullptr));
        } else {
            OCSB_TRY(run_limiter(op));
        }
        oldp = out;
        olds = out_stride;
    }
    return OCSB_OK;
}

}  // namespace

extern "C" int32_t ocsb_hfun_pass(
    ocsb_ctx *ctx, const ocsb_win *win, const void *dem, int32_t dem_dtype,
    int64_t dem_stride, const float *hfun_in, int64_t in_stride,
    float *hfun_out, int64_t out_stride, double dx, double dy, double ghmin,
    double ghmax, const ocsb_hfun_op *ops, int32_t n_ops, int32_t band_rows) {
    CTX_ENTER("ocsb_hfun_pass");
    OCSB_TRY(check_win(win, "ocsb_hfun_pass"));
    OCSB_REQUIRE(ops && n_ops > 0, "ocsb_hfun_pass: empty op list");
    OCSB_REQUIRE(hfun_out && out_stride >= win->width &&
                     (!hfun_in || in_stride >= win->width),
                 "ocsb_hfun_pass: bad hfun buffers");
    OCSB_REQUIRE(band_rows >= 0, "ocsb_hfun_pass: negative band_rows");
    const int w = win->width, h = win->height;
    int rb, re;
    OCSB_TRY(win_band(win, "ocsb_hfun_pass", &rb, &re));
    bool need_dem = false;
    for (int i = 0; i < n_ops; ++i) {
        const ocsb_hfun_op &op = ops[i];
        if (op.kind == OCSB_OP_FEATURE) {
            OCSB_REQUIRE(op.verts_xy && op.nv > 0 && op.nv < (int64_t)2147483647,
                         "ocsb_hfun_pass: op %d: bad vertex set", i);
            OCSB_REQUIRE(!ocsb_is_none(op.target_size) && op.target_size > 0,
                         "Argument target_size must be greater than zero.");
            OCSB_REQUIRE(!ocsb_is_none(op.expansion_rate),
                         "ocsb_hfun_pass: op %d: expansion_rate is NaN", i);
            if (!ocsb_is_none(ghmax) && ghmax != 0.0)
                OCSB_REQUIRE(op.expansion_rate * op.target_size != 0.0,
                             "ocsb_hfun_pass: op %d: float division by zero "
                             "(expansion_rate * target_size == 0 with hmax set)", i);
        } else if (op.kind == OCSB_OP_LIMITER) {
            need_dem = true;
        } else {
            OCSB_REQUIRE(false, "ocsb_hfun_pass: op %d: unknown kind %d", i, op.kind);
        }
    }
    if (need_dem) {
        OCSB_REQUIRE(dem && dem_stride >= w &&
                         (dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64),
                     "ocsb_hfun_pass: a limiter op needs the DEM window");
        OCSB_REQUIRE(w >= 2 && h >= 2,
                     "Shape of array too small to calculate a numerical gradient, "
                     "at least (edge_order + 1) elements are required.");
        OCSB_REQUIRE(!ocsb_is_none(dx) && !ocsb_is_none(dy),
                     "ocsb_hfun_pass: dx/dy is NaN");
    }

    // vertex sets: staged once, whole
    std::vector<Staged> sv(n_ops);
    std::vector<const double *> verts(n_ops, nullptr);
    for (int i = 0; i < n_ops; ++i)
        if (ops[i].kind == OCSB_OP_FEATURE) {
            OCSB_TRY(sv[i].init(ctx, ops[i].verts_xy, 1, ops[i].nv * 2,
                                ops[i].nv * 2, 8, STAGE_IN));
            verts[i] = (const double *)sv[i].dev;
        }

    PassBuf bz, bi, bo;
    bz.user = (const char *)dem; bz.stride = dem_stride;
    bz.elem = dem_dtype == OCSB_F64 ? 8 : 4;
    // memory kind of the first row that exists (see Staged::init)
    auto first_row = [](const void *p, int64_t row, int64_t stride, size_t elem) {
        return (const void *)((const char *)p + (size_t)row * (size_t)stride * elem);
    };
    const int zrow0 = rb > 0 ? rb - 1 : 0;
    bz.host = need_dem && !ocsb_is_device_ptr(first_row(dem, zrow0, dem_stride, bz.elem));
    bi.user = (const char *)hfun_in; bi.stride = in_stride;
    bi.host = hfun_in && !ocsb_is_device_ptr(first_row(hfun_in, rb, in_stride, 4));
    bo.user = (const char *)hfun_out; bo.stride = out_stride;
    bo.host = !ocsb_is_device_ptr(first_row(hfun_out, rb, out_stride, 4));

    if (!bz.host && !bi.host && !bo.host) {
        // everything resident: one band, no copies
        return pass_ops(ctx, win, rb, re, dem, dem_dtype, dem_stride, hfun_in,
                        in_stride, hfun_out, out_stride, dx, dy, ghmin, ghmax,
                        ops, n_ops, verts, nullptr);
    }

    // ---- streamed: H2D(k+1, k+2) | ops(k) | D2H(k-1) -------------------------
    // Nothing below waits on the host until the final drain: the candidate-pool
    // overflow flags of the feature ops stay on the device and are checked
    // once at the end (a band that overflowed is redone synchronously).
    OCSB_TRY(pass_streams(ctx));
    int rows = band_rows;
    if (rows == 0) {
        rows = (int)((80ll << 20) / (4ll * w));
        rows = rows / 64 * 64;          // whole 64-px tiles of the K1 pyramid
        if (rows < 64) rows = 64;
    }
    if (rows > re - rb) rows = re - rb;
    // band schedule: the first H2D and the last D2H cannot overlap with
    // anything, so the bands ramp up from / down to 128 rows at both ends
    std::vector<int> cuts;   // band k = rows [cuts[k], cuts[k+1])
    {
        std::vector<int> head, tail;
        int lo = rb, hi = re;
        for (int sz = 128; sz < rows && hi - lo > 4 * rows; sz *= 2) {
            head.push_back(lo); lo += sz;
            hi -= sz; tail.push_back(hi);
        }
        cuts = head;
        for (int a = lo; a < hi; a += rows) cuts.push_back(a);
        for (size_t i = tail.size(); i-- > 0;) cuts.push_back(tail[i]);
        cuts.push_back(re);
    }
    const int nb = (int)cuts.size() - 1;
    // an aliased host in/out window travels through ONE device buffer
    const bool alias = hfun_in && (const void *)hfun_in == (const void *)hfun_out &&
                       in_stride == out_stride;
    ScratchGuard sg(ctx);
    constexpr int NS = OCSB_PASS_SETS;
    int *flags = nullptr;
    OCSB_TRY(sg.get(&flags, (size_t)nb * n_ops));
    OCSB_CUDA(cudaMemsetAsync(flags, 0, (size_t)nb * n_ops * sizeof(int), ctx->stream));
    for (int s = 0; s < (nb < NS ? nb : NS); ++s) {
        if (bz.host) OCSB_TRY(sg.get(&bz.band[s], (size_t)(rows + 2) * w * bz.elem));
        if (bo.host) OCSB_TRY(sg.get(&bo.band[s], (size_t)rows * w * 4));
        if (bi.host && !alias) OCSB_TRY(sg.get(&bi.band[s], (size_t)rows * w * 4));
    }
    OCSB_CUDA(cudaEventRecord(ctx->ev_alloc, ctx->stream));
    OCSB_CUDA(cudaStreamWaitEvent(ctx->s_h2d, ctx->ev_alloc, 0));
    OCSB_CUDA(cudaStreamWaitEvent(ctx->s_d2h, ctx->ev_alloc, 0));

    auto band_of = [&](int k, int *a, int *b) {
        *a = cuts[k];
        *b = cuts[k + 1];
    };
    auto issue_h2d = [&](int k) -> int32_t {
        const int s = k % NS;
        int a, b;
        band_of(k, &a, &b);
        // the buffers of set s were last used by band k-NS: its D2H (which
        // follows its kernels) must have drained
        if (k >= NS) OCSB_CUDA(cudaStreamWaitEvent(ctx->s_h2d, ctx->ev_out[s], 0));
        if (bz.host) {
            const int za = a > 0 ? a - 1 : 0, zb = b < h ? b + 1 : h;
            OCSB_CUDA(cudaMemcpy2DAsync(
                bz.band[s], (size_t)w * bz.elem,
                bz.user + (size_t)za * bz.stride * bz.elem, (size_t)bz.stride * bz.elem,
                (size_t)w * bz.elem, (size_t)(zb - za), cudaMemcpyHostToDevice,
                ctx->s_h2d));
        }
        if (bi.host) {
            char *dst = alias ? bo.band[s] : bi.band[s];
            OCSB_CUDA(cudaMemcpy2DAsync(
                dst, (size_t)w * 4, bi.user + (size_t)a * bi.stride * 4,
                (size_t)bi.stride * 4, (size_t)w * 4, (size_t)(b - a),
                cudaMemcpyHostToDevice, ctx->s_h2d));
        }
        OCSB_CUDA(cudaEventRecord(ctx->ev_in[s], ctx->s_h2d));
        return OCSB_OK;
    };

    int32_t status = issue_h2d(0);
    if (status == OCSB_OK && nb > 1) status = issue_h2d(1);
    for (int k = 0; k < nb && status == OCSB_OK; ++k) {
        const int s = k % NS;
        int a, b;
        band_of(k, &a, &b);
        if (k + 2 < nb) status = issue_h2d(k + 2);
        if (status != OCSB_OK) break;
        // (virtual) addresses of window row 0 of this band's buffers
        const int za = a > 0 ? a - 1 : 0;
        const void *zp = dem;
        int64_t zs = dem_stride;
        if (bz.host) { zp = bz.band[s] - (size_t)za * w * bz.elem; zs = w; }
        float *op_ = hfun_out;
        int64_t os_ = out_stride;
        if (bo.host) { op_ = (float *)(bo.band[s] - (size_t)a * w * 4); os_ = w; }
        const float *ip = hfun_in;
        int64_t is_ = in_stride;
        if (bi.host) {
            ip = alias ? op_ : (const float *)(bi.band[s] - (size_t)a * w * 4);
            is_ = w;
        }
        cudaError_t e = cudaStreamWaitEvent(ctx->stream, ctx->ev_in[s], 0);
        // device-resident output written in place: band k's kernels must not
        // overtake nothing -- they run in stream order on ctx->stream
        if (e != cudaSuccess) { status = OCSB_ECUDA; ocsb_set_error("cudaStreamWaitEvent -> %s", cudaGetErrorString(e)); break; }
        status = pass_ops(ctx, win, a, b, zp, dem_dtype, zs, ip, is_, op_, os_, dx,
                          dy, ghmin, ghmax, ops, n_ops, verts,
                          flags + (size_t)k * n_ops);
        if (status != OCSB_OK) break;
        cudaEventRecord(ctx->ev_comp[s], ctx->stream);
        cudaStreamWaitEvent(ctx->s_d2h, ctx->ev_comp[s], 0);
        if (bo.host) {
            e = cudaMemcpy2DAsync((char *)hfun_out + (size_t)a * out_stride * 4,
                                  (size_t)out_stride * 4, bo.band[s], (size_t)w * 4,
                                  (size_t)w * 4, (size_t)(b - a),
                                  cudaMemcpyDeviceToHost, ctx->s_d2h);
            if (e != cudaSuccess) { status = OCSB_ECUDA; ocsb_set_error("D2H -> %s", cudaGetErrorString(e)); break; }
        }
        cudaEventRecord(ctx->ev_out[s], ctx->s_d2h);
    }
    // drain all three streams before the scratch is released / the caller reads
    cudaStreamSynchronize(ctx->s_h2d);
    cudaStreamSynchronize(ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->s_d2h);
    if (status == OCSB_OK && e != cudaSuccess) {
        ocsb_set_error("ocsb_hfun_pass: %s", cudaGetErrorString(e));
        status = OCSB_ECUDA;
    }
    if (status != OCSB_OK) return status;

    // ---- bands whose candidate pools overflowed: redo them synchronously ------
    std::vector<int> hflags((size_t)nb * n_ops, 0);
    OCSB_CUDA(cudaMemcpy(hflags.data(), flags, hflags.size() * sizeof(int),
                         cudaMemcpyDeviceToHost));
    for (int k = 0; k < nb; ++k) {
        bool bad = false;
        for (int i = 0; i < n_ops; ++i) bad = bad || hflags[(size_t)k * n_ops + i];
        if (!bad) continue;
        int a, b;
        band_of(k, &a, &b);
        // the single-band call: everything resident, or one streamed band
        // (rows == b - a) through the synchronous retrying path
        if (!bz.host && !bi.host && !bo.host) {
            OCSB_TRY(pass_ops(ctx, win, a, b, dem, dem_dtype, dem_stride, hfun_in,
                              in_stride, hfun_out, out_stride, dx, dy, ghmin,
                              ghmax, ops, n_ops, verts, nullptr));
        } else {
            Staged sz, si, so;
            const int za = a > 0 ? a - 1 : 0, zb = b < h ? b + 1 : h;
            if (need_dem)
                OCSB_TRY(sz.init(ctx, dem, zb - za, w, dem_stride, bz.elem, STAGE_IN, za));
            const bool al = hfun_in && (const void *)hfun_in == (const void *)hfun_out &&
                            in_stride == out_stride;
            if (al) {
                OCSB_TRY(so.init(ctx, hfun_out, b - a, w, out_stride, 4, STAGE_INOUT, a));
            } else {
                if (hfun_in) OCSB_TRY(si.init(ctx, hfun_in, b - a, w, in_stride, 4, STAGE_IN, a));
                OCSB_TRY(so.init(ctx, hfun_out, b - a, w, out_stride, 4, STAGE_OUT, a));
            }
            const float *ip = al ? (const float *)so.dev : (const float *)si.dev;
            const int64_t is_ = al ? so.dev_stride : si.dev_stride;
            OCSB_TRY(pass_ops(ctx, win, a, b, sz.dev, dem_dtype,
                              need_dem ? sz.dev_stride : 0, ip, is_,
                              (float *)so.dev, so.dev_stride, dx, dy, ghmin,
                              ghmax, ops, n_ops, verts, nullptr));
            OCSB_TRY(finish_call(ctx, {&sz, &si, &so}));
        }
    }
    return OCSB_OK;
}


// ---------------------------------------------------------------------------
// The op list over a LIST of semantic windows in one tiled sweep, every buffer
// resident (see ocsb200.h).  The window table is a unit-of-work table of the
// SAME launches: no per-window launch, no host loop.
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_hfun_pass_windows(
    ocsb_ctx *ctx, const ocsb_win *wins, const int64_t *dem_off,
    const int64_t *hfun_off, const int32_t *keep_wh, int32_t n_wins,
    const void *dem, int32_t dem_dtype, int64_t dem_stride, int32_t dem_rows,
    int32_t dem_cols, const float *hfun_in, int64_t in_stride, float *hfun_out,
    int64_t out_stride, double dx, double dy, double ghmin, double ghmax,
    const ocsb_hfun_op *ops, int32_t n_ops, int32_t *status_dev) {
    CTX_ENTER("ocsb_hfun_pass_windows");
    OCSB_REQUIRE(n_wins >= 0, "ocsb_hfun_pass_windows: negative window count");
    if (n_wins == 0) return OCSB_OK;
    OCSB_REQUIRE(wins && hfun_off && ops && n_ops > 0 && hfun_out,
                 "ocsb_hfun_pass_windows: NULL arguments");
    OCSB_REQUIRE(ocsb_is_device_ptr(hfun_out) && (!hfun_in || ocsb_is_device_ptr(hfun_in)),
                 "ocsb_hfun_pass_windows: the size raster must be device memory");
    OCSB_REQUIRE(!hfun_in || in_stride == out_stride,
                 "ocsb_hfun_pass_windows: hfun_in and hfun_out must share one layout");
    OCSB_REQUIRE(!status_dev || ocsb_is_device_ptr(status_dev),
                 "ocsb_hfun_pass_windows: status_dev must be device memory");
    bool need_dem = false;
    std::vector<PassFeat> pf;
    std::vector<PassLim> pl;
    for (int i = 0; i < n_ops; ++i) {
        const ocsb_hfun_op &op = ops[i];
        if (op.kind == OCSB_OP_FEATURE) {
            OCSB_REQUIRE(op.verts_xy && op.nv > 0 && op.nv < (int64_t)2147483647,
                         "ocsb_hfun_pass_windows: op %d: bad vertex set", i);
            OCSB_REQUIRE(ocsb_is_device_ptr(op.verts_xy),
                         "ocsb_hfun_pass_windows: op %d: vertices must be device memory", i);
            OCSB_REQUIRE(!ocsb_is_none(op.target_size) && op.target_size > 0,
                         "Argument target_size must be greater than zero.");
            OCSB_REQUIRE(!ocsb_is_none(op.expansion_rate),
                         "ocsb_hfun_pass_windows: op %d: expansion_rate is NaN", i);
            if (!ocsb_is_none(ghmax) && ghmax != 0.0)
                OCSB_REQUIRE(op.expansion_rate * op.target_size != 0.0,
                             "ocsb_hfun_pass_windows: op %d: float division by zero "
                             "(expansion_rate * target_size == 0 with hmax set)", i);
            PassFeat f;
            f.verts = op.verts_xy; f.nv = op.nv;
            f.rate_target = op.expansion_rate * op.target_size;
            f.target = op.target_size;
            f.r = (!ocsb_is_none(ghmax) && ghmax != 0.0)
                      ? (ghmax - op.target_size) / f.rate_target
                      : std::numeric_limits<double>::infinity();
            f.fuse_clamp = op.fuse_clamp;
            pf.push_back(f);
        } else if (op.kind == OCSB_OP_LIMITER) {
            need_dem = true;
            pl.push_back(PassLim{op.hmin, op.hmax, op.lower_bound, op.upper_bound,
                                 op.fuse_clamp});
        } else {
            OCSB_REQUIRE(false, "ocsb_hfun_pass_windows: op %d: unknown kind %d", i, op.kind);
        }
    }
    if (need_dem) {
        OCSB_REQUIRE(dem && dem_off && ocsb_is_device_ptr(dem) && dem_stride > 0 &&
                         dem_rows > 0 && dem_cols > 0 && dem_cols <= dem_stride &&
                         (dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64),
                     "ocsb_hfun_pass_windows: a limiter op needs the resident DEM block");
        OCSB_REQUIRE(!ocsb_is_none(dx) && !ocsb_is_none(dy),
                     "ocsb_hfun_pass_windows: dx/dy is NaN");
    }
    std::vector<PassWin> pw(n_wins);
    for (int k = 0; k < n_wins; ++k) {
        OCSB_TRY(check_win(&wins[k], "ocsb_hfun_pass_windows"));
        int rb, re;
        OCSB_TRY(win_band(&wins[k], "ocsb_hfun_pass_windows", &rb, &re));
        const int w = wins[k].width, h = wins[k].height;
        if (need_dem)
            OCSB_REQUIRE(w >= 2 && h >= 2,
                         "Shape of array too small to calculate a numerical gradient, "
                         "at least (edge_order + 1) elements are required.");
        PassWin &p = pw[k];
        win_axes(&wins[k], &p.ax, &p.ay);
        p.rb = rb; p.re = re;
        p.keep_w = keep_wh ? keep_wh[2 * k] : w;
        p.keep_h = keep_wh ? keep_wh[2 * k + 1] : h;
        OCSB_REQUIRE(p.keep_w >= 0 && p.keep_w <= w && p.keep_h >= 0 && p.keep_h <= h,
                     "ocsb_hfun_pass_windows: bad keep size of window %d", k);
        p.hoff = hfun_off[k];
        p.dem_c = 0; p.dem_r = 0;
        if (need_dem) {
            // floor division: a window may start above the stored block
            int64_t r = dem_off[k] / dem_stride, c = dem_off[k] - r * dem_stride;
            if (c < 0) { c += dem_stride; --r; }
            p.dem_c = (int)c; p.dem_r = (int)r;
        }
    }
    PassDesc d;
    d.wins = pw.data(); d.n_wins = n_wins;
    if (need_dem) {
        d.dem = dem; d.dem_dtype = dem_dtype; d.dem_stride = dem_stride;
        d.dem_rows = dem_rows; d.dem_cols = dem_cols;
    }
    d.in = hfun_in; d.in_stride = in_stride;
    d.out = hfun_out; d.out_stride = out_stride;
    d.dx = dx; d.dy = dy; d.ghmin = ghmin; d.ghmax = ghmax;
    d.feats = pf.data(); d.nf = (int)pf.size();
    d.lims = pl.data(); d.nl = (int)pl.size();
    d.ext_overflow = status_dev;
    d.sticky = status_dev != nullptr;
    if (k_pass_supported(d)) return k_pass_tiled(ctx, d);
    // ---- op lists the tiled sweep does not take (float64 DEM, mixed clamp
    //      fusing, more than 4 ops of a kind): window by window, op by op ------
    OCSB_REQUIRE(!status_dev, "ocsb_hfun_pass_windows: this op list only runs "
                              "synchronously (status_dev must be NULL)");
    std::vector<const double *> verts(n_ops, nullptr);
    for (int i = 0; i < n_ops; ++i) verts[i] = ops[i].verts_xy;
    const size_t es = dem_dtype == OCSB_F64 ? 8 : 4;
    for (int k = 0; k < n_wins; ++k) {
        const PassWin &p = pw[k];
        OCSB_REQUIRE(p.keep_w == wins[k].width && p.keep_h == wins[k].height,
                     "ocsb_hfun_pass_windows: partial windows need the tiled sweep");
        const char *zp = need_dem ? (const char *)dem + dem_off[k] * (int64_t)es : nullptr;
        OCSB_TRY(pass_ops(ctx, &wins[k], p.rb, p.re, zp, dem_dtype, dem_stride,
                          hfun_in ? hfun_in + hfun_off[k] : nullptr, in_stride,
                          hfun_out + hfun_off[k], out_stride, dx, dy, ghmin, ghmax,
                          ops, n_ops, verts, nullptr));
    }
    return OCSB_OK;
}

// ---------------------------------------------------------------------------
// One op over a LIST of semantic windows (the window loops of
// hfun/raster.py:1212-1260 and :1393-1436), buffers resident on the device.
// Windows are independent units of work (own coordinates, own search radius,
// own one-sided borders), so they are spread over OCSB_WIN_STREAMS worker
// streams: the small per-window launches (8 pyramid kernels + evaluation, or
// one stencil launch) of different windows overlap instead of each running
// alone on a fraction of the SMs.
// ---------------------------------------------------------------------------
extern "C" int32_t ocsb_hfun_op_windows(
    ocsb_ctx *ctx, const ocsb_win *wins, const int64_t *dem_off,
    const int64_t *hfun_off, int32_t n_wins, const void *dem, int32_t dem_dtype,
    int64_t dem_stride, const float *hfun_old, int64_t old_stride,
    float *hfun_new, int64_t new_stride, double dx, double dy, double ghmin,
    double ghmax, const ocsb_hfun_op *op, int32_t in_order) {
    CTX_ENTER("ocsb_hfun_op_windows");
    OCSB_REQUIRE(n_wins >= 0, "ocsb_hfun_op_windows: negative window count");
    if (n_wins == 0) return OCSB_OK;
    OCSB_REQUIRE(wins && hfun_off && op && hfun_new,
                 "ocsb_hfun_op_windows: NULL arguments");
    OCSB_REQUIRE(op->kind == OCSB_OP_FEATURE || op->kind == OCSB_OP_LIMITER,
                 "ocsb_hfun_op_windows: unknown op kind %d", op->kind);
    OCSB_REQUIRE(ocsb_is_device_ptr(hfun_new) &&
                     (!hfun_old || ocsb_is_device_ptr(hfun_old)),
                 "ocsb_hfun_op_windows: the size raster must be device memory");
    const bool lim = op->kind == OCSB_OP_LIMITER;
    if (lim) {
        OCSB_REQUIRE(dem && dem_off && ocsb_is_device_ptr(dem) &&
                         (dem_dtype == OCSB_F32 || dem_dtype == OCSB_F64),
                     "ocsb_hfun_op_windows: a limiter op needs the device DEM");
        OCSB_REQUIRE(!ocsb_is_none(dx) && !ocsb_is_none(dy),
                     "ocsb_hfun_op_windows: dx/dy is NaN");
    } else {
        OCSB_REQUIRE(op->verts_xy && op->nv > 0 && op->nv < (int64_t)2147483647,
                     "ocsb_hfun_op_windows: bad vertex set");
        OCSB_REQUIRE(!ocsb_is_none(op->target_size) && op->target_size > 0,
                     "Argument target_size must be greater than zero.");
        OCSB_REQUIRE(!ocsb_is_none(op->expansion_rate),
                     "ocsb_hfun_op_windows: expansion_rate is NaN");
        if (!ocsb_is_none(ghmax) && ghmax != 0.0)
            OCSB_REQUIRE(op->expansion_rate * op->target_size != 0.0,
                         "ocsb_hfun_op_windows: float division by zero "
                         "(expansion_rate * target_size == 0 with hmax set)");
    }
    std::vector<int> rb(n_wins), re(n_wins);
    for (int32_t k = 0; k < n_wins; ++k) {
        OCSB_TRY(check_win(&wins[k], "ocsb_hfun_op_windows"));
        OCSB_TRY(win_band(&wins[k], "ocsb_hfun_op_windows", &rb[k], &re[k]));
        OCSB_REQUIRE(new_stride >= wins[k].width &&
                         (!hfun_old || old_stride >= wins[k].width) &&
                         (!lim || dem_stride >= wins[k].width),
                     "ocsb_hfun_op_windows: window %d wider than a stride", k);
        if (lim)
            OCSB_REQUIRE(wins[k].width >= 2 && wins[k].height >= 2,
                         "Shape of array too small to calculate a numerical gradient, "
                         "at least (edge_order + 1) elements are required.");
    }
    // vertices: staged once on the context's stream, before the fork
    Staged sv;
    std::vector<const double *> verts(1, nullptr);
    if (!lim) {
        OCSB_TRY(sv.init(ctx, op->verts_xy, 1, op->nv * 2, op->nv * 2, 8, STAGE_IN));
        verts[0] = (const double *)sv.dev;
    }
    // overlapping windows write shared pixels: list order must decide there, as
    // in the reference's sequential loop.  Only windows that really share
    // pixels are ordered (window k waits for the event of every earlier window
    // it overlaps); the rest still run side by side, a wavefront over the grid.
    const int ns = n_wins < OCSB_WIN_STREAMS ? n_wins : OCSB_WIN_STREAMS;
    struct Rect { int64_t r0, r1, c0, c1; };
    std::vector<Rect> rect;
    struct EventList {
        std::vector<cudaEvent_t> ev;
        ~EventList() { for (auto e : ev) if (e) cudaEventDestroy(e); }
    } done;
    if (in_order && n_wins > 1) {
        rect.resize(n_wins);
        done.ev.assign(n_wins, nullptr);
        for (int32_t k = 0; k < n_wins; ++k) {
            // (virtual) row 0 / column 0 of the window in the hfun buffer
            int64_t row = hfun_off[k] / new_stride, col = hfun_off[k] % new_stride;
            if (col < 0) { col += new_stride; --row; }
            rect[k] = {row + rb[k], row + re[k], col, col + wins[k].width};
            OCSB_CUDA(cudaEventCreateWithFlags(&done.ev[k], cudaEventDisableTiming));
        }
    }
    for (int i = 0; i < ns; ++i) {
        if (!ctx->s_win[i])
            OCSB_CUDA(cudaStreamCreateWithFlags(&ctx->s_win[i], cudaStreamNonBlocking));
        if (!ctx->ev_win[i])
            OCSB_CUDA(cudaEventCreateWithFlags(&ctx->ev_win[i], cudaEventDisableTiming));
    }
    OCSB_TRY(pass_streams(ctx));     // ev_alloc doubles as the fork event
    ScratchGuard sg(ctx);
    int *flags = nullptr;
    OCSB_TRY(sg.get(&flags, (size_t)n_wins));
    OCSB_CUDA(cudaMemsetAsync(flags, 0, (size_t)n_wins * sizeof(int), ctx->stream));
    OCSB_CUDA(cudaEventRecord(ctx->ev_alloc, ctx->stream));
    for (int i = 0; i < ns; ++i)
        OCSB_CUDA(cudaStreamWaitEvent(ctx->s_win[i], ctx->ev_alloc, 0));
    const size_t es = dem_dtype == OCSB_F64 ? 8 : 4;
    auto ptrs = [&](int32_t k, const void **zp, const float **ip, float **op_) {
        *zp = lim ? (const void *)((const char *)dem + (size_t)dem_off[k] * es) : nullptr;
        *ip = hfun_old ? hfun_old + hfun_off[k] : nullptr;
        *op_ = hfun_new + hfun_off[k];
    };
    cudaStream_t main_stream = ctx->stream;
    int32_t status = OCSB_OK;
    ctx->keep_ok = false;       // the windows' passes run side by side: own scratch each
    for (int32_t k = 0; k < n_wins && status == OCSB_OK; ++k) {
        const void *zp; const float *ip; float *opp;
        ptrs(k, &zp, &ip, &opp);
        ctx->stream = ctx->s_win[k % ns];        // everything below enqueues here
        if (!rect.empty())
            for (int32_t j = 0; j < k; ++j)
                if (rect[j].r0 < rect[k].r1 && rect[k].r0 < rect[j].r1 &&
                    rect[j].c0 < rect[k].c1 && rect[k].c0 < rect[j].c1 &&
                    (j % ns) != (k % ns))        // same stream: already ordered
                    cudaStreamWaitEvent(ctx->stream, done.ev[j], 0);
        status = pass_ops(ctx, &wins[k], rb[k], re[k], zp, dem_dtype, dem_stride, ip,
                          old_stride, opp, new_stride, dx, dy, ghmin, ghmax, op, 1,
                          verts, flags + k);
        if (!rect.empty() && status == OCSB_OK)
            cudaEventRecord(done.ev[k], ctx->stream);
    }
    ctx->stream = main_stream;
    ctx->keep_ok = true;
    for (int i = 0; i < ns; ++i) {               // join
        cudaEventRecord(ctx->ev_win[i], ctx->s_win[i]);
        cudaStreamWaitEvent(ctx->stream, ctx->ev_win[i], 0);
    }
    if (status != OCSB_OK) {
        cudaStreamSynchronize(ctx->stream);
        return status;
    }
    if (!lim) {
        // windows whose candidate pools overflowed left their pixels untouched:
        // redo them through the synchronous, retrying path
        std::vector<int> hflags((size_t)n_wins, 0);
        OCSB_CUDA(cudaMemcpyAsync(hflags.data(), flags, hflags.size() * sizeof(int),
                                  cudaMemcpyDeviceToHost, ctx->stream));
        OCSB_CUDA(cudaStreamSynchronize(ctx->stream));
        bool redo_rest = false;
        for (int32_t k = 0; k < n_wins; ++k) {
            // ordered lists: a redone window rewrites pixels it shares with
            // later windows, so everything after the first overflow is redone
            // in list order (old / new are distinct buffers: idempotent)
            if (!hflags[k] && !redo_rest) continue;
            if (!rect.empty()) redo_rest = true;
            const void *zp; const float *ip; float *opp;
            ptrs(k, &zp, &ip, &opp);
            OCSB_TRY(pass_ops(ctx, &wins[k], rb[k], re[k], zp, dem_dtype, dem_stride,
                              ip, old_stride, opp, new_stride, dx, dy, ghmin, ghmax,
                              op, 1, verts, nullptr));
        }
    }
    return finish_call(ctx, {&sv});
}
