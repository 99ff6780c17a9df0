// Band exchange over peer memory: the per-step exchange of a
// row-band run -- every rank's contour vertices (the all-gather of the north
// star) and its first / last owned DEM row (the 1-row halos of the gradient
// stencil, ocsmesh/hfun/raster.py:1399) -- as stores into the PEERS' buffers
// through NVLink instead of a NCCL collective between two launches.
//
//   push: packs this rank's block [first row | last row | vertices padded to cap]
//         straight into slot `rank` of every rank's exchange buffer (its own
//         included), then the last CTA to finish publishes the step's epoch in
//         every rank's flag word `rank` (fence.sys + st.release.sys);
//   pull: waits until all `world` flag words of its OWN buffer carry the epoch
//         (ld.acquire.sys, bounded), then unpacks: the dense vertex array the
//         pass reads and the two halo rows of the stored band.
//
// The buffer holds two generations of slots (epoch parity): a rank may be one
// step ahead of a peer, never two (its pull of step k+1 waits for every peer's
// push of k+1, which follows that peer's pull of k in stream order), so a push
// never lands in slots a peer is still unpacking.
//
// Buffers come from ocsb_p2p_alloc (cudaMalloc + CUDA IPC handle) and are
// mapped by the peers with ocsb_p2p_open; the reference is a single process and
// has no counterpart of any of this.
#include "ocsb_common.cuh"

namespace {

constexpr int kMaxWorld = 32;
constexpr long long kFlagBytes = 256;       // 32 x 8-byte epochs

__host__ __device__ inline long long p2p_row_bytes(int w) {
    long long b = 2ll * w * 4;
    return b + ((16 - b % 16) % 16);        // the float64 pairs behind it stay aligned
}
__host__ __device__ inline long long p2p_blk_bytes(long long cap, int w) {
    return p2p_row_bytes(w) + cap * 16;
}

struct PushArgs {
    const double2 *verts;
    long long n, cap;
    const float *row_first, *row_last;
    int w, world, rank;
    long long epoch;
    double far;
    unsigned char *const *peers;        // device array: the ranks' exchange buffers
    unsigned long long *done;           // local counter of finished CTAs (monotone)
};

__global__ void __launch_bounds__(256) band_push_kernel(PushArgs A) {
    const int p = blockIdx.y;
    const long long blk = p2p_blk_bytes(A.cap, A.w);
    const long long gen = (A.epoch & 1) * (long long)A.world * blk;
    unsigned char *dst = A.peers[p] + gen + (long long)A.rank * blk;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 2ll * A.w)
        reinterpret_cast<float *>(dst)[i] = i < A.w ? A.row_first[i] : A.row_last[i - A.w];
    if (i < A.cap)
        reinterpret_cast<double2 *>(dst + p2p_row_bytes(A.w))[i] =
            A.n > 0 ? A.verts[i < A.n ? i : A.n - 1] : make_double2(A.far, A.far);
    // every store of this CTA is visible system-wide before its arrival counts
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned long long total = (unsigned long long)gridDim.x * gridDim.y;
        const unsigned long long old = atomicAdd(A.done, 1ull);
        if (old + 1 == total * (unsigned long long)A.epoch) {
            __threadfence_system();
            for (int q = 0; q < A.world; ++q) {
                unsigned long long *flag = reinterpret_cast<unsigned long long *>(
                    A.peers[q] + 2ll * A.world * blk) + A.rank;
                asm volatile("st.release.sys.global.u64 [%0], %1;"
                             ::"l"(flag), "l"((unsigned long long)A.epoch) : "memory");
            }
        }
    }
}

struct PullArgs {
    const unsigned char *own;
    long long cap, epoch;
    int w, world, rank;
    double2 *verts_out;
    float *halo_above, *halo_below;
    int *status;                        // set to 1 when a peer never showed up
};

__global__ void __launch_bounds__(256) band_pull_kernel(PullArgs A) {
    const long long blk = p2p_blk_bytes(A.cap, A.w);
    if ((int)threadIdx.x < A.world) {
        const unsigned long long *flag = reinterpret_cast<const unsigned long long *>(
            A.own + 2ll * A.world * blk) + threadIdx.x;
        unsigned long long v = 0;
        long long spins = 0;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
            if (v >= (unsigned long long)A.epoch) break;
            if (++spins > 20000000ll) {      // ~2 s: a peer died; do not hang the GPU
                if (A.status) atomicExch(A.status, 1);
                break;
            }
            __nanosleep(100);
        }
    }
    __syncthreads();
    const unsigned char *gen = A.own + (A.epoch & 1) * (long long)A.world * blk;
    const long long rowb = p2p_row_bytes(A.w);
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)A.world * A.cap) {
        const long long r = i / A.cap, k = i - r * A.cap;
        A.verts_out[i] = __ldcg(reinterpret_cast<const double2 *>(gen + r * blk + rowb) + k);
    }
    if (i < A.w) {
        if (A.halo_above)       // last owned row of the rank above
            A.halo_above[i] = __ldcg(
                reinterpret_cast<const float *>(gen + (long long)(A.rank - 1) * blk) + A.w + i);
        if (A.halo_below)       // first owned row of the rank below
            A.halo_below[i] = __ldcg(
                reinterpret_cast<const float *>(gen + (long long)(A.rank + 1) * blk) + i);
    }
}

}  // namespace

extern "C" int64_t ocsb_band_p2p_bytes(int32_t world, int64_t cap, int32_t width) {
    if (world < 1 || world > kMaxWorld || cap < 1 || width < 1) return -1;
    return 2ll * world * p2p_blk_bytes(cap, width) + kFlagBytes;
}

extern "C" int32_t ocsb_p2p_alloc(ocsb_ctx *ctx, int64_t bytes, void **ptr, uint8_t *handle64) {
    OCSB_REQUIRE(ctx && ptr && handle64 && bytes > 0, "ocsb_p2p_alloc: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    OCSB_CUDA(cudaSetDevice(ctx->device));
    void *p = nullptr;
    OCSB_CUDA(cudaMalloc(&p, (size_t)bytes));
    cudaError_t e = cudaMemset(p, 0, (size_t)bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        cudaFree(p);
        cudaGetLastError();
        ocsb_set_error("ocsb_p2p_alloc: %s", cudaGetErrorString(e));
        return OCSB_ECUDA;
    }
    memcpy(handle64, &h, 64);
    *ptr = p;
    return OCSB_OK;
}

extern "C" int32_t ocsb_p2p_open(ocsb_ctx *ctx, const uint8_t *handle64, void **ptr) {
    OCSB_REQUIRE(ctx && ptr && handle64, "ocsb_p2p_open: bad arguments");
    OCSB_CUDA(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void *p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        ocsb_set_error("ocsb_p2p_open: %s", cudaGetErrorString(e));
        return OCSB_ECUDA;
    }
    *ptr = p;
    return OCSB_OK;
}

extern "C" int32_t ocsb_p2p_close(ocsb_ctx *ctx, void *ptr) {
    OCSB_REQUIRE(ctx && ptr, "ocsb_p2p_close: bad arguments");
    OCSB_CUDA(cudaIpcCloseMemHandle(ptr));
    return OCSB_OK;
}

extern "C" int32_t ocsb_p2p_free(ocsb_ctx *ctx, void *ptr) {
    OCSB_REQUIRE(ctx && ptr, "ocsb_p2p_free: bad arguments");
    OCSB_CUDA(cudaFree(ptr));
    return OCSB_OK;
}

extern "C" int32_t ocsb_band_push(ocsb_ctx *ctx, const double *verts, int64_t n, int64_t cap,
                                  const float *row_first, const float *row_last,
                                  int32_t width, const void *const *peer_bufs_dev,
                                  int32_t world, int32_t rank, int64_t epoch, double far,
                                  unsigned long long *done_counter_dev) {
    OCSB_REQUIRE(ctx, "ctx is NULL");
    OCSB_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world && cap >= 1 &&
                     width >= 1 && n >= 0 && n <= cap && epoch >= 1 && row_first && row_last &&
                     peer_bufs_dev && done_counter_dev && (n == 0 || verts),
                 "ocsb_band_push: bad arguments");
    OCSB_REQUIRE(n == 0 || ((uintptr_t)verts & 15) == 0,
                 "ocsb_band_push: verts must be 16-byte aligned");
    PushArgs A;
    A.verts = reinterpret_cast<const double2 *>(verts);
    A.n = n; A.cap = cap;
    A.row_first = row_first; A.row_last = row_last;
    A.w = width; A.world = world; A.rank = rank;
    A.epoch = epoch; A.far = far;
    A.peers = (unsigned char *const *)peer_bufs_dev;
    A.done = done_counter_dev;
    const long long items = std::max<long long>(2ll * width, cap);
    const dim3 grid((unsigned)((items + 255) / 256), (unsigned)world);
    OCSB_TIME(ctx, KC_COORDS);
    band_push_kernel<<<grid, 256, 0, ctx->stream>>>(A);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}

extern "C" int32_t ocsb_band_pull(ocsb_ctx *ctx, const void *own_buf, int32_t world,
                                  int32_t rank, int64_t cap, int32_t width, int64_t epoch,
                                  double *verts_out, float *halo_above, float *halo_below,
                                  int32_t *status_dev) {
    OCSB_REQUIRE(ctx, "ctx is NULL");
    OCSB_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world && cap >= 1 &&
                     width >= 1 && epoch >= 1 && own_buf && verts_out,
                 "ocsb_band_pull: bad arguments");
    OCSB_REQUIRE((halo_above == nullptr || rank > 0) && (halo_below == nullptr || rank < world - 1),
                 "ocsb_band_pull: a halo row was asked for beyond the first / last rank");
    OCSB_REQUIRE(((uintptr_t)verts_out & 15) == 0,
                 "ocsb_band_pull: verts_out must be 16-byte aligned");
    PullArgs A;
    A.own = reinterpret_cast<const unsigned char *>(own_buf);
    A.cap = cap; A.epoch = epoch;
    A.w = width; A.world = world; A.rank = rank;
    A.verts_out = reinterpret_cast<double2 *>(verts_out);
    A.halo_above = halo_above; A.halo_below = halo_below;
    A.status = status_dev;
    const long long items = std::max<long long>((long long)world * cap, width);
    OCSB_TIME(ctx, KC_COORDS);
    band_pull_kernel<<<(unsigned)((items + 255) / 256), 256, 0, ctx->stream>>>(A);
    OCSB_LAUNCHED(ctx);
    OCSB_CHECK_LAUNCH();
    return OCSB_OK;
}
