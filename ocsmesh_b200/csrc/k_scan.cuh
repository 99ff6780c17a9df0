// Shared device helpers for the stream-compaction passes (mask boundary
// pixels in k_constraint.cu, contour segments in k_contour.cu): per-CTA counts
// -> exclusive offsets (two-level scan) -> ordered emit.  Included inside each
// translation unit (internal linkage).
#pragma once

namespace {

// exclusive scan of counts[n] in place by ONE CTA; total -> *total
__global__ void __launch_bounds__(1024)
scan_counts_kernel(int *__restrict__ counts, int64_t n,
                   unsigned long long *__restrict__ total) {
    __shared__ int s_warp[32];
    __shared__ unsigned long long s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const int v = i < n ? counts[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_warp[wid] = x;
        __syncthreads();
        if (wid == 0) {
            int t = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t += y;
            }
            s_warp[lane] = t;
        }
        __syncthreads();
        const unsigned long long carry = s_carry;
        const int before = (wid ? s_warp[wid - 1] : 0) + (x - v);
        // offsets stay below 2^31: the caller bounds the vertex count
        if (i < n) counts[i] = (int)(carry + (unsigned long long)before);
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + (unsigned long long)(before + v);
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = s_carry;
}

// ---- two-level exclusive scan for long count arrays -------------------------
// (1) every CTA scans its own 1024 counts in place and writes its total,
// (2) ONE CTA scans the totals (scan_counts_kernel), (3) the totals' offsets
// are added back.  400 Mpx / 1024 = 390 k counts: 3 short launches instead of
// 381 serial rounds of one CTA.
__global__ void __launch_bounds__(1024)
scan_block_kernel(int *__restrict__ counts, int64_t n, int *__restrict__ sums) {
    __shared__ int s_warp[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    const int v = i < n ? counts[i] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[wid] = x;
    __syncthreads();
    if (wid == 0) {
        int t = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, t, o);
            if (lane >= o) t += y;
        }
        s_warp[lane] = t;
    }
    __syncthreads();
    if (i < n) counts[i] = (wid ? s_warp[wid - 1] : 0) + (x - v);
    if (threadIdx.x == 0) sums[blockIdx.x] = s_warp[31];
}

__global__ void __launch_bounds__(1024)
scan_add_kernel(int *__restrict__ counts, int64_t n, const int *__restrict__ sums) {
    const int64_t i = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    if (i < n) counts[i] += sums[blockIdx.x];
}

// counts[n] -> exclusive offsets in place, grand total -> *total (device).
// `sums` is scratch of (n + 1023) / 1024 ints.  3 launches on `stream`.
inline void scan_counts(int *counts, int64_t n, int *sums,
                        unsigned long long *total, cudaStream_t stream) {
    const int64_t nb = (n + 1023) / 1024;
    scan_block_kernel<<<(unsigned)nb, 1024, 0, stream>>>(counts, n, sums);
    scan_counts_kernel<<<1, 1024, 0, stream>>>(sums, nb, total);
    scan_add_kernel<<<(unsigned)nb, 1024, 0, stream>>>(counts, n, sums);
}

}  // namespace
