// Shared device helpers for the stream-compaction passes (mask boundary
// pixels in k_constraint.cu, contour segments in k_contour.cu): per-CTA counts
// -> exclusive offsets by ONE CTA -> ordered emit.  Included inside each
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

}  // namespace
