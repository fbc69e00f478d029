// Fixed-order block reductions shared by the reduction kernel (cs_diag.cu) and the stage kernels that produce the
// conservation sums themselves (cs_stage.cu).  No floating-point atomics anywhere: a result depends on the launch
// geometry only, never on which block or warp finishes first.
#pragma once

#include <cuda_runtime.h>

__device__ __forceinline__ double cs_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// Sum of two values over a block of THREADS threads (a multiple of 32); valid in thread 0.  Ends with a barrier,
// so it can be called twice in a row.
template <int THREADS>
__device__ __forceinline__ void cs_block_sum2(double& a, double& b) {
    __shared__ double sa[THREADS / 32], sb[THREADS / 32];
    a = cs_warp_sum(a);
    b = cs_warp_sum(b);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sa[w] = a, sb[w] = b;
    __syncthreads();
    if (w == 0) {
        a = l < THREADS / 32 ? sa[l] : 0.0;
        b = l < THREADS / 32 ? sb[l] : 0.0;
        a = cs_warp_sum(a);
        b = cs_warp_sum(b);
    }
    __syncthreads();
}

// Every block has written partials[2 * block + {0, 1}]; the block that draws the last ticket adds all `count`
// pairs in index order and stores the result.  Call from all threads of the block.
template <int THREADS>
__device__ __forceinline__ void cs_finish_partials(double* partials, unsigned int* ticket, int block, int count, double a, double b,
                                                   double* out) {
    __shared__ bool last;
    if (threadIdx.x == 0) {
        partials[2 * block] = a;
        partials[2 * block + 1] = b;
        __threadfence();
        last = atomicAdd(ticket, 1u) == (unsigned int)count - 1u;
    }
    __syncthreads();
    if (last) {
        __threadfence();
        double x = 0.0, y = 0.0;
        for (int k = threadIdx.x; k < count; k += THREADS) {
            x += __ldcg(partials + 2 * k);                // written by other blocks: read through L2, never a stale L1 line
            y += __ldcg(partials + 2 * k + 1);
        }
        cs_block_sum2<THREADS>(x, y);
        if (threadIdx.x == 0) {
            out[0] = x, out[1] = y;
            *ticket = 0;                                  // ready for the next launch on this stream
        }
    }
}
