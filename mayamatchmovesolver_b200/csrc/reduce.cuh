// Deterministic warp / block reductions shared by the LM kernels.
#pragma once
#include <cuda_runtime.h>

namespace mmb {

constexpr int kMaxNA = 64;     // largest per-frame eliminated block

// ---- reductions --------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// deterministic block reduction (blockDim multiple of 32, <= 1024)
__device__ __forceinline__ double block_sum(double v, double *sh /*32*/) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    double r = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.0;
    if (w == 0) r = warp_sum(r);
    if (threadIdx.x == 0) sh[0] = r;
    __syncthreads();
    r = sh[0];
    return r;
}

}  // namespace mmb
