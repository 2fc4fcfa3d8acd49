// Activation dtype helpers: kernels are templated on T in {float, __nv_bfloat16}, compute in fp32.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "common.cuh"

namespace p2p {

__device__ __forceinline__ float ldf(const float* p) { return __ldg(p); }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// Dispatch a templated launcher on the C-ABI dtype code.
#define P2P_DISPATCH_DTYPE(dtype, T, ...)                                          \
    do {                                                                           \
        if ((dtype) == P2P_F32) { using T = float; __VA_ARGS__; }                  \
        else if ((dtype) == P2P_BF16) { using T = __nv_bfloat16; __VA_ARGS__; }    \
        else return p2p::fail(P2P_EINVAL, "unknown dtype code %d", (int)(dtype));  \
    } while (0)

#define P2P_DISPATCH_DTYPE2(dta, TA, dtb, TB, ...)                                  \
    do {                                                                           \
        if ((dta) == P2P_F32 && (dtb) == P2P_F32) { using TA = float; using TB = float; __VA_ARGS__; }                          \
        else if ((dta) == P2P_F32 && (dtb) == P2P_BF16) { using TA = float; using TB = __nv_bfloat16; __VA_ARGS__; }            \
        else if ((dta) == P2P_BF16 && (dtb) == P2P_F32) { using TA = __nv_bfloat16; using TB = float; __VA_ARGS__; }            \
        else if ((dta) == P2P_BF16 && (dtb) == P2P_BF16) { using TA = __nv_bfloat16; using TB = __nv_bfloat16; __VA_ARGS__; }   \
        else return p2p::fail(P2P_EINVAL, "unknown dtype codes %d, %d", (int)(dta), (int)(dtb));                                \
    } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide sum for blockDim.x <= 1024 (result valid in every thread)
__device__ __forceinline__ float block_sum(float v, float* scratch /* >= 33 floats */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    if (warp == 0) {
        float t = lane < nw ? scratch[lane] : 0.f;
        t = warp_sum(t);
        if (lane == 0) scratch[32] = t;
    }
    __syncthreads();
    return scratch[32];
}

}  // namespace p2p
