// Activation dtype helpers: kernels are templated on T in {float, __nv_bfloat16}, compute in fp32.
#pragma once
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "common.cuh"

namespace p2p {

__device__ __forceinline__ float ldf(const float* p) { return __ldg(p); }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// four consecutive elements (16 B of fp32 / 8 B of bf16; the address must be aligned to that)
__device__ __forceinline__ void ld4(const float* p, float (&v)[4]) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void ld4(const __nv_bfloat16* p, float (&v)[4]) {
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(p));
    const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&t.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&t.y));
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st4(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void st4(__nv_bfloat16* p, const float (&v)[4]) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 t;
    t.x = *reinterpret_cast<const uint32_t*>(&a);
    t.y = *reinterpret_cast<const uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = t;
}
inline bool vec4_ok(const void* p, size_t elem) { return p == nullptr || (((uintptr_t)p) % (4 * elem)) == 0; }

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
