// Bandwidth-bound layer kernels of the pix2pixHD path (NHWC): InstanceNorm(+activation, +residual),
// AvgPool 3x3/s2, tanh, reflection-pad adjoint, bias gradient, LSGAN / L1 loss reductions, fused Adam.
// Reference call sites: models/networks.py:18-25 (InstanceNorm2d affine=False), :166,309 (AvgPool2d
// count_include_pad=False), :68-110 (GANLoss), pix2pixHD_model.py:390-398 (feature matching), train.py:164-184.
#include "dtype.cuh"
#include <mutex>

namespace p2p {
namespace {

constexpr int kEw = 256;   // threads of elementwise kernels

__device__ __forceinline__ float act_fwd(float x, int act, float slope) {
    if (act == 1) return x > 0.f ? x : 0.f;
    if (act == 2) return x > 0.f ? x : slope * x;
    return x;
}
__device__ __forceinline__ float act_grad(float x, int act, float slope) {
    if (act == 1) return x > 0.f ? 1.f : 0.f;
    if (act == 2) return x > 0.f ? 1.f : slope;
    return 1.f;
}

// ---- InstanceNorm statistics: raw[n,c] = (sum(y - shift), sum((y - shift)^2)), shift = y[n,0,c] -------------
// grid (ceil(C/32), N, S), block (32, 8): rows of the HW chunk are strided over threadIdx.y
template <typename T>
__global__ void in_stats_kernel(const T* __restrict__ y, float* __restrict__ raw, long long HW, int C, long long chunk) {
    __shared__ float s1[8][33], s2[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x, n = blockIdx.y;
    const long long r0 = (long long)blockIdx.z * chunk, r1 = min(HW, r0 + chunk);
    float a = 0.f, b = 0.f;
    if (c < C) {
        const T* base = y + (long long)n * HW * C + c;
        const float shift = ldf(base);
        for (long long r = r0 + threadIdx.y; r < r1; r += 8) {
            const float v = ldf(base + r * C) - shift;
            a += v; b += v * v;
        }
    }
    s1[threadIdx.y][threadIdx.x] = a; s2[threadIdx.y][threadIdx.x] = b;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
#pragma unroll
        for (int i = 1; i < 8; ++i) { a += s1[i][threadIdx.x]; b += s2[i][threadIdx.x]; }
        atomicAdd(raw + ((long long)n * C + c) * 2, a);
        atomicAdd(raw + ((long long)n * C + c) * 2 + 1, b);
    }
}

template <typename T>
__global__ void in_finalize_kernel(const T* __restrict__ y, float* __restrict__ stats, long long HW, int C, int NC, float eps) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= NC) return;
    const int n = i / C, c = i % C;
    const float shift = ldf(y + (long long)n * HW * C + c);
    const float m = stats[2 * i] / (float)HW;
    const float var = fmaxf(stats[2 * i + 1] / (float)HW - m * m, 0.f);   // biased variance (PyTorch instance_norm)
    stats[2 * i] = m + shift;
    stats[2 * i + 1] = rsqrtf(var + eps);
}

template <typename TY, typename T>
__global__ void in_apply_kernel(const TY* __restrict__ y, const T* __restrict__ res, T* __restrict__ z,
                                const float* __restrict__ stats, long long HW, int C, long long total, int norm,
                                int act, float slope) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        float v = ldf(y + i);
        if (norm) {
            const long long n = i / (HW * C);
            const float2 st = *reinterpret_cast<const float2*>(stats + (n * C + c) * 2);
            v = (v - st.x) * st.y;
        }
        v = act_fwd(v, act, slope);
        if (res) v += ldf(res + i);
        stf(z + i, v);
    }
}

// backward reductions: work[n,c] = (sum g, sum g*xhat), g = dz * act'(xhat)
template <typename TY, typename T>
__global__ void in_bwd_stats_kernel(const TY* __restrict__ y, const float* __restrict__ stats, const T* __restrict__ dz,
                                    float* __restrict__ work, long long HW, int C, long long chunk, int act, float slope) {
    __shared__ float s1[8][33], s2[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x, n = blockIdx.y;
    const long long r0 = (long long)blockIdx.z * chunk, r1 = min(HW, r0 + chunk);
    float a = 0.f, b = 0.f;
    if (c < C) {
        const long long off = (long long)n * HW * C + c;
        const float mean = stats[((long long)n * C + c) * 2], rstd = stats[((long long)n * C + c) * 2 + 1];
        for (long long r = r0 + threadIdx.y; r < r1; r += 8) {
            const float xh = (ldf(y + off + r * C) - mean) * rstd;
            const float g = ldf(dz + off + r * C) * act_grad(xh, act, slope);
            a += g; b += g * xh;
        }
    }
    s1[threadIdx.y][threadIdx.x] = a; s2[threadIdx.y][threadIdx.x] = b;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
#pragma unroll
        for (int i = 1; i < 8; ++i) { a += s1[i][threadIdx.x]; b += s2[i][threadIdx.x]; }
        atomicAdd(work + ((long long)n * C + c) * 2, a);
        atomicAdd(work + ((long long)n * C + c) * 2 + 1, b);
    }
}

template <typename TY, typename T>
__global__ void in_bwd_apply_kernel(const TY* __restrict__ y, const float* __restrict__ stats, const T* __restrict__ dz,
                                    const float* __restrict__ work, T* __restrict__ dy, long long HW, int C,
                                    long long total, int norm, int act, float slope) {
    const float inv = 1.f / (float)HW;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const float v = ldf(y + i), g0 = ldf(dz + i);
        if (!norm) { stf(dy + i, g0 * act_grad(v, act, slope)); continue; }
        const int c = (int)(i % C);
        const long long nc = (i / (HW * C)) * C + c;
        const float2 st = *reinterpret_cast<const float2*>(stats + nc * 2);
        const float2 wk = *reinterpret_cast<const float2*>(work + nc * 2);
        const float xh = (v - st.x) * st.y;
        const float g = g0 * act_grad(xh, act, slope);
        stf(dy + i, st.y * (g - wk.x * inv - xh * wk.y * inv));
    }
}


// ---- four-channel (vector) forms of the InstanceNorm kernels: C % 4 == 0, aligned pointers, < 2^31 vectors ----------
// grid (ceil(C4/32), N, S), block (32, 8)
template <typename T>
__global__ void __launch_bounds__(256) in_stats_v4_kernel(const T* __restrict__ y, float* __restrict__ raw, int HW, int C, int chunk) {
    __shared__ float sm[8][32][9];
    const int cg = blockIdx.x * 32 + threadIdx.x, n = blockIdx.y, C4 = C >> 2;
    const int r0 = blockIdx.z * chunk, r1 = min(HW, r0 + chunk);
    float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
    if (cg < C4) {
        const T* base = y + (long long)n * HW * C + cg * 4;
        float shift[4];
        ld4(base, shift);
        int r = r0 + threadIdx.y;
        for (; r + 24 < r1; r += 32) {           // four rows in flight per thread
            float v[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) ld4(base + (long long)(r + 8 * u) * C, v[u]);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < 4; ++i) { const float d = v[u][i] - shift[i]; a[i] += d; b[i] += d * d; }
        }
        for (; r < r1; r += 8) {
            float v[4];
            ld4(base + (long long)r * C, v);
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float d = v[i] - shift[i]; a[i] += d; b[i] += d * d; }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { sm[threadIdx.y][threadIdx.x][i] = a[i]; sm[threadIdx.y][threadIdx.x][4 + i] = b[i]; }
    __syncthreads();
    // 256 threads: one (column group, component) pair each
    const int t = threadIdx.y * 32 + threadIdx.x, g = t >> 3, k = t & 7, cgo = blockIdx.x * 32 + g;
    if (cgo < C4) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += sm[i][g][k];
        atomicAdd(raw + ((long long)n * C + cgo * 4 + (k & 3)) * 2 + (k >> 2), s);
    }
}

// Single-launch statistics (round 2): the CTAs of one (image, 128-channel block) add their partial sums into a zeroed
// workspace; the LAST of them to finish (self-resetting atomicInc counter) turns the sums into (mean, rstd) -- or, for the
// backward pass, copies them out -- and zeroes the workspace again.  Replaces memset + reduction + finalize (3 graph nodes).
// ws: [N][C][2] floats, cnt: [N][gridDim.x] counters, both zero between calls (one stream: calls are ordered).
__device__ __forceinline__ bool last_cta_of_group(unsigned* cnt, unsigned S) {
    __shared__ unsigned s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0 && threadIdx.y == 0) s_last = atomicInc(cnt, S - 1) == S - 1;
    __syncthreads();
    return s_last != 0;
}

template <typename T>
__global__ void __launch_bounds__(256) in_stats_fused_v4_kernel(const T* __restrict__ y, float* __restrict__ stats, float* __restrict__ ws,
                                                                unsigned* __restrict__ cnt, int HW, int C, int chunk, float eps) {
    pdl_trigger();
    pdl_wait();
    // LX channel groups x RY rows of threads: 32 x 8 on a full channel block; a block with fewer than 32 groups (C < 128, or
    // the last block of C % 128 != 0) spreads its 256 threads over more rows instead of leaving lanes idle
    __shared__ float sm[256 * 9];
    const int n = blockIdx.y, C4 = C >> 2;
    const int t = threadIdx.y * 32 + threadIdx.x;
    const int LX = min(32, C4 - (int)blockIdx.x * 32), RY = 256 / LX;
    const int cx = t % LX, ry = t / LX, cg = blockIdx.x * 32 + cx;
    const int r0 = blockIdx.z * chunk, r1 = min(HW, r0 + chunk);
    float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
    if (ry < RY) {
        const T* base = y + (long long)n * HW * C + cg * 4;
        float shift[4];
        ld4(base, shift);
        int r = r0 + ry;
        for (; r + 3 * RY < r1; r += 4 * RY) {           // four rows in flight per thread
            float v[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) ld4(base + (long long)(r + RY * u) * C, v[u]);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < 4; ++i) { const float d = v[u][i] - shift[i]; a[i] += d; b[i] += d * d; }
        }
        for (; r < r1; r += RY) {
            float v[4];
            ld4(base + (long long)r * C, v);
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float d = v[i] - shift[i]; a[i] += d; b[i] += d * d; }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { sm[t * 9 + i] = a[i]; sm[t * 9 + 4 + i] = b[i]; }
    __syncthreads();
    const int g = t >> 3, k = t & 7, cgo = blockIdx.x * 32 + g;
    const long long wi = ((long long)n * C + cgo * 4 + (k & 3)) * 2 + (k >> 2);
    if (g < LX) {
        float s = 0.f;
        for (int i = 0; i < RY; ++i) s += sm[(i * LX + g) * 9 + k];
        if (gridDim.z > 1) atomicAdd(ws + wi, s); else ws[wi] = s;
    }
    if (!last_cta_of_group(cnt + (long long)n * gridDim.x + blockIdx.x, gridDim.z)) return;
    // finalize this (image, channel block): thread t owns (channel group g, component k); pair lanes k and k + 4 of a channel
    float tot = 0.f;
    if (cgo < C4) { tot = __ldcg(ws + wi); ws[wi] = 0.f; }
    const float other = __shfl_xor_sync(0xffffffffu, tot, 4);            // sum <-> sum of squares of the same channel
    if (cgo < C4 && k < 4) {
        const int c = cgo * 4 + k;
        const float shift = ldf(y + (long long)n * HW * C + c);
        const float m = tot / (float)HW;
        const float var = fmaxf(other / (float)HW - m * m, 0.f);          // biased variance (PyTorch instance_norm)
        *reinterpret_cast<float2*>(stats + ((long long)n * C + c) * 2) = make_float2(m + shift, rsqrtf(var + eps));
    }
}

template <typename TY, typename T>
__global__ void __launch_bounds__(256) in_bwd_stats_fused_v4_kernel(const TY* __restrict__ y, const float* __restrict__ stats,
                                                                    const T* __restrict__ dz, float* __restrict__ work,
                                                                    float* __restrict__ ws, unsigned* __restrict__ cnt, int HW, int C,
                                                                    int chunk, int act, float slope) {
    pdl_trigger();
    pdl_wait();
    __shared__ float sm[256 * 9];                // thread layout: see in_stats_fused_v4_kernel
    const int n = blockIdx.y, C4 = C >> 2;
    const int t = threadIdx.y * 32 + threadIdx.x;
    const int LX = min(32, C4 - (int)blockIdx.x * 32), RY = 256 / LX;
    const int cx = t % LX, ry = t / LX, cg = blockIdx.x * 32 + cx;
    const int r0 = blockIdx.z * chunk, r1 = min(HW, r0 + chunk);
    float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
    if (ry < RY) {
        const long long off = (long long)n * HW * C + cg * 4;
        const float4* sp = reinterpret_cast<const float4*>(stats + ((long long)n * C + cg * 4) * 2);
        const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1);
        const float mean[4] = {s0.x, s0.z, s1.x, s1.z}, rstd[4] = {s0.y, s0.w, s1.y, s1.w};
        int r = r0 + ry;
        for (; r + 3 * RY < r1; r += 4 * RY) {           // four rows in flight per thread
            float v[4][4], g[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                ld4(y + off + (long long)(r + RY * u) * C, v[u]);
                ld4(dz + off + (long long)(r + RY * u) * C, g[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float xh = (v[u][i] - mean[i]) * rstd[i];
                    const float gg = g[u][i] * act_grad(xh, act, slope);
                    a[i] += gg; b[i] += gg * xh;
                }
        }
        for (; r < r1; r += RY) {
            float v[4], g[4];
            ld4(y + off + (long long)r * C, v);
            ld4(dz + off + (long long)r * C, g);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float xh = (v[i] - mean[i]) * rstd[i];
                const float gg = g[i] * act_grad(xh, act, slope);
                a[i] += gg; b[i] += gg * xh;
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { sm[t * 9 + i] = a[i]; sm[t * 9 + 4 + i] = b[i]; }
    __syncthreads();
    const int g = t >> 3, k = t & 7, cgo = blockIdx.x * 32 + g;
    const long long wi = ((long long)n * C + cgo * 4 + (k & 3)) * 2 + (k >> 2);
    if (g < LX) {
        float s = 0.f;
        for (int i = 0; i < RY; ++i) s += sm[(i * LX + g) * 9 + k];
        if (gridDim.z > 1) atomicAdd(ws + wi, s); else ws[wi] = s;
    }
    if (!last_cta_of_group(cnt + (long long)n * gridDim.x + blockIdx.x, gridDim.z)) return;
    if (g < LX) { work[wi] = __ldcg(ws + wi); ws[wi] = 0.f; }
}

template <typename TY, typename T>
__global__ void __launch_bounds__(256) in_apply_v4_kernel(const TY* __restrict__ y, const T* __restrict__ res, T* __restrict__ z,
                                                          const float* __restrict__ stats, unsigned HWC4, unsigned C4,
                                                          unsigned total4, int norm, int act, float slope) {
    pdl_trigger();
    pdl_wait();
#pragma unroll 2
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
        float v[4];
        ld4(y + 4ll * i, v);
        if (norm) {
            const unsigned n = i / HWC4, c4 = i % C4;
            const float4* sp = reinterpret_cast<const float4*>(stats + ((long long)n * C4 + c4) * 8);
            const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1);
            v[0] = (v[0] - s0.x) * s0.y; v[1] = (v[1] - s0.z) * s0.w;
            v[2] = (v[2] - s1.x) * s1.y; v[3] = (v[3] - s1.z) * s1.w;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = act_fwd(v[k], act, slope);
        if (res) {
            float r[4];
            ld4(res + 4ll * i, r);
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] += r[k];
        }
        st4(z + 4ll * i, v);
    }
}

template <typename TY, typename T>
__global__ void __launch_bounds__(256) in_bwd_stats_v4_kernel(const TY* __restrict__ y, const float* __restrict__ stats,
                                                              const T* __restrict__ dz, float* __restrict__ work, int HW, int C,
                                                              int chunk, int act, float slope) {
    __shared__ float sm[8][32][9];
    const int cg = blockIdx.x * 32 + threadIdx.x, n = blockIdx.y, C4 = C >> 2;
    const int r0 = blockIdx.z * chunk, r1 = min(HW, r0 + chunk);
    float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
    if (cg < C4) {
        const long long off = (long long)n * HW * C + cg * 4;
        const float4* sp = reinterpret_cast<const float4*>(stats + ((long long)n * C + cg * 4) * 2);
        const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1);
        const float mean[4] = {s0.x, s0.z, s1.x, s1.z}, rstd[4] = {s0.y, s0.w, s1.y, s1.w};
        int r = r0 + threadIdx.y;
        for (; r + 24 < r1; r += 32) {           // four rows in flight per thread
            float v[4][4], g[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                ld4(y + off + (long long)(r + 8 * u) * C, v[u]);
                ld4(dz + off + (long long)(r + 8 * u) * C, g[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float xh = (v[u][i] - mean[i]) * rstd[i];
                    const float gg = g[u][i] * act_grad(xh, act, slope);
                    a[i] += gg; b[i] += gg * xh;
                }
        }
        for (; r < r1; r += 8) {
            float v[4], g[4];
            ld4(y + off + (long long)r * C, v);
            ld4(dz + off + (long long)r * C, g);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float xh = (v[i] - mean[i]) * rstd[i];
                const float gg = g[i] * act_grad(xh, act, slope);
                a[i] += gg; b[i] += gg * xh;
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { sm[threadIdx.y][threadIdx.x][i] = a[i]; sm[threadIdx.y][threadIdx.x][4 + i] = b[i]; }
    __syncthreads();
    const int t = threadIdx.y * 32 + threadIdx.x, g = t >> 3, k = t & 7, cgo = blockIdx.x * 32 + g;
    if (cgo < C4) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += sm[i][g][k];
        atomicAdd(work + ((long long)n * C + cgo * 4 + (k & 3)) * 2 + (k >> 2), s);
    }
}

template <typename TY, typename T>
__global__ void __launch_bounds__(256) in_bwd_apply_v4_kernel(const TY* __restrict__ y, const float* __restrict__ stats,
                                                              const T* __restrict__ dz, const float* __restrict__ work,
                                                              T* __restrict__ dy, unsigned HWC4, unsigned C4, unsigned total4,
                                                              float inv, int norm, int act, float slope) {
    pdl_trigger();
    pdl_wait();
#pragma unroll 2
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
        float v[4], g[4], o[4];
        ld4(y + 4ll * i, v);
        ld4(dz + 4ll * i, g);
        if (!norm) {
#pragma unroll
            for (int k = 0; k < 4; ++k) o[k] = g[k] * act_grad(v[k], act, slope);
        } else {
            const unsigned n = i / HWC4, c4 = i % C4;
            const long long so = ((long long)n * C4 + c4) * 8;
            const float4 s0 = __ldg(reinterpret_cast<const float4*>(stats + so)), s1 = __ldg(reinterpret_cast<const float4*>(stats + so) + 1);
            const float4 w0 = __ldg(reinterpret_cast<const float4*>(work + so)), w1 = __ldg(reinterpret_cast<const float4*>(work + so) + 1);
            const float mean[4] = {s0.x, s0.z, s1.x, s1.z}, rstd[4] = {s0.y, s0.w, s1.y, s1.w};
            const float wa[4] = {w0.x, w0.z, w1.x, w1.z}, wb[4] = {w0.y, w0.w, w1.y, w1.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float xh = (v[k] - mean[k]) * rstd[k];
                const float gg = g[k] * act_grad(xh, act, slope);
                o[k] = rstd[k] * (gg - wa[k] * inv - xh * wb[k] * inv);
            }
        }
        st4(dy + 4ll * i, o);
    }
}

// in_bwd_apply_v4_kernel writing dy straight into the layout the stride-1 gradient kernels read (ops._dy_wrap): a grid of
// `rows` pixel rows at the pitch of the PADDED input, image n at row n*Hh*Wh, dy[n,h,w] at +h*Wh+w, zeros everywhere else.
template <typename TY, typename T>
__global__ void __launch_bounds__(256) in_bwd_apply_wrap_v4_kernel(const TY* __restrict__ y, const float* __restrict__ stats,
                                                                   const T* __restrict__ dz, const float* __restrict__ work,
                                                                   T* __restrict__ grid, int N, int H, int W, unsigned C4, int Hh, int Wh,
                                                                   unsigned total4, float inv, int norm, int act, float slope) {
    pdl_trigger();
    pdl_wait();
    const unsigned plane = (unsigned)(Hh * Wh);
#pragma unroll 2
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
        const unsigned c4 = i % C4, r = i / C4;
        const unsigned n = r / plane, rem = r - n * plane;
        const unsigned h = rem / (unsigned)Wh, w = rem - h * (unsigned)Wh;
        float o[4] = {0.f, 0.f, 0.f, 0.f};
        if (n < (unsigned)N && h < (unsigned)H && w < (unsigned)W) {
            const long long src = ((((long long)n * H + h) * W + w) * C4 + c4) * 4;
            float v[4], g[4];
            ld4(y + src, v);
            ld4(dz + src, g);
            if (!norm) {
#pragma unroll
                for (int k = 0; k < 4; ++k) o[k] = g[k] * act_grad(v[k], act, slope);
            } else {
                const long long so = ((long long)n * C4 + c4) * 8;
                const float4 s0 = __ldg(reinterpret_cast<const float4*>(stats + so)), s1 = __ldg(reinterpret_cast<const float4*>(stats + so) + 1);
                const float4 w0 = __ldg(reinterpret_cast<const float4*>(work + so)), w1 = __ldg(reinterpret_cast<const float4*>(work + so) + 1);
                const float mean[4] = {s0.x, s0.z, s1.x, s1.z}, rstd[4] = {s0.y, s0.w, s1.y, s1.w};
                const float wa[4] = {w0.x, w0.z, w1.x, w1.z}, wb[4] = {w0.y, w0.w, w1.y, w1.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float xh = (v[k] - mean[k]) * rstd[k];
                    const float gg = g[k] * act_grad(xh, act, slope);
                    o[k] = rstd[k] * (gg - wa[k] * inv - xh * wb[k] * inv);
                }
            }
        }
        st4(grid + 4ll * i, o);
    }
}

// Round 2 (second half): the same arithmetic with one (image, channel group) per THREAD.  grid (gx, N), 256 threads, 256*gx a
// multiple of C4: thread t = blockIdx.x*256 + threadIdx.x keeps channel group t % C4 for its whole life and walks the rows
// t / C4, + 256*gx/C4, ...; the four (mean, rstd) and (sum g, sum g*xhat) pairs are loaded once instead of per element, and the
// per-element divisions are gone.  WRAP: rows are those of the zero-margined grid at the padded input's pitch (see
// in_bwd_apply_wrap_v4_kernel), margin rows are written as zeros; the slice blockIdx.y == N zeroes the tail of the grid.
template <typename TY, typename T, bool WRAP>
__global__ void __launch_bounds__(256) in_bwd_apply_rows_v4_kernel(const TY* __restrict__ y, const float* __restrict__ stats,
                                                                   const T* __restrict__ dz, const float* __restrict__ work,
                                                                   T* __restrict__ out, int N, int H, int W, unsigned C4, int Hh, int Wh,
                                                                   long long grid_rows, float inv, int norm, int act, float slope) {
    pdl_trigger();
    pdl_wait();
    const unsigned t = blockIdx.x * 256u + threadIdx.x, c4 = t % C4, row0 = t / C4, rstep = gridDim.x * 256u / C4;
    const unsigned n = blockIdx.y;
    const float zero4[4] = {0.f, 0.f, 0.f, 0.f};
    if (WRAP && n == (unsigned)N) {                      // tail of the grid beyond the last image
        for (long long q = (long long)N * Hh * Wh + row0; q < grid_rows; q += rstep) st4(out + (q * C4 + c4) * 4, zero4);
        return;
    }
    float mean[4] = {0.f, 0.f, 0.f, 0.f}, rstd[4] = {1.f, 1.f, 1.f, 1.f}, wa[4] = {0.f, 0.f, 0.f, 0.f}, wb[4] = {0.f, 0.f, 0.f, 0.f};
    if (norm) {
        const long long so = ((long long)n * C4 + c4) * 8;
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(stats + so)), s1 = __ldg(reinterpret_cast<const float4*>(stats + so) + 1);
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(work + so)), w1 = __ldg(reinterpret_cast<const float4*>(work + so) + 1);
        mean[0] = s0.x; mean[1] = s0.z; mean[2] = s1.x; mean[3] = s1.z;
        rstd[0] = s0.y; rstd[1] = s0.w; rstd[2] = s1.y; rstd[3] = s1.w;
        wa[0] = w0.x; wa[1] = w0.z; wa[2] = w1.x; wa[3] = w1.z;
        wb[0] = w0.y; wb[1] = w0.w; wb[2] = w1.y; wb[3] = w1.w;
    }
    const unsigned HW = (unsigned)(H * W), rows = WRAP ? (unsigned)(Hh * Wh) : HW;
    const TY* yb = y + ((long long)n * HW * C4 + c4) * 4;
    const T* zb = dz + ((long long)n * HW * C4 + c4) * 4;
    T* ob = out + ((long long)n * rows * C4 + c4) * 4;
#pragma unroll 2
    for (unsigned q = row0; q < rows; q += rstep) {
        unsigned r = q;
        if (WRAP) {
            const unsigned h = q / (unsigned)Wh, w = q - h * (unsigned)Wh;
            if (h >= (unsigned)H || w >= (unsigned)W) { st4(ob + (long long)q * C4 * 4, zero4); continue; }
            r = h * (unsigned)W + w;
        }
        float v[4], g[4], o[4];
        ld4(yb + (long long)r * C4 * 4, v);
        ld4(zb + (long long)r * C4 * 4, g);
        if (!norm) {
#pragma unroll
            for (int k = 0; k < 4; ++k) o[k] = g[k] * act_grad(v[k], act, slope);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float xh = (v[k] - mean[k]) * rstd[k];
                const float gg = g[k] * act_grad(xh, act, slope);
                o[k] = rstd[k] * (gg - wa[k] * inv - xh * wb[k] * inv);
            }
        }
        st4(ob + (long long)q * C4 * 4, o);
    }
}

// grid.x for the kernel above: ~16 CTAs per SM in total, a multiple of C4 / gcd(C4, 256), no more threads than (row, group) pairs
inline unsigned rows_grid_x(unsigned C4, long long rows, int slices) {
    unsigned a = C4, b = 256;
    while (b) { const unsigned r = a % b; a = b; b = r; }
    const unsigned m = C4 / a;                                       // 256 * gx % C4 == 0  <=>  gx % m == 0
    long long want = (148LL * 16 + slices - 1) / slices;
    const long long cap = (rows * C4 + 255) / 256;
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    return (unsigned)((want + m - 1) / m * m);
}

// adjoint of ReflectionPad2d, four channels per thread
template <typename TI, typename T>
__global__ void __launch_bounds__(256) reflect_fold_v4_kernel(const TI* __restrict__ dxp, T* __restrict__ dx, int H, int W, int C4,
                                                              int p, unsigned total4) {
    pdl_trigger();
    pdl_wait();
    const int HP = H + 2 * p, WP = W + 2 * p;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
        const unsigned c4 = i % C4;
        unsigned r = i / C4;
        const int w = (int)(r % W); r /= W;
        const int h = (int)(r % H);
        const long long n = r / H;
        int hs[3], ws[3], nh = 0, nw = 0;
        hs[nh++] = h + p;
        if (h >= 1 && h <= p) hs[nh++] = p - h;
        if (h >= H - 1 - p && h <= H - 2) hs[nh++] = 2 * (H - 1) + p - h;
        ws[nw++] = w + p;
        if (w >= 1 && w <= p) ws[nw++] = p - w;
        if (w >= W - 1 - p && w <= W - 2) ws[nw++] = 2 * (W - 1) + p - w;
        float s[4] = {0.f, 0.f, 0.f, 0.f};
        for (int a = 0; a < nh; ++a)
            for (int b = 0; b < nw; ++b) {
                float v[4];
                ld4(dxp + (((n * HP + hs[a]) * WP + ws[b]) * C4 + c4) * 4, v);
#pragma unroll
                for (int k = 0; k < 4; ++k) s[k] += v[k];
            }
        st4(dx + 4ll * i, s);
    }
}

// ---- AvgPool2d(3, stride 2, pad 1, count_include_pad=False) -------------------------------------------
template <typename T>
__global__ void avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int H, int W, int C, int OH, int OW,
                                   long long total) {
    pdl_trigger();
    pdl_wait();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        long long r = i / C;
        const int ow = (int)(r % OW); r /= OW;
        const int oh = (int)(r % OH);
        const long long n = r / OH;
        const int h0 = max(2 * oh - 1, 0), h1 = min(2 * oh + 1, H - 1), w0 = max(2 * ow - 1, 0), w1 = min(2 * ow + 1, W - 1);
        float s = 0.f;
        for (int h = h0; h <= h1; ++h)
            for (int w = w0; w <= w1; ++w) s += ldf(x + ((n * H + h) * W + w) * C + c);
        stf(y + i, s / (float)((h1 - h0 + 1) * (w1 - w0 + 1)));
    }
}

template <typename T, typename I>        // I: index type (unsigned when the tensor has < 2^31 elements: 64-bit divisions are slow)
__global__ void avgpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int H, int W, int C, int OH, int OW,
                                   long long total) {
    pdl_trigger();
    pdl_wait();
    for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < (I)total; i += (I)gridDim.x * blockDim.x) {
        const int c = (int)(i % (I)C);
        I r = i / (I)C;
        const int w = (int)(r % (I)W); r /= (I)W;
        const int h = (int)(r % (I)H);
        const long long n = (long long)(r / (I)H);
        // outputs whose window [2o-1, 2o+1] contains h: o in [ceil((h-1)/2), floor((h+1)/2)]
        float s = 0.f;
        for (int oh = h / 2; oh <= min((h + 1) / 2, OH - 1); ++oh) {
            const int ch = min(2 * oh + 1, H - 1) - max(2 * oh - 1, 0) + 1;
            for (int ow = w / 2; ow <= min((w + 1) / 2, OW - 1); ++ow) {
                const int cw = min(2 * ow + 1, W - 1) - max(2 * ow - 1, 0) + 1;
                s += ldf(dy + ((n * OH + oh) * OW + ow) * C + c) / (float)(ch * cw);
            }
        }
        stf(dx + i, s);
    }
}

// ---- elementwise ------------------------------------------------------------------------------------------
template <typename T>
__global__ void tanh_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long n) {
    pdl_trigger();
    pdl_wait();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        stf(y + i, tanhf(ldf(x + i)));
}
template <typename T>
__global__ void tanh_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, long long n) {
    pdl_trigger();
    pdl_wait();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float t = ldf(y + i);
        stf(dx + i, ldf(dy + i) * (1.f - t * t));
    }
}

template <typename TI, typename T>
__global__ void reflect_fold_kernel(const TI* __restrict__ dxp, T* __restrict__ dx, int H, int W, int C, int p,
                                    long long total) {
    const int HP = H + 2 * p, WP = W + 2 * p;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        long long r = i / C;
        const int w = (int)(r % W); r /= W;
        const int h = (int)(r % H);
        const long long n = r / H;
        int hs[3], ws[3], nh = 0, nw = 0;
        hs[nh++] = h + p;
        if (h >= 1 && h <= p) hs[nh++] = p - h;
        if (h >= H - 1 - p && h <= H - 2) hs[nh++] = 2 * (H - 1) + p - h;
        ws[nw++] = w + p;
        if (w >= 1 && w <= p) ws[nw++] = p - w;
        if (w >= W - 1 - p && w <= W - 2) ws[nw++] = 2 * (W - 1) + p - w;
        float s = 0.f;
        for (int a = 0; a < nh; ++a)
            for (int b = 0; b < nw; ++b) s += ldf(dxp + ((n * HP + hs[a]) * WP + ws[b]) * C + c);
        stf(dx + i, s);
    }
}

// dbias[j] += column sums; grid (ceil(J/32), S), block (32, 8)
template <typename T>
__global__ void bias_grad_kernel(const T* __restrict__ dout, float* __restrict__ dbias, long long rows, int J, long long chunk) {
    pdl_trigger();
    pdl_wait();
    __shared__ float s1[8][33];
    const int j = blockIdx.x * 32 + threadIdx.x;
    const long long r0 = (long long)blockIdx.y * chunk, r1 = min(rows, r0 + chunk);
    float a = 0.f;
    if (j < J)
        for (long long r = r0 + threadIdx.y; r < r1; r += 8) a += ldf(dout + r * J + j);
    s1[threadIdx.y][threadIdx.x] = a;
    __syncthreads();
    if (threadIdx.y == 0 && j < J) {
#pragma unroll
        for (int i = 1; i < 8; ++i) a += s1[i][threadIdx.x];
        atomicAdd(dbias + j, a);
    }
}

// ---- losses ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void mse_const_kernel(const T* __restrict__ x, float target, float wgt, float* __restrict__ loss,
                                 T* __restrict__ dx, float gcoef, const float* __restrict__ gdev, long long n) {
    pdl_trigger();
    pdl_wait();
    __shared__ float scratch[33];
    float acc = 0.f;
    if (gdev) gcoef *= __ldg(gdev);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float d = ldf(x + i) - target;
        acc += d * d;
        if (dx) stf(dx + i, gcoef * d);
    }
    if (loss) { acc = block_sum(acc, scratch); if (threadIdx.x == 0) atomicAdd(loss, acc * wgt / (float)n); }
}

template <typename T>
__global__ void l1_kernel(const T* __restrict__ a, const T* __restrict__ b, float wgt, float* __restrict__ loss,
                          T* __restrict__ da, float gcoef, const float* __restrict__ gdev, long long n) {
    pdl_trigger();
    pdl_wait();
    __shared__ float scratch[33];
    float acc = 0.f;
    if (gdev) gcoef *= __ldg(gdev);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float d = ldf(a + i) - ldf(b + i);
        acc += fabsf(d);
        if (da) stf(da + i, d > 0.f ? gcoef : (d < 0.f ? -gcoef : 0.f));
    }
    if (loss) { acc = block_sum(acc, scratch); if (threadIdx.x == 0) atomicAdd(loss, acc * wgt / (float)n); }
}

__global__ void adam_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            __nv_bfloat16* __restrict__ wb, long long n, float step_size, float beta1, float beta2,
                            float inv_bc2_sqrt, float eps) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float gi = g[i];
        const float mi = beta1 * m[i] + (1.f - beta1) * gi;
        const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
        m[i] = mi; v[i] = vi;
        const float wi = w[i] - step_size * mi / (sqrtf(vi) * inv_bc2_sqrt + eps);
        w[i] = wi;
        if (wb) wb[i] = __float2bfloat16_rn(wi);
    }
}

// device-resident step counter and learning rate: the whole optimisation step can live in one CUDA graph
__global__ void adam_prepare_kernel(int* step, const float* lr, float beta1, float beta2, float* scratch) {
    const int t = ++(*step);
    scratch[0] = (float)((double)lr[0] / (1.0 - pow((double)beta1, (double)t)));
    scratch[1] = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)t)));
}

__global__ void adam_dev_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                                __nv_bfloat16* __restrict__ wb, long long n, const float* __restrict__ scratch, float beta1,
                                float beta2, float eps, float gscale) {
    const float step_size = scratch[0], inv_bc2_sqrt = scratch[1];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float gi = g[i] * gscale;
        const float mi = beta1 * m[i] + (1.f - beta1) * gi;
        const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
        m[i] = mi; v[i] = vi;
        const float wi = w[i] - step_size * mi / (sqrtf(vi) * inv_bc2_sqrt + eps);
        w[i] = wi;
        if (wb) wb[i] = __float2bfloat16_rn(wi);
    }
}

__global__ void __launch_bounds__(128) adam_dev_v4_kernel(float4* __restrict__ w, const float4* __restrict__ g, float4* __restrict__ m,
                                                          float4* __restrict__ v, uint2* __restrict__ wb, long long n4,
                                                          const float* __restrict__ scratch, float beta1, float beta2, float eps,
                                                          float gscale) {
    const float step_size = scratch[0], inv_bc2_sqrt = scratch[1];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 gq = g[i], mq = m[i], vq = v[i], wq = w[i];
        const float gi[4] = {gq.x * gscale, gq.y * gscale, gq.z * gscale, gq.w * gscale};
        const float mo[4] = {mq.x, mq.y, mq.z, mq.w}, vo[4] = {vq.x, vq.y, vq.z, vq.w}, wo[4] = {wq.x, wq.y, wq.z, wq.w};
        float mi[4], vi[4], wi[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            mi[k] = beta1 * mo[k] + (1.f - beta1) * gi[k];
            vi[k] = beta2 * vo[k] + (1.f - beta2) * gi[k] * gi[k];
            wi[k] = wo[k] - step_size * mi[k] / (sqrtf(vi[k]) * inv_bc2_sqrt + eps);
        }
        m[i] = make_float4(mi[0], mi[1], mi[2], mi[3]);
        v[i] = make_float4(vi[0], vi[1], vi[2], vi[3]);
        w[i] = make_float4(wi[0], wi[1], wi[2], wi[3]);
        if (wb) {
            const __nv_bfloat162 a = __floats2bfloat162_rn(wi[0], wi[1]), b = __floats2bfloat162_rn(wi[2], wi[3]);
            uint2 t;
            t.x = *reinterpret_cast<const uint32_t*>(&a);
            t.y = *reinterpret_cast<const uint32_t*>(&b);
            wb[i] = t;
        }
    }
}

inline int ew_grid(long long n) {
    long long b = (n + kEw - 1) / kEw;
    const long long cap = 148LL * 16;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
inline bool bwd_rows_form() {          // P2P_IN_BWD_ROWS=0: the per-element kernels of round 1
    static const bool on = [] { const char* e = getenv("P2P_IN_BWD_ROWS"); return !(e && e[0] == '0'); }();
    return on;
}
inline int split_for(long long rows, int base_ctas) {      // chunks of the row dimension so that ~4 CTAs/SM exist
    if (deterministic()) return 1;                         // one writer per statistic
    long long s = (148LL * 4 + base_ctas - 1) / base_ctas;
    const long long maxs = (rows + 63) / 64;
    if (s > maxs) s = maxs;
    if (s < 1) s = 1;
    if (s > 65535) s = 65535;
    return (int)s;
}

}  // namespace
}  // namespace p2p

using namespace p2p;

namespace p2p {
float* g_stats_ws = nullptr;        // [>= N*C*2] floats, zero between calls (p2p_instnorm_workspace)
unsigned* g_stats_cnt = nullptr;    // [>= N * ceil(C/128)] counters, zero between calls
long long g_stats_ws_floats = 0, g_stats_cnt_words = 0;
int g_stats_ws_device = -1;
inline bool ws_fits(int N, int C) {
    if (!g_stats_ws || (long long)N * C * 2 > g_stats_ws_floats || (long long)N * ((C / 4 + 31) / 32) > g_stats_cnt_words) return false;
    int dev = -1;
    cudaGetDevice(&dev);
    return dev == g_stats_ws_device;
}
}  // namespace p2p

// Caller-owned, ZEROED device scratch for the single-launch InstanceNorm reductions (the kernels leave it zeroed): ws_floats
// floats at ws, cnt_words 32-bit words at cnt.  Without it (or when a tensor does not fit) the three-node form runs.
// One workspace serves every layer: the calls of a step are ordered on one stream.
// Scratch query: bytes of the two regions p2p_instnorm_workspace wants for tensors of up to N images x C channels
// (*ws_bytes: fp32 partial sums, *cnt_bytes: 32-bit group counters).  Every other entry point takes its scratch as explicit
// arguments sized by the formulas in this header; this is the only library-held workspace.
extern "C" int p2p_workspace_bytes(int N, int C, long long* ws_bytes, long long* cnt_bytes) {
    if (N < 0 || C < 0 || !ws_bytes || !cnt_bytes) return fail(P2P_EINVAL, "bad argument");
    *ws_bytes = 2LL * N * C * (long long)sizeof(float);
    *cnt_bytes = (long long)N * ((C / 4 + 31) / 32) * (long long)sizeof(unsigned);
    return P2P_OK;
}

extern "C" int p2p_instnorm_workspace(float* ws, long long ws_floats, unsigned* cnt, long long cnt_words) {
    g_stats_ws = ws; g_stats_ws_floats = ws ? ws_floats : 0;
    g_stats_cnt = cnt; g_stats_cnt_words = cnt ? cnt_words : 0;
    g_stats_ws_device = -1;
    if (ws && cnt) cudaGetDevice(&g_stats_ws_device);       // the workspace serves the device that is current now
    return P2P_OK;
}


extern "C" int p2p_instnorm_act_fwd(const void* y, const void* res, void* z, float* stats, int N, long long HW, int C,
                                    int norm, int act, float slope, float eps, int y_dtype, int dtype, void* stream) {
    if (N < 0 || HW < 0 || C < 0) return fail(P2P_EINVAL, "negative size");
    const long long total = (long long)N * HW * C;
    if (total == 0) return P2P_OK;
    if (!y || !z || (norm && !stats)) return fail(P2P_EINVAL, "null pointer argument");
    if (N > 65535) return fail(P2P_ENOTSUP, "N=%d too large", N);
    cudaStream_t st = (cudaStream_t)stream;
    const bool v4 = C % 4 == 0 && total / 4 < 2000000000LL && HW < 2000000000LL;
    P2P_DISPATCH_DTYPE2(y_dtype, TY, dtype, T, {
        const bool vec = v4 && vec4_ok(y, sizeof(TY)) && vec4_ok(res, sizeof(T)) && vec4_ok(z, sizeof(T)) && vec4_ok(stats, 4);
        if (norm && vec && ws_fits(N, C)) {
            const int cb = (C / 4 + 31) / 32, S = split_for(HW, cb * N);
            const int chunk = (int)((HW + S - 1) / S);
            launch_k(in_stats_fused_v4_kernel<TY>, dim3(cb, N, S), dim3(32, 8), 0, st, (const TY*)y, stats, g_stats_ws, g_stats_cnt, (int)HW, C, chunk, eps);
        } else if (norm) {
            if (cudaMemsetAsync(stats, 0, sizeof(float) * 2 * N * C, st) != cudaSuccess) return fail(P2P_ECUDA, "memset failed");
            if (vec) {
                const int cb = (C / 4 + 31) / 32, S = split_for(HW, cb * N);
                const int chunk = (int)((HW + S - 1) / S);
                in_stats_v4_kernel<TY><<<dim3(cb, N, S), dim3(32, 8), 0, st>>>((const TY*)y, stats, (int)HW, C, chunk);
            } else {
                const int cb = (C + 31) / 32, S = split_for(HW, cb * N);
                const long long chunk = (HW + S - 1) / S;
                in_stats_kernel<TY><<<dim3(cb, N, S), dim3(32, 8), 0, st>>>((const TY*)y, stats, HW, C, chunk);
            }
            in_finalize_kernel<TY><<<(N * C + 255) / 256, 256, 0, st>>>((const TY*)y, stats, HW, C, N * C, eps);
        }
        if (vec)
            launch_k(in_apply_v4_kernel<TY, T>, dim3((unsigned)(ew_grid(total / 4))), dim3(kEw), 0, st, (const TY*)y, (const T*)res, (T*)z, stats, (unsigned)(HW * (C / 4)),
                                                                          (unsigned)(C / 4), (unsigned)(total / 4), norm, act, slope);
        else
            in_apply_kernel<TY, T><<<ew_grid(total), kEw, 0, st>>>((const TY*)y, (const T*)res, (T*)z, stats, HW, C, total, norm, act, slope);
    });
    return check_launch("instnorm_act_fwd");
}

// stats [N,C,2] = (mean, rstd) of y [N,HW,C] alone (the reduction half of p2p_instnorm_act_fwd), for callers that apply the
// normalisation with p2p_instnorm_act_planes
extern "C" int p2p_instnorm_stats(const void* y, float* stats, int N, long long HW, int C, float eps, int y_dtype, void* stream) {
    if (N < 0 || HW < 0 || C < 0) return fail(P2P_EINVAL, "negative size");
    if ((long long)N * HW * C == 0) return P2P_OK;
    if (!y || !stats) return fail(P2P_EINVAL, "null pointer argument");
    if (N > 65535) return fail(P2P_ENOTSUP, "N=%d too large", N);
    cudaStream_t st = (cudaStream_t)stream;
    if (C % 4 == 0 && HW < 2000000000LL && vec4_ok(y, y_dtype == P2P_BF16 ? 2 : 4) && vec4_ok(stats, 4) && ws_fits(N, C)) {
        const int cb = (C / 4 + 31) / 32, S = split_for(HW, cb * N);
        const int chunk = (int)((HW + S - 1) / S);
        P2P_DISPATCH_DTYPE(y_dtype, TY, launch_k(in_stats_fused_v4_kernel<TY>, dim3(cb, N, S), dim3(32, 8), 0, st, 
            (const TY*)y, stats, g_stats_ws, g_stats_cnt, (int)HW, C, chunk, eps));
        return check_launch("instnorm_stats (single launch)");
    }
    if (cudaMemsetAsync(stats, 0, sizeof(float) * 2 * N * C, st) != cudaSuccess) return fail(P2P_ECUDA, "memset failed");
    P2P_DISPATCH_DTYPE(y_dtype, TY, {
        if (C % 4 == 0 && HW < 2000000000LL && vec4_ok(y, sizeof(TY)) && vec4_ok(stats, 4)) {
            const int cb = (C / 4 + 31) / 32, S = split_for(HW, cb * N);
            const int chunk = (int)((HW + S - 1) / S);
            in_stats_v4_kernel<TY><<<dim3(cb, N, S), dim3(32, 8), 0, st>>>((const TY*)y, stats, (int)HW, C, chunk);
        } else {
            const int cb = (C + 31) / 32, S = split_for(HW, cb * N);
            const long long chunk = (HW + S - 1) / S;
            in_stats_kernel<TY><<<dim3(cb, N, S), dim3(32, 8), 0, st>>>((const TY*)y, stats, HW, C, chunk);
        }
        in_finalize_kernel<TY><<<(N * C + 255) / 256, 256, 0, st>>>((const TY*)y, stats, HW, C, N * C, eps);
    });
    return check_launch("instnorm_stats");
}

// ---- allocation-free variant of p2p_instnorm_stats (no counterpart in the reference; added
// for callers that cannot allocate: graph capture, serving loops) ------------------------------------------------
// The partial-sum workspace and the group counters of the single-launch reduction live in ONE file-scope device array
// instead of caller-allocated memory: kStaticSums floats of sums, then kStaticCnt 32-bit counters (520 KB per GPU that loads
// the library, held until exit).  Every call shares it; the region a call uses is cleared on the caller's stream first.
namespace p2p {
namespace {
constexpr long long kStaticSums = 131072;      // floats: 2 * N * C  -> N * C <= 65536
constexpr long long kStaticCnt = 2048;         // words:  N * ceil(C / 128)
static __device__ __align__(256) float g_static_stats_scratch[kStaticSums + kStaticCnt];

// device address of the array on the CURRENT device (the symbol's name is only a host-side handle), looked up once per device
float* static_scratch_ptr() {
    static float* ptr[64] = {};
    static std::mutex mu;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { fail(P2P_ECUDA, "cudaGetDevice failed"); return nullptr; }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_stats_scratch);
        if (e != cudaSuccess || !p) { fail(P2P_ECUDA, "cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return nullptr; }
        ptr[dev] = static_cast<float*>(p);
    }
    return ptr[dev];
}
}  // namespace
}  // namespace p2p

extern "C" int p2p_instnorm_stats_static(const void* y, float* stats, int N, long long HW, int C, float eps, int y_dtype,
                                         void* stream) {
    if (N < 0 || HW < 0 || C < 0) return fail(P2P_EINVAL, "negative size");
    if ((long long)N * HW * C == 0) return P2P_OK;
    if (!y || !stats) return fail(P2P_EINVAL, "null pointer argument");
    if (y_dtype != P2P_F32 && y_dtype != P2P_BF16) return fail(P2P_EINVAL, "y_dtype must be P2P_F32 or P2P_BF16");
    const int cb = (C / 4 + 31) / 32;
    if (C % 4 != 0 || HW >= 2000000000LL || N > 65535 || !vec4_ok(y, y_dtype == P2P_BF16 ? 2 : 4) || !vec4_ok(stats, 4))
        return fail(P2P_ENOTSUP, "p2p_instnorm_stats_static needs C %% 4 == 0, HW < 2e9, N <= 65535 and 16-byte aligned y / stats "
                                 "(C=%d HW=%lld N=%d)", C, HW, N);
    if (2LL * N * C > kStaticSums || (long long)N * cb > kStaticCnt)
        return fail(P2P_ENOTSUP, "p2p_instnorm_stats_static: N*C=%lld exceeds the built-in scratch (N*C <= %lld and N*ceil(C/128) <= %lld); "
                                 "use p2p_instnorm_stats", (long long)N * C, kStaticSums / 2, kStaticCnt);
    float* ws = static_scratch_ptr();
    if (!ws) return P2P_ECUDA;
    unsigned* cnt = reinterpret_cast<unsigned*>(ws + kStaticSums);
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps whatever the last call (possibly an aborted or overlapping one) left: clear what must start at zero
    if (cudaMemsetAsync(ws, 0, sizeof(float) * 2 * N * C, st) != cudaSuccess ||
        cudaMemsetAsync(cnt, 0, sizeof(unsigned) * N * cb, st) != cudaSuccess)
        return fail(P2P_ECUDA, "memset of the built-in scratch failed");
    const int S = split_for(HW, cb * N);
    const int chunk = (int)((HW + S - 1) / S);
    P2P_DISPATCH_DTYPE(y_dtype, TY, (in_stats_fused_v4_kernel<TY><<<dim3(cb, N, S), dim3(32, 8), 0, st>>>(
        (const TY*)y, stats, ws, cnt, (int)HW, C, chunk, eps)));
    return check_launch("instnorm_stats_static");
}

extern "C" int p2p_instnorm_act_bwd(const void* y, const float* stats, const void* dz, void* dy, float* work, int N,
                                    long long HW, int C, int norm, int act, float slope, int y_dtype, int dtype, void* stream) {
    if (N < 0 || HW < 0 || C < 0) return fail(P2P_EINVAL, "negative size");
    const long long total = (long long)N * HW * C;
    if (total == 0) return P2P_OK;
    if (!y || !dz || !dy || (norm && (!stats || !work))) return fail(P2P_EINVAL, "null pointer argument");
    if (N > 65535) return fail(P2P_ENOTSUP, "N=%d too large", N);
    cudaStream_t st = (cudaStream_t)stream;
    const bool v4 = C % 4 == 0 && total / 4 < 2000000000LL && HW < 2000000000LL;
    P2P_DISPATCH_DTYPE2(y_dtype, TY, dtype, T, {
        const bool vec = v4 && vec4_ok(y, sizeof(TY)) && vec4_ok(dz, sizeof(T)) && vec4_ok(dy, sizeof(T)) && vec4_ok(stats, 4) && vec4_ok(work, 4);
        if (norm && vec && ws_fits(N, C)) {
            const int cb = (C / 4 + 31) / 32, S = split_for(HW, cb * N);
            const int chunk = (int)((HW + S - 1) / S);
            launch_k(in_bwd_stats_fused_v4_kernel<TY, T>, dim3(cb, N, S), dim3(32, 8), 0, st, (const TY*)y, stats, (const T*)dz, work, g_stats_ws,
                                                                                         g_stats_cnt, (int)HW, C, chunk, act, slope);
        } else if (norm) {
            if (cudaMemsetAsync(work, 0, sizeof(float) * 2 * N * C, st) != cudaSuccess) return fail(P2P_ECUDA, "memset failed");
            if (vec) {
                const int cb = (C / 4 + 31) / 32, S = split_for(HW, cb * N);
                const int chunk = (int)((HW + S - 1) / S);
                in_bwd_stats_v4_kernel<TY, T><<<dim3(cb, N, S), dim3(32, 8), 0, st>>>((const TY*)y, stats, (const T*)dz, work, (int)HW, C, chunk, act, slope);
            } else {
                const int cb = (C + 31) / 32, S = split_for(HW, cb * N);
                const long long chunk = (HW + S - 1) / S;
                in_bwd_stats_kernel<TY, T><<<dim3(cb, N, S), dim3(32, 8), 0, st>>>((const TY*)y, stats, (const T*)dz, work, HW, C, chunk, act, slope);
            }
        }
        if (vec && bwd_rows_form() && HW * (C / 4) < 2000000000LL)
            launch_k(in_bwd_apply_rows_v4_kernel<TY, T, false>, dim3(rows_grid_x((unsigned)(C / 4), HW, N), N), dim3(256), 0, st, (const TY*)y, stats,
                     (const T*)dz, (const float*)work, (T*)dy, N, (int)HW, 1, (unsigned)(C / 4), (int)HW, 1, 0LL, 1.f / (float)HW, norm, act, slope);
        else if (vec)
            launch_k(in_bwd_apply_v4_kernel<TY, T>, dim3((unsigned)(ew_grid(total / 4))), dim3(kEw), 0, st, (const TY*)y, stats, (const T*)dz, work, (T*)dy, (unsigned)(HW * (C / 4)),
                                                                              (unsigned)(C / 4), (unsigned)(total / 4), 1.f / (float)HW, norm, act, slope);
        else
            in_bwd_apply_kernel<TY, T><<<ew_grid(total), kEw, 0, st>>>((const TY*)y, stats, (const T*)dz, work, (T*)dy, HW, C, total, norm, act, slope);
    });
    return check_launch("instnorm_act_bwd");
}

// p2p_instnorm_act_bwd with dy written at the pitch of the padded input (see in_bwd_apply_wrap_v4_kernel): grid [rows, C],
// rows >= N*Hh*Wh, Hh >= H, Wh >= W.  Vector shapes only (C % 4 == 0, 16-byte aligned pointers); P2P_ENOTSUP otherwise.
extern "C" int p2p_instnorm_act_bwd_wrap(const void* y, const float* stats, const void* dz, void* grid, float* work, int N, int H,
                                         int W, int C, int Hh, int Wh, long long rows, int norm, int act, float slope, int y_dtype,
                                         int dtype, void* stream) {
    if (N < 1 || H < 1 || W < 1 || C < 1 || Hh < H || Wh < W || rows < (long long)N * Hh * Wh) return fail(P2P_EINVAL, "bad size");
    if (!y || !dz || !grid || (norm && (!stats || !work))) return fail(P2P_EINVAL, "null pointer argument");
    const long long HW = (long long)H * W, total = rows * C;
    if (N > 65535 || C % 4 != 0 || total / 4 >= 2000000000LL) return fail(P2P_ENOTSUP, "instnorm_act_bwd_wrap: shape not covered (C=%d)", C);
    cudaStream_t st = (cudaStream_t)stream;
    P2P_DISPATCH_DTYPE2(y_dtype, TY, dtype, T, {
        if (!(vec4_ok(y, sizeof(TY)) && vec4_ok(dz, sizeof(T)) && vec4_ok(grid, sizeof(T)) && vec4_ok(stats, 4) && vec4_ok(work, 4)))
            return fail(P2P_ENOTSUP, "instnorm_act_bwd_wrap: pointers must be 16-byte aligned");
        const int cb = (C / 4 + 31) / 32, S = split_for(HW, cb * N);
        const int chunk = (int)((HW + S - 1) / S);
        if (norm && ws_fits(N, C)) {
            launch_k(in_bwd_stats_fused_v4_kernel<TY, T>, dim3(cb, N, S), dim3(32, 8), 0, st, (const TY*)y, stats, (const T*)dz, work, g_stats_ws,
                     g_stats_cnt, (int)HW, C, chunk, act, slope);
        } else if (norm) {
            if (cudaMemsetAsync(work, 0, sizeof(float) * 2 * N * C, st) != cudaSuccess) return fail(P2P_ECUDA, "memset failed");
            in_bwd_stats_v4_kernel<TY, T><<<dim3(cb, N, S), dim3(32, 8), 0, st>>>((const TY*)y, stats, (const T*)dz, work, (int)HW, C, chunk, act, slope);
        }
        if (bwd_rows_form() && (long long)Hh * Wh * (C / 4) < 2000000000LL)
            launch_k(in_bwd_apply_rows_v4_kernel<TY, T, true>, dim3(rows_grid_x((unsigned)(C / 4), (long long)Hh * Wh, N + 1), N + 1), dim3(256), 0, st,
                     (const TY*)y, stats, (const T*)dz, (const float*)work, (T*)grid, N, H, W, (unsigned)(C / 4), Hh, Wh, rows, 1.f / (float)HW, norm, act, slope);
        else
            launch_k(in_bwd_apply_wrap_v4_kernel<TY, T>, dim3((unsigned)(ew_grid(total / 4))), dim3(kEw), 0, st, (const TY*)y, stats, (const T*)dz,
                     (const float*)work, (T*)grid, N, H, W, (unsigned)(C / 4), Hh, Wh, (unsigned)(total / 4), 1.f / (float)HW, norm, act, slope);
    });
    return check_launch("instnorm_act_bwd_wrap");
}

extern "C" int p2p_avgpool3s2_fwd(const void* x, void* y, int N, int H, int W, int C, int dtype, void* stream) {
    if (N < 0 || H < 1 || W < 1 || C < 0) return fail(P2P_EINVAL, "bad size");
    const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
    const long long total = (long long)N * OH * OW * C;
    if (total == 0) return P2P_OK;
    if (!x || !y) return fail(P2P_EINVAL, "null pointer argument");
    P2P_DISPATCH_DTYPE(dtype, T, launch_k(avgpool_fwd_kernel<T>, dim3((unsigned)(ew_grid(total))), dim3(kEw), 0, (cudaStream_t)stream, (const T*)x, (T*)y, H, W, C, OH, OW, total));
    return check_launch("avgpool3s2_fwd");
}

extern "C" int p2p_avgpool3s2_bwd(const void* dy, void* dx, int N, int H, int W, int C, int dtype, void* stream) {
    if (N < 0 || H < 1 || W < 1 || C < 0) return fail(P2P_EINVAL, "bad size");
    const int OH = (H - 1) / 2 + 1, OW = (W - 1) / 2 + 1;
    const long long total = (long long)N * H * W * C;
    if (total == 0) return P2P_OK;
    if (!dx || !dy) return fail(P2P_EINVAL, "null pointer argument");
    if (total < 2000000000LL)
        P2P_DISPATCH_DTYPE(dtype, T, launch_k(avgpool_bwd_kernel<T, unsigned>, dim3((unsigned)(ew_grid(total))), dim3(kEw), 0, (cudaStream_t)stream, (const T*)dy, (T*)dx, H, W, C, OH, OW, total));
    else
        P2P_DISPATCH_DTYPE(dtype, T, launch_k(avgpool_bwd_kernel<T, long long>, dim3((unsigned)(ew_grid(total))), dim3(kEw), 0, (cudaStream_t)stream, (const T*)dy, (T*)dx, H, W, C, OH, OW, total));
    return check_launch("avgpool3s2_bwd");
}

extern "C" int p2p_tanh_fwd(const void* x, void* y, long long n, int dtype, void* stream) {
    if (n <= 0) return n == 0 ? P2P_OK : fail(P2P_EINVAL, "negative size");
    P2P_DISPATCH_DTYPE(dtype, T, launch_k(tanh_fwd_kernel<T>, dim3((unsigned)(ew_grid(n))), dim3(kEw), 0, (cudaStream_t)stream, (const T*)x, (T*)y, n));
    return check_launch("tanh_fwd");
}
extern "C" int p2p_tanh_bwd(const void* y, const void* dy, void* dx, long long n, int dtype, void* stream) {
    if (n <= 0) return n == 0 ? P2P_OK : fail(P2P_EINVAL, "negative size");
    P2P_DISPATCH_DTYPE(dtype, T, launch_k(tanh_bwd_kernel<T>, dim3((unsigned)(ew_grid(n))), dim3(kEw), 0, (cudaStream_t)stream, (const T*)y, (const T*)dy, (T*)dx, n));
    return check_launch("tanh_bwd");
}

extern "C" int p2p_reflect_pad_fold(const void* dxp, void* dx, int N, int H, int W, int C, int p, int dtype, void* stream) {
    if (N < 0 || H < 1 || W < 1 || C < 0 || p < 0) return fail(P2P_EINVAL, "bad size");
    if (p >= H || p >= W) return fail(P2P_EINVAL, "reflection pad %d must be smaller than H=%d, W=%d", p, H, W);
    const long long total = (long long)N * H * W * C;
    if (total == 0) return P2P_OK;
    const bool v4 = C % 4 == 0 && total / 4 < 2000000000LL;
    P2P_DISPATCH_DTYPE(dtype, T, {
        if (v4 && vec4_ok(dxp, sizeof(T)) && vec4_ok(dx, sizeof(T)))
            launch_k(reflect_fold_v4_kernel<T, T>, dim3((unsigned)(ew_grid(total / 4))), dim3(kEw), 0, (cudaStream_t)stream, (const T*)dxp, (T*)dx, H, W, C / 4, p, (unsigned)(total / 4));
        else
            reflect_fold_kernel<T, T><<<ew_grid(total), kEw, 0, (cudaStream_t)stream>>>((const T*)dxp, (T*)dx, H, W, C, p, total);
    });
    return check_launch("reflect_pad_fold");
}

// the same with an fp32 padded gradient (the split-K window pass's red.add output) folded into a tensor of `dtype`
extern "C" int p2p_reflect_pad_fold_f32(const float* dxp, void* dx, int N, int H, int W, int C, int p, int dtype, void* stream) {
    if (N < 0 || H < 1 || W < 1 || C < 0 || p < 0) return fail(P2P_EINVAL, "bad size");
    if (p >= H || p >= W) return fail(P2P_EINVAL, "reflection pad %d must be smaller than H=%d, W=%d", p, H, W);
    const long long total = (long long)N * H * W * C;
    if (total == 0) return P2P_OK;
    if (!dxp || !dx) return fail(P2P_EINVAL, "null pointer argument");
    const bool v4 = C % 4 == 0 && total / 4 < 2000000000LL;
    P2P_DISPATCH_DTYPE(dtype, T, {
        if (v4 && vec4_ok(dxp, 4) && vec4_ok(dx, sizeof(T)))
            launch_k(reflect_fold_v4_kernel<float, T>, dim3((unsigned)(ew_grid(total / 4))), dim3(kEw), 0, (cudaStream_t)stream, dxp, (T*)dx, H, W, C / 4, p, (unsigned)(total / 4));
        else
            reflect_fold_kernel<float, T><<<ew_grid(total), kEw, 0, (cudaStream_t)stream>>>(dxp, (T*)dx, H, W, C, p, total);
    });
    return check_launch("reflect_pad_fold_f32");
}

extern "C" int p2p_bias_grad(const void* dout, float* dbias, long long rows, int J, int dtype, void* stream) {
    if (rows < 0 || J < 0) return fail(P2P_EINVAL, "negative size");
    if (J == 0) return P2P_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(dbias, 0, sizeof(float) * J, st) != cudaSuccess) return fail(P2P_ECUDA, "memset failed");
    if (rows == 0) return P2P_OK;
    const int jb = (J + 31) / 32, S = split_for(rows, jb);
    const long long chunk = (rows + S - 1) / S;
    P2P_DISPATCH_DTYPE(dtype, T, launch_k(bias_grad_kernel<T>, dim3(jb, S), dim3(32, 8), 0, st, (const T*)dout, dbias, rows, J, chunk));
    return check_launch("bias_grad");
}

extern "C" int p2p_loss_mse_const(const void* x, float target, float weight, float* loss, void* dx, float gscale,
                                  const float* gscale_dev, long long n, int dtype, void* stream) {
    if (n <= 0) return n == 0 ? P2P_OK : fail(P2P_EINVAL, "negative size");
    const float gcoef = gscale * weight * 2.f / (float)n;
    const int grid = (loss && deterministic()) ? 1 : (ew_grid(n) > 592 ? 592 : ew_grid(n));
    P2P_DISPATCH_DTYPE(dtype, T, launch_k(mse_const_kernel<T>, dim3((unsigned)(grid)), dim3(kEw), 0, (cudaStream_t)stream, (const T*)x, target, weight, loss, (T*)dx, gcoef, gscale_dev, n));
    return check_launch("loss_mse_const");
}

extern "C" int p2p_loss_l1(const void* a, const void* b, float weight, float* loss, void* da, float gscale,
                           const float* gscale_dev, long long n, int dtype, void* stream) {
    if (n <= 0) return n == 0 ? P2P_OK : fail(P2P_EINVAL, "negative size");
    const float gcoef = gscale * weight / (float)n;
    const int grid = (loss && deterministic()) ? 1 : (ew_grid(n) > 592 ? 592 : ew_grid(n));
    P2P_DISPATCH_DTYPE(dtype, T, launch_k(l1_kernel<T>, dim3((unsigned)(grid)), dim3(kEw), 0, (cudaStream_t)stream, (const T*)a, (const T*)b, weight, loss, (T*)da, gcoef, gscale_dev, n));
    return check_launch("loss_l1");
}

extern "C" int p2p_adam_step_dev(float* w, const float* g, float* m, float* v, void* w_bf16, long long n,
                                 const float* lr_dev, float beta1, float beta2, float eps, float grad_scale, int* step_dev,
                                 float* scratch2, void* stream) {
    if (n <= 0) return n == 0 ? P2P_OK : fail(P2P_EINVAL, "negative size");
    if (!w || !g || !m || !v || !lr_dev || !step_dev || !scratch2) return fail(P2P_EINVAL, "null pointer argument");
    adam_prepare_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev, lr_dev, beta1, beta2, scratch2);
    if (n % 4 == 0 && vec4_ok(w, 4) && vec4_ok(g, 4) && vec4_ok(m, 4) && vec4_ok(v, 4) && vec4_ok(w_bf16, 2))
        // 128-thread blocks: small enough (registers) to share an SM with a resident conv_window_kernel CTA -- train_step runs
        // the generator's Adam on a side stream during the discriminator's backward pass
        adam_dev_v4_kernel<<<(unsigned)((n / 4 + 127) / 128 < 148 * 12 ? (n / 4 + 127) / 128 : 148 * 12), 128, 0, (cudaStream_t)stream>>>((float4*)w, (const float4*)g, (float4*)m, (float4*)v,
                                                                            (uint2*)w_bf16, n / 4, scratch2, beta1, beta2, eps, grad_scale);
    else
        adam_dev_kernel<<<ew_grid(n), kEw, 0, (cudaStream_t)stream>>>(w, g, m, v, (__nv_bfloat16*)w_bf16, n, scratch2, beta1, beta2,
                                                                       eps, grad_scale);
    return check_launch("adam_step_dev");
}

extern "C" int p2p_adam_step(float* w, const float* g, float* m, float* v, void* w_bf16, long long n, float lr,
                             float beta1, float beta2, float eps, int step, void* stream) {
    if (n <= 0) return n == 0 ? P2P_OK : fail(P2P_EINVAL, "negative size");
    if (step < 1) return fail(P2P_EINVAL, "step must be >= 1");
    const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    adam_kernel<<<ew_grid(n), kEw, 0, (cudaStream_t)stream>>>(w, g, m, v, (__nv_bfloat16*)w_bf16, n, (float)(lr / bc1),
                                                               beta1, beta2, (float)(1.0 / sqrt(bc2)), eps);
    return check_launch("adam_step");
}
