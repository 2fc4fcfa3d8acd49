// Spectrogram encode / decode around the networks (reference: models/pix2pixHD_model.py:142-249).
//   encode:    MDCT coefficients [B,F,M] -> dB channels [B,M,F,C] (channels-last image, H = bins, W = frames),
//              sign map, and batch-global min / max / sum / sum-of-squares
//   normalise: (x - min) / (max - min) in place, high-band rows replaced by the caller's noise (or zeros)
//   decode:    network output -> |x| (max-min) + min -> 10 * 10^(x/20) - min_value -> explicit phase decode
//              -> coefficients [B,F,M] for IMDCT4
#include <cfloat>
#include "dtype.cuh"

namespace p2p {
namespace {

__device__ __forceinline__ int float_key(float f) {
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__device__ __forceinline__ float amp_to_db(float x, float amin) { return 20.f * log10f(fmaxf(x, amin)) - 20.f; }

struct SpectroStats {   // device-side accumulator, 32 bytes
    int min_key, max_key;
    double sum, sumsq;
    long long pad;
};

__global__ void spectro_stats_init(SpectroStats* st) {
    st->min_key = float_key(FLT_MAX);
    st->max_key = float_key(-FLT_MAX);
    st->sum = 0.0;
    st->sumsq = 0.0;
}

// grid (M/32, ceil(F/32), B), block (32, 8)
template <int C>
__global__ void __launch_bounds__(256) spectro_encode_kernel(const float* __restrict__ spec, float* __restrict__ logs,
                                                             float* __restrict__ pha, SpectroStats* st, int F, int M,
                                                             float alpha, float amin) {
    __shared__ float tile[C][32][33];
    __shared__ float tsgn[32][33];
    __shared__ float red[40];
    const int b = blockIdx.z, k0 = blockIdx.x * 32, f0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    float vmin = FLT_MAX, vmax = -FLT_MAX, s1 = 0.f, s2 = 0.f;
    for (int r = ty; r < 32; r += 8) {
        const int f = f0 + r, k = k0 + tx;
        if (f < F && k < M) {
            const float s = spec[((long long)b * F + f) * M + k];
            float d0, d1 = 0.f;
            if (C == 2) {   // explicit +/- encoding (pix2pixHD_model.py:149-158)
                const float neg = 0.5f * (fabsf(s) - s), pos = s + neg;
                d0 = amp_to_db(alpha * pos + (1.f - alpha) * neg, amin);
                d1 = amp_to_db((1.f - alpha) * pos + alpha * neg, amin);
                tile[C - 1][r][tx] = d1;
                vmin = fminf(vmin, fminf(d0, d1));
                vmax = fmaxf(vmax, fmaxf(d0, d1));
                s1 += d0 + d1;
                s2 += d0 * d0 + d1 * d1;
            } else {        // magnitude only (:160-162)
                d0 = amp_to_db(fabsf(s) + amin, amin);
                vmin = fminf(vmin, d0);
                vmax = fmaxf(vmax, d0);
                s1 += d0;
                s2 += d0 * d0;
            }
            tile[0][r][tx] = d0;
            tsgn[r][tx] = s > 0.f ? 1.f : (s < 0.f ? -1.f : 0.f);
        }
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int k = k0 + r, f = f0 + tx;
        if (f < F && k < M) {
            const long long o = ((long long)b * M + k) * F + f;
            if (C == 2) reinterpret_cast<float2*>(logs)[o] = make_float2(tile[0][tx][r], tile[1][tx][r]);
            else logs[o] = tile[0][tx][r];
            pha[o] = tsgn[tx][r];
        }
    }
    // block reduction of the statistics
    const int lane = tx, warp = ty;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        vmin = fminf(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
        vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    if (lane == 0) { red[warp] = vmin; red[8 + warp] = vmax; red[16 + warp] = s1; red[24 + warp] = s2; }
    __syncthreads();
    if (warp == 0 && lane == 0) {
        double a1 = 0.0, a2 = 0.0;
        for (int i = 0; i < 8; ++i) {
            vmin = fminf(vmin, red[i]);
            vmax = fmaxf(vmax, red[8 + i]);
            a1 += red[16 + i];
            a2 += red[24 + i];
        }
        atomicMin(&st->min_key, float_key(vmin));
        atomicMax(&st->max_key, float_key(vmax));
        atomicAdd(&st->sum, a1);
        atomicAdd(&st->sumsq, a2);
    }
}

// one thread per (b, k, f): all C channels
template <int C>
__global__ void spectro_normalize_kernel(float* __restrict__ logs, const SpectroStats* __restrict__ st,
                                         float* __restrict__ norm_out, const float* __restrict__ noise, int B, int M,
                                         int F, int mask_rows, int use_mask) {
    const float vmin = key_float(st->min_key), vmax = key_float(st->max_key);
    const long long n = (long long)B * M * F;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        const double cnt = (double)n * C, mean = st->sum / cnt;
        const double var = cnt > 1 ? (st->sumsq - cnt * mean * mean) / (cnt - 1) : 0.0;
        norm_out[0] = vmax;
        norm_out[1] = vmin;
        norm_out[2] = (float)mean;
        norm_out[3] = (float)sqrt(var > 0 ? var : 0.0);
    }
    if (i >= n) return;
    const int f = (int)(i % F);
    const int k = (int)((i / F) % M);
    const int b = (int)(i / ((long long)F * M));
    const float inv = 1.f / (vmax - vmin);
    const bool masked = use_mask && k >= M - mask_rows;
#pragma unroll
    for (int c = 0; c < C; ++c) {
        float v;
        if (masked) v = noise ? noise[(((long long)b * mask_rows + (k - (M - mask_rows))) * F + f) * C + c] : 0.f;
        else v = (logs[i * C + c] - vmin) * inv;
        logs[i * C + c] = v;
    }
}

// grid (M/32, ceil(F/32), B), block (32, 8): x[b, c, k, f] (strided) -> coef[b, f, k]
template <typename T>
__global__ void __launch_bounds__(256) spectro_decode_kernel(const T* __restrict__ x, long long sb, long long sc,
                                                             long long sh, long long sw, const float* __restrict__ norm,
                                                             const float* __restrict__ pha, float* __restrict__ coef,
                                                             int C, int F, int M, float alpha, float amin, float gain) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z, k0 = blockIdx.x * 32, f0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    const float vmax = norm[0], vmin = norm[1];
    for (int r = ty; r < 32; r += 8) {
        const int k = k0 + r, f = f0 + tx;
        if (k < M && f < F) {
            const T* p = x + (long long)b * sb + (long long)k * sh + (long long)f * sw;
            // denormalize (:229-232): |x| (max - min) + min, then DB_to_amplitude(., 10, 0.5) - min_value
            const float a0 = 10.f * exp10f(0.05f * (fabsf(ldf(p)) * (vmax - vmin) + vmin)) - amin;
            float v;
            if (C == 2) {
                const float a1 = 10.f * exp10f(0.05f * (fabsf(ldf(p + sc)) * (vmax - vmin) + vmin)) - amin;
                v = (a0 - a1) / (2.f * alpha - 1.f);
            } else {
                v = pha ? a0 * pha[((long long)b * M + k) * F + f] : a0;
            }
            tile[r][tx] = gain * v;
        }
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int f = f0 + r, k = k0 + tx;
        if (k < M && f < F) coef[((long long)b * F + f) * M + k] = tile[tx][r];
    }
}

}  // namespace
}  // namespace p2p

extern "C" int p2p_spectro_encode(const float* spec, float* logs, float* pha, void* stats_work, int B, int F, int M,
                                  int channels, float alpha, float min_value, void* stream) {
    using namespace p2p;
    if (!spec || !logs || !pha || !stats_work) return fail(P2P_EINVAL, "null pointer argument");
    if (channels != 1 && channels != 2) return fail(P2P_EINVAL, "channels must be 1 or 2, got %d", channels);
    if (B < 1 || F < 1 || M < 1 || B > 65535) return fail(P2P_EINVAL, "bad spectrogram geometry");
    cudaStream_t st = (cudaStream_t)stream;
    spectro_stats_init<<<1, 1, 0, st>>>((SpectroStats*)stats_work);
    dim3 grid((M + 31) / 32, (F + 31) / 32, B), block(32, 8);
    if (channels == 2) spectro_encode_kernel<2><<<grid, block, 0, st>>>(spec, logs, pha, (SpectroStats*)stats_work, F, M, alpha, min_value);
    else spectro_encode_kernel<1><<<grid, block, 0, st>>>(spec, logs, pha, (SpectroStats*)stats_work, F, M, alpha, min_value);
    return check_launch("spectro_encode_kernel");
}

extern "C" int p2p_spectro_normalize(float* logs, const void* stats_work, float* norm_out, const float* noise, int B, int F,
                                     int M, int channels, int mask_rows, int use_mask, void* stream) {
    using namespace p2p;
    if (!logs || !stats_work || !norm_out) return fail(P2P_EINVAL, "null pointer argument");
    if (channels != 1 && channels != 2) return fail(P2P_EINVAL, "channels must be 1 or 2, got %d", channels);
    if (mask_rows < 0 || mask_rows > M) return fail(P2P_EINVAL, "mask_rows %d outside [0, %d]", mask_rows, M);
    const long long n = (long long)B * M * F;
    const unsigned grid = (unsigned)((n + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (channels == 2) spectro_normalize_kernel<2><<<grid, 256, 0, st>>>(logs, (const SpectroStats*)stats_work, norm_out, noise, B, M, F, mask_rows, use_mask);
    else spectro_normalize_kernel<1><<<grid, 256, 0, st>>>(logs, (const SpectroStats*)stats_work, norm_out, noise, B, M, F, mask_rows, use_mask);
    return check_launch("spectro_normalize_kernel");
}

extern "C" int p2p_spectro_decode(const void* x, long long sb, long long sc, long long sh, long long sw, const float* norm,
                                  const float* pha, float* coef, int B, int channels, int F, int M, float alpha,
                                  float min_value, float gain, int dtype, void* stream) {
    using namespace p2p;
    if (!x || !norm || !coef) return fail(P2P_EINVAL, "null pointer argument");
    if (channels != 1 && channels != 2) return fail(P2P_EINVAL, "channels must be 1 or 2, got %d", channels);
    if (B < 1 || F < 1 || M < 1 || B > 65535) return fail(P2P_EINVAL, "bad spectrogram geometry");
    dim3 grid((M + 31) / 32, (F + 31) / 32, B), block(32, 8);
    cudaStream_t st = (cudaStream_t)stream;
    P2P_DISPATCH_DTYPE(dtype, T, spectro_decode_kernel<T><<<grid, block, 0, st>>>((const T*)x, sb, sc, sh, sw, norm, pha, coef, channels, F, M, alpha, min_value, gain));
    return check_launch("spectro_decode_kernel");
}
