// MDCT4 / IMDCT4 as two fused sm_100a kernels (analysis, synthesis).
//
// Replaces models/mdct.py:486-513 (MDCT4.forward) and :542-566 (IMDCT4.forward) of the reference and,
// with the roles swapped, their autograd.  Instead of the reference's
// N-point complex128 FFT with half the output discarded, each frame is folded to M = N/2 reals and
// transformed by a DCT-IV computed with ONE L = M/2-point complex FFT held in shared memory
// (Appendix A.3-A.5; tests/fast_model.py is the numpy statement of the same index formulas).
//
// Data layout: signal x [B,T] fp32 row-major; spectrogram [B,F,M] fp32 row-major (bins fastest).
// A CTA owns a run of consecutive frames of one batch row, so every signal sample is fetched from
// HBM once (the hop overlap is served from shared memory) and every coefficient is written once.
#include "common.cuh"
#include "mdct4_fast.cuh"
#include "smem_fft.cuh"
#include <cmath>
#include <vector>

namespace p2p {
namespace {

using smemfft::kThreads;
using smemfft::cmul;
using smemfft::fft_rows_inplace;

struct Tables {           // views into the 7L-word table buffer
    const float2* pre;    // e^{-j pi (n+1/4)/M}
    const float2* post;   // e^{-j pi k/M}
    const float2* roots;  // e^{-2 pi j m/L}
    const int* perm;      // position p after the in-place DIF passes holds FFT bin perm[p]
};

__device__ __forceinline__ Tables make_tables(const void* t, int L) {
    const float* f = reinterpret_cast<const float*>(t);
    Tables r;
    r.pre = reinterpret_cast<const float2*>(f);
    r.post = reinterpret_cast<const float2*>(f + 2 * L);
    r.roots = reinterpret_cast<const float2*>(f + 4 * L);
    r.perm = reinterpret_cast<const int*>(f + 6 * L);
    return r;
}

// ---- analysis: frames -> window -> fold -> DCT-IV -> coefficients ---------------------------------
// grid (ceil(F/fpb), B).  smem: sig[(fpb-1)*hop+win] | win_s[win] | z[fpb*L] (float2) | o[fpb*M]
__global__ void __launch_bounds__(kThreads)
mdct4_analysis_kernel(const float* __restrict__ x, float* __restrict__ spec,
                      const float* __restrict__ window, const void* __restrict__ tables,
                      long long T, int F, int N, int hop, int win, long long offset, float scale,
                      int fpb) {
    extern __shared__ __align__(16) float smem[];
    const int M = N >> 1, L = N >> 2;
    const int f0 = blockIdx.x * fpb;
    const int nfr = min(fpb, F - f0);
    const int span = (nfr - 1) * hop + win;
    const int span_max = (fpb - 1) * hop + win;
    float* sig = smem;
    float* win_s = sig + ((span_max + 3) & ~3);
    float2* z = reinterpret_cast<float2*>(win_s + ((win + 3) & ~3));
    float* o = reinterpret_cast<float*>(z + fpb * L);
    const Tables tb = make_tables(tables, L);

    const float* xb = x + (long long)blockIdx.y * T;
    const long long t0 = (long long)f0 * hop - offset;
    for (int i = threadIdx.x; i < span; i += kThreads) {
        const long long t = t0 + i;
        sig[i] = (t >= 0 && t < T) ? __ldg(xb + t) : 0.f;
    }
    for (int i = threadIdx.x; i < win; i += kThreads) win_s[i] = __ldg(window + i);
    __syncthreads();

    // fold + pre-twiddle: z[n] = (v[2n] + j v[M-1-2n]) * pre[n]
    const int hM = M >> 1;
    for (int w = threadIdx.x; w < nfr * L; w += kThreads) {
        const int fl = w / L;
        const int n = w - fl * L;
        const float* s = sig + fl * hop;
        auto u = [&](int i) -> float { return i < win ? s[i] * win_s[i] : 0.f; };
        float a, b;
        if (2 * n < hM) {
            a = -u(3 * hM - 1 - 2 * n) - u(3 * hM + 2 * n);
            b = u(hM - 1 - 2 * n) - u(hM + 2 * n);
        } else {
            a = u(2 * n - hM) - u(3 * hM - 1 - 2 * n);
            b = -u(hM + 2 * n) - u(5 * hM - 1 - 2 * n);
        }
        z[w] = cmul(make_float2(a, b), tb.pre[n]);
    }
    __syncthreads();

    fft_rows_inplace(z, nfr, L, tb.roots);

    // post-twiddle + de-interleave: X[2k] = Re Y[k], X[M-1-2k] = -Im Y[k]
    for (int w = threadIdx.x; w < nfr * L; w += kThreads) {
        const int fl = w / L;
        const int p = w - fl * L;
        const int k = tb.perm[p];
        const float2 y = cmul(z[w], tb.post[k]);
        o[fl * M + 2 * k] = scale * y.x;
        o[fl * M + M - 1 - 2 * k] = -scale * y.y;
    }
    __syncthreads();

    float* dst = spec + ((long long)blockIdx.y * F + f0) * M;
    const int total = nfr * M;
    if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        const float4* o4 = reinterpret_cast<const float4*>(o);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (int i = threadIdx.x; i < (total >> 2); i += kThreads) d4[i] = o4[i];
    } else {
        for (int i = threadIdx.x; i < total; i += kThreads) dst[i] = o[i];
    }
}

// ---- synthesis: coefficients -> DCT-IV -> unfold -> window -> overlap-add (gather) -> crop ---------
// grid (ceil(T_out/chunk), B).  Each CTA owns `chunk` output samples and transforms every frame that
// overlaps them (nfmax frames at most), so the overlap-add is a deterministic gather: no atomics, each
// output sample written exactly once.  smem: win_s[win] | z[nfmax*L] (float2) | t[nfmax*M]
__global__ void __launch_bounds__(kThreads)
mdct4_synthesis_kernel(const float* __restrict__ spec, float* __restrict__ out,
                       const float* __restrict__ window, const void* __restrict__ tables,
                       int F, long long T_out, int N, int hop, int win, long long crop_left,
                       float scale, int chunk, int nfmax) {
    extern __shared__ __align__(16) float smem[];
    const int M = N >> 1, L = N >> 2, hM = N >> 2;
    float* win_s = smem;
    float2* z = reinterpret_cast<float2*>(win_s + ((win + 3) & ~3));
    float* t = reinterpret_cast<float*>(z + nfmax * L);
    const Tables tb = make_tables(tables, L);

    const long long s_lo = (long long)blockIdx.x * chunk + crop_left;            // OLA coordinates
    const long long s_hi = min((long long)(blockIdx.x + 1) * chunk, T_out) + crop_left;  // exclusive
    long long fl_ = (s_lo - win + 1 + hop - 1);
    fl_ = fl_ > 0 ? fl_ / hop : 0;                                               // ceil, clamped at 0
    const int f_lo = (int)min(fl_, (long long)F);
    const int f_hi = (int)min((s_hi - 1) / hop, (long long)F - 1);              // inclusive
    const int nfr = max(0, f_hi - f_lo + 1);

    for (int i = threadIdx.x; i < win; i += kThreads) win_s[i] = __ldg(window + i);
    {   // coalesced load of the coefficient rows into t
        const float* src = spec + ((long long)blockIdx.y * F + f_lo) * M;
        const int total = nfr * M;
        if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
            const float4* s4 = reinterpret_cast<const float4*>(src);
            float4* t4 = reinterpret_cast<float4*>(t);
            for (int i = threadIdx.x; i < (total >> 2); i += kThreads) t4[i] = __ldg(s4 + i);
        } else {
            for (int i = threadIdx.x; i < total; i += kThreads) t[i] = __ldg(src + i);
        }
    }
    __syncthreads();

    for (int w = threadIdx.x; w < nfr * L; w += kThreads) {
        const int fl = w / L;
        const int n = w - fl * L;
        const float* X = t + fl * M;
        z[w] = cmul(make_float2(X[2 * n], X[M - 1 - 2 * n]), tb.pre[n]);
    }
    __syncthreads();

    fft_rows_inplace(z, nfr, L, tb.roots);

    for (int w = threadIdx.x; w < nfr * L; w += kThreads) {
        const int fl = w / L;
        const int p = w - fl * L;
        const int k = tb.perm[p];
        const float2 y = cmul(z[w], tb.post[k]);
        t[fl * M + 2 * k] = y.x;
        t[fl * M + M - 1 - 2 * k] = -y.y;
    }
    __syncthreads();

    float* ob = out + (long long)blockIdx.y * T_out;
    const int n_out = (int)(s_hi - s_lo);
    for (int i = threadIdx.x; i < n_out; i += kThreads) {
        const long long s = s_lo + i;
        // frames with 0 <= s - f*hop < win
        long long a = s - win + 1 + hop - 1;
        int fa = (int)(a > 0 ? a / hop : 0);
        fa = max(fa, f_lo);
        const int fb = (int)min(s / hop, (long long)f_hi);
        float acc = 0.f;
        for (int f = fa; f <= fb; ++f) {
            const int n = (int)(s - (long long)f * hop);
            const float* tf = t + (f - f_lo) * M;
            float y;
            if (n < hM) y = tf[hM + n];
            else if (n < 3 * hM) y = -tf[3 * hM - 1 - n];
            else y = -tf[n - 3 * hM];
            acc += win_s[n] * y;
        }
        ob[s - crop_left] = scale * acc;
    }
}

}  // namespace
}  // namespace p2p

// ---- host side ---------------------------------------------------------------------------------
namespace {

int check_geometry(int n_fft, int hop, int win) {
    using namespace p2p;
    if (!is_pow2(n_fft) || n_fft < 16 || n_fft > 4096)
        return fail(P2P_ENOTSUP, "n_fft must be a power of two in [16, 4096], got %d", n_fft);
    if (win < 1 || win > n_fft) return fail(P2P_EINVAL, "win_length %d must be in [1, n_fft=%d]", win, n_fft);
    if (hop < 1 || hop > win) return fail(P2P_EINVAL, "hop_length %d must be in [1, win_length=%d]", hop, win);
    return P2P_OK;
}

constexpr int kMaxSmem = 227 * 1024;

// (V,P) register/lane split of the L-point FFT used by the fast path; 0 when L has no fast path
void fast_shape(int L, int* V, int* P) {
    switch (L) {
        case 64: *V = 8; *P = 8; break;
        case 128: *V = 16; *P = 8; break;
        case 256: *V = 16; *P = 16; break;
        case 512: *V = 32; *P = 16; break;
        case 1024: *V = 32; *P = 32; break;
        default: *V = 0; *P = 0;
    }
}

int sm_count_and_blocks(const void* kernel, size_t smem, int* blocks_per_sm) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, kernel, p2p::fast::kThreads, smem) != cudaSuccess)
        return -1;
    return sms;
}

template <class C>
int launch_analysis_fast(const float* x, float* spec, const float* window, const void* tables, int B, long long T,
                         int F, long long offset, float scale, cudaStream_t stream) {
    using namespace p2p;
    auto kernel = fast::mdct4_analysis_fast<C>;
    const size_t smem = C::SMEM_BYTES;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail(P2P_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int bps = 0;
    const int sms = sm_count_and_blocks((const void*)kernel, smem, &bps);
    if (sms <= 0 || bps <= 0) { cudaGetLastError(); return fail(P2P_ECUDA, "occupancy query failed (smem %zu)", smem); }
    const int tiles_per_row = (F + C::FPT - 1) / C::FPT;
    const long long total = (long long)tiles_per_row * B;
    const long long grid = total < (long long)sms * bps ? total : (long long)sms * bps;
    const int aligned = ((reinterpret_cast<uintptr_t>(x) & 15) == 0) && (T % 4 == 0) && (offset % 4 == 0);
    kernel<<<(unsigned)grid, fast::kThreads, smem, stream>>>(x, spec, window, tables, T, F, offset, scale,
                                                             tiles_per_row, total, aligned);
    return check_launch("mdct4_analysis_fast");
}

template <typename K>
int ensure_smem(K kernel, size_t bytes) {
    if (bytes > (size_t)kMaxSmem) return p2p::fail(P2P_ENOTSUP, "needs %zu B of shared memory", bytes);
    if (bytes > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e != cudaSuccess) return p2p::fail(P2P_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    return P2P_OK;
}

template <class C>
int launch_synthesis_fast(const float* spec, float* out, const float* window, const void* tables, int B, int F,
                          long long T_out, long long crop_left, float scale, cudaStream_t stream) {
    using namespace p2p;
    using S = fast::SynCfg<C>;
    auto kernel = fast::mdct4_synthesis_fast<C>;
    const size_t smem = S::SMEM_BYTES;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail(P2P_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int bps = 0;
    const int sms = sm_count_and_blocks((const void*)kernel, smem, &bps);
    if (sms <= 0 || bps <= 0) { cudaGetLastError(); return fail(P2P_ECUDA, "occupancy query failed (smem %zu)", smem); }
    const long long hf = crop_left / C::M, hl = (crop_left + T_out - 1) / C::M;
    if (hl > 2147483000LL) return fail(P2P_ENOTSUP, "T_out too large");
    const int tiles_per_row = (int)((hl - hf + 1 + C::FPT - 2) / (C::FPT - 1));
    const long long total = (long long)tiles_per_row * B;
    const long long grid = total < (long long)sms * bps ? total : (long long)sms * bps;
    const int aligned_in = (reinterpret_cast<uintptr_t>(spec) & 15) == 0;
    const int vec_out = ((reinterpret_cast<uintptr_t>(out) & 15) == 0) && (T_out % 4 == 0) && (crop_left % 4 == 0);
    kernel<<<(unsigned)grid, fast::kThreads, smem, stream>>>(spec, out, window, tables, F, T_out, crop_left, scale,
                                                             (int)hf, (int)hl, tiles_per_row, total, aligned_in, vec_out);
    return check_launch("mdct4_synthesis_fast");
}

}  // namespace

static int g_force_general = 0;
extern "C" int p2p_mdct4_force_general(int on) { const int old = g_force_general; g_force_general = on; return old; }

extern "C" int p2p_mdct4_table_words(int n_fft) {
    if (!p2p::is_pow2(n_fft) || n_fft < 16 || n_fft > 4096)
        return p2p::fail(P2P_ENOTSUP, "n_fft must be a power of two in [16, 4096], got %d", n_fft);
    return 9 * (n_fft / 4);
}

extern "C" int p2p_mdct4_fill_tables(int n_fft, void* host_tables) {
    const int words = p2p_mdct4_table_words(n_fft);
    if (words < 0) return words;
    if (!host_tables) return p2p::fail(P2P_EINVAL, "host_tables is null");
    const int M = n_fft / 2, L = n_fft / 4;
    float* f = static_cast<float*>(host_tables);
    int* perm = reinterpret_cast<int*>(f + 6 * L);
    const double pi = 3.14159265358979323846264338327950288;
    for (int n = 0; n < L; ++n) {
        const double a = -pi * (n + 0.25) / M, b = -pi * n / M, c = -2.0 * pi * n / L;
        f[2 * n] = (float)cos(a);           f[2 * n + 1] = (float)sin(a);
        f[2 * L + 2 * n] = (float)cos(b);   f[2 * L + 2 * n + 1] = (float)sin(b);
        f[4 * L + 2 * n] = (float)cos(c);   f[4 * L + 2 * n + 1] = (float)sin(c);
    }
    // position -> bin map of the in-place DIF passes (radix 4,...,4[,2]): the pass over sub-size S
    // sends output digit k_i to block k_i of size S/r, and bin = k_1 + r_1 (k_2 + r_2 (...)).
    for (int p = 0; p < L; ++p) {
        int S = L, rem = p, k = 0, weight = 1;
        while (S >= 4) { const int q = S / 4; k += (rem / q) * weight; rem %= q; weight *= 4; S = q; }
        if (S == 2) { k += rem * weight; }
        perm[p] = k;
    }
    // fast-path exchange twiddles xtw[k1*P + j] = e^{-2 pi j (j*k1)/L} for the (V,P) split of L
    int V = 0, P = 0;
    fast_shape(L, &V, &P);
    float* xt = f + 7 * L;
    for (int i = 0; i < 2 * L; ++i) xt[i] = 0.f;
    if (V) {
        for (int k1 = 0; k1 < V; ++k1)
            for (int j = 0; j < P; ++j) {
                const double a = -2.0 * pi * (double)((long long)j * k1 % L) / L;
                xt[2 * (k1 * P + j)] = (float)cos(a);
                xt[2 * (k1 * P + j) + 1] = (float)sin(a);
            }
    }
    return P2P_OK;
}

extern "C" int p2p_mdct4_analysis(const float* x, float* spec, const float* window, const void* tables,
                                  int B, int64_t T, int F, int n_fft, int hop, int win, int64_t offset,
                                  float scale, void* stream) {
    using namespace p2p;
    if (int rc = check_geometry(n_fft, hop, win)) return rc;
    if (B < 0 || T < 0 || F < 0) return fail(P2P_EINVAL, "negative size B=%d T=%lld F=%d", B, (long long)T, F);
    if (B == 0 || F == 0) return P2P_OK;
    if (!x || !spec || !window || !tables) return fail(P2P_EINVAL, "null pointer argument");
    if (B > 65535) return fail(P2P_ENOTSUP, "B=%d exceeds 65535 rows per launch", B);
    const int M = n_fft / 2, L = n_fft / 4;
    if (!g_force_general && win == n_fft && hop == M && L >= 64 && (reinterpret_cast<uintptr_t>(spec) & 7) == 0) {
        cudaStream_t st = (cudaStream_t)stream;
        switch (L) {
            case 64: return launch_analysis_fast<fast::Cfg<8, 8>>(x, spec, window, tables, B, T, F, offset, scale, st);
            case 128: return launch_analysis_fast<fast::Cfg<16, 8>>(x, spec, window, tables, B, T, F, offset, scale, st);
            case 256: return launch_analysis_fast<fast::Cfg<16, 16>>(x, spec, window, tables, B, T, F, offset, scale, st);
            case 512: return launch_analysis_fast<fast::Cfg<32, 16>>(x, spec, window, tables, B, T, F, offset, scale, st);
            case 1024: return launch_analysis_fast<fast::Cfg<32, 32>>(x, spec, window, tables, B, T, F, offset, scale, st);
        }
    }
    int fpb = 4096 / M; if (fpb < 1) fpb = 1; if (fpb > 32) fpb = 32; if (fpb > F) fpb = F;
    const int span_max = (fpb - 1) * hop + win;
    const size_t smem = 4ull * (((span_max + 3) & ~3) + ((win + 3) & ~3) + 2 * fpb * L + fpb * M);
    if (int rc = ensure_smem(mdct4_analysis_kernel, smem)) return rc;
    dim3 grid((F + fpb - 1) / fpb, B);
    mdct4_analysis_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(
        x, spec, window, tables, (long long)T, F, n_fft, hop, win, (long long)offset, scale, fpb);
    return check_launch("mdct4_analysis_kernel");
}

extern "C" int p2p_mdct4_synthesis(const float* spec, float* out, const float* window, const void* tables,
                                   int B, int F, int64_t T_out, int n_fft, int hop, int win,
                                   int64_t crop_left, float scale, void* stream) {
    using namespace p2p;
    if (int rc = check_geometry(n_fft, hop, win)) return rc;
    if (B < 0 || T_out < 0 || F < 0 || crop_left < 0)
        return fail(P2P_EINVAL, "negative size B=%d T_out=%lld F=%d crop_left=%lld", B, (long long)T_out, F, (long long)crop_left);
    if (B == 0 || T_out == 0) return P2P_OK;
    if (!out || !window || !tables || (!spec && F > 0)) return fail(P2P_EINVAL, "null pointer argument");
    if (B > 65535) return fail(P2P_ENOTSUP, "B=%d exceeds 65535 rows per launch", B);
    const int M = n_fft / 2, L = n_fft / 4;
    if (!g_force_general && win == n_fft && hop == M && L >= 64 && F > 0) {
        cudaStream_t st = (cudaStream_t)stream;
        switch (L) {
            case 64: return launch_synthesis_fast<fast::Cfg<8, 8>>(spec, out, window, tables, B, F, T_out, crop_left, scale, st);
            case 128: return launch_synthesis_fast<fast::Cfg<16, 8>>(spec, out, window, tables, B, F, T_out, crop_left, scale, st);
            case 256: return launch_synthesis_fast<fast::Cfg<16, 16>>(spec, out, window, tables, B, F, T_out, crop_left, scale, st);
            case 512: return launch_synthesis_fast<fast::Cfg<32, 16>>(spec, out, window, tables, B, F, T_out, crop_left, scale, st);
            case 1024: return launch_synthesis_fast<fast::Cfg<32, 32>>(spec, out, window, tables, B, F, T_out, crop_left, scale, st);
        }
    }
    const int overlap = (win + hop - 1) / hop;            // frames covering one sample
    int nf_target = 8192 / M; if (nf_target < 2) nf_target = 2;
    int hpb = nf_target - overlap; if (hpb < 1) hpb = 1;
    const int chunk = hpb * hop;
    int nfmax = (chunk + win - 2) / hop + 1; if (nfmax > F) nfmax = F; if (nfmax < 1) nfmax = 1;
    const size_t smem = 4ull * (((win + 3) & ~3) + 2 * nfmax * L + nfmax * M);
    if (int rc = ensure_smem(mdct4_synthesis_kernel, smem)) return rc;
    const long long nchunks = (T_out + chunk - 1) / chunk;
    if (nchunks > 2147483647LL) return fail(P2P_ENOTSUP, "T_out too large");
    dim3 grid((unsigned)nchunks, B);
    mdct4_synthesis_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(
        spec, out, window, tables, F, (long long)T_out, n_fft, hop, win, (long long)crop_left, scale,
        chunk, nfmax);
    return check_launch("mdct4_synthesis_kernel");
}
