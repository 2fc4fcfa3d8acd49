// MDCT2 / IMDCT2 (the DCT-II lapped transform the shipped Pix2PixHDModel wires) as two fused kernels.
//
// Replaces models/mdct.py:377-398 (MDCT2.forward: pad, frame, window, zero-extend, DCT_2N_native) and :432-454
// (IMDCT2.forward: IDCT_2N_native / 2, window, fold, crop) plus dct/dct_native.py:7-67.  What the reference computes
// (verified numerically against it, tests/golden/make_golden_mdct2.py):
//     DCT_2N_native(u)[k]  = (2/N) sum_n u[n] cos(pi (2n+1) k / (2N))
//     IDCT_2N_native(X)[n] = X[0] + 2 sum_{k>=1} X[k] cos(pi (2n+1) k / (2N))        (so IDCT(DCT(u)) = 2u)
// with a 2N-point rfft / irfft of a zero-extended row.  Here (Makhoul): one N-point complex FFT in shared memory on the
// even/odd-reordered row v[n] = u[2n], v[N-1-n] = u[2n+1]:  sum_n u[n] cos(pi (2n+1) k / (2N)) = Re(e^{-j pi k/(2N)} FFT_N(v)[k]);
// the inverse runs the same forward FFT on c_k X[k] e^{-j pi k/(2N)} and keeps the real part.  Both directions take
// (scale, c0): c0 multiplies bin 0, which makes each kernel the other's adjoint (autograd) as well:
//     analysis(scale, c0):  X[k] = scale * c0_k * (2/N) * sum_n w[n] x[f hop + n] cos(.)
//     synthesis(scale, c0): out  = scale * crop(OLA_hop(w[n] * 2 sum_k c0_k X[k] cos(.)))     c0_0 = c0, c0_k = 1 (k > 0)
//   MDCT2 = analysis(1, 1); IMDCT2 = synthesis(1/2, 1/2); d MDCT2 / dx = synthesis(1/N, 1); d IMDCT2 / dX = analysis(N/2, 1/2).
// Layout: signal [B,T] fp32, coefficients [B,F,N] fp32 (bins fastest).  Bound: HBM (4 B in + 8 B out per sample at
// hop = N/2).  Two real frames share one complex FFT (real / imaginary lanes).  The canonical geometry (win == n_fft,
// hop == n_fft/2, n_fft 64..1024) runs the register-FFT + TMA-prefetch kernels of mdct2_fast.cuh; the kernels in this
// file are the general path (any hop / win, n_fft up to 4096).
#include "common.cuh"
#include "smem_fft.cuh"
#include "mdct2_fast.cuh"
#include <cmath>
#include <cstdint>

namespace p2p {
namespace {

using smemfft::kThreads;
using smemfft::cmul;
using smemfft::fft_rows_inplace;

struct Tables2 {          // 6N words: tw[N] complex | roots[N] complex | perm[N] int | inv[N] int
    const float2* tw;     // e^{-j pi k / (2N)}
    const float2* roots;  // e^{-2 pi j m / N}
    const int* perm;      // position p after the in-place DIF passes holds FFT bin perm[p]
    const int* inv;       // FFT bin k sits at position inv[k]
};

__device__ __forceinline__ Tables2 make_tables2(const void* t, int N) {
    const float* f = reinterpret_cast<const float*>(t);
    Tables2 r;
    r.tw = reinterpret_cast<const float2*>(f);
    r.roots = reinterpret_cast<const float2*>(f + 2 * N);
    r.perm = reinterpret_cast<const int*>(f + 4 * N);
    r.inv = reinterpret_cast<const int*>(f + 5 * N);
    return r;
}

// Two real frames share one complex FFT (frame A on the real lane, frame B on the imaginary lane):
//   FFT(vA + j vB)[k] = VA[k] + j VB[k],  VA[k] = (Z[k] + conj Z[N-k]) / 2,  VB[k] = -j (Z[k] - conj Z[N-k]) / 2.
// grid (ceil(F/fpb), B), fpb even.  smem: sig[(fpb-1)*hop+win] | win_s[win] | z[(fpb/2)*N] (float2)
__global__ void __launch_bounds__(kThreads)
mdct2_analysis_kernel(const float* __restrict__ x, float* __restrict__ spec, const float* __restrict__ window,
                      const void* __restrict__ tables, long long T, int F, int N, int lgN, int hop, int win,
                      long long offset, float scale, float c0, int fpb) {
    extern __shared__ __align__(16) float smem[];
    const int f0 = blockIdx.x * fpb;
    const int nfr = min(fpb, F - f0);
    const int npair = (nfr + 1) >> 1;
    const int span = (nfr - 1) * hop + win;
    const int span_max = (fpb - 1) * hop + win;
    float* sig = smem;
    float* win_s = sig + ((span_max + 3) & ~3);
    float2* z = reinterpret_cast<float2*>(win_s + ((win + 3) & ~3));
    const Tables2 tb = make_tables2(tables, N);
    const int mask = N - 1;

    const float* xb = x + (long long)blockIdx.y * T;
    const long long t0 = (long long)f0 * hop - offset;
    for (int i = threadIdx.x; i < span; i += kThreads) {
        const long long t = t0 + i;
        sig[i] = (t >= 0 && t < T) ? __ldg(xb + t) : 0.f;
    }
    for (int i = threadIdx.x; i < win; i += kThreads) win_s[i] = __ldg(window + i);
    __syncthreads();

    // v[p] = u[2p] (p < N/2), u[2(N-1-p)+1] otherwise; u = window * frame, zero beyond win
    const int hN = N >> 1;
    for (int w = threadIdx.x; w < npair * N; w += kThreads) {
        const int pr = w >> lgN;
        const int p = w & mask;
        const int i = p < hN ? 2 * p : 2 * (N - 1 - p) + 1;
        float a = 0.f, b = 0.f;
        if (i < win) {
            const float wv = win_s[i];
            a = sig[(2 * pr) * hop + i] * wv;
            if (2 * pr + 1 < nfr) b = sig[(2 * pr + 1) * hop + i] * wv;
        }
        z[w] = make_float2(a, b);
    }
    __syncthreads();

    fft_rows_inplace(z, npair, N, tb.roots);

    // natural-order bins: coalesced table reads and stores, the digit reversal is a shared-memory gather
    const float sc = scale / (float)N;               // (2/N) * the 1/2 of the split
    float* dst = spec + ((long long)blockIdx.y * F + f0) * N;
    for (int w = threadIdx.x; w < npair * N; w += kThreads) {
        const int pr = w >> lgN;
        const int k = w & mask;
        const float2 a = z[(pr << lgN) + __ldg(tb.inv + k)];
        const float2 m = z[(pr << lgN) + __ldg(tb.inv + ((N - k) & mask))];
        const float2 t = __ldg(tb.tw + k);
        const float s = k == 0 ? sc * c0 : sc;
        // VA = (a + conj m), VB = -j (a - conj m)   (the 1/2 lives in sc)
        const float2 va = make_float2(a.x + m.x, a.y - m.y);
        const float2 vb = make_float2(a.y + m.y, m.x - a.x);
        dst[(long long)(2 * pr) * N + k] = s * (va.x * t.x - va.y * t.y);
        if (2 * pr + 1 < nfr) dst[(long long)(2 * pr + 1) * N + k] = s * (vb.x * t.x - vb.y * t.y);
    }
}

// grid (ceil(T_out/chunk), B).  Each CTA owns `chunk` output samples and transforms every frame that overlaps them, so
// the overlap-add is a deterministic gather (no atomics, each sample written once).  Frames are paired on one complex
// FFT: h[k] = tw[k] (X[k] + j X[N-k]) / 2 (h[0] = c0 X[0]) is Hermitian, so FFT(hA + j hB) = vA + j vB with both real.
// smem: win_s[win] | z[ceil(nfmax/2)*N] (float2) | t[nfmax*N]
__global__ void __launch_bounds__(kThreads)
mdct2_synthesis_kernel(const float* __restrict__ spec, float* __restrict__ out, const float* __restrict__ window,
                       const void* __restrict__ tables, int F, long long T_out, int N, int lgN, int hop, int win,
                       long long crop_left, float scale, float c0, int chunk, int nfmax) {
    extern __shared__ __align__(16) float smem[];
    float* win_s = smem;
    float2* z = reinterpret_cast<float2*>(win_s + ((win + 3) & ~3));
    float* t = reinterpret_cast<float*>(z + ((nfmax + 1) >> 1) * N);
    const Tables2 tb = make_tables2(tables, N);
    const int mask = N - 1;

    const long long s_lo = (long long)blockIdx.x * chunk + crop_left;            // OLA coordinates
    const long long s_hi = min((long long)(blockIdx.x + 1) * chunk, T_out) + crop_left;  // exclusive
    long long fl_ = (s_lo - win + 1 + hop - 1);
    fl_ = fl_ > 0 ? fl_ / hop : 0;                                               // ceil, clamped at 0
    const int f_lo = (int)min(fl_, (long long)F);
    const int f_hi = (int)min((s_hi - 1) / hop, (long long)F - 1);              // inclusive
    const int nfr = max(0, f_hi - f_lo + 1);
    const int npair = (nfr + 1) >> 1;

    for (int i = threadIdx.x; i < win; i += kThreads) win_s[i] = __ldg(window + i);
    {   // coalesced load of the coefficient rows into t
        const float* src = spec + ((long long)blockIdx.y * F + f_lo) * N;
        const int total = nfr * N;
        if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
            const float4* s4 = reinterpret_cast<const float4*>(src);
            float4* t4 = reinterpret_cast<float4*>(t);
            for (int i = threadIdx.x; i < (total >> 2); i += kThreads) t4[i] = __ldg(s4 + i);
        } else {
            for (int i = threadIdx.x; i < total; i += kThreads) t[i] = __ldg(src + i);
        }
    }
    __syncthreads();

    for (int w = threadIdx.x; w < npair * N; w += kThreads) {
        const int pr = w >> lgN;
        const int k = w & mask;
        const float* ta = t + ((2 * pr) << lgN);
        const bool hb = 2 * pr + 1 < nfr;
        const float* tbp = ta + N;
        float2 ha, hbv;
        if (k == 0) {
            ha = make_float2(c0 * ta[0], 0.f);
            hbv = make_float2(hb ? c0 * tbp[0] : 0.f, 0.f);
        } else {
            const float2 tw = __ldg(tb.tw + k);
            const float ak = ta[k], am = ta[N - k];
            ha = make_float2(0.5f * (tw.x * ak - tw.y * am), 0.5f * (tw.x * am + tw.y * ak));
            const float bk = hb ? tbp[k] : 0.f, bm = hb ? tbp[N - k] : 0.f;
            hbv = make_float2(0.5f * (tw.x * bk - tw.y * bm), 0.5f * (tw.x * bm + tw.y * bk));
        }
        z[w] = make_float2(ha.x - hbv.y, ha.y + hbv.x);
    }
    __syncthreads();

    fft_rows_inplace(z, npair, N, tb.roots);

    // position p holds v[n], n = perm[p]; y[2n] = 2 v[n] (n < N/2), y[2(N-1-n)+1] = 2 v[n] otherwise
    const int hN = N >> 1;
    for (int w = threadIdx.x; w < npair * N; w += kThreads) {
        const int pr = w >> lgN;
        const int n = __ldg(tb.perm + (w & mask));
        const int i = n < hN ? 2 * n : 2 * (N - 1 - n) + 1;
        const float2 v = z[w];
        t[((2 * pr) << lgN) + i] = 2.f * v.x;
        if (2 * pr + 1 < nfr) t[((2 * pr + 1) << lgN) + i] = 2.f * v.y;
    }
    __syncthreads();

    float* ob = out + (long long)blockIdx.y * T_out;
    const int n_out = (int)(s_hi - s_lo);
    for (int i = threadIdx.x; i < n_out; i += kThreads) {
        const long long s = s_lo + i;
        long long a = s - win + 1 + hop - 1;                                     // frames with 0 <= s - f*hop < win
        int fa = (int)(a > 0 ? a / hop : 0);
        fa = max(fa, f_lo);
        const int fb = (int)min(s / hop, (long long)f_hi);
        float acc = 0.f;
        for (int f = fa; f <= fb; ++f) {
            const int n = (int)(s - (long long)f * hop);
            acc += win_s[n] * t[((f - f_lo) << lgN) + n];
        }
        ob[s - crop_left] = scale * acc;
    }
}

inline int ilog2i(int v) { int l = 0; while ((1 << l) < v) ++l; return l; }

constexpr int kMaxSmem2 = 227 * 1024;

int check_geometry2(int n_fft, int hop, int win) {
    if (!is_pow2(n_fft) || n_fft < 4 || n_fft > 4096)
        return fail(P2P_ENOTSUP, "n_fft must be a power of two in [4, 4096], got %d", n_fft);
    if (win < 1 || win > n_fft) return fail(P2P_EINVAL, "win_length %d must be in [1, n_fft=%d]", win, n_fft);
    if (hop < 1 || hop > win) return fail(P2P_EINVAL, "hop_length %d must be in [1, win_length=%d]", hop, win);
    return P2P_OK;
}

template <typename K>
int ensure_smem2(K kernel, size_t bytes) {
    if (bytes > (size_t)kMaxSmem2) return fail(P2P_ENOTSUP, "needs %zu B of shared memory", bytes);
    if (bytes > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e != cudaSuccess) return fail(P2P_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    return P2P_OK;
}

int sm_count_and_blocks2(const void* kernel, size_t smem, int* blocks_per_sm) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, kernel, fast2::kThreads, smem) != cudaSuccess) return -1;
    return sms;
}

template <class C>
int launch_analysis2_fast(const float* x, float* spec, const float* window, const void* tables, int B, long long T, int F,
                          long long offset, float scale, float c0, cudaStream_t stream) {
    auto kernel = fast2::mdct2_analysis_fast<C>;
    const size_t smem = C::A_SMEM;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail(P2P_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int bps = 0;
    const int sms = sm_count_and_blocks2((const void*)kernel, smem, &bps);
    if (sms <= 0 || bps <= 0) { cudaGetLastError(); return fail(P2P_ECUDA, "occupancy query failed (smem %zu)", smem); }
    const int tiles_per_row = (F + C::FPT - 1) / C::FPT;
    const long long total = (long long)tiles_per_row * B;
    const long long grid = total < (long long)sms * bps ? total : (long long)sms * bps;
    const int aligned = ((reinterpret_cast<uintptr_t>(x) & 15) == 0) && (T % 4 == 0) && (offset % 4 == 0);
    kernel<<<(unsigned)grid, fast2::kThreads, smem, stream>>>(x, spec, window, tables, T, F, offset, scale, c0, tiles_per_row,
                                                              total, aligned);
    return check_launch("mdct2_analysis_fast");
}

template <class C>
int launch_synthesis2_fast(const float* spec, float* out, const float* window, const void* tables, int B, int F,
                           long long T_out, long long crop_left, float scale, float c0, cudaStream_t stream) {
    auto kernel = fast2::mdct2_synthesis_fast<C>;
    const size_t smem = C::S_SMEM;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail(P2P_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int bps = 0;
    const int sms = sm_count_and_blocks2((const void*)kernel, smem, &bps);
    if (sms <= 0 || bps <= 0) { cudaGetLastError(); return fail(P2P_ECUDA, "occupancy query failed (smem %zu)", smem); }
    const long long hf = crop_left / C::M, hl = (crop_left + T_out - 1) / C::M;
    if (hl > 2147483000LL) return fail(P2P_ENOTSUP, "T_out too large");
    const int tiles_per_row = (int)((hl - hf + 1 + C::FPT - 2) / (C::FPT - 1));
    const long long total = (long long)tiles_per_row * B;
    const long long grid = total < (long long)sms * bps ? total : (long long)sms * bps;
    const int aligned_in = (reinterpret_cast<uintptr_t>(spec) & 15) == 0;
    const int vec_out = ((reinterpret_cast<uintptr_t>(out) & 15) == 0) && (T_out % 4 == 0) && (crop_left % 4 == 0);
    kernel<<<(unsigned)grid, fast2::kThreads, smem, stream>>>(spec, out, window, tables, F, T_out, crop_left, scale, c0,
                                                              (int)hf, (int)hl, tiles_per_row, total, aligned_in, vec_out);
    return check_launch("mdct2_synthesis_fast");
}

// n_fft 512 synthesis: one coefficient buffer and two CTAs per SM (default) or two buffers and one CTA (P2P_MDCT2_NBUF=2)
int synthesis_nbuf_512() {
    static const int n = [] { const char* e = getenv("P2P_MDCT2_NBUF"); return (e && e[0] == '2') ? 2 : 1; }();
    return n;
}

int g_force_general2 = 0;
bool fast2_enabled() {
    static const bool on = [] { const char* e = getenv("P2P_MDCT2_FAST"); return !(e && e[0] == '0'); }();
    return on && !g_force_general2;
}

}  // namespace
}  // namespace p2p

extern "C" int p2p_mdct2_force_general(int on) { const int old = p2p::g_force_general2; p2p::g_force_general2 = on; return old; }

extern "C" int p2p_mdct2_table_words(int n_fft) {
    if (!p2p::is_pow2(n_fft) || n_fft < 4 || n_fft > 4096)
        return p2p::fail(P2P_ENOTSUP, "n_fft must be a power of two in [4, 4096], got %d", n_fft);
    return 6 * n_fft;
}

extern "C" int p2p_mdct2_fill_tables(int n_fft, void* host_tables) {
    const int words = p2p_mdct2_table_words(n_fft);
    if (words < 0) return words;
    if (!host_tables) return p2p::fail(P2P_EINVAL, "host_tables is null");
    const int N = n_fft;
    float* f = static_cast<float*>(host_tables);
    int* perm = reinterpret_cast<int*>(f + 4 * N);
    int* inv = perm + N;
    const double pi = 3.14159265358979323846264338327950288;
    for (int k = 0; k < N; ++k) {
        const double a = -pi * k / (2.0 * N), c = -2.0 * pi * k / N;
        f[2 * k] = (float)cos(a);           f[2 * k + 1] = (float)sin(a);
        f[2 * N + 2 * k] = (float)cos(c);   f[2 * N + 2 * k + 1] = (float)sin(c);
        perm[k] = p2p::smemfft::dif_position_to_bin(k, N);
    }
    for (int p = 0; p < N; ++p) inv[perm[p]] = p;
    return P2P_OK;
}

extern "C" int p2p_mdct2_analysis(const float* x, float* spec, const float* window, const void* tables, int B, int64_t T,
                                  int F, int n_fft, int hop, int win, int64_t offset, float scale, float c0, void* stream) {
    using namespace p2p;
    if (int rc = check_geometry2(n_fft, hop, win)) return rc;
    if (B < 0 || T < 0 || F < 0) return fail(P2P_EINVAL, "negative size B=%d T=%lld F=%d", B, (long long)T, F);
    if (B == 0 || F == 0) return P2P_OK;
    if (!spec || !window || !tables || (!x && T > 0)) return fail(P2P_EINVAL, "null pointer argument");
    if (B > 65535) return fail(P2P_ENOTSUP, "B=%d exceeds 65535 rows per launch", B);
    if (fast2_enabled() && win == n_fft && hop * 2 == n_fft && x) {
        cudaStream_t st = (cudaStream_t)stream;
        switch (n_fft) {
            case 64: return launch_analysis2_fast<fast2::Cfg2<8, 8, 2>>(x, spec, window, tables, B, T, F, offset, scale, c0, st);
            case 128: return launch_analysis2_fast<fast2::Cfg2<16, 8, 2>>(x, spec, window, tables, B, T, F, offset, scale, c0, st);
            case 256: return launch_analysis2_fast<fast2::Cfg2<16, 16, 2>>(x, spec, window, tables, B, T, F, offset, scale, c0, st);
            case 512: return launch_analysis2_fast<fast2::Cfg2<32, 16, 1>>(x, spec, window, tables, B, T, F, offset, scale, c0, st);
            case 1024: return launch_analysis2_fast<fast2::Cfg2<32, 32, 2>>(x, spec, window, tables, B, T, F, offset, scale, c0, st);
        }
    }
    int fpb = 8192 / n_fft; if (fpb < 2) fpb = 2; if (fpb > 32) fpb = 32;      // even: frames are transformed in pairs
    const int span_max = (fpb - 1) * hop + win;
    const size_t smem = 4ull * (((span_max + 3) & ~3) + ((win + 3) & ~3) + 2 * (fpb / 2) * n_fft);
    if (int rc = ensure_smem2(mdct2_analysis_kernel, smem)) return rc;
    dim3 grid((F + fpb - 1) / fpb, B);
    mdct2_analysis_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(x, spec, window, tables, (long long)T, F, n_fft,
                                                                         ilog2i(n_fft), hop, win, (long long)offset, scale, c0, fpb);
    return check_launch("mdct2_analysis_kernel");
}

extern "C" int p2p_mdct2_synthesis(const float* spec, float* out, const float* window, const void* tables, int B, int F,
                                   int64_t T_out, int n_fft, int hop, int win, int64_t crop_left, float scale, float c0,
                                   void* stream) {
    using namespace p2p;
    if (int rc = check_geometry2(n_fft, hop, win)) return rc;
    if (B < 0 || T_out < 0 || F < 0 || crop_left < 0)
        return fail(P2P_EINVAL, "negative size B=%d T_out=%lld F=%d crop_left=%lld", B, (long long)T_out, F, (long long)crop_left);
    if (B == 0 || T_out == 0) return P2P_OK;
    if (!out || !window || !tables || (!spec && F > 0)) return fail(P2P_EINVAL, "null pointer argument");
    if (B > 65535) return fail(P2P_ENOTSUP, "B=%d exceeds 65535 rows per launch", B);
    if (fast2_enabled() && win == n_fft && hop * 2 == n_fft && F > 0) {
        cudaStream_t st = (cudaStream_t)stream;
        switch (n_fft) {
            case 64: return launch_synthesis2_fast<fast2::Cfg2<8, 8, 2>>(spec, out, window, tables, B, F, T_out, crop_left, scale, c0, st);
            case 128: return launch_synthesis2_fast<fast2::Cfg2<16, 8, 2>>(spec, out, window, tables, B, F, T_out, crop_left, scale, c0, st);
            case 256: return launch_synthesis2_fast<fast2::Cfg2<16, 16, 2>>(spec, out, window, tables, B, F, T_out, crop_left, scale, c0, st);
            case 512:
                if (synthesis_nbuf_512() == 2)
                    return launch_synthesis2_fast<fast2::Cfg2<32, 16, 2>>(spec, out, window, tables, B, F, T_out, crop_left, scale, c0, st);
                return launch_synthesis2_fast<fast2::Cfg2<32, 16, 1>>(spec, out, window, tables, B, F, T_out, crop_left, scale, c0, st);
            case 1024: return launch_synthesis2_fast<fast2::Cfg2<32, 32, 2>>(spec, out, window, tables, B, F, T_out, crop_left, scale, c0, st);
        }
    }
    const int overlap = (win + hop - 1) / hop;            // frames covering one sample
    int nf_target = 8192 / n_fft; if (nf_target < 2) nf_target = 2;
    int hpb = nf_target - overlap; if (hpb < 1) hpb = 1;
    const int chunk = hpb * hop;
    int nfmax = (chunk + win - 2) / hop + 1; if (nfmax > F) nfmax = F; if (nfmax < 1) nfmax = 1;
    const size_t smem = 4ull * (((win + 3) & ~3) + 2 * ((nfmax + 1) / 2) * n_fft + nfmax * n_fft);
    if (int rc = ensure_smem2(mdct2_synthesis_kernel, smem)) return rc;
    const long long nchunks = (T_out + chunk - 1) / chunk;
    if (nchunks > 2147483647LL) return fail(P2P_ENOTSUP, "T_out too large");
    dim3 grid((unsigned)nchunks, B);
    mdct2_synthesis_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(spec, out, window, tables, F, (long long)T_out, n_fft,
                                                                          ilog2i(n_fft), hop, win, (long long)crop_left, scale, c0, chunk, nfmax);
    return check_launch("mdct2_synthesis_kernel");
}
