// Fast path of the MDCT4 analysis / synthesis kernels for the canonical geometry
// win == n_fft, hop == n_fft/2 (what MDCT4/IMDCT4 are used with: README.md:93-111 of the reference).
//
// L = n_fft/4 complex points per frame are held in registers as V values on each of P lanes
// (L = V*P): radix-V FFT in registers -> twiddle -> one exchange through warp-private shared memory
// -> V/P radix-P FFTs in registers.  A warp works on 32/P frames at once and never needs a CTA-wide
// barrier inside a tile.  A persistent CTA walks (row, tile) work items; the signal span of the NEXT
// tile is fetched by TMA bulk copies (cp.async.bulk + mbarrier) while the current one is transformed.
#pragma once
#include "reg_fft.cuh"
#include <cstdint>

namespace p2p {
namespace fast {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;

template <int V_, int P_>
struct Cfg {
    static constexpr int V = V_, P = P_;
    static constexpr int L = V * P, M = 2 * L, N = 4 * L;
    static constexpr int G = 32 / P;                 // frames per warp
    static constexpr int FPT = kWarps * G;           // frames per tile
    static constexpr int Q = V / P;                  // radix-P FFTs per lane in stage 2
    static constexpr int ROWPAD = (P == 8) ? 16 : 0; // floats; de-conflicts the 4 frames of a warp
    static constexpr int ROW = M + ROWPAD;           // one hop row in shared memory
    static constexpr int SIG = (FPT + 1) * ROW;      // floats per signal buffer
    static constexpr int XW = 2 * V * (P + 1);       // words of one exchange region (row stride P+1)
    static constexpr int XREG = XW + ((P == 8 && XW % 32 == 0) ? 16 : 0);
    static constexpr bool TABLES_IN_SMEM = (L <= 512);   // pre/post twiddles; else read through L1
    static constexpr bool XTW_IN_REGS = (V <= 16);       // exchange twiddles live in registers
    static constexpr int XV = XTW_IN_REGS ? V : 1;
    static constexpr int WS = N + ROWPAD;            // signed window, same padded layout as a frame
    static_assert(V >= P && V % P == 0, "V must be a multiple of P");
    // shared memory layout (floats): sig[2][SIG] | ws[WS] | exch[kWarps*G][XREG] | tables[4L]? | mbar
    static constexpr int OFF_WS = 2 * SIG;
    static constexpr int OFF_X = OFF_WS + ((WS + 3) & ~3);
    static constexpr int OFF_TB = OFF_X + kWarps * G * XREG;
    static constexpr int OFF_BAR = OFF_TB + (TABLES_IN_SMEM ? 4 * L : 0);
    static constexpr size_t SMEM_BYTES = 4ull * OFF_BAR + 16;
};

// ---- PTX helpers: mbarrier + TMA bulk copy ---------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct TileCoord { int row; int f0; int nfr; };

template <class C>
__device__ __forceinline__ TileCoord tile_coord(long long w, int tiles_per_row, int F) {
    TileCoord t;
    t.row = (int)(w / tiles_per_row);
    t.f0 = (int)(w - (long long)t.row * tiles_per_row) * C::FPT;
    t.nfr = min(C::FPT, F - t.f0);
    return t;
}

// Fetch the (nfr+1) hop rows of a tile into `sig`: in-range 16B-aligned rows by TMA bulk copy signalling
// `bar`, the rest (zero padding outside [0,T), unaligned input) by ordinary loads from all threads.
template <class C>
__device__ __forceinline__ void fetch_tile(float* sig, uint64_t* bar, const float* __restrict__ x, long long T,
                                           long long offset, const TileCoord tc, bool aligned) {
    const float* xb = x + (long long)tc.row * T;
    const long long t_first = (long long)tc.f0 * C::M - offset;
    const int rows = tc.nfr + 1;
    if (threadIdx.x == 0) {
        uint32_t bytes = 0;
        for (int r = 0; r < rows; ++r) {
            const long long t = t_first + (long long)r * C::M;
            if (aligned && t >= 0 && t + C::M <= T) bytes += C::M * 4;
        }
        mbar_expect_tx(bar, bytes);
        for (int r = 0; r < rows; ++r) {
            const long long t = t_first + (long long)r * C::M;
            if (aligned && t >= 0 && t + C::M <= T) bulk_g2s(sig + r * C::ROW, xb + t, C::M * 4, bar);
        }
    }
    for (int r = 0; r < rows; ++r) {
        const long long t = t_first + (long long)r * C::M;
        if (aligned && t >= 0 && t + C::M <= T) continue;
        for (int i = threadIdx.x; i < C::M; i += kThreads) {
            const long long tt = t + i;
            sig[r * C::ROW + i] = (tt >= 0 && tt < T) ? __ldg(xb + tt) : 0.f;
        }
    }
}

template <bool IN_SMEM>
__device__ __forceinline__ float2 ldt(const float2* p) {
    if constexpr (IN_SMEM) return *p; else return __ldg(p);
}

template <int R>
__device__ __forceinline__ int brev(int k) { return (int)(__brev((unsigned)k) >> (32 - ilog2(R))); }

// One warp transforms its G frames of the current tile: fold + signed window + pre-twiddle ->
// radix-V -> exchange -> radix-P -> post-twiddle -> pair up (X[2k], X[2k+1]) by one shuffle -> STG.64.
//   sig : tile signal buffer (rows of C::ROW floats), ws: signed, scaled, padded window
//   ex  : this warp's exchange regions, pre/post/xtw: twiddle tables (shared or global)
template <class C>
__device__ __forceinline__ void analysis_warp(const float* __restrict__ sig, const float* __restrict__ ws,
                                              float* __restrict__ ex_warp, const float2* pre, const float2* post,
                                              const float2* xtw, const float2 (&xr)[C::XV],
                                              float* __restrict__ spec_tile, int nfr, int warp, int lane) {
    constexpr int V = C::V, P = C::P, L = C::L, PAD = C::ROWPAD;
    const int j = lane % P, g = lane / P;
    const bool flip = (lane >> 4) & 1;
    const int fl = warp * C::G + g;
    const float* sf = sig + fl * C::ROW;
    const int step = flip ? -2 * P : 2 * P;
    const int e2j = 2 * j;
    // (first, second) base offsets of the four fold pairs; see DESIGN.md "fold with a signed window"
    auto b1 = [&](int ce, int co) { return flip ? co - e2j : ce + e2j; };
    auto b2 = [&](int ce, int co) { return flip ? ce + e2j : co - e2j; };
    const int aRe1 = b1(3 * L + PAD, 3 * L - 1 + PAD), aRe2 = b2(3 * L + PAD, 3 * L - 1 + PAD);
    const int aIm1 = b1(L, L - 1), aIm2 = b2(L, L - 1);
    const int bRe1 = b1(0, 2 * L - 1), bRe2 = b2(0, 2 * L - 1);
    const int bIm1 = b1(2 * L + PAD, 4 * L - 1 + PAD), bIm2 = b2(2 * L + PAD, 4 * L - 1 + PAD);

    float2 z[V];
#pragma unroll
    for (int n1 = 0; n1 < V; ++n1) {
        const bool caseA = n1 < V / 2;
        const int i = caseA ? n1 : n1 - V / 2;
        const int r1 = (caseA ? aRe1 : bRe1) + step * i, r2 = (caseA ? aRe2 : bRe2) - step * i;
        const int i1 = (caseA ? aIm1 : bIm1) + step * i, i2 = (caseA ? aIm2 : bIm2) - step * i;
        const float re = ws[r1] * sf[r1] + ws[r2] * sf[r2];
        const float im = ws[i1] * sf[i1] + ws[i2] * sf[i2];
        z[n1] = cmulf(make_float2(re, im), ldt<C::TABLES_IN_SMEM>(pre + P * n1 + j));
    }
    fft_reg<V>(z);

    float2* ex = reinterpret_cast<float2*>(ex_warp + g * C::XREG);
#pragma unroll
    for (int k1 = 0; k1 < V; ++k1) {
        float2 v = z[brev<V>(k1)];
        if (k1 > 0) {
            if constexpr (C::XTW_IN_REGS) v = cmulf(v, xr[k1]);
            else v = cmulf(v, __ldg(xtw + k1 * P + j));
        }
        ex[k1 * (P + 1) + j] = v;
    }
    __syncwarp();

    float re_[C::Q][P], nim[C::Q][P];
#pragma unroll
    for (int m = 0; m < C::Q; ++m) {
        float2 b[P];
        const int k1 = j + P * m;
#pragma unroll
        for (int n2 = 0; n2 < P; ++n2) b[n2] = ex[k1 * (P + 1) + n2];
        fft_reg<P>(b);
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float2 y = cmulf(b[brev<P>(k2)], ldt<C::TABLES_IN_SMEM>(post + k1 + V * k2));
            re_[m][k2] = y.x;
            nim[m][k2] = -y.y;
        }
    }
    __syncwarp();   // exchange regions are reused by the next tile

    float* sp = spec_tile + (long long)fl * C::M;
    const bool valid = fl < nfr;
#pragma unroll
    for (int m = 0; m < C::Q; ++m) {
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float other = __shfl_xor_sync(0xffffffffu, nim[C::Q - 1 - m][P - 1 - k2], P - 1);
            if (valid) {
                const int k = j + P * m + V * k2;
                *reinterpret_cast<float2*>(sp + 2 * k) = make_float2(re_[m][k2], other);
            }
        }
    }
}

template <class C>
__global__ void __launch_bounds__(kThreads)
mdct4_analysis_fast(const float* __restrict__ x, float* __restrict__ spec, const float* __restrict__ window,
                    const void* __restrict__ tables, long long T, int F, long long offset, float scale,
                    int tiles_per_row, long long total_tiles, int aligned) {
    extern __shared__ __align__(128) float smem[];
    constexpr int L = C::L, M = C::M, N = C::N;
    float* sig0 = smem;
    float* ws = smem + C::OFF_WS;
    float* exch = smem + C::OFF_X;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BAR);
    const float* gt = reinterpret_cast<const float*>(tables);
    const float2* pre = reinterpret_cast<const float2*>(gt);
    const float2* post = reinterpret_cast<const float2*>(gt + 2 * L);
    const float2* xtw = reinterpret_cast<const float2*>(gt + 7 * L);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
    for (int p = threadIdx.x; p < N; p += kThreads) {
        const float w = __ldg(window + p) * scale;
        ws[p + (p >= 2 * L ? C::ROWPAD : 0)] = p < L ? w : -w;
    }
    if constexpr (C::TABLES_IN_SMEM) {
        float* tb = smem + C::OFF_TB;
        for (int i = threadIdx.x; i < 2 * L; i += kThreads) {
            tb[i] = __ldg(gt + i); tb[2 * L + i] = __ldg(gt + 2 * L + i);
        }
        pre = reinterpret_cast<const float2*>(tb);
        post = reinterpret_cast<const float2*>(tb + 2 * L);
    }
    float2 xr[C::XV];
    if constexpr (C::XTW_IN_REGS) {
#pragma unroll
        for (int k1 = 0; k1 < C::V; ++k1) xr[k1] = __ldg(xtw + k1 * C::P + (lane % C::P));
    } else {
        xr[0] = make_float2(1.f, 0.f);
    }
    __syncthreads();

    long long w = blockIdx.x;
    if (w < total_tiles) fetch_tile<C>(sig0, &bars[0], x, T, offset, tile_coord<C>(w, tiles_per_row, F), aligned);
    __syncthreads();
    uint32_t phase[2] = {0, 0};
    int buf = 0;
    for (; w < total_tiles; w += gridDim.x) {
        const TileCoord tc = tile_coord<C>(w, tiles_per_row, F);
        const long long wn = w + gridDim.x;
        if (wn < total_tiles)
            fetch_tile<C>(sig0 + (buf ^ 1) * C::SIG, &bars[buf ^ 1], x, T, offset, tile_coord<C>(wn, tiles_per_row, F), aligned);
        mbar_wait(&bars[buf], phase[buf]);
        phase[buf] ^= 1;
        if (warp * C::G < tc.nfr)
            analysis_warp<C>(sig0 + buf * C::SIG, ws, exch + warp * C::G * C::XREG, pre, post, xtw, xr,
                             spec + ((long long)tc.row * F + tc.f0) * M, tc.nfr, warp, lane);
        __syncthreads();
        buf ^= 1;
    }
}

// ---- synthesis ------------------------------------------------------------------------------------
// A tile = FPT consecutive coefficient rows (frames fA .. fA+FPT-1) of one batch row; it produces the
// FPT-1 output hop slots that lie between them (slot h = first half of frame h + second half of frame
// h-1), so consecutive tiles overlap by one frame and every output sample is written exactly once.
template <class C>
struct SynCfg {
    static constexpr int ROWS = C::FPT;
    static constexpr int BUF = ROWS * C::ROW;                       // floats per coefficient buffer
    static constexpr int OFF_X = 2 * BUF;
    static constexpr int OFF_TB = OFF_X + kWarps * C::G * C::XREG;
    static constexpr int OFF_BAR = OFF_TB + (C::TABLES_IN_SMEM ? 4 * C::L : 0);
    static constexpr size_t SMEM_BYTES = 4ull * OFF_BAR + 16;
    static constexpr int G4 = C::M / 4;                             // float4 groups per hop slot
    static constexpr int NQ = G4 > kThreads ? G4 / kThreads : 1;    // window register sets per thread
    static constexpr int SS = G4 < kThreads ? kThreads / G4 : 1;    // slots covered per pass
};

template <class C>
__device__ __forceinline__ void fetch_rows(float* buf, uint64_t* bar, const float* __restrict__ spec_row, int fA,
                                           int F, bool aligned) {
    if (threadIdx.x == 0) {
        uint32_t bytes = 0;
        for (int r = 0; r < C::FPT; ++r) {
            const int f = fA + r;
            if (aligned && f >= 0 && f < F) bytes += C::M * 4;
        }
        mbar_expect_tx(bar, bytes);
        for (int r = 0; r < C::FPT; ++r) {
            const int f = fA + r;
            if (aligned && f >= 0 && f < F) bulk_g2s(buf + r * C::ROW, spec_row + (long long)f * C::M, C::M * 4, bar);
        }
    }
    if (!aligned) {
        for (int r = 0; r < C::FPT; ++r) {
            const int f = fA + r;
            if (f < 0 || f >= F) continue;
            for (int i = threadIdx.x; i < C::M; i += kThreads) buf[r * C::ROW + i] = __ldg(spec_row + (long long)f * C::M + i);
        }
    }
}

// One warp: coefficient rows of its G frames -> DCT-IV -> t rows written back IN PLACE (same rows).
template <class C>
__device__ __forceinline__ void synthesis_warp(float* __restrict__ buf, float* __restrict__ ex_warp, const float2* pre,
                                               const float2* post, const float2* xtw, const float2 (&xr)[C::XV],
                                               int fA, int F, int warp, int lane) {
    constexpr int V = C::V, P = C::P, M = C::M;
    const int j = lane % P, g = lane / P;
    const bool flip = (lane >> 4) & 1;
    const int fl = warp * C::G + g;
    float* row = buf + fl * C::ROW;
    const bool valid = (fA + fl) >= 0 && (fA + fl) < F;
    const int step = flip ? -2 * P : 2 * P;
    const int first = flip ? M - 1 - 2 * j : 2 * j, second = flip ? 2 * j : M - 1 - 2 * j;

    float2 z[V];
#pragma unroll
    for (int n1 = 0; n1 < V; ++n1) {
        const float a = row[first + step * n1], b = row[second - step * n1];
        z[n1] = cmulf(make_float2(flip ? b : a, flip ? a : b), ldt<C::TABLES_IN_SMEM>(pre + P * n1 + j));
    }
    fft_reg<V>(z);

    float2* ex = reinterpret_cast<float2*>(ex_warp + g * C::XREG);
#pragma unroll
    for (int k1 = 0; k1 < V; ++k1) {
        float2 v = z[brev<V>(k1)];
        if (k1 > 0) {
            if constexpr (C::XTW_IN_REGS) v = cmulf(v, xr[k1]);
            else v = cmulf(v, __ldg(xtw + k1 * P + j));
        }
        ex[k1 * (P + 1) + j] = v;
    }
    __syncwarp();

    float re_[C::Q][P], nim[C::Q][P];
#pragma unroll
    for (int m = 0; m < C::Q; ++m) {
        float2 b[P];
        const int k1 = j + P * m;
#pragma unroll
        for (int n2 = 0; n2 < P; ++n2) b[n2] = ex[k1 * (P + 1) + n2];
        fft_reg<P>(b);
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float2 y = cmulf(b[brev<P>(k2)], ldt<C::TABLES_IN_SMEM>(post + k1 + V * k2));
            re_[m][k2] = y.x;
            nim[m][k2] = -y.y;
        }
    }
    __syncwarp();

#pragma unroll
    for (int m = 0; m < C::Q; ++m) {
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float other = __shfl_xor_sync(0xffffffffu, nim[C::Q - 1 - m][P - 1 - k2], P - 1);
            const int k = j + P * m + V * k2;
            *reinterpret_cast<float2*>(row + 2 * k) = valid ? make_float2(re_[m][k2], other) : make_float2(0.f, 0.f);
        }
    }
}

template <class C>
__global__ void __launch_bounds__(kThreads)
mdct4_synthesis_fast(const float* __restrict__ spec, float* __restrict__ out, const float* __restrict__ window,
                     const void* __restrict__ tables, int F, long long T_out, long long crop_left, float scale,
                     int h_first, int h_last, int tiles_per_row, long long total_tiles, int aligned_in, int vec_out) {
    using S = SynCfg<C>;
    extern __shared__ __align__(128) float smem[];
    constexpr int L = C::L, M = C::M;
    float* buf0 = smem;
    float* exch = smem + S::OFF_X;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::OFF_BAR);
    const float* gt = reinterpret_cast<const float*>(tables);
    const float2* pre = reinterpret_cast<const float2*>(gt);
    const float2* post = reinterpret_cast<const float2*>(gt + 2 * L);
    const float2* xtw = reinterpret_cast<const float2*>(gt + 7 * L);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
    if constexpr (C::TABLES_IN_SMEM) {
        float* tb = smem + S::OFF_TB;
        for (int i = threadIdx.x; i < 2 * L; i += kThreads) { tb[i] = __ldg(gt + i); tb[2 * L + i] = __ldg(gt + 2 * L + i); }
        pre = reinterpret_cast<const float2*>(tb);
        post = reinterpret_cast<const float2*>(tb + 2 * L);
    }
    float2 xr[C::XV];
    if constexpr (C::XTW_IN_REGS) {
#pragma unroll
        for (int k1 = 0; k1 < C::V; ++k1) xr[k1] = __ldg(xtw + k1 * C::P + (lane % C::P));
    } else {
        xr[0] = make_float2(1.f, 0.f);
    }
    // overlap-add role of this thread: float4 group(s) q of every SS-th hop slot; window (with the
    // unfold signs and the output scale folded in) stays in registers
    const int q0 = threadIdx.x % S::G4;
    const int sl0 = (S::G4 < kThreads) ? threadIdx.x / S::G4 : 0;
    float w1[S::NQ][4], w2[S::NQ][4];
#pragma unroll
    for (int qi = 0; qi < S::NQ; ++qi) {
        const int i = 4 * (q0 + kThreads * qi);
        const bool lo = i < L;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float a = __ldg(window + i + c) * scale, b = __ldg(window + 2 * L + i + c) * scale;
            w1[qi][c] = lo ? a : -a;
            w2[qi][c] = -b;
        }
    }
    __syncthreads();

    auto coord = [&](long long w, int& row, int& h0, int& ns) {
        row = (int)(w / tiles_per_row);
        h0 = h_first + (int)(w - (long long)row * tiles_per_row) * (C::FPT - 1);
        ns = min(C::FPT - 1, h_last + 1 - h0);
    };
    long long w = blockIdx.x;
    int row, h0, ns;
    if (w < total_tiles) {
        coord(w, row, h0, ns);
        fetch_rows<C>(buf0, &bars[0], spec + (long long)row * F * M, h0 - 1, F, aligned_in);
    }
    __syncthreads();
    uint32_t phase[2] = {0, 0};
    int b = 0;
    for (; w < total_tiles; w += gridDim.x) {
        coord(w, row, h0, ns);
        const long long wn = w + gridDim.x;
        if (wn < total_tiles) {
            int rown, h0n, nsn;
            coord(wn, rown, h0n, nsn);
            fetch_rows<C>(buf0 + (b ^ 1) * S::BUF, &bars[b ^ 1], spec + (long long)rown * F * M, h0n - 1, F, aligned_in);
        }
        mbar_wait(&bars[b], phase[b]);
        phase[b] ^= 1;
        float* buf = buf0 + b * S::BUF;
        if (warp * C::G <= ns)       // frames fA .. fA+ns are needed (ns+1 rows)
            synthesis_warp<C>(buf, exch + warp * C::G * C::XREG, pre, post, xtw, xr, h0 - 1, F, warp, lane);
        __syncthreads();

        float* orow = out + (long long)row * T_out;
        for (int sl = sl0; sl < ns; sl += S::SS) {
            const float* tp = buf + sl * C::ROW;          // frame h-1
            const float* th = tp + C::ROW;                // frame h
#pragma unroll
            for (int qi = 0; qi < S::NQ; ++qi) {
                const int i = 4 * (q0 + kThreads * qi);
                const bool lo = i < L;
                const float4 A = *reinterpret_cast<const float4*>(th + (lo ? L + i : 3 * L - 4 - i));
                const float4 Bv = *reinterpret_cast<const float4*>(tp + (lo ? L - 4 - i : i - L));
                float4 o;
                o.x = w1[qi][0] * (lo ? A.x : A.w) + w2[qi][0] * (lo ? Bv.w : Bv.x);
                o.y = w1[qi][1] * (lo ? A.y : A.z) + w2[qi][1] * (lo ? Bv.z : Bv.y);
                o.z = w1[qi][2] * (lo ? A.z : A.y) + w2[qi][2] * (lo ? Bv.y : Bv.z);
                o.w = w1[qi][3] * (lo ? A.w : A.x) + w2[qi][3] * (lo ? Bv.x : Bv.w);
                const long long s = (long long)(h0 + sl) * M + i - crop_left;
                if (vec_out && s >= 0 && s + 3 < T_out) {
                    *reinterpret_cast<float4*>(orow + s) = o;
                } else {
                    if (s >= 0 && s < T_out) orow[s] = o.x;
                    if (s + 1 >= 0 && s + 1 < T_out) orow[s + 1] = o.y;
                    if (s + 2 >= 0 && s + 2 < T_out) orow[s + 2] = o.z;
                    if (s + 3 >= 0 && s + 3 < T_out) orow[s + 3] = o.w;
                }
            }
        }
        __syncthreads();
        b ^= 1;
    }
}

}  // namespace fast
}  // namespace p2p
