// Fast path of the MDCT2 / IMDCT2 kernels for the canonical geometry win == n_fft, hop == n_fft/2
// (models/mdct.py:352-454 as the shipped Pix2PixHDModel wires them), n_fft = N in {64 .. 1024}.
//
// Two consecutive frames A, B ride on ONE N-point complex FFT (A on the real lane, B on the imaginary lane); the
// N = V*P points of a pair are held in registers, V per lane on P lanes (index logic modelled in numpy and checked
// against the DCT-II definition before it was written down here):
//   radix-V FFT in registers -> twiddle -> exchange through a pair-private shared-memory region, P rows at a time
//   -> radix-P FFT(s) in registers.  For V = 2P a lane takes the rows k1 = j and V - j (lane 0: 0 and P), so that
//   the bins k and N - k the real/imaginary split needs sit in the SAME lane; for V = P one shuffle fetches them.
// A warp works on 32/P pairs; a persistent CTA walks (row, tile) work items and the next tile's signal span /
// coefficient rows arrive by TMA bulk copies (cp.async.bulk + mbarrier) while the current one is transformed.
// Lanes 16-31 hold their points rotated by V/2 in analysis (they read the odd-position samples while lanes 0-15
// read the even ones: no bank conflict on the stride-2 reads); FFT_V of the rotated sequence is (-1)^k1 times the
// plain one, which is folded into the lane's exchange twiddle.
#pragma once
#include "mdct4_fast.cuh"      // mbarrier / bulk-copy helpers, brev, kWarps, reg_fft

namespace p2p {
namespace fast2 {

using fast::kWarps;
using fast::kThreads;
using fast::brev;
using fast::mbar_init;
using fast::mbar_fence_init;
using fast::mbar_expect_tx;
using fast::mbar_wait;
using fast::bulk_g2s;

template <int V_, int P_, int NBUF_S_>
struct Cfg2 {
    static constexpr int V = V_, P = P_;
    static constexpr int N = V * P, M = N / 2;
    static constexpr int Q = V / P;                       // exchange passes = rows per lane in stage 2
    static constexpr int G = 32 / P;                      // frame pairs per warp
    static constexpr int PAIRS = kWarps * G;
    static constexpr int FPT = 2 * PAIRS;                 // frames per tile
    static constexpr int XW = 2 * P * (P + 1);            // words of one pair's exchange region (P rows of P+1 float2)
    static constexpr bool XTW_IN_REGS = (V <= 16);
    static constexpr int XV = XTW_IN_REGS ? V : 1;
    // V = 32: the exchange twiddles w_N^(j k1) come from a [k1][j] table in shared memory (conflict-free, immediate
    // offsets) - a gather from roots[] touches up to 16 cache lines per load.  The (-1)^k1 of flipped lanes is part of
    // the table when P = 32 (one lane per column); for P < 32 a whole pair is flipped and the sign moves to the output.
    static constexpr int XT = XTW_IN_REGS ? 0 : 2 * N;    // words
    static constexpr bool SIGN_AT_OUTPUT = !XTW_IN_REGS && P < 32;
    static_assert(Q == 1 || Q == 2, "V must be P or 2P");
    static constexpr int MINB = (N <= 512) ? 2 : 1;       // CTAs per SM the register allocation must allow
    // analysis: sig[2][ASIG] | ws[N + APAD] | exch[PAIRS][XW] | tw[2N] | mbar
    static constexpr int APAD = (P == 8) ? 8 : 0;         // floats per hop row; shifts the pairs of a warp by 16 banks
    static constexpr int AROW = M + APAD;
    static constexpr int ASIG = (FPT + 1) * AROW;
    static constexpr int A_OFF_WS = 2 * ASIG;
    static constexpr int A_OFF_X = A_OFF_WS + N + APAD;
    static constexpr int A_OFF_TW = A_OFF_X + PAIRS * XW;
    static constexpr int A_OFF_XT = A_OFF_TW + 2 * N;
    static constexpr int A_OFF_BAR = A_OFF_XT + XT;
    static constexpr size_t A_SMEM = 4ull * A_OFF_BAR + 16;
    // synthesis: buf[NBUF][SBUF] | exch | tw | mbar
    static constexpr int NBUF = NBUF_S_;
    static constexpr int SPAD = (P < 32) ? P / 2 : 0;     // floats per coefficient row (same purpose)
    static constexpr int SROW = N + SPAD;
    static constexpr int SBUF = FPT * SROW + 4;           // + slack: row[N - 0] of the last row is read (and discarded)
    static constexpr int S_OFF_X = NBUF * SBUF;
    static constexpr int S_OFF_TW = S_OFF_X + PAIRS * XW;
    static constexpr int S_OFF_XT = S_OFF_TW + 2 * N;
    static constexpr int S_OFF_BAR = S_OFF_XT + XT;
    static constexpr size_t S_SMEM = 4ull * S_OFF_BAR + 16;
    static constexpr int G4 = M / 4;                      // float4 groups per hop slot (overlap-add pass)
    static constexpr int NQ = G4 > kThreads ? G4 / kThreads : 1;
    static constexpr int SS = G4 < kThreads ? kThreads / G4 : 1;
};

__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// index into roots[] (e^{-2 pi i m / N}) of lane j's exchange twiddle for first-stage bin k1
template <class C>
__device__ __forceinline__ int xtw_index(int j, int k1, bool flip) {
    return (j * k1 + ((flip && (k1 & 1)) ? C::N / 2 : 0)) & (C::N - 1);
}

// z: V points of column j (slot i holds n1 = i, or (i + V/2) % V on a flipped lane)  ->  X[h][k2] = FFT_N bin K1(h) + V k2,
// K1(0) = j, K1(1) = (j == 0 ? P : V - j).
template <class C>
__device__ __forceinline__ void fft_core2(float2 (&z)[C::V], float2* __restrict__ ex, const float2* __restrict__ xt_lane,
                                          const float2 (&xr)[C::XV], int j, float2 (&X)[C::Q][C::P]) {
    constexpr int V = C::V, P = C::P;
    fft_reg<V>(z);
#pragma unroll
    for (int h = 0; h < C::Q; ++h) {
#pragma unroll
        for (int r = 0; r < P; ++r) {
            const int k1 = h * P + r;
            float2 v = z[brev<V>(k1)];
            if (k1 > 0) {
                if constexpr (C::XTW_IN_REGS) v = cmulf(v, xr[k1]);
                else v = cmulf(v, xt_lane[k1 * P]);
            }
            ex[r * (P + 1) + j] = v;
        }
        __syncwarp();
        const float2* row = ex + (h == 0 ? j : ((P - j) & (P - 1))) * (P + 1);
        float2 b[P];
#pragma unroll
        for (int n2 = 0; n2 < P; ++n2) b[n2] = row[n2];
        __syncwarp();                  // the region is rewritten by the next pass / the next tile
        fft_reg<P>(b);
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) X[h][k2] = b[brev<P>(k2)];
    }
}

// fills the shared-memory table (V = 32); `signed_cols`: columns j >= 16 carry the (-1)^k1 of flipped lanes (analysis, P = 32)
template <class C>
__device__ __forceinline__ void fill_xt(float2* xt, const float2* roots, bool signed_cols) {
    if constexpr (!C::XTW_IN_REGS) {
        for (int i = threadIdx.x; i < C::N; i += kThreads) {
            const int k1 = i / C::P, j = i % C::P;
            xt[i] = __ldg(roots + xtw_index<C>(j, k1, signed_cols && C::P == 32 && j >= 16));
        }
    }
}

template <class C>
__device__ __forceinline__ void load_xtw(float2 (&xr)[C::XV], const float2* roots, int j, bool flip) {
    if constexpr (C::XTW_IN_REGS) {
#pragma unroll
        for (int k1 = 0; k1 < C::V; ++k1) xr[k1] = __ldg(roots + xtw_index<C>(j, k1, flip));
    } else {
        xr[0] = make_float2(1.f, 0.f);
    }
}

// ---- analysis -------------------------------------------------------------------------------------------
struct Tile2 { int row; int f0; int nfr; };

template <class C>
__device__ __forceinline__ Tile2 tile2(long long w, int tiles_per_row, int F) {
    Tile2 t;
    t.row = (int)(w / tiles_per_row);
    t.f0 = (int)(w - (long long)t.row * tiles_per_row) * C::FPT;
    t.nfr = min(C::FPT, F - t.f0);
    return t;
}

// The (nfr+1) hop rows of a tile: in-range, 16-byte aligned rows by TMA bulk copy (ONE copy when rows are not padded),
// the rest (zero padding outside [0,T), unaligned input) by ordinary loads from all threads.
template <class C>
__device__ __forceinline__ void fetch_span(float* sig, uint64_t* bar, const float* __restrict__ x, long long T,
                                           long long offset, const Tile2 tc, bool aligned) {
    const float* xb = x + (long long)tc.row * T;
    const long long t_first = (long long)tc.f0 * C::M - offset;
    const int rows = tc.nfr + 1;
    int r_lo = 0, r_hi = 0;
    if (aligned) {
        r_lo = t_first >= 0 ? 0 : (int)min((long long)rows, (-t_first + C::M - 1) / C::M);
        const long long fit = (T - t_first) / C::M;
        r_hi = (int)max((long long)r_lo, min((long long)rows, fit));
    }
    if (threadIdx.x == 0) {
        fence_proxy_async_smem();
        mbar_expect_tx(bar, (uint32_t)(r_hi - r_lo) * C::M * 4);
        if constexpr (C::APAD == 0) {
            if (r_hi > r_lo)
                bulk_g2s(sig + r_lo * C::AROW, xb + t_first + (long long)r_lo * C::M, (uint32_t)(r_hi - r_lo) * C::M * 4, bar);
        } else {
#pragma unroll 1
            for (int r = r_lo; r < r_hi; ++r)
                bulk_g2s(sig + r * C::AROW, xb + t_first + (long long)r * C::M, C::M * 4, bar);
        }
    }
#pragma unroll 1
    for (int r = 0; r < rows; ++r) {
        if (r >= r_lo && r < r_hi) { r = r_hi - 1; continue; }
        const long long t = t_first + (long long)r * C::M;
        for (int i = threadIdx.x; i < C::M; i += kThreads) {
            const long long tt = t + i;
            sig[r * C::AROW + i] = (tt >= 0 && tt < T) ? __ldg(xb + tt) : 0.f;
        }
    }
}

// out[k] of both frames of a pair from Zk = FFT bin k, Zm = FFT bin N-k, t = e^{-j pi k / (2N)}:
//   A: Re((Zk + conj Zm) t),  B: Re(-j (Zk - conj Zm) t)          (scale / N lives in the window)
__device__ __forceinline__ void split_pair(float2 Zk, float2 Zm, float2 t, float& a, float& b) {
    a = (Zk.x + Zm.x) * t.x - (Zk.y - Zm.y) * t.y;
    b = (Zk.y + Zm.y) * t.x - (Zm.x - Zk.x) * t.y;
}

template <class C>
__device__ __forceinline__ void analysis2_warp(const float* __restrict__ sig, const float* __restrict__ ws,
                                               float* __restrict__ ex_warp, const float2* __restrict__ tw,
                                               const float2* __restrict__ xt, const float2 (&xr)[C::XV],
                                               float* __restrict__ spec_tile, int nfr, int warp, int lane, float c0) {
    constexpr int V = C::V, P = C::P, N = C::N, H = V / 2;
    const int j = lane % P, g = lane / P;
    const bool flip = (lane >> 4) & 1;
    const int pr = warp * C::G + g;                        // pair of the tile: frames 2 pr, 2 pr + 1
    const bool vA = 2 * pr < nfr, vB = 2 * pr + 1 < nfr;
    const float* sA = sig + (2 * pr) * C::AROW;
    // v[n] = u[2n] (n < N/2), u[2(N-1-n)+1] otherwise, n = P n1 + j; see the header for the rotation of flipped lanes
    const int e = 2 * j;
    const int a_lo = flip ? N - 1 - e : e, a_hi = flip ? e : N - 1 - e;
    const int st = flip ? -2 * P : 2 * P;
    float2 z[V];
#pragma unroll
    for (int i = 0; i < V; ++i) {
        int idx = i < H ? a_lo + st * i : a_hi - st * (i - H);
        if constexpr (C::APAD != 0) {                      // samples of the second hop row sit APAD floats further
            const bool up = (i < H) ? ((i >= V / 4) != flip) : (((i - H) < V / 4) != flip);
            idx += up ? C::APAD : 0;
        }
        const float w = ws[idx];
        const float a = sA[idx], b = sA[idx + C::AROW];
        z[i] = make_float2(vA ? w * a : 0.f, vB ? w * b : 0.f);
    }
    float2 X[C::Q][P];
    fft_core2<C>(z, reinterpret_cast<float2*>(ex_warp + g * C::XW), xt + j, xr, j, X);

    float* rowA = spec_tile + (long long)(2 * pr) * N;
    if constexpr (C::Q == 2) {
        const int k1b = (j == 0) ? P : V - j;
        // all bins of a flipped pair's odd columns come out negated (rows j and V - j have the parity of j)
        const float sg = (C::SIGN_AT_OUTPUT && flip && (j & 1)) ? -1.f : 1.f;
        float* pA = rowA + j;                              // bins j + V k2
        float* qA = rowA + k1b;                            // bins k1b + V k2
        const float2* tp = tw + j;
        const float2* tq = tw + k1b;
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float2 Zp = X[0][k2], Zs = X[1][P - 1 - k2];
            const float2 mp = (j == 0) ? X[0][(P - k2) % P] : Zs;
            const float2 ms = (j == 0) ? X[1][k2] : Zp;
            float2 t1 = tp[V * k2], t2 = tq[V * (P - 1 - k2)];
            if constexpr (C::SIGN_AT_OUTPUT) { t1.x *= sg; t1.y *= sg; t2.x *= sg; t2.y *= sg; }
            float ap, bp, as, bs;
            split_pair(Zp, mp, t1, ap, bp);
            split_pair(Zs, ms, t2, as, bs);
            if (k2 == 0 && j == 0) { ap *= c0; bp *= c0; }
            if (vA) { pA[V * k2] = ap; qA[V * (P - 1 - k2)] = as; }
            if (vB) { pA[N + V * k2] = bp; qA[N + V * (P - 1 - k2)] = bs; }
        }
    } else {
        float* rowB = rowA + N;
        const int src = (lane & ~(P - 1)) | ((P - j) & (P - 1));
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float2 pub = (j == 0) ? X[0][(P - k2) % P] : X[0][P - 1 - k2];
            float2 m;
            m.x = __shfl_sync(0xffffffffu, pub.x, src);
            m.y = __shfl_sync(0xffffffffu, pub.y, src);
            const int k = j + V * k2;
            float a, b;
            split_pair(X[0][k2], m, tw[k], a, b);
            if (k2 == 0 && j == 0) { a *= c0; b *= c0; }
            if (vA) rowA[k] = a;
            if (vB) rowB[k] = b;
        }
    }
}

template <class C>
__global__ void __launch_bounds__(kThreads, C::MINB)
mdct2_analysis_fast(const float* __restrict__ x, float* __restrict__ spec, const float* __restrict__ window,
                    const void* __restrict__ tables, long long T, int F, long long offset, float scale, float c0,
                    int tiles_per_row, long long total_tiles, int aligned) {
    extern __shared__ __align__(128) float smem[];
    constexpr int N = C::N, M = C::M;
    float* sig0 = smem;
    float* ws = smem + C::A_OFF_WS;
    float* exch = smem + C::A_OFF_X;
    float2* tw = reinterpret_cast<float2*>(smem + C::A_OFF_TW);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::A_OFF_BAR);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float2* gtw = reinterpret_cast<const float2*>(tables);
    const float2* roots = gtw + N;

    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
    const float sc = scale / (float)N;
    for (int p = threadIdx.x; p < N; p += kThreads) {
        ws[p + (p >= M ? C::APAD : 0)] = __ldg(window + p) * sc;
        tw[p] = __ldg(gtw + p);
    }
    float2 xr[C::XV];
    load_xtw<C>(xr, roots, lane % C::P, (lane >> 4) & 1);
    float2* xt = reinterpret_cast<float2*>(smem + C::A_OFF_XT);
    fill_xt<C>(xt, roots, true);
    __syncthreads();

    long long w = blockIdx.x;
    if (w < total_tiles) fetch_span<C>(sig0, &bars[0], x, T, offset, tile2<C>(w, tiles_per_row, F), aligned);
    __syncthreads();
    uint32_t phase[2] = {0, 0};
    int buf = 0;
    for (; w < total_tiles; w += gridDim.x) {
        const Tile2 tc = tile2<C>(w, tiles_per_row, F);
        const long long wn = w + gridDim.x;
        if (wn < total_tiles)
            fetch_span<C>(sig0 + (buf ^ 1) * C::ASIG, &bars[buf ^ 1], x, T, offset, tile2<C>(wn, tiles_per_row, F), aligned);
        mbar_wait(&bars[buf], phase[buf]);
        phase[buf] ^= 1;
        if (warp * C::G * 2 < tc.nfr)
            analysis2_warp<C>(sig0 + buf * C::ASIG, ws, exch + warp * C::G * C::XW, tw, xt, xr,
                              spec + ((long long)tc.row * F + tc.f0) * N, tc.nfr, warp, lane, c0);
        __syncthreads();
        buf ^= 1;
    }
}

// ---- synthesis ------------------------------------------------------------------------------------------
// A tile = FPT consecutive coefficient rows (frames fA .. fA+FPT-1) of one batch row; it produces the FPT-1 output
// hop slots between them (slot h = first half of frame h + second half of frame h-1), so consecutive tiles overlap
// by one frame and every output sample is written exactly once (no atomics).
template <class C>
__device__ __forceinline__ void fetch_rows2(float* buf, uint64_t* bar, const float* __restrict__ spec_row, int fA,
                                            int F, bool aligned) {
    const int r_lo = max(0, -fA), r_hi = max(r_lo, min(C::FPT, F - fA));     // rows holding real frames
    if (aligned) {
        if (threadIdx.x == 0) {
            fence_proxy_async_smem();
            mbar_expect_tx(bar, (uint32_t)(r_hi - r_lo) * C::N * 4);
#pragma unroll 1
            for (int r = r_lo; r < r_hi; ++r)
                bulk_g2s(buf + r * C::SROW, spec_row + (long long)(fA + r) * C::N, C::N * 4, bar);
        }
    } else {
        if (threadIdx.x == 0) mbar_expect_tx(bar, 0);
#pragma unroll 1
        for (int r = r_lo; r < r_hi; ++r)
            for (int i = threadIdx.x; i < C::N; i += kThreads)
                buf[r * C::SROW + i] = __ldg(spec_row + (long long)(fA + r) * C::N + i);
    }
    // rows of frames outside [0, F) (first / last tile of a batch row) are zero coefficients
#pragma unroll 1
    for (int r = 0; r < C::FPT; ++r) {
        if (r >= r_lo && r < r_hi) { r = r_hi - 1; continue; }
        for (int i = threadIdx.x; i < C::N; i += kThreads) buf[r * C::SROW + i] = 0.f;
    }
}

// One warp: the coefficient rows of its G pairs -> y = 2 sum_k c_k X[k] cos(pi (2n+1) k / (2N)) written back IN PLACE.
// h[k] = tw[k] (X[k] + j X[N-k]) (h[0] = 2 c0 X[0]) is Hermitian-symmetric in the sense that FFT_N(h) is real, so
// FFT_N(hA + j hB) carries frame A on the real and frame B on the imaginary lane:  v[n] -> y[2n] (n < N/2), y[2(N-1-n)+1].
template <class C>
__device__ __forceinline__ void synthesis2_warp(float* __restrict__ buf, float* __restrict__ ex_warp,
                                                const float2* __restrict__ tw, const float2* __restrict__ xt,
                                                const float2 (&xr)[C::XV], int fA, int F, int warp, int lane, float c0x2) {
    constexpr int V = C::V, P = C::P, N = C::N;
    const int j = lane % P, g = lane / P;
    const int pr = warp * C::G + g;
    float* rA = buf + (2 * pr) * C::SROW;
    const float* rB = rA + C::SROW;
    const int fa = fA + 2 * pr;
    const bool vA = fa >= 0 && fa < F, vB = fa + 1 >= 0 && fa + 1 < F;

    float2 z[V];
#pragma unroll
    for (int i = 0; i < V; ++i) {
        const int k = P * i + j;
        const float ak = rA[k], am = rA[N - k], bk = rB[k], bm = rB[N - k];
        const float2 t = tw[k];
        float2 hA = make_float2(t.x * ak - t.y * am, t.x * am + t.y * ak);
        float2 hB = make_float2(t.x * bk - t.y * bm, t.x * bm + t.y * bk);
        if (i == 0 && j == 0) { hA = make_float2(c0x2 * ak, 0.f); hB = make_float2(c0x2 * bk, 0.f); }
        z[i] = make_float2(hA.x - hB.y, hA.y + hB.x);      // rows of frames outside [0, F) hold zeros (fetch_rows2)
    }
    float2 X[C::Q][P];
    fft_core2<C>(z, reinterpret_cast<float2*>(ex_warp + g * C::XW), xt + j, xr, j, X);

    const bool all_valid = __all_sync(0xffffffffu, vA && vB);
#pragma unroll
    for (int h = 0; h < C::Q; ++h) {
        const int k1 = h == 0 ? j : (j == 0 ? P : V - j);
        float* lo = rA + 2 * k1;                           // y[2n], n = k1 + V k2 < N/2
        float* hi = rA + 2 * N - 1 - 2 * k1;               // y[2(N-1-n)+1] otherwise
        if (all_valid) {
#pragma unroll
            for (int k2 = 0; k2 < P; ++k2) {
                float* d = k2 < P / 2 ? lo + 2 * V * k2 : hi - 2 * V * k2;
                d[0] = X[h][k2].x;
                d[C::SROW] = X[h][k2].y;
            }
        } else {                                           // a zero row must stay exactly zero: its partner's rounding noise
#pragma unroll
            for (int k2 = 0; k2 < P; ++k2) {
                float* d = k2 < P / 2 ? lo + 2 * V * k2 : hi - 2 * V * k2;
                d[0] = vA ? X[h][k2].x : 0.f;
                d[C::SROW] = vB ? X[h][k2].y : 0.f;
            }
        }
    }
}

template <class C>
__global__ void __launch_bounds__(kThreads, C::MINB)
mdct2_synthesis_fast(const float* __restrict__ spec, float* __restrict__ out, const float* __restrict__ window,
                     const void* __restrict__ tables, int F, long long T_out, long long crop_left, float scale, float c0,
                     int h_first, int h_last, int tiles_per_row, long long total_tiles, int aligned_in, int vec_out) {
    extern __shared__ __align__(128) float smem[];
    constexpr int N = C::N, M = C::M, NB = C::NBUF;
    float* buf0 = smem;
    float* exch = smem + C::S_OFF_X;
    float2* tw = reinterpret_cast<float2*>(smem + C::S_OFF_TW);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::S_OFF_BAR);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float2* gtw = reinterpret_cast<const float2*>(tables);
    const float2* roots = gtw + N;

    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
    for (int p = threadIdx.x; p < N; p += kThreads) tw[p] = __ldg(gtw + p);
    float2 xr[C::XV];
    load_xtw<C>(xr, roots, lane % C::P, false);
    float2* xt = reinterpret_cast<float2*>(smem + C::S_OFF_XT);
    fill_xt<C>(xt, roots, false);
    // overlap-add role of this thread: float4 group(s) q of every SS-th hop slot; window * scale stays in registers
    const int q0 = threadIdx.x % C::G4;
    const int sl0 = (C::G4 < kThreads) ? threadIdx.x / C::G4 : 0;
    float w1[C::NQ][4], w2[C::NQ][4];
#pragma unroll
    for (int qi = 0; qi < C::NQ; ++qi) {
        const int i = 4 * (q0 + kThreads * qi);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            w1[qi][c] = __ldg(window + i + c) * scale;
            w2[qi][c] = __ldg(window + M + i + c) * scale;
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
        fetch_rows2<C>(buf0, &bars[0], spec + (long long)row * F * N, h0 - 1, F, aligned_in);
    }
    __syncthreads();
    uint32_t phase[2] = {0, 0};
    int b = 0;
    for (; w < total_tiles; w += gridDim.x) {
        coord(w, row, h0, ns);
        const long long wn = w + gridDim.x;
        int rown = 0, h0n = 0, nsn = 0;
        if (wn < total_tiles) coord(wn, rown, h0n, nsn);
        if (NB == 2 && wn < total_tiles)
            fetch_rows2<C>(buf0 + (b ^ 1) * C::SBUF, &bars[b ^ 1], spec + (long long)rown * F * N, h0n - 1, F, aligned_in);
        mbar_wait(&bars[b], phase[b]);
        phase[b] ^= 1;
        float* buf = buf0 + b * C::SBUF;
        if (warp * C::G * 2 <= ns)       // frames fA .. fA+ns are needed (ns+1 rows)
            synthesis2_warp<C>(buf, exch + warp * C::G * C::XW, tw, xt, xr, h0 - 1, F, warp, lane, 2.f * c0);
        __syncthreads();

        float* orow = out + (long long)row * T_out;
        for (int sl = sl0; sl < ns; sl += C::SS) {
            const float* tp = buf + sl * C::SROW;         // frame h-1
            const float* th = tp + C::SROW;               // frame h
#pragma unroll
            for (int qi = 0; qi < C::NQ; ++qi) {
                const int i = 4 * (q0 + kThreads * qi);
                const float4 A = *reinterpret_cast<const float4*>(th + i);
                const float4 Bv = *reinterpret_cast<const float4*>(tp + M + i);
                float4 o;
                o.x = w1[qi][0] * A.x + w2[qi][0] * Bv.x;
                o.y = w1[qi][1] * A.y + w2[qi][1] * Bv.y;
                o.z = w1[qi][2] * A.z + w2[qi][2] * Bv.z;
                o.w = w1[qi][3] * A.w + w2[qi][3] * Bv.w;
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
        if (NB == 1) {
            if (wn < total_tiles)
                fetch_rows2<C>(buf0, &bars[0], spec + (long long)rown * F * N, h0n - 1, F, aligned_in);
            __syncthreads();      // the unaligned path stages rows by ordinary stores
        } else {
            b ^= 1;
        }
    }
}

}  // namespace fast2
}  // namespace p2p
