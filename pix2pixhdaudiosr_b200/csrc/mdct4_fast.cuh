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
    // (V,P) = (16,8) (n_fft 512): with the post table in shared memory the analysis kernel needs 115 920 B per CTA, 208 B
    // more than two CTAs per SM allow; its 1 KB post table is read through L1 instead (16 instead of 8 warps per SM)
    static constexpr bool POST_IN_SMEM = TABLES_IN_SMEM && !(V == 16 && P == 8);
    static constexpr int TBF = TABLES_IN_SMEM ? (POST_IN_SMEM ? 6 * L : 4 * L) : 0;   // floats of tables in shared memory
    static constexpr bool XTW_IN_REGS = (V <= 16);       // exchange twiddles live in registers
    static constexpr int XV = XTW_IN_REGS ? V : 1;
    static constexpr int WS = N + ROWPAD;            // signed window, same padded layout as a frame
    static_assert(V >= P && V % P == 0, "V must be a multiple of P");
    // shared memory layout (floats): sig[2][SIG] | ws[WS] | exch[kWarps*G][XREG] | tables[6L]? (pre, pre_f, post) | mbar
    static constexpr int OFF_WS = 2 * SIG;
    static constexpr int OFF_X = OFF_WS + ((WS + 3) & ~3);
    static constexpr int OFF_TB = OFF_X + kWarps * G * XREG;
    static constexpr int OFF_BAR = OFF_TB + TBF;
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
    // bulk (TMA) rows are [r_lo, r_hi): in range and aligned; the rest is staged by ordinary loads
    int r_lo = 0, r_hi = 0;
    if (aligned) {
        r_lo = t_first >= 0 ? 0 : (int)min((long long)rows, (-t_first + C::M - 1) / C::M);
        const long long fit = (T - t_first) / C::M;             // rows r < fit end inside [0,T)
        r_hi = (int)max((long long)r_lo, min((long long)rows, fit));
    }
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, (uint32_t)(r_hi - r_lo) * C::M * 4);
#pragma unroll 1
        for (int r = r_lo; r < r_hi; ++r)
            bulk_g2s(sig + r * C::ROW, xb + t_first + (long long)r * C::M, C::M * 4, bar);
    }
#pragma unroll 1
    for (int r = 0; r < rows; ++r) {
        if (r >= r_lo && r < r_hi) { r = r_hi - 1; continue; }
        const long long t = t_first + (long long)r * C::M;
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

// ---- lane roles -------------------------------------------------------------------------------------
// Lane l: j = l % P (column of the L = V*P point FFT), g = l / P (frame of the warp), flip = (l>>4)&1.
// Fold/de-interleave reads are stride-2 floats, so lanes 0-15 would collide with lanes 16-31 on the same
// banks.  "Flipped" lanes therefore read the ODD-position operand of every pair first, and - so that all
// shared-memory offsets stay compile-time immediates - hold their V points in REVERSED order:
//   register n1 of a flipped lane holds z[P*pi(n1) + j],  pi(n1) = (c - n1) mod V
// (c = V/2-1 for analysis, V-1 for synthesis).  FFT_V of the reversed sequence is w_V^{c k} Z[-k mod V]; the
// phase is folded into the lane's exchange twiddle and the row reversal into the stage-2 row pointer.
template <class C>
struct LaneRole {
    int j, g;
    bool flip;
    __device__ __forceinline__ LaneRole(int lane) : j(lane % C::P), g(lane / C::P), flip((lane >> 4) & 1) {}
    // index into roots[] (e^{-2 pi i m/L}) of this lane's exchange twiddle for first-stage output slot k1
    __device__ __forceinline__ int xtw_index(int k1, int c) const {
        if (!flip) return j * k1;
        const int kappa = (C::V - k1) % C::V;
        return (j * kappa - c * k1 * C::P) & (C::L - 1);
    }
};

template <class C>
__device__ __forceinline__ int pi_rev(int n1, int c) { return (c - n1 + 2 * C::V) % C::V; }

// stages 1b..2 shared by analysis and synthesis: z (V registers, lane order as above) -> Y[k1 + V k2] * post
// returns re_[m][k2] = Re, nim[m][k2] = -Im for k = j + P m + V k2.
// SWZ: the exchange region is V rows of exactly P float2 (no padding column), element (row r, column c) at
// r*P + (c ^ (r % P)) -- conflict-free like the padded form, and 2*V*P floats = one frame row, so the synthesis kernel can
// lay it over the frame's own (already consumed) coefficient row; `xstride` = floats between the regions of a warp's frames.
template <class C, bool SWZ = false>
__device__ __forceinline__ void fft_core(float2 (&z)[C::V], float* __restrict__ ex_warp, const float2* post,
                                         const float2* roots, const float2 (&xr)[C::XV], const LaneRole<C>& ln, int c,
                                         float (&re_)[C::Q][C::P], float (&nim)[C::Q][C::P], int xstride = C::XREG) {
    constexpr int V = C::V, P = C::P, PS = SWZ ? P : P + 1;
    fft_reg<V>(z);
    float2* ex = reinterpret_cast<float2*>(ex_warp + ln.g * xstride);
#pragma unroll
    for (int k1 = 0; k1 < V; ++k1) {
        float2 v = z[brev<V>(k1)];
        if (k1 > 0) {
            if constexpr (C::XTW_IN_REGS) v = cmulf(v, xr[k1]);
            else v = cmulf(v, __ldg(roots + ln.xtw_index(k1, c)));
        }
        ex[k1 * PS + (SWZ ? (ln.j ^ (k1 % P)) : ln.j)] = v;          // flipped lanes: physical row k1 holds logical row (V-k1)%V
    }
    __syncwarp();
#pragma unroll
    for (int m = 0; m < C::Q; ++m) {
        float2 b[P];
        const int k1 = ln.j + P * m;
        const int rN = k1, rF = (V - k1) % V;
        const float2* rowN = ex + rN * PS;
        const float2* rowF = ex + rF * PS;
        const int rL = ln.flip ? rF : rN;
        const float2* rowL = ln.flip ? rowF : rowN;
#pragma unroll
        for (int n2 = 0; n2 < P; ++n2) {
            if constexpr (P == 32) b[n2] = (n2 >= 16 ? rowF : rowN)[SWZ ? (n2 ^ ((n2 >= 16 ? rF : rN) % P)) : n2];   // columns 16.. were written by flipped lanes
            else b[n2] = rowL[SWZ ? (n2 ^ (rL % P)) : n2];
        }
        fft_reg<P>(b);
        const float2* pk = post + k1;
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float2 y = cmulf(b[brev<P>(k2)], ldt<C::POST_IN_SMEM>(pk + V * k2));
            re_[m][k2] = y.x;
            nim[m][k2] = -y.y;
        }
    }
    __syncwarp();   // exchange regions are reused by the next tile
}

// One warp transforms its G frames of the current tile: fold + signed window + pre-twiddle ->
// radix-V -> exchange -> radix-P -> post-twiddle -> pair up (X[2k], X[2k+1]) by one shuffle -> STG.64.
//   sig : tile signal buffer (rows of C::ROW floats), ws: signed, scaled, padded window
//   pre : this lane's pre-twiddle row base (pre + j, or the reversed-order copy for flipped lanes)
template <class C>
__device__ __forceinline__ void analysis_warp(const float* __restrict__ sig, const float* __restrict__ ws,
                                              float* __restrict__ ex_warp, const float2* pre_lane, const float2* post,
                                              const float2* roots, const float2 (&xr)[C::XV],
                                              float* __restrict__ spec_tile, int nfr, int warp, const LaneRole<C>& ln) {
    constexpr int V = C::V, P = C::P, L = C::L, PAD = C::ROWPAD, H = V / 2, CREV = V / 2 - 1;
    const int fl = warp * C::G + ln.g;
    const float* sf = sig + fl * C::ROW;
    const int e2j = 2 * ln.j;
    // pair (Ce, Co): even-position operand at Ce + 2j + 2P i, odd-position operand at Co - 2j - 2P i.
    // first load = base F + 2P i, second load = base S - 2P i for BOTH lane kinds (see LaneRole).
    auto F = [&](int ce, int co) { return ln.flip ? co - e2j - 2 * P * (H - 1) : ce + e2j; };
    auto S = [&](int ce, int co) { return ln.flip ? ce + e2j + 2 * P * (H - 1) : co - e2j; };
    const int fo[4] = {F(3 * L + PAD, 3 * L - 1 + PAD), F(L, L - 1), F(0, 2 * L - 1), F(2 * L + PAD, 4 * L - 1 + PAD)};
    const int so[4] = {S(3 * L + PAD, 3 * L - 1 + PAD), S(L, L - 1), S(0, 2 * L - 1), S(2 * L + PAD, 4 * L - 1 + PAD)};
    const float *sF[4], *sS[4], *wF[4], *wS[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) { sF[t] = sf + fo[t]; sS[t] = sf + so[t]; wF[t] = ws + fo[t]; wS[t] = ws + so[t]; }

    float2 z[V];
#pragma unroll
    for (int n1 = 0; n1 < V; ++n1) {
        constexpr int dummy = 0; (void)dummy;
        const int tA = n1 < H ? 0 : 2;                 // pair types: A.re, A.im, B.re, B.im
        const int i = n1 < H ? n1 : n1 - H;
        const float re = wF[tA][2 * P * i] * sF[tA][2 * P * i] + wS[tA][-2 * P * i] * sS[tA][-2 * P * i];
        const float im = wF[tA + 1][2 * P * i] * sF[tA + 1][2 * P * i] + wS[tA + 1][-2 * P * i] * sS[tA + 1][-2 * P * i];
        float2 tw;
        if constexpr (C::TABLES_IN_SMEM) tw = pre_lane[P * n1];
        else tw = __ldg(pre_lane + P * (ln.flip ? pi_rev<C>(n1, CREV) : n1));
        z[n1] = cmulf(make_float2(re, im), tw);
    }
    float re_[C::Q][P], nim[C::Q][P];
    fft_core<C>(z, ex_warp, post, roots, xr, ln, CREV, re_, nim);

    float* sp = spec_tile + (long long)fl * C::M + 2 * ln.j;
    const bool valid = fl < nfr;
#pragma unroll
    for (int m = 0; m < C::Q; ++m) {
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float other = __shfl_xor_sync(0xffffffffu, nim[C::Q - 1 - m][P - 1 - k2], P - 1);
            if (valid) *reinterpret_cast<float2*>(sp + 2 * (P * m + V * k2)) = make_float2(re_[m][k2], other);
        }
    }
}

// Prologue shared by both kernels: pre / pre_f / post tables into shared memory (when they fit), the lane's
// pre-twiddle base pointer and its exchange twiddles (registers when V <= 16).
template <class C>
__device__ __forceinline__ void setup_tables(float* tb, const float* gt, int c, const LaneRole<C>& ln,
                                             const float2*& pre_lane, const float2*& post, const float2*& roots,
                                             float2 (&xr)[C::XV]) {
    constexpr int L = C::L;
    const float2* gpre = reinterpret_cast<const float2*>(gt);
    post = reinterpret_cast<const float2*>(gt + 2 * L);
    roots = reinterpret_cast<const float2*>(gt + 4 * L);
    pre_lane = gpre + ln.j;
    if constexpr (C::TABLES_IN_SMEM) {
        float2* t2 = reinterpret_cast<float2*>(tb);
        for (int i = threadIdx.x; i < L; i += kThreads) {
            const int n1 = i / C::P, jj = i % C::P;
            t2[i] = __ldg(gpre + i);                                          // pre
            t2[L + i] = __ldg(gpre + C::P * pi_rev<C>(n1, c) + jj);            // pre_f: reversed register order
            if constexpr (C::POST_IN_SMEM) t2[2 * L + i] = __ldg(post + i);   // post
        }
        pre_lane = t2 + (ln.flip ? L : 0) + ln.j;
        if constexpr (C::POST_IN_SMEM) post = t2 + 2 * L;
    }
    if constexpr (C::XTW_IN_REGS) {
#pragma unroll
        for (int k1 = 0; k1 < C::V; ++k1) xr[k1] = __ldg(roots + ln.xtw_index(k1, c));
    } else {
        xr[0] = make_float2(1.f, 0.f);
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
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const LaneRole<C> ln(lane);

    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
    for (int p = threadIdx.x; p < N; p += kThreads) {
        const float w = __ldg(window + p) * scale;
        ws[p + (p >= 2 * L ? C::ROWPAD : 0)] = p < L ? w : -w;
    }
    const float2 *pre_lane, *post, *roots;
    float2 xr[C::XV];
    setup_tables<C>(smem + C::OFF_TB, reinterpret_cast<const float*>(tables), C::V / 2 - 1, ln, pre_lane, post, roots, xr);
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
            analysis_warp<C>(sig0 + buf * C::SIG, ws, exch + warp * C::G * C::XREG, pre_lane, post, roots, xr,
                             spec + ((long long)tc.row * F + tc.f0) * M, tc.nfr, warp, ln);
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
    // (16,16), n_fft 1024: the exchange region of a frame is laid over the frame's own coefficient row (swizzled form,
    // 2*V*P floats = M = the row), which brings the CTA from 106 KB to 72 KB of shared memory: three CTAs per SM
    static constexpr bool ALIAS = (C::V == 16 && C::P == 16 && C::ROWPAD == 0);
    static constexpr int MINB = ALIAS ? 3 : 0;                      // 0: no occupancy request (as before)
    static constexpr int OFF_X = 2 * BUF;
    static constexpr int OFF_TB = OFF_X + (ALIAS ? 0 : kWarps * C::G * C::XREG);
    static constexpr int OFF_BAR = OFF_TB + C::TBF;
    static constexpr size_t SMEM_BYTES = 4ull * OFF_BAR + 16;
    static constexpr int G4 = C::M / 4;                             // float4 groups per hop slot
    static constexpr int NQ = G4 > kThreads ? G4 / kThreads : 1;    // window register sets per thread
    static constexpr int SS = G4 < kThreads ? kThreads / G4 : 1;    // slots covered per pass
};

template <class C>
__device__ __forceinline__ void fetch_rows(float* buf, uint64_t* bar, const float* __restrict__ spec_row, int fA,
                                           int F, bool aligned) {
    const int r_lo = max(0, -fA), r_hi = max(r_lo, min(C::FPT, F - fA));     // rows holding real frames
    if (aligned) {
        if (threadIdx.x == 0) {
            mbar_expect_tx(bar, (uint32_t)(r_hi - r_lo) * C::M * 4);
#pragma unroll 1
            for (int r = r_lo; r < r_hi; ++r)
                bulk_g2s(buf + r * C::ROW, spec_row + (long long)(fA + r) * C::M, C::M * 4, bar);
        }
    } else {
        if (threadIdx.x == 0) mbar_expect_tx(bar, 0);
#pragma unroll 1
        for (int r = r_lo; r < r_hi; ++r)
            for (int i = threadIdx.x; i < C::M; i += kThreads)
                buf[r * C::ROW + i] = __ldg(spec_row + (long long)(fA + r) * C::M + i);
    }
}

// One warp: coefficient rows of its G frames -> DCT-IV -> t rows written back IN PLACE (same rows).
template <class C>
__device__ __forceinline__ void synthesis_warp(float* __restrict__ buf, float* __restrict__ ex_warp,
                                               const float2* pre_lane, const float2* post, const float2* roots,
                                               const float2 (&xr)[C::XV], int fA, int F, int warp,
                                               const LaneRole<C>& ln) {
    constexpr int V = C::V, P = C::P, M = C::M, CREV = V - 1;
    const int fl = warp * C::G + ln.g;
    float* row = buf + fl * C::ROW;
    const bool valid = (fA + fl) >= 0 && (fA + fl) < F;
    // z[n] = X[2n] + j X[M-1-2n]: even operand at 2j + 2P n1, odd operand at M-1-2j - 2P n1; flipped lanes
    // read the odd one first and hold their points in reversed order (LaneRole)
    const float* pF = row + (ln.flip ? M - 1 - 2 * ln.j - 2 * P * (V - 1) : 2 * ln.j);
    const float* pS = row + (ln.flip ? 2 * ln.j + 2 * P * (V - 1) : M - 1 - 2 * ln.j);

    float2 z[V];
#pragma unroll
    for (int n1 = 0; n1 < V; ++n1) {
        const float a = pF[2 * P * n1], b = pS[-2 * P * n1];
        float2 tw;
        if constexpr (C::TABLES_IN_SMEM) tw = pre_lane[P * n1];
        else tw = __ldg(pre_lane + P * (ln.flip ? pi_rev<C>(n1, CREV) : n1));
        z[n1] = cmulf(make_float2(ln.flip ? b : a, ln.flip ? a : b), tw);
    }
    float re_[C::Q][P], nim[C::Q][P];
    if constexpr (SynCfg<C>::ALIAS) {
        __syncwarp();                       // every lane of the frame holds its coefficients: the row becomes the exchange region
        fft_core<C, true>(z, buf + warp * C::G * C::ROW, post, roots, xr, ln, CREV, re_, nim, C::ROW);
    } else {
        fft_core<C>(z, ex_warp, post, roots, xr, ln, CREV, re_, nim);
    }

    float* tp = row + 2 * ln.j;
#pragma unroll
    for (int m = 0; m < C::Q; ++m) {
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
            const float other = __shfl_xor_sync(0xffffffffu, nim[C::Q - 1 - m][P - 1 - k2], P - 1);
            re_[m][k2] = valid ? re_[m][k2] : 0.f;
            *reinterpret_cast<float2*>(tp + 2 * (P * m + V * k2)) = make_float2(re_[m][k2], valid ? other : 0.f);
        }
    }
}

template <class C>
__global__ void __launch_bounds__(kThreads, SynCfg<C>::MINB)
mdct4_synthesis_fast(const float* __restrict__ spec, float* __restrict__ out, const float* __restrict__ window,
                     const void* __restrict__ tables, int F, long long T_out, long long crop_left, float scale,
                     int h_first, int h_last, int tiles_per_row, long long total_tiles, int aligned_in, int vec_out) {
    using S = SynCfg<C>;
    extern __shared__ __align__(128) float smem[];
    constexpr int L = C::L, M = C::M;
    float* buf0 = smem;
    float* exch = smem + S::OFF_X;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::OFF_BAR);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const LaneRole<C> ln(lane);

    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_fence_init(); }
    const float2 *pre_lane, *post, *roots;
    float2 xr[C::XV];
    setup_tables<C>(smem + S::OFF_TB, reinterpret_cast<const float*>(tables), C::V - 1, ln, pre_lane, post, roots, xr);
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
            synthesis_warp<C>(buf, exch + warp * C::G * C::XREG, pre_lane, post, roots, xr, h0 - 1, F, warp, ln);
        __syncthreads();

        float* orow = out + (long long)row * T_out;
        for (int sl = sl0; sl < ns; sl += S::SS) {
            const float* tp = buf + sl * C::ROW;          // frame h-1
            const float* th = tp + C::ROW;                // frame h
#pragma unroll
            for (int qi = 0; qi < S::NQ; ++qi) {
                const int i = 4 * (q0 + kThreads * qi);
                const bool lo = i < L;
                float4 o;
                if (lo) {
                    const float4 A = *reinterpret_cast<const float4*>(th + L + i);
                    const float4 Bv = *reinterpret_cast<const float4*>(tp + L - 4 - i);
                    o.x = w1[qi][0] * A.x + w2[qi][0] * Bv.w;
                    o.y = w1[qi][1] * A.y + w2[qi][1] * Bv.z;
                    o.z = w1[qi][2] * A.z + w2[qi][2] * Bv.y;
                    o.w = w1[qi][3] * A.w + w2[qi][3] * Bv.x;
                } else {
                    const float4 A = *reinterpret_cast<const float4*>(th + 3 * L - 4 - i);
                    const float4 Bv = *reinterpret_cast<const float4*>(tp + i - L);
                    o.x = w1[qi][0] * A.w + w2[qi][0] * Bv.x;
                    o.y = w1[qi][1] * A.z + w2[qi][1] * Bv.y;
                    o.z = w1[qi][2] * A.y + w2[qi][2] * Bv.z;
                    o.w = w1[qi][3] * A.x + w2[qi][3] * Bv.w;
                }
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
