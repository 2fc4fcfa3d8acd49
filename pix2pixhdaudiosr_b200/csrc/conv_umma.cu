// Implicit-GEMM convolution passes on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
//
// One persistent, warp-specialised kernel family covers every bf16 layer of models/networks.py whose gathered
// channel count is a multiple of 16:
//   MODE 0  forward-type pass, weights K-major      (Conv2d forward, ConvTranspose2d parity passes)
//   MODE 1  forward-type pass, weights MN-major     (input gradients: the same weight buffer read transposed)
//   MODE 2  weight gradient: dW_t[j,k] = sum_m dout[m,j] * in[gather(m,t),k]   (both operands MN-major)
//
// Layout in shared memory (all modes): blocks of [rows][64 bf16] with the 128-byte swizzle, i.e. the 16-byte
// chunk c of row r lives at r*128 + ((c ^ (r & 7)) << 4) inside a 1024-byte aligned block.  Read as a K-major
// operand the rows are GEMM rows/columns; read as an MN-major operand the rows are the reduction index and the
// 64 contiguous elements are GEMM rows/columns -- the memory image is the same, only the descriptor differs.
//
// Warp roles (288 threads, 1 CTA per SM):  warps 0-3 gather operands with cp.async (im2col on the fly: zero
// or reflection padding, strided gathers), warps 4-7 drain TMEM (bias, dtype conversion, scatter / split-K
// reduction), warp 8 allocates TMEM and issues tcgen05.mma.  Two accumulators (2 x 256 columns) let the
// epilogue of one tile overlap the main loop of the next.
#include <cuda.h>
#include <cstdint>
#include <cstdlib>
#include "dtype.cuh"

namespace p2p {
namespace {

constexpr int BM = 128;                  // GEMM rows per tile = TMEM lanes
constexpr int kStages = 4;
constexpr int kLook = kStages - 1;       // cp.async groups in flight before a stage is published
constexpr int kProducerThreads = 128;
constexpr int kThreads = 288;
constexpr int A_BYTES = 128 * 128;       // 16 KB: [128 rows][64 bf16] or 2 x [64 rows][64 bf16]
constexpr int B_BYTES = 256 * 128;       // 32 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int SMEM_BYTES = kStages * STAGE_BYTES + 1024 /* alignment slack */ + 256 /* barriers */ + 2048 /* staged bias */;
constexpr int BLOCK64 = 64 * 128;        // bytes of one [64 rows][64 bf16] block

struct UmmaArgs {
    p2p_conv_pass P;
    const __nv_bfloat16* in;     // gathered tensor
    const __nv_bfloat16* w;      // weights (MODE 0/1) or dout (MODE 2)
    const float* bias;
    void* out;                   // out (MODE 0/1: fp32 or bf16) or dW (MODE 2: fp32)
    long long M;                 // GEMM rows (MODE 0/1) / reduction length (MODE 2)
    int out_bf16;
    int BN;                      // tile width (multiple of 16, <= 256)
    int m_tiles, n_tiles;        // MODE 2: m_tiles = j tiles (128 wide), n_tiles = k tiles (BN wide)
    int nsplit, kb_per_split, total_kb;
    int cpb;                     // 64-channel chunks per tap (MODE 0/1)
    int atomic;                  // epilogue reduces with red.add.f32
    long long nwork;
};

struct RowCoord { int n, y, x; bool ok; };

__device__ __forceinline__ RowCoord decode_row(const p2p_conv_pass& P, long long m, long long M) {
    RowCoord r;
    r.ok = m < M;
    const long long mm = r.ok ? m : 0;
    r.x = (int)(mm % P.gx);
    const long long q = mm / P.gx;
    r.y = (int)(q % P.gy);
    r.n = (int)(q / P.gy);
    return r;
}

__device__ __forceinline__ long long gather_pixel(const p2p_conv_pass& P, const RowCoord& rc, int t) {
    int ih = rc.y * P.sy + P.dh[t], iw = rc.x * P.sx + P.dw[t];
    if (P.pad_mode == P2P_PAD_REFLECT) {
        ih = ih < 0 ? -ih : (ih >= P.H ? 2 * (P.H - 1) - ih : ih);
        iw = iw < 0 ? -iw : (iw >= P.W ? 2 * (P.W - 1) - iw : iw);
    }
    if (ih < 0 || ih >= P.H || iw < 0 || iw >= P.W) return -1;
    return ((long long)rc.n * P.H + ih) * P.W + iw;
}

__device__ __forceinline__ long long out_pixel(const p2p_conv_pass& P, const RowCoord& rc) {
    return ((long long)rc.n * P.OH + (rc.y * P.oys + P.oyo)) * P.OW + (rc.x * P.oxs + P.oxo);
}

// ---- PTX -------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {     // one lane of a converged warp
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc),
        "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// shared-memory matrix descriptor, 128-byte swizzle (layout_type 2), descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: D fp32, A/B bf16, M = 128, N = n
__device__ __forceinline__ uint32_t make_idesc(int n, int a_mn_major, int b_mn_major) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

struct Work { int mt, nt, sp, tap; };

template <int MODE>
__device__ __forceinline__ Work decode_work(const UmmaArgs& a, long long w) {
    Work k;
    k.sp = (int)(w % a.nsplit);
    w /= a.nsplit;
    k.nt = (int)(w % a.n_tiles);
    w /= a.n_tiles;
    if (MODE == 2) {
        k.mt = (int)(w % a.m_tiles);
        k.tap = (int)(w / a.m_tiles);
    } else {
        k.mt = (int)w;
        k.tap = 0;
    }
    return k;
}

template <int MODE>
__global__ void __launch_bounds__(kThreads, 1) conv_umma_kernel(const __grid_constant__ UmmaArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = base + kStages * STAGE_BYTES;
    // barrier slots (8 bytes each): full[kStages], empty[kStages], tfull[2], tempty[2]; then the TMEM base word
    auto full = [&](int s) { return bars + 8u * s; };
    auto empty = [&](int s) { return bars + 8u * (kStages + s); };
    auto tfull = [&](int i) { return bars + 8u * (2 * kStages + i); };
    auto tempty = [&](int i) { return bars + 8u * (2 * kStages + 2 + i); };
    const uint32_t tmem_slot = bars + 8u * (2 * kStages + 4);

    const p2p_conv_pass& P = a.P;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int K = P.K, J = P.J, BN = a.BN;

    if (warp == 8) {
        if (lane == 0) {
            for (int s = 0; s < kStages; ++s) { mbar_init(full(s), kProducerThreads); mbar_init(empty(s), 1); }
            for (int i = 0; i < 2; ++i) { mbar_init(tfull(i), 1); mbar_init(tempty(i), 128); }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_trigger();
    pdl_wait();

    if (warp < 4) {
        // ================================ producers: im2col gather with cp.async ================================
        int it = 0;
        for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x) {
            const Work wk = decode_work<MODE>(a, w);
            const int kb0 = wk.sp * a.kb_per_split, kb1 = min(a.total_kb, kb0 + a.kb_per_split);
            if (MODE != 2) {
                const RowCoord rc = decode_row(P, (long long)wk.mt * BM + tid, a.M);
                const int j0 = wk.nt * BN;
                int cur_tap = -1;
                long long px = -1;
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const int t = kb / a.cpb, c0 = (kb - t * a.cpb) * 64;
                    const int nch = min(64, K - c0) >> 3;
                    if (t != cur_tap) { px = rc.ok ? gather_pixel(P, rc, t) : -1; cur_tap = t; }
                    const int s = it % kStages;
                    if (it >= kStages) mbar_wait(empty(s), (uint32_t)((it / kStages - 1) & 1));
                    const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_BYTES;
                    {   // A: this thread's GEMM row
                        const __nv_bfloat16* src = a.in + (px >= 0 ? px * K + c0 : 0);
                        const uint32_t dst = sa + tid * 128;
                        const int sz = px >= 0 ? 16 : 0, sw = tid & 7;
#pragma unroll
                        for (int ch = 0; ch < 8; ++ch)
                            if (ch < nch) cp_async16(dst + ((ch ^ sw) << 4), src + ch * 8, sz);
                    }
                    const __nv_bfloat16* wt = a.w + (long long)P.wtap[t] * P.w_tap;
                    if (MODE == 0) {    // B rows = output channels, 64 gathered channels contiguous
                        for (int j = tid; j < BN; j += kProducerThreads) {
                            if (j0 + j < J) {
                                const __nv_bfloat16* src = wt + (long long)(j0 + j) * P.w_j + c0;
                                const uint32_t dst = sb + j * 128;
                                const int sw = j & 7;
#pragma unroll
                                for (int ch = 0; ch < 8; ++ch)
                                    if (ch < nch) cp_async16(dst + ((ch ^ sw) << 4), src + ch * 8, 16);
                            }
                        }
                    } else {            // B rows = gathered channels, output channels contiguous (64 per block)
                        const int kv = nch << 3, nchunk = 64 * (BN >> 3);
                        for (int idx = tid; idx < nchunk; idx += kProducerThreads) {
                            const int k = idx & 63, cj = idx >> 6;
                            if (k < kv && j0 + cj * 8 < J) {
                                const __nv_bfloat16* src = wt + (long long)(c0 + k) * P.w_k + j0 + cj * 8;
                                cp_async16(sb + (cj >> 3) * BLOCK64 + k * 128 + (((cj & 7) ^ (k & 7)) << 4), src, 16);
                            }
                        }
                    }
                    cp_async_commit();
                    if (it >= kLook) {
                        cp_async_wait<kLook>();
                        fence_proxy_async();
                        mbar_arrive(full((it - kLook) % kStages));
                    }
                }
            } else {
                const int t = wk.tap, j0 = wk.mt * BM, k0 = wk.nt * BN;
                const int row = tid & 63, half = tid >> 6, sw = row & 7;
                const __nv_bfloat16* dout = a.w;
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    const RowCoord rc = decode_row(P, (long long)kb * 64 + row, a.M);
                    const long long px = rc.ok ? gather_pixel(P, rc, t) : -1;
                    const long long opx = rc.ok ? out_pixel(P, rc) : 0;
                    const int s = it % kStages;
                    if (it >= kStages) mbar_wait(empty(s), (uint32_t)((it / kStages - 1) & 1));
                    const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_BYTES;
                    {   // A = dout rows (reduction index = pixel), block `half` of the 128 output channels
                        const __nv_bfloat16* src = dout + opx * J + j0 + half * 64;
                        const uint32_t dst = sa + half * BLOCK64 + row * 128;
                        const int sz = rc.ok ? 16 : 0;
#pragma unroll
                        for (int ch = 0; ch < 8; ++ch)
                            if (j0 + half * 64 + ch * 8 < J) cp_async16(dst + ((ch ^ sw) << 4), src + ch * 8, sz);
                    }
                    {   // B = gathered input rows, chunk columns half, half+2, ...
                        const __nv_bfloat16* src = a.in + (px >= 0 ? px * K : 0) + k0;
                        const int sz = px >= 0 ? 16 : 0;
                        for (int cj = half; cj < (BN >> 3); cj += 2)
                            if (k0 + cj * 8 < K)
                                cp_async16(sb + (cj >> 3) * BLOCK64 + row * 128 + (((cj & 7) ^ sw) << 4), src + cj * 8, sz);
                    }
                    cp_async_commit();
                    if (it >= kLook) {
                        cp_async_wait<kLook>();
                        fence_proxy_async();
                        mbar_arrive(full((it - kLook) % kStages));
                    }
                }
            }
        }
        cp_async_wait<0>();
        fence_proxy_async();
        for (int d = it > kLook ? it - kLook : 0; d < it; ++d) mbar_arrive(full(d % kStages));
    } else if (warp == 8) {
        // ================================ MMA issuer (warp-uniform schedule, one elected lane issues) ==========
        {
            const uint32_t idesc = make_idesc(BN, MODE == 2, MODE != 0);
            // loop-invariant descriptor halves; the k16 advance is an add on the start-address field
            const uint64_t a_hi = MODE == 2 ? make_desc(0, BLOCK64, 1024) : make_desc(0, 16, 1024);
            const uint64_t b_hi = MODE == 0 ? make_desc(0, 16, 1024) : make_desc(0, BLOCK64, 1024);
            const uint32_t a_k16 = MODE == 2 ? (2048u >> 4) : (32u >> 4), b_k16 = MODE == 0 ? (32u >> 4) : (2048u >> 4);
            int it = 0, li = 0;
            for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x, ++li) {
                const Work wk = decode_work<MODE>(a, w);
                const int kb0 = wk.sp * a.kb_per_split, kb1 = min(a.total_kb, kb0 + a.kb_per_split);
                const int acc = li & 1, v = li >> 1;
                if (v >= 1) mbar_wait(tempty(acc), (uint32_t)((v - 1) & 1));
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * 256;
                int cc = MODE != 2 ? kb0 % a.cpb : 0;        // channel chunk inside the tap (the last one may be short)
                for (int kb = kb0; kb < kb1; ++kb, ++it) {
                    int nk16 = 4;
                    if (MODE != 2) {
                        nk16 = min(64, K - cc * 64) >> 4;
                        if (++cc == a.cpb) cc = 0;
                    }
                    const int s = it % kStages;
                    const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_BYTES;
                    const uint64_t ad = a_hi | (uint64_t)(sa >> 4), bd = b_hi | (uint64_t)(sb >> 4);
                    mbar_wait(full(s), (uint32_t)((it / kStages) & 1));
                    tc_fence_after();
                    if (elect_one()) {
                        umma_bf16(d_tmem, ad, bd, idesc, kb > kb0 ? 1u : 0u);
                        if (nk16 > 1) umma_bf16(d_tmem, ad + a_k16, bd + b_k16, idesc, 1u);
                        if (nk16 > 2) umma_bf16(d_tmem, ad + 2 * a_k16, bd + 2 * b_k16, idesc, 1u);
                        if (nk16 > 3) umma_bf16(d_tmem, ad + 3 * a_k16, bd + 3 * b_k16, idesc, 1u);
                        umma_commit(empty(s));
                    }
                    __syncwarp();
                }
                if (elect_one()) umma_commit(tfull(acc));
                __syncwarp();
            }
        }
        __syncwarp();
    } else {
        // ================================ epilogue: TMEM -> registers -> global ================================
        const int ew = warp - 4, row = ew * 32 + lane;
        int li = 0;
        for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x, ++li) {
            const Work wk = decode_work<MODE>(a, w);
            const int acc = li & 1, v = li >> 1;
            long long obase;            // element offset of this thread's output row
            bool rok;
            int c_begin, c_count;       // first output column index and number of valid columns of the tile
            if (MODE != 2) {
                const RowCoord rc = decode_row(P, (long long)wk.mt * BM + row, a.M);
                rok = rc.ok;
                c_begin = wk.nt * BN;
                c_count = min(BN, J - c_begin);
                obase = out_pixel(P, rc) * J + c_begin;
            } else {
                const int j = wk.mt * BM + row;
                rok = j < J;
                c_begin = wk.nt * BN;
                c_count = min(BN, K - c_begin);
                obase = (long long)P.wtap[wk.tap] * P.w_tap + (long long)j * P.w_j + c_begin;
            }
            const bool add_bias = MODE != 2 && a.bias != nullptr && wk.sp == 0;
            float* bs_ = reinterpret_cast<float*>(smem_raw + (bars + 256 - smem_u32(smem_raw))) + acc * 256;
            if (MODE != 2) {
                for (int c = row; c < BN; c += 128) bs_[c] = (add_bias && c < c_count) ? __ldg(a.bias + c_begin + c) : 0.f;
                asm volatile("bar.sync 1, 128;" ::: "memory");      // double-buffered by accumulator slot
            }
            mbar_wait(tfull(acc), (uint32_t)(v & 1));
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + acc * 256;
            for (int c2 = 0; c2 < BN; c2 += 32) {
              uint32_t r2[2][16];
              tmem_ld16(taddr + c2, r2[0]);
              if (c2 + 16 < BN) tmem_ld16(taddr + c2 + 16, r2[1]);
              tmem_ld_wait();
#pragma unroll
              for (int hh = 0; hh < 2; ++hh) {
                const int c = c2 + hh * 16;
                if (!rok || c >= c_count) continue;
                float f[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(r2[hh][i]);
                if (MODE != 2) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 bv = *reinterpret_cast<const float4*>(bs_ + c + 4 * i);
                        f[4 * i] += bv.x; f[4 * i + 1] += bv.y; f[4 * i + 2] += bv.z; f[4 * i + 3] += bv.w;
                    }
                }
                const bool fullv = c + 16 <= c_count;
                if (a.out_bf16) {
                    __nv_bfloat16* o = (__nv_bfloat16*)a.out + obase + c;
                    if (fullv && (((uintptr_t)o) & 15) == 0) {
                        uint32_t pk[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
                            pk[i] = *reinterpret_cast<uint32_t*>(&h);
                        }
                        reinterpret_cast<uint4*>(o)[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        reinterpret_cast<uint4*>(o)[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (c + i < c_count) o[i] = __float2bfloat16_rn(f[i]);
                    }
                } else {
                    float* o = (float*)a.out + obase + c;
                    if (a.atomic) {
                        if (fullv && (((uintptr_t)o) & 15) == 0) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) red_add_v4(o + 4 * i, f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; ++i)
                                if (c + i < c_count) atomicAdd(o + i, f[i]);
                        }
                    } else if (fullv && (((uintptr_t)o) & 15) == 0) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            reinterpret_cast<float4*>(o)[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (c + i < c_count) o[i] = f[i];
                    }
                }
              }
            }
            tc_fence_before();
            mbar_arrive(tempty(acc));
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

int g_force_simple = -1;     // -1: read P2P_CONV_SIMPLE from the environment on first use
int g_sm_count = 0;
long long g_launches = 0;    // tcgen05 kernel launches so far (tests assert the tensor-core path really ran)

bool aligned16(const void* p) { return (((uintptr_t)p) & 15) == 0; }

int sm_count() {
    if (!g_sm_count) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    return g_sm_count;
}

template <int MODE>
int launch(const UmmaArgs& a, cudaStream_t st, const char* what) {
    static bool configured[64] = {};            // per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        if (cudaFuncSetAttribute(conv_umma_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess)
            return check_launch("cudaFuncSetAttribute(conv_umma_kernel)");
        configured[dev] = true;
    }
    const long long grid = a.nwork < sm_count() ? a.nwork : sm_count();
    launch_k(conv_umma_kernel<MODE>, dim3((unsigned)grid), dim3(kThreads), SMEM_BYTES, st, a);
    ++g_launches;
    return check_launch(what);
}

int pick_bn(int cols) {
    const int nt = (cols + 255) / 256;
    int bn = ((cols + nt - 1) / nt + 15) / 16 * 16;
    return bn < 16 ? 16 : bn;
}

}  // namespace

void note_umma_launch() { ++g_launches; }     // the window kernel (conv_window.cu) is a tcgen05 kernel too

bool conv_umma_disabled() {
    if (g_force_simple < 0) {
        const char* e = getenv("P2P_CONV_SIMPLE");
        g_force_simple = (e && e[0] == '1') ? 1 : 0;
    }
    return g_force_simple == 1;
}

// Returns 1 when the tcgen05 path does not cover this call (caller falls through to the CUDA-core kernels),
// 0 on success, negative on error.
int conv_forward_umma(const p2p_conv_pass* P, const void* in, const void* weight, const float* bias, void* out, int dtype,
                      int out_dtype, cudaStream_t st) {
    if (conv_umma_disabled() || dtype != P2P_BF16) return 1;
    const long long M = (long long)P->N * P->gy * P->gx;
    if (M <= 0 || P->K % 16 != 0 || !aligned16(in) || !aligned16(weight) || !aligned16(out)) return 1;
    int mode;
    if (P->w_k == 1 && P->w_j % 8 == 0 && P->w_tap % 8 == 0) mode = 0;
    else if (P->w_j == 1 && P->w_k % 8 == 0 && P->w_tap % 8 == 0 && P->J % 8 == 0) mode = 1;
    else return 1;
    UmmaArgs a;
    a.P = *P;
    a.in = (const __nv_bfloat16*)in;
    a.w = (const __nv_bfloat16*)weight;
    a.bias = bias;
    a.out = out;
    a.M = M;
    a.out_bf16 = out_dtype == P2P_BF16;
    const long long mt = (M + BM - 1) / BM;
    a.cpb = (P->K + 63) / 64;
    a.total_kb = P->ntaps * a.cpb;
    // Tile width and reduction split: these layers are operand-load bound when few tiles exist, so model a work
    // item's cost as k-blocks x (A rows + B rows) and minimise rounds x cost over BN (multiples of 16) and, for fp32
    // outputs scattered densely (reduced with red.add into a zeroed buffer), over the split of the reduction.
    const bool can_split = !a.out_bf16 && P->oys == 1 && P->oxs == 1 && P->gy == P->OH && P->gx == P->OW;
    const int sms = sm_count();
    double best = 1e300;
    int best_bn = 16, best_split = 1;
    for (int bn = 16; bn <= 256; bn += 16) {
        const long long nt = (P->J + bn - 1) / bn;
        if (bn > 16 && nt == (P->J + bn - 17) / (bn - 16)) continue;     // a narrower tile covers J equally
        const int max_split = (can_split && !deterministic()) ? (a.total_kb >= 8 ? a.total_kb / 4 : 1) : 1;
        for (int sp = 1; sp <= max_split; sp = sp < 4 ? sp + 1 : sp * 2) {
            const long long items = mt * nt * sp;
            const long long rounds = (items + sms - 1) / sms;
            const int kbs = (a.total_kb + sp - 1) / sp;
            const double cost = (double)rounds * (kbs * (double)(128 + bn) + 600.0 + (sp > 1 ? 4.0 * bn : 0.0));
            if (cost < best) { best = cost; best_bn = bn; best_split = sp; }
        }
    }
    a.BN = best_bn;
    a.n_tiles = (P->J + a.BN - 1) / a.BN;
    if (mt * a.n_tiles > 2000000000LL) return 1;
    a.m_tiles = (int)mt;
    const long long tiles = mt * a.n_tiles;
    a.kb_per_split = (a.total_kb + best_split - 1) / best_split;
    a.nsplit = (a.total_kb + a.kb_per_split - 1) / a.kb_per_split;
    a.atomic = a.nsplit > 1;
    a.nwork = tiles * a.nsplit;
    if (a.atomic && cudaMemsetAsync(out, 0, (size_t)M * P->J * sizeof(float), st) != cudaSuccess)
        return check_launch("cudaMemsetAsync(split-K output)");
    return mode == 0 ? launch<0>(a, st, "conv_umma_kernel<0>") : launch<1>(a, st, "conv_umma_kernel<1>");
}

int conv_wgrad_umma(const p2p_conv_pass* P, const void* in, const void* dout, float* dw, int dtype, cudaStream_t st) {
    if (conv_umma_disabled() || dtype != P2P_BF16) return 1;
    const long long M = (long long)P->N * P->gy * P->gx;
    if (M <= 0 || P->K % 16 != 0 || P->J % 8 != 0 || P->w_k != 1 || !aligned16(in) || !aligned16(dout) || !aligned16(dw))
        return 1;
    if (P->J < 16) return 1;          // one or two output channels: 126 of 128 tensor rows would idle
    UmmaArgs a;
    a.P = *P;
    a.in = (const __nv_bfloat16*)in;
    a.w = (const __nv_bfloat16*)dout;
    a.bias = nullptr;
    a.out = dw;
    a.M = M;
    a.out_bf16 = 0;
    a.BN = pick_bn(P->K);
    a.n_tiles = (P->K + a.BN - 1) / a.BN;
    a.m_tiles = (P->J + BM - 1) / BM;
    a.cpb = 1;
    a.total_kb = (int)((M + 63) / 64);
    const long long tiles = (long long)P->ntaps * a.m_tiles * a.n_tiles;
    int nsplit = 1;
    if (tiles < 2LL * sm_count()) {
        nsplit = (int)((2LL * sm_count() + tiles - 1) / tiles);
        const int max_split = a.total_kb / 4;
        if (nsplit > max_split) nsplit = max_split;
        if (nsplit < 1 || deterministic()) nsplit = 1;
    }
    a.kb_per_split = (a.total_kb + nsplit - 1) / nsplit;
    a.nsplit = (a.total_kb + a.kb_per_split - 1) / a.kb_per_split;
    a.atomic = 1;                 // dW += : the caller may hand in a live gradient buffer (FusedAdam's flat grads)
    a.nwork = tiles * a.nsplit;
    return launch<2>(a, st, "conv_umma_kernel<2>");
}

}  // namespace p2p

extern "C" long long p2p_conv_umma_launches(void) { return p2p::g_launches; }

extern "C" int p2p_conv_force_simple(int on) {
    const int old = p2p::conv_umma_disabled() ? 1 : 0;
    if (on >= 0) p2p::g_force_simple = on ? 1 : 0;      // negative: query only
    return old;
}
