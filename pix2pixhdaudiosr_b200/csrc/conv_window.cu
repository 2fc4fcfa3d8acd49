// Stride-1 convolution as a TMA-fed implicit GEMM with *shifted-window* operand descriptors (tcgen05 / TMEM).
//
// The input is a materialised-padding NHWC tensor xp [N, Hp, Wp, C] (zero or reflection padding written once by
// p2p_pad_nhwc), so a "valid" convolution  out[n,y,x,:] = sum_t W_t . xp[n, y+dh_t, x+dw_t, :]  reads, for GEMM row
// q = (n*Hp + y)*Wp + x, the pixel row q + dh_t*Wp + dw_t: every tap is the SAME shared-memory window read through a
// descriptor whose start address is shifted by dh_t*Wp + dw_t rows.  One TMA box load per 64-channel chunk
// therefore feeds all KH*KW taps (9x / 16x / 49x less operand traffic than gathering im2col rows per tap), and no
// thread touches the activations.  Rows whose (y, x) fall in the padding margin are computed and discarded
// ((Hp*Wp)/(OH*OW) - 1 overhead: 3 % at 256x64, 6 % at 512x128 with 7x7).
//
// Window layout in shared memory: rows of 64 bf16 (128 bytes), 128-byte swizzle keyed on the absolute address (what
// TMA writes and what the UMMA descriptor's base-offset field describes for starts that are not 1024-byte aligned).
//
// Warp roles (192 threads, 1 CTA / SM, persistent over (row tile, column tile) work items):
//   warp 0: TMA producer -- window boxes (2 buffers) and per-tap weight tiles (4-stage ring)
//   warp 1: TMEM allocation + tcgen05.mma issue
//   warps 2-5: epilogue (TMEM -> registers -> bias -> global), double-buffered accumulators
#include <cuda.h>
#include <cstdint>
#include "dtype.cuh"

namespace p2p {
namespace {

constexpr int WBM = 128;
constexpr int kBStagesMax = 16;
constexpr int kWinRowsMax = 312;                    // 39 KB per window buffer
constexpr int WIN_BYTES = kWinRowsMax * 128;
constexpr int WB_BYTES = 256 * 128;                 // one weight tile: up to 256 output channels x 64 channels
constexpr int WRING_BYTES = 4 * WB_BYTES;             // 128 KB of weight tiles in flight, cut into b_stage_bytes stages
constexpr int EPI_STAGE_BYTES = 4 * 4096;           // one 32-row x 128-byte transposition tile per epilogue warp
constexpr int kWinThreads = 320;                    // conv_window_kernel: TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter)
constexpr int WEPI_STAGE_BYTES = 8 * 4096;
constexpr int WDATA_BYTES = 2 * WIN_BYTES + WRING_BYTES - 16384;     // two window buffers + the weight ring, carved per launch (win_bytes)
constexpr int WSMEM = WDATA_BYTES + 1024 + 512 + 32 * kBStagesMax + 2048 + WEPI_STAGE_BYTES;     // + barriers + stage records + 2 x 256 staged bias values
// stage-record flags (written by the producer, read by the MMA warp)
constexpr uint32_t REC_TILE_FIRST = 1u << 8, REC_CHUNK_FIRST = 1u << 9, REC_CHUNK_LAST = 1u << 10, REC_TILE_LAST = 1u << 11, REC_LAST = 1u << 12;
static_assert(WSMEM <= 232448, "window kernel shared memory");
constexpr int kWThreads = 192;

struct WindowArgs {
    CUtensorMap tm_x;        // xp viewed [Q = N*Hp*Wp rows][C]: box [64 channels][rows_per_box]
    CUtensorMap tm_w;        // weights: K-major [taps*J rows][C] box [64][BN]; MN-major [taps*K rows][J] box [64][64]
    const float* bias;
    void* out;
    float* stats;            // optional [N][J][2] fp32, zeroed by the host: per (image, channel) sum and sum of squares of the
                             // bias-free accumulators over the valid output pixels (InstanceNorm statistics from the epilogue)
    long long Q;             // GEMM rows = padded pixels of one plane
    long long Qp;            // rows between consecutive planes (>= Q)
    int N, Hp, Wp, C, J;
    int OHf, OWf, oys, oxs;  // output tensor extent and scatter stride: pixel (y*oys+oyo, x*oxs+oxo)
    int ntaps, nplanes, nsub;
    int gy[4], gx[4], oyo[4], oxo[4];   // per sub-pass: valid rows / columns of the padded grid, scatter offsets
    int pbeg[4][5];          // sub-pass u: taps [pbeg[u][p], pbeg[u][p+1]) read plane p
    int win_rows, rows_per_box, nboxes;
    int BN, n_tiles, m_tiles, cpb;
    int out_bf16, b_mn_major, out_bf16_hint;
    int base_off_mode;
    int tps, b_tap_bytes;          // taps per ring stage (amortises the issuing thread's per-stage overhead), bytes of one tap tile
    int b_stages, b_stage_bytes;   // weight ring: stages of tps tap tiles, as many as fit WRING_BYTES
    int fast_epi;                  // output rows 16-byte aligned: staged, coalesced epilogue stores
    long long* dbg;                // P2P_WINDOW_TIMING builds: per CTA {wait A, wait B, issue, wait TMEM, stages, total} cycles of the MMA warp
    int mrep;                      // 128-row MMAs per weight tile (1 or 2): a work item covers 128*mrep padded pixels
    int win_bytes;                 // bytes of one window buffer (multiple of 1024); the ring gets WDATA_BYTES - 2*win_bytes
    int acc_slots, half_cols;      // TMEM: accumulator slots of 256 (or 512) columns, column stride between the two row halves
    int resident;                  // the weight ring holds ALL (chunk, tap) tiles of a work item: loaded once, re-used by every item
    int nsplit, ci_per_split;      // split of the (channel chunk, plane) loop across CTAs; > 1 => fp32 red.add epilogue   // weight-tile ring: as many stages as fit 128 KB (TMA latency needs bytes in flight)
    short shift[P2P_MAX_TAPS];       // dh*Wp + dw
    unsigned char wtap[P2P_MAX_TAPS];
    long long nwork;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    // bounded: a protocol error traps (an error the host sees) instead of hanging the GPU.  One try_wait suspends the thread
    // for up to the hardware's time limit, so 2^22 failed attempts are seconds, far beyond any legitimate wait here.
    uint32_t tries = 0;
    for (;;) {
        uint32_t ok;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return;
        if (++tries > (1u << 22)) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
        "l"(tm), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
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
// ---- CTA pairs (tcgen05 cta_group::2): one MMA covers 2 x 128 rows, each CTA holds half of the B tile ------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t saddr, uint32_t rank) {      // shared::cluster address of `saddr` in CTA `rank`
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cbar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cbar) : "memory");
}
// TMA load issued by the pair's leader into the shared memory (same offset `dst`) of the CTAs in `mask`; completion is
// signalled on the leader's barrier `bar` (.cta_group::2: the barrier address selects the CTA of the pair that receives it)
__device__ __forceinline__ void tma_load_2d_to(uint32_t dst, const CUtensorMap* tm, int c0, int c1, uint32_t bar, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%2, %3}], [%4], %5;"
        ::"r"(dst), "l"(tm), "r"(c0), "r"(c1), "r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc),
        "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma2_commit(uint32_t bar, uint16_t mask) {     // arrives on `bar` (same offset) in the CTAs of `mask`
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"(mask) : "memory");
}
template <int CG> __device__ __forceinline__ void umma_cg(uint32_t d, uint64_t ad, uint64_t bd, uint32_t idesc, uint32_t acc) {
    if constexpr (CG == 2) umma2_bf16(d, ad, bd, idesc, acc); else umma_bf16(d, ad, bd, idesc, acc);
}
template <int CG> __device__ __forceinline__ void commit_cg(uint32_t bar, uint16_t mask) {
    if constexpr (CG == 2) umma2_commit(bar, mask); else umma_commit(bar);
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


// ---- coalesced epilogue stores --------------------------------------------------------------------------------------------
// After tcgen05.ld a lane owns one accumulator ROW (32 lanes = 32 rows), so a direct v4 store touches 32 different 128-byte
// lines with 16 bytes each and the LSU, not HBM, bounds the epilogue.  Each warp therefore transposes its 32-row chunk through
// a private 4 KB shared-memory tile (16-byte pieces XOR-swizzled against bank conflicts) and stores rows cooperatively:
// P lanes cover one row chunk of P*16 contiguous bytes, 32/P rows per instruction.
template <int P> __device__ __forceinline__ int piece_swz(int row) {
    return P == 8 ? (row & 7) : (P == 4 ? ((row >> 1) & 3) : (P == 2 ? ((row >> 2) & 1) : 0));
}
enum { EPI_ST = 0, EPI_RED = 1 };
// regs: this lane's row chunk as P 16-byte pieces.  row_off: byte offset of this lane's row start (chunk column 0) in out;
// okmask: ballot of valid rows; valid_bytes: bytes of the chunk inside the tensor (multiple of esz).
// st: statistics target of this chunk's first column (nullptr: none), &stats[(n*J + column) * 2], valid when all valid rows of
// the warp belong to image n; bias4: shared-memory address of the chunk's staged bias values (statistics are taken of v - bias).
template <int P>
__device__ __forceinline__ void stage_store_chunk(uint32_t stage, int lane, const uint32_t* regs, long long row_off, unsigned okmask,
                                                  char* out, int valid_bytes, int mode, int esz, float* st = nullptr,
                                                  const float* bias4 = nullptr) {
    __syncwarp();                               // the previous chunk's reads of the tile are done
#pragma unroll
    for (int p = 0; p < P; ++p)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stage + (uint32_t)(lane * P + (p ^ piece_swz<P>(lane))) * 16u),
                     "r"(regs[4 * p]), "r"(regs[4 * p + 1]), "r"(regs[4 * p + 2]), "r"(regs[4 * p + 3]) : "memory");
    __syncwarp();
    const int p = lane % P, r0 = lane / P;
    float sa[4] = {0.f, 0.f, 0.f, 0.f}, sq[4] = {0.f, 0.f, 0.f, 0.f}, bb[4] = {0.f, 0.f, 0.f, 0.f};
    if (st) {
        const float4 b4 = *reinterpret_cast<const float4*>(bias4 + p * 4);
        bb[0] = b4.x; bb[1] = b4.y; bb[2] = b4.z; bb[3] = b4.w;
    }
#pragma unroll
    for (int k = 0; k < P; ++k) {
        const int row = r0 + (32 / P) * k;
        uint32_t v0, v1, v2, v3;
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v0), "=r"(v1), "=r"(v2), "=r"(v3)
                     : "r"(stage + (uint32_t)(row * P + (p ^ piece_swz<P>(row))) * 16u));
        const long long ro = __shfl_sync(0xffffffffu, row_off, row);
        if (!((okmask >> row) & 1u)) continue;
        if (st) {           // fp32 pieces: this lane sees the same 4 columns in every row it reads
            const float d0 = __uint_as_float(v0) - bb[0], d1 = __uint_as_float(v1) - bb[1], d2 = __uint_as_float(v2) - bb[2],
                        d3 = __uint_as_float(v3) - bb[3];
            sa[0] += d0; sa[1] += d1; sa[2] += d2; sa[3] += d3;
            sq[0] += d0 * d0; sq[1] += d1 * d1; sq[2] += d2 * d2; sq[3] += d3 * d3;
        }
        char* dst = out + ro + p * 16;
        if (p * 16 + 16 <= valid_bytes) {
            if (mode == EPI_RED) red_add_v4(reinterpret_cast<float*>(dst), __uint_as_float(v0), __uint_as_float(v1), __uint_as_float(v2), __uint_as_float(v3));
            else *reinterpret_cast<uint4*>(dst) = make_uint4(v0, v1, v2, v3);
        } else if (p * 16 < valid_bytes) {      // ragged last piece of the tensor row
            const uint32_t vv[4] = {v0, v1, v2, v3};
            if (esz == 4) {
                for (int i = 0; i < 4 && p * 16 + 4 * i < valid_bytes; ++i) {
                    if (mode == EPI_RED) atomicAdd(reinterpret_cast<float*>(dst) + i, __uint_as_float(vv[i]));
                    else reinterpret_cast<uint32_t*>(dst)[i] = vv[i];
                }
            } else {
                for (int i = 0; i < 8 && p * 16 + 2 * i < valid_bytes; ++i)
                    reinterpret_cast<uint16_t*>(dst)[i] = (uint16_t)(vv[i >> 1] >> ((i & 1) * 16));
            }
        }
    }
    if (st) {               // lanes p, p + P, ... hold partial sums of the same 4 columns: fold them, one lane per piece adds
#pragma unroll
        for (int o = P; o < 32; o <<= 1) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                sa[i] += __shfl_xor_sync(0xffffffffu, sa[i], o);
                sq[i] += __shfl_xor_sync(0xffffffffu, sq[i], o);
            }
        }
        if (r0 == 0 && p * 16 + 16 <= valid_bytes) {
            float* t = st + p * 8;          // [column][sum, sum of squares]
            red_add_v4(t, sa[0], sq[0], sa[1], sq[1]);
            red_add_v4(t + 4, sa[2], sq[2], sa[3], sq[3]);
        }
    }
}
__device__ __forceinline__ void stage_store(int P, uint32_t stage, int lane, const uint32_t* regs, long long row_off, unsigned okmask,
                                            char* out, int valid_bytes, int mode, int esz, float* st = nullptr,
                                            const float* bias4 = nullptr) {
    if (P == 8) stage_store_chunk<8>(stage, lane, regs, row_off, okmask, out, valid_bytes, mode, esz, st, bias4);
    else if (P == 4) stage_store_chunk<4>(stage, lane, regs, row_off, okmask, out, valid_bytes, mode, esz, st, bias4);
    else stage_store_chunk<2>(stage, lane, regs, row_off, okmask, out, valid_bytes, mode, esz);
}

// 128-byte-swizzle matrix descriptor; base_off = (start >> 7) & 7 for starts inside a 1024-byte swizzle atom
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t base_off) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46) | ((uint64_t)(base_off & 7) << 49) | (2ull << 61);
}

// the MMAs of one weight-ring stage: up to 4 taps x up to 4 k16 steps, each against MREP 128-row halves of the window (the
// halves share the weight tile, so it is written to shared memory once per 128*MREP pixels)
template <int MREP, int CG>
__device__ __forceinline__ void issue_stage(uint32_t d_tmem, uint32_t half_cols, uint32_t a_hi, uint32_t b_hi, uint32_t al0, uint32_t al1,
                                            uint32_t al2, uint32_t al3, uint32_t bl, uint32_t b_k16, uint32_t b_tap, uint32_t idesc,
                                            uint32_t meta, uint32_t accumulate) {
    auto mk = [](uint32_t hi, uint32_t lo) { return ((uint64_t)hi << 32) | lo; };
    constexpr uint32_t A_HALF = (128u * 128u) >> 4;         // descriptor advance to the second 128 window rows
    const uint32_t al[4] = {al0, al1, al2, al3};
    if ((meta & 63u) == (4u | (4u << 3))) {                 // the common stage: 4 taps x 64 channels, no predicates
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t bk = bl + (uint32_t)k * b_tap;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
#pragma unroll
                for (int h = 0; h < MREP; ++h)
                    umma_cg<CG>(d_tmem + h * half_cols, mk(a_hi, al[k] + 2 * j + h * A_HALF), mk(b_hi, bk + j * b_k16), idesc,
                              (k | j) == 0 ? accumulate : 1u);
            }
        }
    } else {
        const uint32_t nt_s = meta & 7u, nk16 = (meta >> 3) & 7u;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if ((uint32_t)k < nt_s) {
                const uint32_t bk = bl + (uint32_t)k * b_tap;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if ((uint32_t)j < nk16) {
#pragma unroll
                        for (int h = 0; h < MREP; ++h)
                            umma_cg<CG>(d_tmem + h * half_cols, mk(a_hi, al[k] + 2 * j + h * A_HALF), mk(b_hi, bk + j * b_k16), idesc,
                                      (k | j) == 0 ? accumulate : 1u);
                    }
                }
            }
        }
    }
}

// CG = 2: the CTAs of a pair (cluster of 2) each own 128*mrep window rows and half of every weight tile; the LEADER's TMA warp
// fills both CTAs' shared memory (multicast-addressed loads, one destination each) and its MMA warp issues
// tcgen05.mma.cta_group::2 (M = 256) for both, so the refill loop never crosses the pair; the other CTA only drains its TMEM.  Measured (profiles/r02_mma_rate_cg2.txt): 46-50 cycles per MMA for
// N <= 96 against 93 for cta_group::1 -- the per-instruction floor that capped the 96-channel layers at half the tensor peak.
template <int CG>
__global__ void __launch_bounds__(kWinThreads, 1) conv_window_kernel(const __grid_constant__ WindowArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
    const uint32_t cid = blockIdx.x / CG, ncl = gridDim.x / CG;          // pair index / number of pairs: the work-item stride
#ifdef P2P_WINDOW_TIMING
    long long g_entry;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_entry));
#endif
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t win_bytes = (uint32_t)a.win_bytes;
    const uint32_t wbase = base + 2 * win_bytes;
    const uint32_t bars = base + WDATA_BYTES;
    const int kBStages = a.b_stages;
    // barriers: afull[2] aempty[2] bfull[16] bempty[16] tfull[2] tempty[2], then the TMEM base word
    auto afull = [&](int i) { return bars + 8u * i; };
    auto aempty = [&](int i) { return bars + 8u * (2 + i); };
    auto bfull = [&](int s) { return bars + 8u * (4 + s); };
    auto bempty = [&](int s) { return bars + 8u * (4 + kBStagesMax + s); };
    auto tfull = [&](int i) { return bars + 8u * (4 + 2 * kBStagesMax + i); };
    auto tempty = [&](int i) { return bars + 8u * (6 + 2 * kBStagesMax + i); };
    const uint32_t tmem_slot = bars + 8u * (8 + 2 * kBStagesMax);
    const uint32_t WB_STRIDE = (uint32_t)a.b_stage_bytes;       // = tps * TAP_STRIDE
    const uint32_t TAP_STRIDE = (uint32_t)a.b_tap_bytes;
    const uint32_t recs = bars + 512;                           // one 32-byte stage record per weight-ring stage
    const uint32_t bias_u32 = recs + 32 * kBStagesMax;
    const uint32_t epi_stage = bias_u32 + 2048;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int BN = a.BN, C = a.C, J = a.J;
    const uint32_t a_bytes = (uint32_t)a.nboxes * a.rows_per_box * 128u;
    const int BNL = BN / CG;                                             // weight-tile columns held by this CTA
    const uint32_t b_bytes = a.b_mn_major ? (uint32_t)((BNL + 63) / 64) * 8192u : (uint32_t)BNL * 128u;

    if (warp == 1) {
        if (lane == 0) {
            // full / empty barriers are used in the leader only; tfull receives the multicast commit in both CTAs; the leader's
            // tempty collects one arrive per epilogue warp of every CTA
            for (int i = 0; i < 2; ++i) { mbar_init(afull(i), 1); mbar_init(aempty(i), 1); mbar_init(tfull(i), 1); mbar_init(tempty(i), 8 * CG); }
            for (int s = 0; s < kBStages; ++s) { mbar_init(bfull(s), 1); mbar_init(bempty(s), 1); }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        if constexpr (CG == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (CG == 2) cluster_sync_all();          // the peer's barriers are initialised before anything arrives on them
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_trigger();          // the next kernel on the stream may start its prologue while this one runs ...
    pdl_wait();             // ... and this one touches global memory only after its predecessor has completed (common.cuh)
    // barrier addresses as seen from this CTA for arrivals that must land in the leader
    auto lead = [&](uint32_t bar) { return CG == 2 ? mapa_rank(bar, 0) : bar; };

    if (warp == 0) {
        // ================================ TMA producer ================================
        // Besides the loads, the producer writes one 32-byte record per weight-ring stage (A/B descriptor low words, tap and
        // k16 counts, tile/chunk boundary flags) so that the MMA warp's loop is flat: wait, read the record, issue.  Measured
        // with -DP2P_WINDOW_TIMING: the nested schedule walk in the issuing warp cost as many cycles as the MMAs themselves.
        if (lane == 0 && rank == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_x) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_w) : "memory");
            const uint32_t a_lo = (uint32_t)make_desc(0, 16, 1024, 0);
            const uint32_t b_lo = (uint32_t)(a.b_mn_major ? make_desc(0, 8192, 1024, 0) : make_desc(0, 16, 1024, 0));
            int ia = 0;                     // window loads issued so far
            uint32_t bs = 0, bph = 1;       // weight ring: stage, parity of the "empty" wait (first lap passes at once)
            const uint32_t nwork = (uint32_t)a.nwork, n_tiles = (uint32_t)a.n_tiles, nsplit = (uint32_t)a.nsplit;
            for (uint32_t w = cid; w < nwork; w += ncl) {
                const uint32_t nt = w % n_tiles, wr = w / n_tiles;
                const uint32_t sp = wr % nsplit, mtu = wr / nsplit;
                const int u = (int)(mtu / (uint32_t)a.m_tiles);
                const int q0 = (int)((mtu - (uint32_t)u * (uint32_t)a.m_tiles) * CG) * (WBM * a.mrep);      // CTA r of a pair: + r * 128 * mrep
                const int j0 = (int)nt * BN;                                                                  // CTA r of a pair: + r * BNL
                const int ci0 = (int)sp * a.ci_per_split, ci1 = min(a.cpb * a.nplanes, ((int)sp + 1) * a.ci_per_split);
                const uint32_t f_last_tile = (w + ncl >= nwork) ? REC_LAST : 0u;
                for (int ci = ci0; ci < ci1; ++ci, ++ia) {
                    const int cc = ci / a.nplanes, pl = ci - cc * a.nplanes;
                    const int c0 = cc * 64;
                    const uint32_t nk16 = (uint32_t)(min(64, C - c0) >> 4);
                    const int ab = ia & 1;
                    const uint32_t win = base + ab * win_bytes;
                    if (ia >= 2) mbar_wait(aempty(ab), (uint32_t)(((ia >> 1) - 1) & 1));
                    if constexpr (CG == 2) {
                        mbar_expect_tx(afull(ab), 2 * a_bytes);
                        for (int r = 0; r < 2; ++r)
                            for (int bx = 0; bx < a.nboxes; ++bx)
                                tma_load_2d_to(win + bx * a.rows_per_box * 128, &a.tm_x, c0,
                                               pl * (int)a.Qp + q0 + r * (WBM * a.mrep) + bx * a.rows_per_box, afull(ab), (uint16_t)(1u << r));
                    } else {
                        mbar_expect_tx(afull(ab), a_bytes);
                        for (int bx = 0; bx < a.nboxes; ++bx)
                            tma_load_2d(win + bx * a.rows_per_box * 128, &a.tm_x, c0, pl * (int)a.Qp + q0 + bx * a.rows_per_box, afull(ab));
                    }
                    const int t_beg = a.pbeg[u][pl], t_end = a.pbeg[u][pl + 1];
                    for (int t0 = t_beg; t0 < t_end; t0 += a.tps) {       // one ring stage = up to tps taps
                        const int nt_s = min(a.tps, t_end - t0);
                        const uint32_t s = bs;
                        uint32_t meta = (uint32_t)nt_s | (nk16 << 3);
                        if (t0 == t_beg) meta |= REC_CHUNK_FIRST | (ci == ci0 ? REC_TILE_FIRST : 0u);
                        if (t0 + a.tps >= t_end) meta |= REC_CHUNK_LAST | (ci == ci1 - 1 ? (REC_TILE_LAST | f_last_tile) : 0u);
                        uint32_t alo[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            alo[k] = a_lo | ((win + (uint32_t)a.shift[min(t0 + k, t_end - 1)] * 128u) >> 4);
                        const uint32_t stage_base = wbase + s * WB_STRIDE;
                        mbar_wait(bempty(s), bph);
                        if (++bs == (uint32_t)kBStages) { bs = 0; bph ^= 1u; }
                        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(recs + 32u * s), "r"(alo[0]), "r"(alo[1]), "r"(alo[2]),
                                     "r"(alo[3]) : "memory");
                        asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(recs + 32u * s + 16u), "r"(meta), "r"(b_lo | (stage_base >> 4)) : "memory");
                        if (a.resident && w != cid) {       // weights already in place (first work item): publish the record only
                            mbar_arrive(bfull(s));
                        } else
                        if constexpr (CG == 2) {
                            mbar_expect_tx(bfull(s), 2 * b_bytes * (uint32_t)nt_s);     // release: the record is visible to the waiter
                            for (int k = 0; k < nt_s; ++k) {
                                const int t = t0 + k;
                                const uint32_t dst = stage_base + (uint32_t)k * TAP_STRIDE;
                                for (int r = 0; r < 2; ++r) {
                                    if (!a.b_mn_major) {
                                        tma_load_2d_to(dst, &a.tm_w, c0, (int)a.wtap[t] * J + j0 + r * BNL, bfull(s), (uint16_t)(1u << r));
                                    } else {
                                        for (int jb = 0; jb * 64 < BNL; ++jb)
                                            tma_load_2d_to(dst + jb * 8192, &a.tm_w, j0 + r * BNL + jb * 64, (int)a.wtap[t] * C + c0, bfull(s),
                                                           (uint16_t)(1u << r));
                                    }
                                }
                            }
                        } else {
                        mbar_expect_tx(bfull(s), b_bytes * (uint32_t)nt_s);         // release: the record is visible to the waiter
                        for (int k = 0; k < nt_s; ++k) {
                            const int t = t0 + k;
                            const uint32_t dst = stage_base + (uint32_t)k * TAP_STRIDE;
                            if (!a.b_mn_major) {
                                tma_load_2d(dst, &a.tm_w, c0, (int)a.wtap[t] * J + j0, bfull(s));
                            } else {
                                for (int jb = 0; jb * 64 < BN; ++jb)
                                    tma_load_2d(dst + jb * 8192, &a.tm_w, j0 + jb * 64, (int)a.wtap[t] * C + c0, bfull(s));
                            }
                        }
                        }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================================ MMA issuer ================================
        // the whole warp walks the records (keeps every value warp-uniform); one elected lane issues
        if (cid < a.nwork && rank == 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a.b_mn_major << 16) | ((uint32_t)(BN >> 3) << 17) |
                                   ((uint32_t)((WBM * CG) >> 4) << 24);
            // descriptors: the high words are loop invariant, the low words come from the stage record; per-k16 advance is an
            // add on the low word (+32 B along K for K-major tiles, +2048 B for MN-major weight tiles)
            const uint32_t a_hi = (uint32_t)(make_desc(0, 16, 1024, 0) >> 32);
            const uint32_t b_hi = (uint32_t)((a.b_mn_major ? make_desc(0, 8192, 1024, 0) : make_desc(0, 16, 1024, 0)) >> 32);
            const uint32_t b_k16 = a.b_mn_major ? (2048u >> 4) : (32u >> 4);
            const uint32_t b_tap = TAP_STRIDE >> 4;
            int ia = 0, li = 0;
            uint32_t bs = 0, bph = 0;       // weight ring: stage, parity of the "full" wait
            uint32_t d_tmem = tmem_base, accumulate = 0;
#ifdef P2P_WINDOW_TIMING
            // running checkpoints: every cycle of the warp between loop entry and exit lands in exactly one bucket
            long long c_wa = 0, c_wb = 0, c_is = 0, c_wt = 0, c_n = 0, c_sh = 0;
            const long long c_t0 = clock64();
            long long c_last = c_t0;
            long long g_loop0;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_loop0));
#define CP(dst) { const long long now_ = clock64(); dst += now_ - c_last; c_last = now_; }
#else
#define CP(dst)
#endif
            const uint32_t slot_mask = (uint32_t)a.acc_slots - 1u, slot_shift = (uint32_t)a.acc_slots >> 1;      // 1 or 2 slots
            const uint32_t half_cols = (uint32_t)a.half_cols;
            for (;;) {
                const uint32_t s = bs;
                CP(c_sh);
                mbar_wait(bfull(s), bph);
                CP(c_wb);
                if (++bs == (uint32_t)kBStages) { bs = 0; bph ^= 1u; }
                uint32_t al0, al1, al2, al3, meta, bl;
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(al0), "=r"(al1), "=r"(al2), "=r"(al3) : "r"(recs + 32u * s) : "memory");
                asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(meta), "=r"(bl) : "r"(recs + 32u * s + 16u) : "memory");
                if (meta & REC_TILE_FIRST) {
                    const uint32_t acc = (uint32_t)li & slot_mask, v = (uint32_t)li >> slot_shift;
                    if (v >= 1) mbar_wait(tempty(acc), (v - 1) & 1u);
                    d_tmem = tmem_base + acc * 256;
                    accumulate = 0;
                    CP(c_wt);
                }
                if (meta & REC_CHUNK_FIRST) {
                    mbar_wait(afull(ia & 1), (uint32_t)((ia >> 1) & 1));
                    CP(c_wa);
                }
                tc_fence_after();
                if (elect_one()) {
                    if (a.mrep == 2) issue_stage<2, CG>(d_tmem, half_cols, a_hi, b_hi, al0, al1, al2, al3, bl, b_k16, b_tap, idesc, meta, accumulate);
                    else issue_stage<1, CG>(d_tmem, half_cols, a_hi, b_hi, al0, al1, al2, al3, bl, b_k16, b_tap, idesc, meta, accumulate);
                    commit_cg<CG>(bempty(s), 1);                                            // the leader refills
                    if (meta & REC_CHUNK_LAST) commit_cg<CG>(aempty(ia & 1), 1);
                    if (meta & REC_TILE_LAST) commit_cg<CG>(tfull((uint32_t)li & slot_mask), 3);    // both CTAs drain their rows
                }
                __syncwarp();
                accumulate = 1;
                if (meta & REC_CHUNK_LAST) ++ia;
                if (meta & REC_TILE_LAST) ++li;
#ifdef P2P_WINDOW_TIMING
                ++c_n;
#endif
                CP(c_is);
                if (meta & REC_LAST) break;
            }
#ifdef P2P_WINDOW_TIMING
            if (a.dbg && lane == 0) {
                long long* o = a.dbg + 8ll * blockIdx.x;
                long long g_loop1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_loop1));
                o[0] = c_wa; o[1] = c_wb; o[2] = c_is; o[3] = c_wt; o[4] = c_n; o[5] = clock64() - c_t0; o[6] = li;
                o[7] = ((g_loop0 - g_entry) << 32) | (g_loop1 - g_loop0);       // ns: entry -> loop start | loop duration
                long long* o2 = a.dbg + 9ll * gridDim.x + 8ll * blockIdx.x;
                o2[0] = 0; o2[1] = 0; o2[2] = c_sh; o2[3] = 0; o2[4] = 0;
            }
#endif
        }
        __syncwarp();
    } else {
        // ================================ epilogue ================================
        // eight warps: two per TMEM lane quarter.  With two row halves (mrep = 2) group g takes half g; with one, the groups
        // take alternate 32-column chunks.
        const int quarter = warp & 3, row = quarter * 32 + lane;
        const int ewarp = warp - 2, g = ewarp >> 2;
        const int et = ewarp * 32 + lane;                 // 0..255 over the epilogue warps
        const int half_lo = a.mrep == 2 ? g : 0, half_hi = a.mrep == 2 ? g + 1 : 1;
        const int c2_lo = a.mrep == 2 ? 0 : 32 * g, c2_step = a.mrep == 2 ? 32 : 64;
        const uint32_t plane = (uint32_t)(a.Hp * a.Wp);     // the host keeps every row index below 2^31: 32-bit divisions throughout
        float* bias_s = reinterpret_cast<float*>(smem_raw + (bias_u32 - smem_u32(smem_raw)));
        int li = 0;
        const uint32_t nwork = (uint32_t)a.nwork, n_tiles = (uint32_t)a.n_tiles, nsplit = (uint32_t)a.nsplit;
        for (uint32_t w = cid; w < nwork; w += ncl, ++li) {
            const int nt = (int)(w % n_tiles);
            const uint32_t wr = w / n_tiles;
            const int sp = (int)(wr % nsplit);
            const uint32_t mtu = wr / nsplit;
            const int u = (int)(mtu / (uint32_t)a.m_tiles);
            const uint32_t mt = (mtu - (uint32_t)u * (uint32_t)a.m_tiles) * CG + rank;
            const int acc = li & (a.acc_slots - 1), v = li >> (a.acc_slots >> 1);
            const int c_begin = nt * BN, c_count = min(BN, J - c_begin);
            float* bs_ = bias_s + (li & 1) * 256;
            for (int c = et; c < BN; c += 256) bs_[c] = (a.bias && sp == 0 && c < c_count) ? __ldg(a.bias + c_begin + c) : 0.f;
            for (int half = half_lo; half < half_hi; ++half) {
            const uint32_t q = (mt * (uint32_t)a.mrep + (uint32_t)half) * WBM + row;
            const int n = (int)(q / plane);
            const int rem = (int)(q - (uint32_t)n * plane);
            const int y = rem / a.Wp, x = rem - y * a.Wp;
            const bool rok = (long long)q < a.Q && y < a.gy[u] && x < a.gx[u];
            const long long obase = (((long long)n * a.OHf + (y * a.oys + a.oyo[u])) * a.OWf + (x * a.oxs + a.oxo[u])) * J + c_begin;
            if (half == half_lo) {
                asm volatile("bar.sync 1, 256;" ::: "memory");      // bias staged (double-buffered by tile parity)
                mbar_wait(tfull(acc), (uint32_t)(v & 1));
                tc_fence_after();
            }
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * 256 + half * a.half_cols;
            if (a.fast_epi) {
                const int esz = a.out_bf16 ? 2 : 4;
                const unsigned okmask = __ballot_sync(0xffffffffu, rok);
                const long long row_off = obase * esz;
                const uint32_t stage = epi_stage + (uint32_t)ewarp * 4096u;
                const int mode = a.nsplit > 1 ? EPI_RED : EPI_ST;
                // InstanceNorm statistics: the warp's valid rows normally lie in one image (n_lo == n_hi): column sums through
                // the staged tile; a warp that straddles two images adds row by row
                int n_lo = 0, n_hi = 0;
                if (a.stats && okmask) {
                    n_lo = __reduce_min_sync(0xffffffffu, rok ? n : 0x7fffffff);
                    n_hi = __reduce_max_sync(0xffffffffu, rok ? n : -1);
                }
                for (int c2 = c2_lo; c2 < BN; c2 += c2_step) {
                    uint32_t r2[32];
                    const bool two = c2 + 16 < BN;
                    tmem_ld16(taddr + c2, *reinterpret_cast<uint32_t(*)[16]>(&r2[0]));
                    if (two) tmem_ld16(taddr + c2 + 16, *reinterpret_cast<uint32_t(*)[16]>(&r2[16]));
                    tmem_ld_wait();
                    if (c2 >= c_count) break;
                    const int ncol = two ? 32 : 16;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        if (4 * i < ncol) {
                            const float4 bv = *reinterpret_cast<const float4*>(bs_ + c2 + 4 * i);
                            r2[4 * i] = __float_as_uint(__uint_as_float(r2[4 * i]) + bv.x);
                            r2[4 * i + 1] = __float_as_uint(__uint_as_float(r2[4 * i + 1]) + bv.y);
                            r2[4 * i + 2] = __float_as_uint(__uint_as_float(r2[4 * i + 2]) + bv.z);
                            r2[4 * i + 3] = __float_as_uint(__uint_as_float(r2[4 * i + 3]) + bv.w);
                        }
                    }
                    int P = ncol / 4;
                    if (a.out_bf16) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(r2[2 * i]), __uint_as_float(r2[2 * i + 1]));
                            r2[i] = *reinterpret_cast<const uint32_t*>(&h);
                        }
                        P = ncol / 8;
                    }
                    const int valid = min(ncol, c_count - c2) * esz;
                    float* st = nullptr;
                    if (a.stats && okmask) {
                        if (n_lo == n_hi) {
                            st = a.stats + ((long long)n_lo * J + c_begin + c2) * 2;
                        } else if (rok) {           // rare: per-row adds (r2 holds acc + bias in fp32 here: stats only with fp32 output)
                            float* t = a.stats + ((long long)n * J + c_begin + c2) * 2;
#pragma unroll
                            for (int i = 0; i < 32; ++i) {          // compile-time register indices: r2 must not move to local memory
                                if (i < ncol && c2 + i < c_count) {
                                    const float d = __uint_as_float(r2[i]) - bs_[c2 + i];
                                    atomicAdd(t + 2 * i, d);
                                    atomicAdd(t + 2 * i + 1, d * d);
                                }
                            }
                        }
                    }
                    stage_store(P, stage, lane, r2, row_off + (long long)c2 * esz, okmask, (char*)a.out, valid, mode, esz, st, bs_ + c2);
                }
            } else
            for (int c2 = c2_lo; c2 < BN; c2 += c2_step) {
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
                    for (int i = 0; i < 4; ++i) {
                        const float4 bv = *reinterpret_cast<const float4*>(bs_ + c + 4 * i);
                        f[4 * i] = __uint_as_float(r2[hh][4 * i]) + bv.x;
                        f[4 * i + 1] = __uint_as_float(r2[hh][4 * i + 1]) + bv.y;
                        f[4 * i + 2] = __uint_as_float(r2[hh][4 * i + 2]) + bv.z;
                        f[4 * i + 3] = __uint_as_float(r2[hh][4 * i + 3]) + bv.w;
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
                    } else if (a.nsplit > 1) {
                        float* o = (float*)a.out + obase + c;
                        if (fullv && (((uintptr_t)o) & 15) == 0) {
#pragma unroll
                            for (int i = 0; i < 4; ++i) red_add_v4(o + 4 * i, f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; ++i)
                                if (c + i < c_count) atomicAdd(o + i, f[i]);
                        }
                    } else {
                        float* o = (float*)a.out + obase + c;
                        if (fullv && (((uintptr_t)o) & 15) == 0) {
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
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {            // one arrival per epilogue warp, on the leader's barrier (its MMA warp reuses the slot)
                if constexpr (CG == 2) mbar_arrive_cluster(lead(tempty(acc)));
                else mbar_arrive(tempty(acc));
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if constexpr (CG == 2) cluster_sync_all();          // no CTA of the pair leaves (or frees TMEM) while the other still signals it
#ifdef P2P_WINDOW_TIMING
    if (a.dbg && tid == 0) {
        long long g_exit;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_exit));
        a.dbg[8ll * gridDim.x + blockIdx.x] = g_exit - g_entry;             // ns: whole CTA lifetime
    }
#endif
    if (warp == 1) {
        tc_fence_after();
        if constexpr (CG == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// zero / reflection padding written once, optionally split into stride x stride parity planes:
//   planes[(a*s + b)][n, y, x, :] = x[n, map(s*y + a - pt), map(s*x + b - pl), :]      (s = 1: one padded tensor)
template <typename T>
__global__ void pad_nhwc_kernel(const T* __restrict__ x, T* __restrict__ xp, int N, int H, int W, int C8, int pt, int pl,
                                int Hp, int Wp, int reflect, int s, long long plane_rows) {
    // C8 = 16-byte vectors per pixel; every plane occupies plane_rows >= N*Hp*Wp pixel rows (the tail is zero-filled)
    const long long plane = plane_rows * C8, used = (long long)N * Hp * Wp * C8;
    const long long total = plane * s * s;
    pdl_trigger();
    pdl_wait();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int pq = (int)(i / plane);
        long long p = i - pq * plane;
        if (p >= used) { reinterpret_cast<uint4*>(xp)[i] = make_uint4(0, 0, 0, 0); continue; }
        const int c = (int)(p % C8);
        p /= C8;
        const int xx = (int)(p % Wp);
        p /= Wp;
        const int yy = (int)(p % Hp);
        const int n = (int)(p / Hp);
        int sy = s * yy + pq / s - pt, sx = s * xx + pq % s - pl;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (reflect) {
            sy = sy < 0 ? -sy : (sy >= H ? 2 * (H - 1) - sy : sy);
            sx = sx < 0 ? -sx : (sx >= W ? 2 * (W - 1) - sx : sx);
        }
        if (sy >= 0 && sy < H && sx >= 0 && sx < W)
            v = reinterpret_cast<const uint4*>(x)[(((long long)n * H + sy) * W + sx) * C8 + c];
        reinterpret_cast<uint4*>(xp)[i] = v;
    }
}


// ====================================================================================================================
// Swapped form for layers with J <= 128 output channels.  Measured (scratch/mma_rate.cu): one tcgen05.mma of M = 128, K = 16
// costs >= 105 cycles whatever N <= 192 is, so with the pixels on the M side a 96-channel layer can reach at most 48/105 of
// the tensor peak.  Here the WEIGHT tile is the A operand (M = 128 output channels, rows >= J are don't-care) and 256 padded
// pixels of the shifted window are the B operand (N = 256, 128 math cycles per instruction): D^T[j][pixel].  The epilogue
// transposes 32 channels x 32 pixels through shared memory and stores 128 contiguous bytes per pixel and warp.
// ====================================================================================================================
constexpr int SBM = 256;                            // pixels per tile = N of the MMA
constexpr int kSwapRowsMax = 400;                   // 256 + halo rows, 50 KB per window buffer
constexpr int SWIN_BYTES = kSwapRowsMax * 128;
constexpr int SW_TAP_BYTES = 16384;                 // [128 j][64 c] (K-major) or 2 x [64 c][64 j] (MN-major) bf16
constexpr int kSwapStages = 6;
constexpr int SWSMEM = 2 * SWIN_BYTES + kSwapStages * SW_TAP_BYTES + 1024 + 512 + 2 * SBM * 8 + EPI_STAGE_BYTES;
static_assert(SWSMEM <= 232448, "swapped window kernel shared memory");

__global__ void __launch_bounds__(kWThreads, 1) conv_window_swap_kernel(const __grid_constant__ WindowArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t wbase = base + 2 * SWIN_BYTES;
    const uint32_t bars = wbase + kSwapStages * SW_TAP_BYTES;
    auto afull = [&](int i) { return bars + 8u * i; };
    auto aempty = [&](int i) { return bars + 8u * (2 + i); };
    auto bfull = [&](int s) { return bars + 8u * (4 + s); };
    auto bempty = [&](int s) { return bars + 8u * (4 + kSwapStages + s); };
    auto tfull = [&](int i) { return bars + 8u * (4 + 2 * kSwapStages + i); };
    auto tempty = [&](int i) { return bars + 8u * (6 + 2 * kSwapStages + i); };
    const uint32_t tmem_slot = bars + 8u * (8 + 2 * kSwapStages);
    const uint32_t table_u32 = bars + 512;                  // long long [2][256]: output element offset of each pixel, -1 = dropped
    const uint32_t epi_stage = table_u32 + 2 * SBM * 8;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int C = a.C, J = a.J;
    const uint32_t a_bytes = (uint32_t)a.nboxes * a.rows_per_box * 128u;

    if (warp == 1) {
        if (lane == 0) {
            for (int i = 0; i < 2; ++i) { mbar_init(afull(i), 1); mbar_init(aempty(i), 1); mbar_init(tfull(i), 1); mbar_init(tempty(i), 128); }
            for (int s = 0; s < kSwapStages; ++s) { mbar_init(bfull(s), 1); mbar_init(bempty(s), 1); }
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

    if (warp == 0) {
        // ================================ TMA producer ================================
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_x) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_w) : "memory");
            int ia = 0;
            uint32_t bs = 0, bph = 1;
            for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x) {
                const int jt = (int)(w % a.n_tiles);
                const long long mtu = w / a.n_tiles;
                const int u = (int)(mtu / a.m_tiles);
                const long long q0 = (mtu - (long long)u * a.m_tiles) * SBM;
                const int j0 = jt * 128;
                for (int ci = 0; ci < a.cpb * a.nplanes; ++ci, ++ia) {
                    const int cc = ci / a.nplanes, pl = ci - cc * a.nplanes;
                    const int c0 = cc * 64;
                    const int ab = ia & 1;
                    if (ia >= 2) mbar_wait(aempty(ab), (uint32_t)(((ia >> 1) - 1) & 1));
                    mbar_expect_tx(afull(ab), a_bytes);
                    for (int bx = 0; bx < a.nboxes; ++bx)
                        tma_load_2d(base + ab * SWIN_BYTES + bx * a.rows_per_box * 128, &a.tm_x, c0,
                                    (int)((long long)pl * a.Qp + q0 + (long long)bx * a.rows_per_box), afull(ab));
                    const int t_end = a.pbeg[u][pl + 1];
                    for (int t = a.pbeg[u][pl]; t < t_end; ++t) {
                        const uint32_t s = bs;
                        mbar_wait(bempty(s), bph);
                        if (++bs == (uint32_t)kSwapStages) { bs = 0; bph ^= 1u; }
                        mbar_expect_tx(bfull(s), SW_TAP_BYTES);
                        const uint32_t dst = wbase + s * SW_TAP_BYTES;
                        if (!a.b_mn_major) {
                            tma_load_2d(dst, &a.tm_w, c0, (int)a.wtap[t] * J + j0, bfull(s));
                        } else {
                            tma_load_2d(dst, &a.tm_w, j0, (int)a.wtap[t] * C + c0, bfull(s));
                            tma_load_2d(dst + 8192, &a.tm_w, j0 + 64, (int)a.wtap[t] * C + c0, bfull(s));
                        }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================================ MMA issuer ================================
        {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a.b_mn_major << 15) | ((uint32_t)(SBM >> 3) << 17) |
                                   ((uint32_t)(128 >> 4) << 24);
            const uint64_t w_hi = a.b_mn_major ? make_desc(0, 8192, 1024, 0) : make_desc(0, 16, 1024, 0);     // weights: the A operand
            const uint32_t w_k16 = a.b_mn_major ? (2048u >> 4) : (32u >> 4);
            const uint64_t x_hi = make_desc(0, 16, 1024, 0);                                                  // window rows: the B operand
            int ia = 0, li = 0;
            uint32_t bs = 0, bph = 0;
            for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x, ++li) {
                const int u = (int)((w / a.n_tiles) / a.m_tiles);
                const int acc = li & 1, v = li >> 1;
                if (v >= 1) mbar_wait(tempty(acc), (uint32_t)((v - 1) & 1));
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * 256;
                uint32_t accumulate = 0;
                for (int ci = 0; ci < a.cpb * a.nplanes; ++ci, ++ia) {
                    const int cc = ci / a.nplanes, pl = ci - cc * a.nplanes;
                    const int nk16 = min(64, C - cc * 64) >> 4;
                    const int ab = ia & 1;
                    mbar_wait(afull(ab), (uint32_t)((ia >> 1) & 1));
                    tc_fence_after();
                    const uint32_t win = base + ab * SWIN_BYTES;
                    const int t_end = a.pbeg[u][pl + 1];
                    for (int t = a.pbeg[u][pl]; t < t_end; ++t) {
                        const uint32_t s = bs;
                        const uint64_t xd = x_hi | (uint64_t)((win + (uint32_t)a.shift[t] * 128u) >> 4);
                        const uint64_t wd = w_hi | (uint64_t)((wbase + s * SW_TAP_BYTES) >> 4);
                        mbar_wait(bfull(s), bph);
                        if (++bs == (uint32_t)kSwapStages) { bs = 0; bph ^= 1u; }
                        tc_fence_after();
                        if (elect_one()) {
                            umma_bf16(d_tmem, wd, xd, idesc, accumulate);
                            if (nk16 > 1) umma_bf16(d_tmem, wd + w_k16, xd + 2, idesc, 1);
                            if (nk16 > 2) umma_bf16(d_tmem, wd + 2 * w_k16, xd + 4, idesc, 1);
                            if (nk16 > 3) umma_bf16(d_tmem, wd + 3 * w_k16, xd + 6, idesc, 1);
                            umma_commit(bempty(s));
                        }
                        __syncwarp();
                        accumulate = 1;
                    }
                    if (elect_one()) umma_commit(aempty(ab));
                    __syncwarp();
                }
                if (elect_one()) umma_commit(tfull(acc));
                __syncwarp();
            }
        }
        __syncwarp();
    } else {
        // ================================ epilogue ================================
        const int quarter = warp & 3;                       // TMEM lanes 32*quarter .. +31 = output channels of this warp
        const int et = (warp - 2) * 32 + lane;
        const long long plane = (long long)a.Hp * a.Wp;
        long long* table = reinterpret_cast<long long*>(smem_raw + (table_u32 - smem_u32(smem_raw)));
        const uint32_t stage = epi_stage + (uint32_t)(warp - 2) * 4096u;
        const int esz = a.out_bf16 ? 2 : 4;
        int li = 0;
        for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x, ++li) {
            const int jt = (int)(w % a.n_tiles);
            const long long mtu = w / a.n_tiles;
            const int u = (int)(mtu / a.m_tiles);
            const long long mt = mtu - (long long)u * a.m_tiles;
            const int acc = li & 1, v = li >> 1;
            long long* tb = table + acc * SBM;
            for (int i = et; i < SBM; i += 128) {           // where each of the tile's 256 padded pixels lands (or -1)
                const long long q = mt * SBM + i;
                long long off = -1;
                if (q < a.Q) {
                    const int n = (int)(q / plane);
                    const int rem = (int)(q - (long long)n * plane);
                    const int y = rem / a.Wp, x = rem - y * a.Wp;
                    if (y < a.gy[u] && x < a.gx[u])
                        off = (((long long)n * a.OHf + (y * a.oys + a.oyo[u])) * a.OWf + (x * a.oxs + a.oxo[u])) * J;
                }
                tb[i] = off;
            }
            const int jch = jt * 128 + quarter * 32 + lane;                 // this lane's output channel
            const float bj = (a.bias && jch < J) ? __ldg(a.bias + jch) : 0.f;
            asm volatile("bar.sync 1, 128;" ::: "memory");
            mbar_wait(tfull(acc), (uint32_t)(v & 1));
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * 256;
            const int jc = jt * 128 + quarter * 32 + (lane & 7) * 4;        // first of the 4 channels this lane stores
            for (int c2 = 0; c2 < SBM; c2 += 32) {
                uint32_t r[32];
                tmem_ld16(taddr + c2, *reinterpret_cast<uint32_t(*)[16]>(&r[0]));
                tmem_ld16(taddr + c2 + 16, *reinterpret_cast<uint32_t(*)[16]>(&r[16]));
                tmem_ld_wait();
                __syncwarp();                       // the previous chunk's reads of the tile are done
#pragma unroll
                for (int i = 0; i < 32; ++i)        // tile[pixel i][channel lane]
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(stage + (uint32_t)(i * 32 + lane) * 4u), "f"(__uint_as_float(r[i]) + bj) : "memory");
                __syncwarp();
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int pix = (lane >> 3) + 4 * k;
                    float4 f;
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f.x), "=f"(f.y), "=f"(f.z), "=f"(f.w)
                                 : "r"(stage + (uint32_t)(pix * 32 + (lane & 7) * 4) * 4u));
                    const long long off = tb[c2 + pix];
                    if (off < 0 || jc >= J) continue;
                    char* dst = (char*)a.out + (off + jc) * esz;
                    if (jc + 4 <= J) {
                        if (a.out_bf16) {
                            const __nv_bfloat162 h0 = __floats2bfloat162_rn(f.x, f.y), h1 = __floats2bfloat162_rn(f.z, f.w);
                            uint2 t2;
                            t2.x = *reinterpret_cast<const uint32_t*>(&h0);
                            t2.y = *reinterpret_cast<const uint32_t*>(&h1);
                            *reinterpret_cast<uint2*>(dst) = t2;
                        } else {
                            *reinterpret_cast<float4*>(dst) = f;
                        }
                    } else {
                        const float fv[4] = {f.x, f.y, f.z, f.w};
                        for (int i = 0; i < 4 && jc + i < J; ++i) {
                            if (a.out_bf16) reinterpret_cast<__nv_bfloat16*>(dst)[i] = __float2bfloat16_rn(fv[i]);
                            else reinterpret_cast<float*>(dst)[i] = fv[i];
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
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// ---- weight gradients in window form -------------------------------------------------------------------------------------
// dW[wtap_t][j][c] += sum_q A[a_plane_t][q][j] * B[b_plane_t][q + shift_t][c]: the reduction index q runs over the padded pixel
// grid (A = dout laid out at the same pitch, zero where no output exists).  Both operands are MN-major (rows = q): A is one
// TMA tile per 64-row block, B one shifted WINDOW shared by all taps of a group; every tap owns BN TMEM columns.
// One tcgen05.mma costs >= 105 cycles whatever its N (scratch/mma_rate.cu), so the channel tile BN is made as wide as possible
// (up to 256: the window is then BN/64 chunk boxes, LBO = bytes of one chunk) and the taps are cut into SETS of <= 512/BN taps;
// a work item = (128 output channels, BN input channels, one tap set, one slice of the pixel blocks).
constexpr int kGStagesMax = 8;
constexpr int GA_BYTES = 2 * 8192;                  // [64 q][128 j]
constexpr int GRING_BYTES = 196608;                 // 192 KB cut into stages of A tile + B window (all chunks)
constexpr int GSMEM = GRING_BYTES + 1024 + 256 + EPI_STAGE_BYTES;
constexpr int kMaxGroups = 8;

struct WgradWinArgs {
    CUtensorMap tm_a;        // dout planes [rows][J]: box [64 j][64 rows]
    CUtensorMap tm_b;        // input planes [rows][C]: box [64 c][b_rows_per_box]
    float* dw;
    long long Qs;            // rows per plane (multiple of 64, zero tail)
    long long w_tap;
    int J, C, BN, j_tiles, c_tiles, nsplit, kb_per_split, total_kb;
    int ntaps, ngroups, atomic;
    int nsets, set_taps;     // taps [set * set_taps, min(ntaps, (set + 1) * set_taps)) share the TMEM columns of one work item
    int nchunks, chunk_bytes;   // 64-channel boxes of the B window and the bytes of one
    int b_nboxes, b_rows_per_box;
    int stages, stage_bytes;
    int fast_epi;
    int gbeg[kMaxGroups + 1];
    short ga[kMaxGroups], gb[kMaxGroups];
    short shift[32];
    unsigned char wtap[32];
    long long nwork;
};

__global__ void __launch_bounds__(kWThreads, 1) conv_wgrad_window_kernel(const __grid_constant__ WgradWinArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = base + GRING_BYTES;
    const int kGStages = a.stages;
    const uint32_t GSTAGE = (uint32_t)a.stage_bytes;
    auto full = [&](int s) { return bars + 8u * s; };
    auto empty = [&](int s) { return bars + 8u * (kGStagesMax + s); };
    const uint32_t tfull = bars + 8u * (2 * kGStagesMax), tempty = tfull + 8, tmem_slot = tfull + 16;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int BN = a.BN;
    const uint32_t b_bytes = (uint32_t)a.b_nboxes * a.b_rows_per_box * 128u;

    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < kGStages; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
            mbar_init(tfull, 1);
            mbar_init(tempty, 128);
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

    if (warp == 0) {
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_a) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_b) : "memory");
            uint32_t rs = 0, rph = 1;       // ring stage, parity of the "empty" wait (first lap passes at once)
            for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x) {
                const int sp = (int)(w % a.nsplit);
                long long r = w / a.nsplit;
                const int set = (int)(r % a.nsets);
                r /= a.nsets;
                const int ct = (int)(r % a.c_tiles), jt = (int)(r / a.c_tiles);
                const int t_lo = set * a.set_taps, t_hi = min(a.ntaps, t_lo + a.set_taps);
                const int nch = min(a.nchunks, (a.C - ct * BN + 63) / 64);      // chunks of this channel tile that exist
                const int kb0 = sp * a.kb_per_split, kb1 = min(a.total_kb, kb0 + a.kb_per_split);
                for (int kb = kb0; kb < kb1; ++kb)
                    for (int g = 0; g < a.ngroups; ++g) {
                        if (max(a.gbeg[g], t_lo) >= min(a.gbeg[g + 1], t_hi)) continue;     // no tap of this set in the group
                        const uint32_t s = rs;
                        mbar_wait(empty(s), rph);
                        if (++rs == (uint32_t)kGStages) { rs = 0; rph ^= 1u; }
                        mbar_expect_tx(full(s), GA_BYTES + b_bytes * (uint32_t)nch);
                        const uint32_t sa = base + s * GSTAGE, sb = sa + GA_BYTES;
                        const long long qa = (long long)a.ga[g] * a.Qs + (long long)kb * 64;
                        const long long qb = (long long)a.gb[g] * a.Qs + (long long)kb * 64;
                        tma_load_2d(sa, &a.tm_a, jt * 128, (int)qa, full(s));
                        tma_load_2d(sa + 8192, &a.tm_a, jt * 128 + 64, (int)qa, full(s));
                        for (int ch = 0; ch < nch; ++ch)
                            for (int bx = 0; bx < a.b_nboxes; ++bx)
                                tma_load_2d(sb + ch * a.chunk_bytes + bx * a.b_rows_per_box * 128, &a.tm_b, ct * BN + ch * 64,
                                            (int)(qb + (long long)bx * a.b_rows_per_box), full(s));
                    }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        {   // whole warp walks the schedule, one elected lane issues
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(BN >> 3) << 17) |
                                   ((uint32_t)(WBM >> 4) << 24);
            const uint64_t d_hi = make_desc(0, 8192, 1024, 0);      // both operands MN-major; +2048 B per k16
            const uint64_t b_hi = make_desc(0, (uint32_t)a.chunk_bytes, 1024, 0);      // 64-channel blocks of the window: one chunk apart
            int li = 0;
            uint32_t rs = 0, rph = 0;
            for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x, ++li) {
                const int sp = (int)(w % a.nsplit);
                const int set = (int)((w / a.nsplit) % a.nsets);
                const int t_lo = set * a.set_taps, t_hi = min(a.ntaps, t_lo + a.set_taps);
                const int kb0 = sp * a.kb_per_split, kb1 = min(a.total_kb, kb0 + a.kb_per_split);
                if (li >= 1) mbar_wait(tempty, (uint32_t)((li - 1) & 1));
                tc_fence_after();
                for (int kb = kb0; kb < kb1; ++kb)
                    for (int g = 0; g < a.ngroups; ++g) {
                        const int t_beg = max(a.gbeg[g], t_lo), t_end = min(a.gbeg[g + 1], t_hi);
                        if (t_beg >= t_end) continue;
                        const uint32_t s = rs;
                        const uint32_t sa = base + s * GSTAGE, sb = sa + GA_BYTES;
                        const uint64_t ad = d_hi | (uint64_t)(sa >> 4);
                        const uint32_t first = kb > kb0 ? 1u : 0u;
                        mbar_wait(full(s), rph);
                        if (++rs == (uint32_t)kGStages) { rs = 0; rph ^= 1u; }
                        tc_fence_after();
                        if (elect_one()) {
                            for (int t = t_beg; t < t_end; ++t) {
                                const uint32_t d_tmem = tmem_base + (t - t_lo) * BN;
                                const uint64_t bd = b_hi | (uint64_t)((sb + (uint32_t)a.shift[t] * 128u) >> 4);
                                umma_bf16(d_tmem, ad, bd, idesc, first);
                                umma_bf16(d_tmem, ad + 128, bd + 128, idesc, 1);
                                umma_bf16(d_tmem, ad + 256, bd + 256, idesc, 1);
                                umma_bf16(d_tmem, ad + 384, bd + 384, idesc, 1);
                            }
                            umma_commit(empty(s));
                        }
                        __syncwarp();
                    }
                if (elect_one()) umma_commit(tfull);
                __syncwarp();
            }
        }
        __syncwarp();
    } else {
        const int quarter = warp & 3, row = quarter * 32 + lane;
        int li = 0;
        for (long long w = blockIdx.x; w < a.nwork; w += gridDim.x, ++li) {
            long long r = w / a.nsplit;
            const int set = (int)(r % a.nsets);
            r /= a.nsets;
            const int ct = (int)(r % a.c_tiles), jt = (int)(r / a.c_tiles);
            const int t_lo = set * a.set_taps, t_hi = min(a.ntaps, t_lo + a.set_taps);
            const int j = jt * 128 + row, c_begin = ct * BN, c_count = min(BN, a.C - c_begin);
            const bool rok = j < a.J;
            mbar_wait(tfull, (uint32_t)(li & 1));
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
            if (a.fast_epi) {
                const unsigned okmask = __ballot_sync(0xffffffffu, rok);
                const uint32_t stage = bars + 256 + (uint32_t)(warp - 2) * 4096u;
                const int mode = a.atomic ? EPI_RED : EPI_ST;
                for (int t = t_lo; t < t_hi; ++t) {
                    const long long row_off = ((long long)a.wtap[t] * a.w_tap + (long long)j * a.C + c_begin) * 4;
                    for (int c2 = 0; c2 < BN; c2 += 32) {
                        uint32_t v2[32];
                        const bool two = c2 + 16 < BN;
                        tmem_ld16(taddr + (t - t_lo) * BN + c2, *reinterpret_cast<uint32_t(*)[16]>(&v2[0]));
                        if (two) tmem_ld16(taddr + (t - t_lo) * BN + c2 + 16, *reinterpret_cast<uint32_t(*)[16]>(&v2[16]));
                        tmem_ld_wait();
                        if (c2 >= c_count) break;
                        const int ncol = two ? 32 : 16;
                        stage_store(ncol / 4, stage, lane, v2, row_off + (long long)c2 * 4, okmask, (char*)a.dw, min(ncol, c_count - c2) * 4, mode, 4);
                    }
                }
            } else
            for (int t = t_lo; t < t_hi; ++t) {
                float* o = a.dw + (long long)a.wtap[t] * a.w_tap + (long long)j * a.C + c_begin;
                for (int c2 = 0; c2 < BN; c2 += 32) {
                    uint32_t v2[2][16];
                    tmem_ld16(taddr + (t - t_lo) * BN + c2, v2[0]);
                    if (c2 + 16 < BN) tmem_ld16(taddr + (t - t_lo) * BN + c2 + 16, v2[1]);
                    tmem_ld_wait();
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
                        const int c = c2 + hh * 16;
                        const uint32_t(&v)[16] = v2[hh];
                        if (!rok || c >= c_count) continue;
                        const bool vec = c + 16 <= c_count && ((((uintptr_t)(o + c)) & 15) == 0);
                        if (!a.atomic) {            // sole writer of a tile the caller declared zero: plain stores
                            if (vec) {
#pragma unroll
                                for (int i = 0; i < 4; ++i)
                                    reinterpret_cast<float4*>(o + c)[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                                                      __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
                            } else {
#pragma unroll
                                for (int i = 0; i < 16; ++i)
                                    if (c + i < c_count) o[c + i] = __uint_as_float(v[i]);
                            }
                        } else if (vec) {           // red.add (a read-add-write epilogue measured 1.7x slower)
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                red_add_v4(o + c + 4 * i, __uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                           __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; ++i)
                                if (c + i < c_count) atomicAdd(o + c + i, __uint_as_float(v[i]));
                        }
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(tempty);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// libcuda is resolved at run time through the runtime API, so the library still loads on a host without a driver
EncodeFn encode_fn() {
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (EncodeFn)p;
    }
    return fn;
}

int encode_2d(CUtensorMap* tm, const void* ptr, unsigned long long inner, unsigned long long rows, unsigned box_inner,
              unsigned box_rows) {
    EncodeFn cuTensorMapEncodeTiled = encode_fn();
    if (!cuTensorMapEncodeTiled) return fail(P2P_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[2] = {inner, rows};
    cuuint64_t strides[1] = {inner * 2};            // bytes, bf16
    cuuint32_t box[2] = {box_inner, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(P2P_ECUDA, "cuTensorMapEncodeTiled failed (%d): inner %llu rows %llu box %u x %u", (int)r, inner, rows, box_inner, box_rows);
    return P2P_OK;
}

long long* g_window_dbg = nullptr;    // P2P_WINDOW_TIMING builds only (p2p_conv_window_debug_buffer)
int g_window_base_off_mode = 0;     // measured on B200: shifted starts need base_offset 0 (absolute-address swizzle)

}  // namespace
void note_umma_launch();
}  // namespace p2p

extern "C" int p2p_conv_window_debug_buffer(long long* buf) {    // 8 x int64 per CTA; only read by -DP2P_WINDOW_TIMING builds
    p2p::g_window_dbg = buf;
    return 0;
}

extern "C" int p2p_conv_window_base_off_mode(int mode) {
    const int old = p2p::g_window_base_off_mode;
    if (mode >= 0) p2p::g_window_base_off_mode = mode;
    return old;
}

extern "C" int p2p_pad_nhwc(const void* x, void* xp, int N, int H, int W, int C, int pt, int pb, int pl, int pr, int reflect,
                            int dtype, void* stream) {
    using namespace p2p;
    if (!x || !xp) return fail(P2P_EINVAL, "null pointer argument");
    const int esz = dtype == P2P_BF16 ? 2 : 4;
    if ((C * esz) % 16 != 0) return fail(P2P_ENOTSUP, "p2p_pad_nhwc needs whole 16-byte vectors per pixel (C * %d bytes)", esz);
    if (reflect && (pt >= H || pb >= H || pl >= W || pr >= W)) return fail(P2P_EINVAL, "reflection padding must be smaller than the input");
    if (N < 1 || H < 1 || W < 1 || pt < 0 || pb < 0 || pl < 0 || pr < 0) return fail(P2P_EINVAL, "bad padding geometry");
    const int Hp = H + pt + pb, Wp = W + pl + pr, C8 = C * esz / 16;
    const long long total = (long long)N * Hp * Wp * C8;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    launch_k(pad_nhwc_kernel<uint4>, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, (const uint4*)x, (uint4*)xp, N, H, W, C8, pt, pl, Hp, Wp, reflect, 1, (long long)N * Hp * Wp);
    return check_launch("pad_nhwc_kernel");
}

// planes [s*s][N, Hh, Wh, C]: plane (a, b) holds the zero-padded input sampled at (s*y + a - pt, s*x + b - pl)
extern "C" int p2p_pad_planes(const void* x, void* planes, int N, int H, int W, int C, int pt, int pl, int Hh, int Wh, int s,
                              long long plane_rows, int reflect, int dtype, void* stream) {
    using namespace p2p;
    if (!x || !planes) return fail(P2P_EINVAL, "null pointer argument");
    const int esz = dtype == P2P_BF16 ? 2 : 4;
    if ((C * esz) % 16 != 0) return fail(P2P_ENOTSUP, "p2p_pad_planes needs whole 16-byte vectors per pixel (C * %d bytes)", esz);
    if (N < 1 || H < 1 || W < 1 || Hh < 1 || Wh < 1 || s < 1 || s > 2) return fail(P2P_EINVAL, "bad plane geometry");
    if (plane_rows < (long long)N * Hh * Wh) return fail(P2P_EINVAL, "plane_rows smaller than N*Hh*Wh");
    if (reflect && s != 1) return fail(P2P_ENOTSUP, "reflection padding of parity planes");
    const int C8 = C * esz / 16;
    const long long total = plane_rows * C8 * s * s;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    launch_k(pad_nhwc_kernel<uint4>, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, (const uint4*)x, (uint4*)planes, N, H, W, C8, pt, pl, Hh, Wh, reflect, s, plane_rows);
    return check_launch("pad_planes_kernel");
}

// General form (include/p2p_b200.h: p2p_conv_window_desc): up to 4 sub-passes (output parities) share the planes.
extern "C" int p2p_conv_window_run_stats(const p2p_conv_window_desc* d, const void* planes, const void* weight, const float* bias,
                                         void* out, int out_dtype, float* stats, void* stream);
extern "C" int p2p_conv_window_run(const p2p_conv_window_desc* d, const void* planes, const void* weight, const float* bias,
                                   void* out, int out_dtype, void* stream) {
    return p2p_conv_window_run_stats(d, planes, weight, bias, out, out_dtype, nullptr, stream);
}

// stats (optional): [N][J][2] fp32, written by this call (zeroed, then accumulated by the epilogue): per (image, output channel)
// sum and sum of squares of (out - bias) over the output pixels -- InstanceNorm's statistics without a second pass over `out`.
// Needs an fp32 output with J % 4 == 0 and enough tiles to fill the chip (no channel-loop split): P2P_ENOTSUP otherwise, with
// nothing launched.
extern "C" int p2p_conv_window_run_stats(const p2p_conv_window_desc* d, const void* planes, const void* weight, const float* bias,
                                         void* out, int out_dtype, float* stats, void* stream) {
    using namespace p2p;
    if (!d || !planes || !weight || !out) return fail(P2P_EINVAL, "null pointer argument");
    const int ntaps = d->ntaps, nplanes = d->nplanes, C = d->C, J = d->J, Hp = d->Hp, Wp = d->Wp;
    if (ntaps < 1 || ntaps > P2P_MAX_TAPS || nplanes < 1 || nplanes > 4 || d->wtaps_total < 1 || d->nsub < 1 || d->nsub > 4)
        return fail(P2P_EINVAL, "bad tap / plane / sub-pass count");
    if (C % 16 != 0) return fail(P2P_ENOTSUP, "window convolution needs C %% 16 == 0, got %d", C);
    WindowArgs a;
    a.N = d->N; a.Hp = Hp; a.Wp = Wp; a.C = C; a.J = J;
    a.OHf = d->OHf; a.OWf = d->OWf; a.oys = d->oys; a.oxs = d->oxs;
    a.ntaps = ntaps; a.nplanes = nplanes; a.nsub = d->nsub;
    a.Q = (long long)d->N * Hp * Wp;
    a.Qp = d->plane_rows > 0 ? d->plane_rows : a.Q;
    if (a.Qp < a.Q) return fail(P2P_EINVAL, "plane_rows smaller than N*Hp*Wp");
    if ((long long)nplanes * a.Qp > 2000000000LL) return fail(P2P_ENOTSUP, "too many rows");
    for (int u = 0; u < 4; ++u) {
        a.gy[u] = a.gx[u] = a.oyo[u] = a.oxo[u] = 0;
        for (int p = 0; p <= 4; ++p) a.pbeg[u][p] = 0;
    }
    for (int u = 0; u < d->nsub; ++u) {
        a.gy[u] = d->gy[u]; a.gx[u] = d->gx[u]; a.oyo[u] = d->oyo[u]; a.oxo[u] = d->oxo[u];
        if (a.gy[u] < 1 || a.gx[u] < 1 || a.gy[u] > Hp || a.gx[u] > Wp || a.oyo[u] < 0 || a.oxo[u] < 0 ||
            (a.gy[u] - 1) * d->oys + a.oyo[u] >= d->OHf || (a.gx[u] - 1) * d->oxs + a.oxo[u] >= d->OWf)
            return fail(P2P_EINVAL, "sub-pass %d scatters outside the output tensor", u);
    }
    int halo = 0, t = 0;
    for (int u = 0; u < d->nsub; ++u) {         // taps grouped by sub-pass, then sorted by plane
        for (int p = 0; p < nplanes; ++p) {
            a.pbeg[u][p] = t;
            while (t < ntaps && d->tap_sub[t] == u && d->tap_plane[t] == p) {
                if (d->tap_dh[t] < 0 || d->tap_dw[t] < 0 ||
                    (!d->wrap && (d->tap_dh[t] + a.gy[u] > Hp || d->tap_dw[t] + a.gx[u] > Wp)))
                    return fail(P2P_EINVAL, "tap %d reads outside its plane", t);
                if (d->tap_w[t] >= d->wtaps_total) return fail(P2P_EINVAL, "weight tap out of range");
                const int sh = d->tap_dh[t] * Wp + d->tap_dw[t];
                if (sh > 32767) return fail(P2P_ENOTSUP, "tap shift too large");
                a.shift[t] = (short)sh;
                a.wtap[t] = d->tap_w[t];
                if (sh > halo) halo = sh;
                ++t;
            }
        }
        for (int p = nplanes; p <= 4; ++p) a.pbeg[u][p] = t;
        if (a.pbeg[u][nplanes] == a.pbeg[u][0]) return fail(P2P_EINVAL, "sub-pass %d has no taps", u);
    }
    if (t != ntaps) return fail(P2P_EINVAL, "taps must be grouped by sub-pass and sorted by plane");
    // swapped form (weights on the M side, 256 pixels on the N side) for J <= 128: see conv_window_swap_kernel
    static int swap_env = -1;
    if (swap_env < 0) { const char* e = getenv("P2P_CONV_SWAP"); swap_env = (e && e[0] == '1') ? 1 : 0; }      // opt-in since the 256-row tiles below: 9.15 vs 9.19 ms per step
    a.Q = (long long)d->N * Hp * Wp;
    const long long tiles256 = ((a.Q + SBM - 1) / SBM) * d->nsub * ((J + 127) / 128);
    // measured per layer (gpurun_out/layers_swap.log vs layers_noswap.log): a win for J = 96 (res blocks: dgrad 253 -> 164 us per
    // 4 calls), a loss for J <= 64 (fixed 256-pixel epilogue, 16 KB weight tiles) and mixed at J = 128
    const bool swap = swap_env && J >= 96 && J < 128 && J % 4 == 0 && SBM + halo <= kSwapRowsMax && tiles256 >= 120 &&
                      (((uintptr_t)out) & 15) == 0;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // column tile: whole J when it fits 256 columns, else balanced multiples of 16 (swapped form: 128 weight rows per MMA)
    {
        const int nt = (J + 255) / 256;
        const int bn = ((J + nt - 1) / nt + 15) / 16 * 16;
        a.BN = swap ? 128 : (bn < 16 ? 16 : bn);
    }
    a.n_tiles = (J + a.BN - 1) / a.BN;
    // CTA pairs (tcgen05 cta_group::2, see conv_window_kernel): opt-in with P2P_CONV_CG=2.  Measured on B200 (profiles/README.md,
    // round 2): the pair form issues MMAs at the same ~78 cycles each as single CTAs in this kernel (the operand reads out of
    // shared memory, not the instruction floor of the microbenchmark, pace it) and loses to them on every G3L2 layer.
    static int cg_env = -1;
    if (cg_env < 0) { const char* e = getenv("P2P_CONV_CG"); cg_env = (e && e[0] == '2') ? 2 : 1; }
    const int CG = (!swap && cg_env == 2 && sms % 2 == 0) ? 2 : 1;
    const int slots = sms / CG;                 // CTAs (CG = 1) or CTA pairs (CG = 2) resident at once
    // Two 128-row MMAs per weight tile (mrep = 2) halve the weight bytes TMA writes into shared memory per pixel.  The kernel
    // is bound by the shared-memory port (MMA operand reads + TMA fills + epilogue staging, 128 B/clk), see DESIGN.md section 6,
    // so this pays whenever the longer tiles still fill the chip: rounds(256-row tiles) * 2 * 0.8 < rounds(128-row tiles).
    static int mrep_env = -1;
    if (mrep_env < 0) { const char* e = getenv("P2P_CONV_MREP"); mrep_env = (e && e[0] == '1') ? 1 : 2; }
    a.mrep = 1;
    a.out_bf16_hint = out_dtype == P2P_BF16;
    if (!swap && mrep_env == 2) {
        const long long w1 = ((a.Q + 128 * CG - 1) / (128 * CG)) * a.n_tiles * d->nsub, w2 = ((a.Q + 256 * CG - 1) / (256 * CG)) * a.n_tiles * d->nsub;
        const long long r1 = (w1 + slots - 1) / slots, r2 = (w2 + slots - 1) / slots;
        const int nb2 = (256 + halo + 255) / 256, rpb2 = ((256 + halo + nb2 - 1) / nb2 + 7) / 8 * 8;
        const int rows2 = (nb2 * rpb2 * 128 + 1023) / 1024 * 1024;         // bytes of one 256-row window buffer, box-rounded
        const int bnl = a.BN / CG;
        const int tap_bytes_k = ((bnl * 128 + 1023) / 1024) * 1024, tap_bytes_mn = ((bnl + 63) / 64) * 8192;
        const int tap_max = tap_bytes_k > tap_bytes_mn ? tap_bytes_k : tap_bytes_mn;
        if (w1 * 2 > slots && r2 * 16 < r1 * 10 && WDATA_BYTES - 2 * rows2 >= 3 * tap_max) a.mrep = 2;
        // few-tile layers (channel loop split across CTAs below): these are bound by the weight bytes TMA pulls out of L2, and
        // the split fills the chip either way -- take the longer tiles whenever at least two of them exist
        const int cpb_ = (C + 63) / 64 * nplanes;
        if (w1 * 2 <= slots && !a.out_bf16_hint && cpb_ >= 4 && a.Q > 256 * CG && w2 * (cpb_ / 2) >= slots / 2 &&
            WDATA_BYTES - 2 * rows2 >= 3 * tap_max && !deterministic()) a.mrep = 2;
    }
    // pairs: a layer whose (half) weight set fits next to 128-row windows but not next to 256-row ones keeps the short tiles
    // and its weights resident (see `resident` below)
    auto win_bytes_for = [&](int mrep) {
        const int rows = WBM * mrep + halo, nb = (rows + 255) / 256, rpb = ((rows + nb - 1) / nb + 7) / 8 * 8;
        const int wb = (rows * 128 + 1023) / 1024 * 1024, bb = (nb * rpb * 128 + 1023) / 1024 * 1024;
        return wb > bb ? wb : bb;
    };
    auto resident_fits = [&](int mrep) {
        if (swap || (J + a.BN - 1) / a.BN != 1 || d->nsub != 1) return false;
        const int bnl = a.BN / CG;
        const bool mn = !(d->w_k == 1);
        const int tapb = mn ? ((bnl + 63) / 64) * 8192 : ((bnl * 128 + 1023) / 1024) * 1024;
        const int ring = WDATA_BYTES - 2 * win_bytes_for(mrep);
        for (int tps = 4; tps >= 1; --tps) {
            int stages = 0;
            for (int p = 0; p < nplanes; ++p) stages += (a.pbeg[0][p + 1] - a.pbeg[0][p] + tps - 1) / tps;
            stages *= (C + 63) / 64;
            if (stages >= 2 && stages <= kBStagesMax && (long long)stages * tps * tapb <= ring) return true;
        }
        return false;
    };
    if (CG == 2 && a.mrep == 2 && !resident_fits(2) && resident_fits(1)) a.mrep = 1;
    const int TR = swap ? SBM : WBM * a.mrep, rows_max = swap ? kSwapRowsMax : (WDATA_BYTES / 2) / 128;
    a.win_rows = TR + halo;
    if (a.win_rows > rows_max || (!swap && a.mrep == 1 && a.win_rows > kWinRowsMax))
        return fail(P2P_ENOTSUP, "window of %d rows exceeds the %d-row buffer", a.win_rows, rows_max);
    a.win_bytes = (a.win_rows * 128 + 1023) / 1024 * 1024;
    a.half_cols = a.BN <= 128 ? 128 : 256;
    a.acc_slots = (a.mrep == 2 && a.BN > 128) ? 1 : 2;
    a.nboxes = (a.win_rows + 255) / 256;
    a.rows_per_box = (a.win_rows + a.nboxes - 1) / a.nboxes;
    a.rows_per_box = (a.rows_per_box + 7) / 8 * 8;          // keeps every box 1024-byte aligned
    if (a.nboxes * a.rows_per_box > rows_max) return fail(P2P_ENOTSUP, "window boxes exceed the buffer");
    if (!swap && a.nboxes * a.rows_per_box * 128 > a.win_bytes) {
        a.win_bytes = (a.nboxes * a.rows_per_box * 128 + 1023) / 1024 * 1024;
    }
    const long long w_tap = d->w_tap, w_k = d->w_k, w_j = d->w_j;
    if (w_k == 1 && w_j % 8 == 0 && w_tap % 8 == 0) a.b_mn_major = 0;
    else if (w_j == 1 && w_k % 8 == 0 && w_tap % 8 == 0 && J % 8 == 0) a.b_mn_major = 1;
    else return fail(P2P_ENOTSUP, "weight strides not 16-byte friendly");
    if (a.b_mn_major ? (w_tap != (long long)C * w_k) : (w_tap != (long long)J * w_j))
        return fail(P2P_ENOTSUP, "weights must be densely packed per tap");
    a.bias = bias;
    a.out = out;
    a.stats = stats;
    a.out_bf16 = out_dtype == P2P_BF16;
    a.cpb = (C + 63) / 64;
    a.base_off_mode = g_window_base_off_mode;
    a.fast_epi = ((long long)J * (a.out_bf16 ? 2 : 4)) % 16 == 0 && (((uintptr_t)out) & 15) == 0;
    a.dbg = g_window_dbg;
    const int BNL = a.BN / CG;                  // weight-tile columns per CTA
    if (CG == 2 && (a.BN % 16 != 0)) return fail(P2P_ENOTSUP, "CTA pairs need a column tile that is a multiple of 16");
    a.b_tap_bytes = a.b_mn_major ? ((BNL + 63) / 64) * 8192 : ((BNL * 128 + 1023) / 1024) * 1024;
    const int ring_bytes = swap ? WRING_BYTES : WDATA_BYTES - 2 * a.win_bytes;
    a.tps = ring_bytes / ((CG == 2 ? 6 : 3) * a.b_tap_bytes);          // keep >= 3 (pairs: 6) stages in flight: the refill of a
    if (a.tps > 4) a.tps = 4;                                           // stage waits for its MMAs to COMPLETE and for TMA's latency
    if (a.tps < 1) a.tps = 1;
    a.b_stage_bytes = a.tps * a.b_tap_bytes;
    a.b_stages = ring_bytes / a.b_stage_bytes;
    if (a.b_stages > kBStagesMax) a.b_stages = kBStagesMax;
    if (a.b_stages < 2) return fail(P2P_ENOTSUP, "weight ring too small for this window");
    // Resident weights: a layer whose whole weight set (this CTA's half of it) fits the ring keeps it there -- every work item
    // after the first re-uses the tiles, so only the windows stream from L2 (the 96-channel layers would otherwise pull
    // ~40 B/clk/SM of weights at the cta_group::2 MMA rate, the L2's whole bandwidth)
    a.resident = 0;
    static int res_env = -1;
    if (res_env < 0) { const char* e = getenv("P2P_CONV_RESIDENT"); res_env = (e && e[0] == '0') ? 0 : 1; }
    const long long nwork0 = ((a.Q + (long long)TR * CG - 1) / ((long long)TR * CG)) * a.n_tiles * d->nsub;
    const bool will_split = !swap && out_dtype != P2P_BF16 && nwork0 * 2 <= slots && ((C + 63) / 64) * nplanes >= 4 && !deterministic();
    if (!swap && res_env && !will_split && nwork0 > slots && (J + a.BN - 1) / a.BN == 1 && d->nsub == 1) {
        for (int tps = 4; tps >= 1 && !a.resident; --tps) {
            int stages = 0;
            for (int p = 0; p < nplanes; ++p) stages += (a.pbeg[0][p + 1] - a.pbeg[0][p] + tps - 1) / tps;
            stages *= (C + 63) / 64;
            if (stages >= 2 && stages <= kBStagesMax && (long long)stages * tps * a.b_tap_bytes <= ring_bytes) {
                a.resident = 1; a.tps = tps; a.b_stage_bytes = tps * a.b_tap_bytes; a.b_stages = stages;
            }
        }
    }
    const long long mt = swap ? (a.Q + TR - 1) / TR : (a.Q + (long long)TR * CG - 1) / ((long long)TR * CG);      // row tiles per CTA (pair)
    if (mt * a.nsub > 2000000000LL / a.n_tiles) return fail(P2P_ENOTSUP, "too many tiles");
    if ((long long)nplanes * a.Qp + halo + (long long)TR * CG > 2000000000LL) return fail(P2P_ENOTSUP, "too many rows for 32-bit row indices");
    a.m_tiles = (int)mt;
    a.nwork = mt * a.n_tiles * a.nsub;
    a.nsplit = 1;
    a.ci_per_split = a.cpb * nplanes;
    const bool may_split = !swap && !a.out_bf16 && a.nwork * 2 <= slots && a.cpb * nplanes >= 4 && !deterministic();
    if (may_split && stats)
        return fail(P2P_ENOTSUP, "epilogue statistics are not available for split channel loops (few-tile layers)");
    if (may_split) {
        // few tiles (the 16x4-pixel bottleneck layers): split the channel loop so that one wave fills the chip; the
        // partial tiles meet in the fp32 output through red.add
        int ns = (int)(slots / a.nwork);
        if (ns > a.cpb * nplanes / 2) ns = a.cpb * nplanes / 2;
        a.ci_per_split = (a.cpb * nplanes + ns - 1) / ns;
        a.nsplit = (a.cpb * nplanes + a.ci_per_split - 1) / a.ci_per_split;
        if (a.nsplit > 1 && a.resident) return fail(P2P_ENOTSUP, "internal: resident weights with a split channel loop");
        if (a.nsplit > 1) {
            const size_t bytes = (size_t)d->N * d->OHf * d->OWf * J * sizeof(float);
            if (cudaMemsetAsync(out, 0, bytes, (cudaStream_t)stream) != cudaSuccess) return check_launch("cudaMemsetAsync(split output)");
            a.nwork *= a.nsplit;
        }
    }
    if (stats) {
        if (swap || a.out_bf16 || !a.fast_epi || a.nsplit > 1 || deterministic())
            return fail(P2P_ENOTSUP, "epilogue statistics need an fp32 output with J %% 4 == 0 and an unsplit channel loop");
        if (cudaMemsetAsync(stats, 0, sizeof(float) * 2 * (size_t)d->N * J, (cudaStream_t)stream) != cudaSuccess)
            return check_launch("cudaMemsetAsync(stats)");
    }
    // wrap mode: the last image's rows are reached through shifts past row Q (the caller's buffer holds Q + halo rows)
    if (int rc = encode_2d(&a.tm_x, planes, (unsigned long long)C, (unsigned long long)(nplanes - 1) * a.Qp + a.Q + (d->wrap ? halo : 0), 64,
                           (unsigned)a.rows_per_box)) return rc;
    if (!a.b_mn_major) {
        if (int rc = encode_2d(&a.tm_w, weight, (unsigned long long)C, (unsigned long long)d->wtaps_total * J, 64, (unsigned)BNL)) return rc;
    } else {
        if (int rc = encode_2d(&a.tm_w, weight, (unsigned long long)J, (unsigned long long)d->wtaps_total * C, 64, 64)) return rc;
    }
    static bool configured[64] = {};            // the attribute is per device
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        if (cudaFuncSetAttribute(conv_window_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, WSMEM) != cudaSuccess ||
            cudaFuncSetAttribute(conv_window_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, WSMEM) != cudaSuccess ||
            cudaFuncSetAttribute(conv_window_swap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SWSMEM) != cudaSuccess)
            return check_launch("cudaFuncSetAttribute(conv_window_kernel)");
        configured[dev] = true;
    }
    const long long grid = (a.nwork < slots ? a.nwork : slots) * CG;
    if (swap) {
        conv_window_swap_kernel<<<(unsigned)grid, kWThreads, SWSMEM, (cudaStream_t)stream>>>(a);
        note_umma_launch();
        return check_launch("conv_window_swap_kernel");
    }
    if (CG == 2) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)grid);
        cfg.blockDim = dim3(kWinThreads);
        cfg.dynamicSmemBytes = WSMEM;
        cfg.stream = (cudaStream_t)stream;
        cudaLaunchAttribute attr[2];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = pdl_enabled() ? 2 : 1;
        if (cudaLaunchKernelEx(&cfg, conv_window_kernel<2>, a) != cudaSuccess) return check_launch("conv_window_kernel<2> (cluster launch)");
    } else {
        launch_k(conv_window_kernel<1>, dim3((unsigned)grid), dim3(kWinThreads), WSMEM, (cudaStream_t)stream, a);
    }
    note_umma_launch();
    return check_launch("conv_window_kernel");
}

// out[n,y,x,j] (+bias) = sum_{dh<KH, dw<KW} sum_c xp[n, y+dh, x+dw, c] * Wt[c, j],  y < OH = Hp-KH+1, x < OW = Wp-KW+1.
// weight tap index = flip ? (KH-1-dh)*KW + (KW-1-dw) : dh*KW + dw; element (tap, c, j) at tap*w_tap + c*w_k + j*w_j with
// either w_k == 1 (K-major) or w_j == 1 (MN-major, the transposed read used for input gradients).
extern "C" int p2p_conv_window_stats(const void* xp, const void* weight, const float* bias, void* out, int N, int Hp, int Wp, int C,
                                     int J, int KH, int KW, int flip, long long w_tap, long long w_k, long long w_j, int out_dtype,
                                     float* stats, void* stream);
extern "C" int p2p_conv_window(const void* xp, const void* weight, const float* bias, void* out, int N, int Hp, int Wp, int C,
                               int J, int KH, int KW, int flip, long long w_tap, long long w_k, long long w_j, int out_dtype,
                               void* stream) {
    return p2p_conv_window_stats(xp, weight, bias, out, N, Hp, Wp, C, J, KH, KW, flip, w_tap, w_k, w_j, out_dtype, nullptr, stream);
}

extern "C" int p2p_conv_window_stats(const void* xp, const void* weight, const float* bias, void* out, int N, int Hp, int Wp, int C,
                                     int J, int KH, int KW, int flip, long long w_tap, long long w_k, long long w_j, int out_dtype,
                                     float* stats, void* stream) {
    using namespace p2p;
    if (KH < 1 || KW < 1 || KH * KW > P2P_MAX_TAPS || Hp < KH || Wp < KW) return fail(P2P_EINVAL, "bad window geometry");
    p2p_conv_window_desc d;
    d.nplanes = 1; d.N = N; d.Hp = Hp; d.Wp = Wp; d.C = C;
    d.OHf = Hp - KH + 1; d.OWf = Wp - KW + 1; d.J = J;
    d.wtaps_total = KH * KW; d.w_tap = w_tap; d.w_k = w_k; d.w_j = w_j;
    d.nsub = 1; d.oys = d.oxs = 1; d.wrap = 0; d.plane_rows = 0;
    for (int u = 0; u < 4; ++u) { d.gy[u] = d.OHf; d.gx[u] = d.OWf; d.oyo[u] = d.oxo[u] = 0; }
    d.ntaps = KH * KW;
    for (int i = 0; i < KH; ++i)
        for (int j = 0; j < KW; ++j) {
            const int t = i * KW + j;
            d.tap_sub[t] = 0; d.tap_plane[t] = 0; d.tap_dh[t] = (short)i; d.tap_dw[t] = (short)j;
            d.tap_w[t] = (unsigned char)(flip ? (KH - 1 - i) * KW + (KW - 1 - j) : t);
        }
    return p2p_conv_window_run_stats(&d, xp, weight, bias, out, out_dtype, stats, stream);
}

// dW[tap_w[t]][j][c] (fp32, [wtaps_total][J][C], caller-zeroed) += sum over the padded grid of
//   aplanes[tap_a[t]][q][j] * bplanes[tap_b[t]][q + tap_dh[t]*Wp + tap_dw[t]][c];
// planes are [plane_rows][channels] bf16 with plane_rows a multiple of 64 (p2p_pad_planes zero-fills the tail);
// taps must be grouped by (tap_a, tap_b).  ntaps * BN <= 512 TMEM columns: ntaps <= 32.
static thread_local int g_last_wgrad_stored = 0;
// 1 when this thread's last p2p_conv_wgrad_window call STORED dW (one writer per tile), 0 when it accumulated with red.add
extern "C" int p2p_conv_wgrad_window_stored(void) { return g_last_wgrad_stored; }

extern "C" int p2p_conv_wgrad_window(const void* aplanes, int n_aplanes, const void* bplanes, int n_bplanes, long long plane_rows,
                                     int Wp, int J, int C, float* dw, int wtaps_total, int ntaps, const int* tap_a,
                                     const int* tap_b, const int* tap_dh, const int* tap_dw, const int* tap_w, int dw_is_zero,
                                     void* stream) {
    using namespace p2p;
    if (!aplanes || !bplanes || !dw || !tap_a || !tap_b || !tap_dh || !tap_dw || !tap_w) return fail(P2P_EINVAL, "null pointer argument");
    if (ntaps < 1 || ntaps > 32 || wtaps_total < 1) return fail(P2P_ENOTSUP, "window wgrad supports 1..32 taps, got %d", ntaps);
    if (plane_rows < 64 || plane_rows % 64 != 0) return fail(P2P_EINVAL, "plane_rows must be a positive multiple of 64");
    if (J % 8 != 0 || C % 16 != 0) return fail(P2P_ENOTSUP, "window wgrad needs J %% 8 == 0 and C %% 16 == 0");
    if ((long long)(n_aplanes > n_bplanes ? n_aplanes : n_bplanes) * plane_rows > 2000000000LL) return fail(P2P_ENOTSUP, "too many rows");
    WgradWinArgs a;
    a.dw = dw; a.Qs = plane_rows; a.J = J; a.C = C; a.ntaps = ntaps; a.w_tap = (long long)J * C;
    int halo = 0, ng = 0;
    for (int t = 0; t < ntaps; ++t) {
        if (tap_a[t] < 0 || tap_a[t] >= n_aplanes || tap_b[t] < 0 || tap_b[t] >= n_bplanes || tap_dh[t] < 0 || tap_dw[t] < 0)
            return fail(P2P_EINVAL, "bad tap %d", t);
        if (tap_w[t] < 0 || tap_w[t] >= wtaps_total) return fail(P2P_EINVAL, "weight tap out of range");
        if (t == 0 || tap_a[t] != tap_a[t - 1] || tap_b[t] != tap_b[t - 1]) {
            if (ng == kMaxGroups) return fail(P2P_ENOTSUP, "too many operand-plane pairs");
            for (int g = 0; g < ng; ++g)
                if (a.ga[g] == tap_a[t] && a.gb[g] == tap_b[t]) return fail(P2P_EINVAL, "taps must be grouped by operand planes");
            a.ga[ng] = (short)tap_a[t]; a.gb[ng] = (short)tap_b[t]; a.gbeg[ng] = t;
            ++ng;
        }
        const int sh = tap_dh[t] * Wp + tap_dw[t];
        if (sh > 32767) return fail(P2P_ENOTSUP, "tap shift too large");
        a.shift[t] = (short)sh;
        a.wtap[t] = (unsigned char)tap_w[t];
        if (sh > halo) halo = sh;
    }
    for (int g = ng; g <= kMaxGroups; ++g) a.gbeg[g] = ntaps;
    a.ngroups = ng;
    const int brows = 64 + halo;
    if (brows > kWinRowsMax) return fail(P2P_ENOTSUP, "window of %d rows exceeds the %d-row buffer", brows, kWinRowsMax);
    a.b_nboxes = (brows + 255) / 256;
    a.b_rows_per_box = ((brows + a.b_nboxes - 1) / a.b_nboxes + 7) / 8 * 8;
    if (a.b_nboxes * a.b_rows_per_box > kWinRowsMax) return fail(P2P_ENOTSUP, "window boxes exceed the buffer");
    a.chunk_bytes = ((a.b_nboxes * a.b_rows_per_box * 128 + 1023) / 1024) * 1024;
    // channel tile: as wide as two ring stages allow (fewer, wider MMAs), balanced over C; then taps per set = 512 / BN columns
    const int Cr = (C + 15) / 16 * 16;
    int max_chunks = (GRING_BYTES / 2 - GA_BYTES) / a.chunk_bytes;
    if (max_chunks > 4) max_chunks = 4;
    if (max_chunks < 1) return fail(P2P_ENOTSUP, "window too large for the stage ring");
    const int bn_cap = max_chunks * 64;
    const int nct = (Cr + bn_cap - 1) / bn_cap;
    int bn = ((Cr + nct - 1) / nct + 15) / 16 * 16;
    if (bn > bn_cap) bn = bn_cap;
    // Channel-tile width by a two-term cost model (round 2, measured constants): per 64-pixel step a work item pulls the dout
    // tile (16 KB) plus BN/64 window chunks through TMA (~24 B/clk/SM) and issues set_taps x 4 MMAs of max(93, BN/2) cycles;
    // the TMEM holds set_taps x BN = 512 accumulator columns.  A wide tile halves the MMA count but doubles the window bytes
    // per accumulated column -- the wide-channel layers (1536-channel bottleneck, D's 256 -> 512) are TMA-bound at BN = 256.
    static int bn_model = -1;
    if (bn_model < 0) { const char* e = getenv("P2P_WGRAD_BN_MODEL"); bn_model = (e && e[0] == '0') ? 0 : 1; }
    if (bn_model && C >= 256) {                 // (measured: the 192-channel layers lose with narrower tiles)
        double best = 1e30;
        int best_bn = bn;
        for (int cand = 64; cand <= bn_cap && cand <= Cr; cand += 64) {
            if (C % cand != 0 && cand != bn) continue;                   // whole tiles only (ragged tiles keep the balanced width)
            const int tmax_c = 512 / cand, nsets_c = (ntaps + tmax_c - 1) / tmax_c, set_c = (ntaps + nsets_c - 1) / nsets_c;
            const double mma = set_c * 4.0 * (cand / 2 > 93 ? cand / 2 : 93);
            const double tma = (GA_BYTES + (double)((cand + 63) / 64) * a.chunk_bytes) / 24.0;
            const double cost = (mma > tma ? mma : tma) * nsets_c / ((double)ntaps * cand);      // cycles per accumulated (tap, column)
            if (cost < best * 0.97) { best = cost; best_bn = cand; }
        }
        bn = best_bn;
    }
    a.BN = bn;
    a.nchunks = (bn + 63) / 64;
    a.j_tiles = (J + 127) / 128;
    a.c_tiles = (C + bn - 1) / bn;
    const int tmax = 512 / bn;
    a.nsets = (ntaps + tmax - 1) / tmax;
    a.set_taps = (ntaps + a.nsets - 1) / a.nsets;
    a.nsets = (ntaps + a.set_taps - 1) / a.set_taps;
    a.stage_bytes = GA_BYTES + a.nchunks * a.chunk_bytes;
    a.stages = GRING_BYTES / a.stage_bytes;
    if (a.stages > kGStagesMax) a.stages = kGStagesMax;
    a.total_kb = (int)(plane_rows / 64);
    const long long items = (long long)a.j_tiles * a.c_tiles * a.nsets;
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int nsplit = (int)(sms / items);                       // at most one wave: every extra split is another pass of atomics over dW
    if (nsplit > a.total_kb / 2) nsplit = a.total_kb / 2;
    if (nsplit < 1 || deterministic() || dw_is_zero == 2) nsplit = 1;     // 2: dW holds stale values and MUST be overwritten
    a.kb_per_split = (a.total_kb + nsplit - 1) / nsplit;
    a.nsplit = (a.total_kb + a.kb_per_split - 1) / a.kb_per_split;
    a.fast_epi = C % 4 == 0 && (((uintptr_t)dw) & 15) == 0;
    a.atomic = !(dw_is_zero && a.nsplit == 1);    // dW += with red.add unless the buffer is known zero and the tile has one writer
    g_last_wgrad_stored = a.atomic ? 0 : 1;
    a.nwork = items * a.nsplit;
    if (int rc = encode_2d(&a.tm_a, aplanes, (unsigned long long)J, (unsigned long long)n_aplanes * plane_rows, 64, 64)) return rc;
    if (int rc = encode_2d(&a.tm_b, bplanes, (unsigned long long)C, (unsigned long long)n_bplanes * plane_rows, 64, (unsigned)a.b_rows_per_box)) return rc;
    static bool configured[64] = {};            // the attribute is per device
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        if (cudaFuncSetAttribute(conv_wgrad_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GSMEM) != cudaSuccess)
            return check_launch("cudaFuncSetAttribute(conv_wgrad_window_kernel)");
        configured[dev] = true;
    }
    const long long grid = a.nwork < sms ? a.nwork : sms;
    launch_k(conv_wgrad_window_kernel, dim3((unsigned)grid), dim3(kWThreads), GSMEM, (cudaStream_t)stream, a);
    note_umma_launch();
    return check_launch("conv_wgrad_window_kernel");
}
