// microbenchmark (round 2): cycles per tcgen05.mma (K=16, kind::f16) issued back to back by one thread
//   (a) cta_group::2, M = 256 (128 rows per CTA of the pair), N = 32..256: does the >= 105-cycle per-instruction cost of the
//       cta_group::1 form (r01_mma_rate_microbench.txt) cover twice the rows?
//   (b) cta_group::1 with the A operand in TMEM (.ts form): does the floor come from the shared-memory read of A?
// smem / TMEM contents are arbitrary (timing only).  One CTA per SM, all 148 SMs busy.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma2_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma1_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma1_ts_bf16(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit2(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void commit1(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// (a) CTA pair: the leader issues M = 256 MMAs; both CTAs wait on their own copy of the barrier (multicast commit)
__global__ void __cluster_dims__(2, 1, 1) k_cg2(long long* out, int N, int iters, int nacc) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t sa = base, sb = base + 64 * 1024, bar = base + 160 * 1024, slot = bar + 64;
    const int tid = threadIdx.x, warp = tid >> 5;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    for (int i = tid; i < 40 * 1024; i += blockDim.x) reinterpret_cast<uint32_t*>(raw)[i] = 0x3c003c00u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        if (tid == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(slot));
    const long long t0 = clock64();
    if (tid == 0) {
        if (rank == 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((256u >> 4) << 24);
            const uint64_t ad0 = make_desc(sa, 16, 1024), bd0 = make_desc(sb, 16, 1024);
            for (int it = 0; it < iters; ++it) {
                umma2_bf16(tmem + (uint32_t)((0 % nacc) * N), ad0, bd0, idesc, 1);
                umma2_bf16(tmem + (uint32_t)((1 % nacc) * N), ad0 + 2, bd0 + 2, idesc, 1);
                umma2_bf16(tmem + (uint32_t)((2 % nacc) * N), ad0 + 4, bd0 + 4, idesc, 1);
                umma2_bf16(tmem + (uint32_t)((3 % nacc) * N), ad0 + 6, bd0 + 6, idesc, 1);
            }
            commit2(bar);
        }
        mbar_wait(bar, 0);
        out[blockIdx.x] = clock64() - t0;
    }
    __syncthreads();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    cluster_sync();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

// (b) one CTA: A from TMEM (columns 256.. of the allocation; K = 16 bf16 = 8 columns per MMA) or from smem
__global__ void k_ts(long long* out, int N, int iters, int from_tmem) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t sa = base, sb = base + 64 * 1024, bar = base + 160 * 1024, slot = bar + 64;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 40 * 1024; i += blockDim.x) reinterpret_cast<uint32_t*>(raw)[i] = 0x3c003c00u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        if (tid == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(slot));
    if (tid == 0) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t ad0 = make_desc(sa, 16, 1024), bd0 = make_desc(sb, 16, 1024);
        const uint32_t at = tmem + 256;
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            if (from_tmem) {
                umma1_ts_bf16(tmem, at, bd0, idesc, 1);
                umma1_ts_bf16(tmem, at + 8, bd0 + 2, idesc, 1);
                umma1_ts_bf16(tmem, at + 16, bd0 + 4, idesc, 1);
                umma1_ts_bf16(tmem, at + 24, bd0 + 6, idesc, 1);
            } else {
                umma1_bf16(tmem, ad0, bd0, idesc, 1);
                umma1_bf16(tmem, ad0 + 2, bd0 + 2, idesc, 1);
                umma1_bf16(tmem, ad0 + 4, bd0 + 4, idesc, 1);
                umma1_bf16(tmem, ad0 + 6, bd0 + 6, idesc, 1);
            }
        }
        commit1(bar);
        mbar_wait(bar, 0);
        out[blockIdx.x] = clock64() - t0;
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

static double run_max(long long* d, int n) {
    long long h[148];
    cudaMemcpy(h, d, sizeof(long long) * n, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < n; ++i) mx = h[i] > mx ? h[i] : mx;
    return (double)mx;
}

int main() {
    long long* d;
    cudaMalloc(&d, 148 * 8);
    cudaFuncSetAttribute(k_cg2, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_ts, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 512;
    printf("(a) cta_group::2, M=256 (2 x 128 rows), K=16: cycles per MMA, %d groups of 4, 74 CTA pairs\n", iters);
    for (int N : {32, 48, 64, 96, 128, 192, 256})
        for (int nacc : {1, 2}) {
            if (nacc * N > 512) continue;
            cudaMemset(d, 0, 148 * 8);
            k_cg2<<<148, 128, 200 * 1024>>>(d, N, iters, nacc);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            printf("cg2 M 256  N %3d  accumulators %d : %7.1f cyc/MMA  (math floor %d per SM)\n", N, nacc, run_max(d, 148) / (iters * 4.0), 128 * N / 256);
        }
    printf("(b) cta_group::1, M=128, K=16: A from shared memory vs A from TMEM\n");
    for (int N : {32, 64, 96, 128, 192, 256})
        for (int ft : {0, 1}) {
            cudaMemset(d, 0, 148 * 8);
            k_ts<<<148, 128, 200 * 1024>>>(d, N, iters, ft);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            printf("cg1 M 128  N %3d  A from %s : %7.1f cyc/MMA  (math floor %d)\n", N, ft ? "TMEM" : "smem", run_max(d, 148) / (iters * 4.0), 128 * N / 256);
        }
    return 0;
}
