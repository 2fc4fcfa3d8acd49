// microbenchmark: cycles per tcgen05.mma (M=128, K=16, kind::f16) issued back to back by one thread, for several N,
// operand major-ness and A start shifts; smem contents are arbitrary (timing only).  One CTA per SM.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46) | (2ull << 61);
}
// mode bit0: A MN-major, bit1: B MN-major.  groups of 4 MMAs (k16 = 0..3) then a commit, `iters` groups; ntmem accumulators rotated
__global__ void k(long long* out, int N, int mode, int iters, int shift_rows, int nacc, int commit_every) {
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
        const int amn = mode & 1, bmn = (mode >> 1) & 1;
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)amn << 15) | ((uint32_t)bmn << 16) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t ad0 = amn ? make_desc(sa + shift_rows * 128, 8192, 1024) : make_desc(sa + shift_rows * 128, 16, 1024);
        const uint64_t bd0 = bmn ? make_desc(sb, 8192, 1024) : make_desc(sb, 16, 1024);
        const uint32_t ak = amn ? 128 : 2, bk = bmn ? 128 : 2;
        const long long t0 = clock64();
        uint32_t ph = 0;
        for (int it = 0; it < iters; ++it) {
            // the 4 MMAs of a group go to accumulators 0..nacc-1 round-robin (nacc = 1: one dependent chain)
            umma_bf16(tmem + (uint32_t)((0 % nacc) * N), ad0, bd0, idesc, 1);
            umma_bf16(tmem + (uint32_t)((1 % nacc) * N), ad0 + ak, bd0 + bk, idesc, 1);
            umma_bf16(tmem + (uint32_t)((2 % nacc) * N), ad0 + 2 * ak, bd0 + 2 * bk, idesc, 1);
            umma_bf16(tmem + (uint32_t)((3 % nacc) * N), ad0 + 3 * ak, bd0 + 3 * bk, idesc, 1);
            if ((it + 1) % commit_every == 0) { umma_commit(bar); mbar_wait(bar, ph); ph ^= 1; }
        }
        if (iters % commit_every != 0) { umma_commit(bar); mbar_wait(bar, ph); }
        const long long t1 = clock64();
        out[blockIdx.x] = t1 - t0;
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}
// W warps (one elected thread each) issue concurrently to disjoint accumulators: is the ~105-cycle cost per issuing thread or per SM?
__global__ void k2(long long* out, int N, int iters, int W) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t sa = base, sb = base + 64 * 1024, bar = base + 160 * 1024, slot = bar + 64;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 40 * 1024; i += blockDim.x) reinterpret_cast<uint32_t*>(raw)[i] = 0x3c003c00u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        if (tid == 0) { for (int i = 0; i < 4; ++i) mbar_init(bar + 8 * i, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tmem;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(slot));
    long long t0 = clock64();
    if (warp < W && (tid & 31) == 0) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t ad0 = make_desc(sa, 16, 1024), bd0 = make_desc(sb, 16, 1024);
        const uint32_t d = tmem + (uint32_t)(warp * N);
        for (int it = 0; it < iters; ++it) {
            umma_bf16(d, ad0, bd0, idesc, 1);
            umma_bf16(d, ad0 + 2, bd0 + 2, idesc, 1);
            umma_bf16(d, ad0 + 4, bd0 + 4, idesc, 1);
            umma_bf16(d, ad0 + 6, bd0 + 6, idesc, 1);
        }
        umma_commit(bar + 8 * warp);
        mbar_wait(bar + 8 * warp, 0);
    }
    __syncthreads();
    if (tid == 0) out[blockIdx.x] = clock64() - t0;
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}
int main() {
    long long* d; cudaMalloc(&d, 148 * 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 512;
    printf("cycles per MMA (M=128,K=16), %d groups of 4, all 148 SMs busy; commit+wait every C groups\n", iters);
    const int Ns[] = {16, 48, 64, 96, 128, 192, 256};
    for (int ce : {512})
    for (int mode : {0, 3})
        for (int N : Ns) {
            for (int nacc : {1, 2, 4}) {
                const int shift = 0;
                if (nacc * N > 512) continue;
                k<<<148, 128, 200 * 1024>>>(d, N, mode, iters, shift, nacc, ce);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
                long long mx = 0; for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
                printf("commit_every %3d  A %s  B %s  N %3d  independent accumulators %d : %7.1f cyc/MMA  (math floor %d)\n", ce, (mode & 1) ? "MN" : "K ", (mode & 2) ? "MN" : "K ", N, nacc,
                       (double)mx / (iters * 4.0), 128 * N / 256);
            }
        }
    printf("\nW issuing warps per CTA, each 512 groups of 4 MMAs to its own accumulator: cycles per MMA counted over ALL warps' MMAs\n");
    cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int N : {48, 96, 128})
        for (int W : {1, 2, 4}) {
            if (W * N > 512) continue;
            k2<<<148, 128, 200 * 1024>>>(d, N, iters, W);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
            long long mx = 0; for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
            printf("N %3d  issuing warps %d : %7.1f cyc/MMA (math floor %d)\n", N, W, (double)mx / (iters * 4.0 * W), 128 * N / 256);
        }
    return 0;
}
