// Shared helpers for the C-ABI translation units (error slot, launch checks).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include "../../include/p2p_b200.h"

namespace p2p {

char* error_slot();   // thread-local 512-byte message buffer (defined in c_abi.cu)

inline int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_slot(), 512, fmt, ap);
    va_end(ap);
    return code;
}

long long& launch_counter();   // kernels launched through the C ABI by this process (defined in c_abi.cu)

inline int check_launch(const char* what) {
    ++launch_counter();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(P2P_ECUDA, "%s: %s", what, cudaGetErrorString(e));
    return P2P_OK;
}

int& deterministic_flag();      // -1 = not read yet (defined in c_abi.cu; env P2P_DETERMINISTIC, p2p_set_deterministic)
// Deterministic mode: no kernel lets several CTAs add floating-point partial results into one address in an order that
// depends on scheduling (split-K / split-pixel red.add, multi-chunk statistics and loss atomics): every reduction runs
// with one writer per address, at lower parallelism.  Results are then bit-reproducible run to run.
inline bool deterministic() {
    int& f = deterministic_flag();
    if (f < 0) { const char* e = getenv("P2P_DETERMINISTIC"); f = (e && e[0] == '1') ? 1 : 0; }
    return f != 0;
}

// ---- programmatic dependent launch (PDL) -----------------------------------------------------------------------------------
// A kernel launched through launch_k() may be scheduled while its predecessor on the stream is still draining (the
// predecessor announces that with pdl_trigger()); it must execute pdl_wait() before its first access to global memory, which
// blocks until the predecessor grid has completed and its writes are visible.  Inside a CUDA graph the attribute becomes a
// programmatic edge: launch latency, the prologue (barrier init, TMEM allocation, table loads) and the CTA ramp-up of kernel
// n+1 overlap the tail of kernel n.  Only kernels that call pdl_wait() may be launched this way.  P2P_PDL=0 turns it off.
inline bool pdl_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("P2P_PDL"); v = (e && e[0] == '0') ? 0 : 1; }
    return v != 0;
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// Kernels without dynamic shared memory ask for the SAME (maximum shared memory) L1 carve-out as the 227 KB convolution
// kernels they alternate with (P2P_CARVEOUT=0 keeps the driver's default): streaming element-wise kernels gain nothing from
// a large L1, and a carve-out change between two launches costs an SM reconfiguration.
inline bool carveout_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("P2P_CARVEOUT"); v = (e && e[0] == '1') ? 1 : 0; }
    return v != 0;
}
void note_kernel_for_carveout(const void* fn);      // c_abi.cu: cudaFuncSetAttribute once per (function, device)

template <typename... KA, typename... A>
inline cudaError_t launch_k(void (*kern)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A&&... args) {
    if (smem == 0 && carveout_enabled()) note_kernel_for_carveout((const void*)kern);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KA>(args)...);
}
#endif

inline bool is_pow2(long long v) { return v > 0 && (v & (v - 1)) == 0; }

}  // namespace p2p
