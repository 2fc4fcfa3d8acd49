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

inline bool is_pow2(long long v) { return v > 0 && (v & (v - 1)) == 0; }

}  // namespace p2p
