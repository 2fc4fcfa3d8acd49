// Shared helpers for the C-ABI translation units (error slot, launch checks).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
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

inline bool is_pow2(long long v) { return v > 0 && (v & (v - 1)) == 0; }

}  // namespace p2p
