// Shared-memory FFT used by the general (any-geometry) MDCT4 kernels and by the MDCT2 / DCT-II kernels.
#pragma once
#include <cuda_runtime.h>

namespace p2p {
namespace smemfft {

constexpr int kThreads = 256;

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

// In-place decimation-in-frequency FFT of `nfr` independent L-point rows in shared memory
// (row stride L), radix-4 passes plus one radix-2 pass when log2(L) is odd.  Natural-order input,
// digit-reversed output (Tables::perm).  All kThreads threads cooperate; ends with __syncthreads().
static __device__ void fft_rows_inplace(float2* z, int nfr, int L, const float2* __restrict__ roots) {
    int S = L;
    while (S >= 4) {
        const int q = S >> 2;
        const int per_row = L >> 2;
        const int tw_step = L / S;
        for (int w = threadIdx.x; w < nfr * per_row; w += kThreads) {
            const int row = w / per_row;
            const int r = w - row * per_row;
            const int blk = r / q;
            const int j = r - blk * q;
            float2* p = z + row * L + blk * S + j;
            const float2 a0 = p[0], a1 = p[q], a2 = p[2 * q], a3 = p[3 * q];
            const float2 s02 = make_float2(a0.x + a2.x, a0.y + a2.y);
            const float2 d02 = make_float2(a0.x - a2.x, a0.y - a2.y);
            const float2 s13 = make_float2(a1.x + a3.x, a1.y + a3.y);
            const float2 d13 = make_float2(a1.x - a3.x, a1.y - a3.y);
            const float2 y0 = make_float2(s02.x + s13.x, s02.y + s13.y);
            const float2 y2 = make_float2(s02.x - s13.x, s02.y - s13.y);
            // y1 = d02 - j*d13 ; y3 = d02 + j*d13
            const float2 y1 = make_float2(d02.x + d13.y, d02.y - d13.x);
            const float2 y3 = make_float2(d02.x - d13.y, d02.y + d13.x);
            p[0] = y0;
            if (j == 0) {
                p[q] = y1; p[2 * q] = y2; p[3 * q] = y3;
            } else {
                p[q] = cmul(y1, roots[tw_step * j]);
                p[2 * q] = cmul(y2, roots[tw_step * 2 * j]);
                p[3 * q] = cmul(y3, roots[tw_step * 3 * j]);
            }
        }
        __syncthreads();
        S = q;
    }
    if (S == 2) {
        const int per_row = L >> 1;
        for (int w = threadIdx.x; w < nfr * per_row; w += kThreads) {
            const int row = w / per_row;
            const int r = w - row * per_row;
            float2* p = z + row * L + 2 * r;
            const float2 a0 = p[0], a1 = p[1];
            p[0] = make_float2(a0.x + a1.x, a0.y + a1.y);
            p[1] = make_float2(a0.x - a1.x, a0.y - a1.y);
        }
        __syncthreads();
    }
}


// position -> bin map of the in-place DIF passes (radix 4,...,4[,2]): the pass over sub-size S sends output digit
// k_i to block k_i of size S/r, and bin = k_1 + r_1 (k_2 + r_2 (...)).  Host side.
inline int dif_position_to_bin(int p, int L) {
    int S = L, rem = p, k = 0, weight = 1;
    while (S >= 4) { const int q = S / 4; k += (rem / q) * weight; rem %= q; weight *= 4; S = q; }
    if (S == 2) { k += rem * weight; }
    return k;
}

}  // namespace smemfft
}  // namespace p2p
