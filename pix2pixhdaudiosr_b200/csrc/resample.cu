// Polyphase windowed-sinc resampling on the device: the arithmetic of torchaudio.functional.resample
// (what data/audio_dataset.py:56-58 calls three times per item on CPU workers): after reducing orig/new by their gcd,
//     out[b, i*new + j] = sum_{k < K} kernel[j, k] * x_pad[b, i*orig + k],   K = 2*width + orig,
// x_pad = x shifted right by `width` zeros and zero-extended; the [new, K] filter bank is built on the host in double
// (pix2pixhdaudiosr_b200/data/resample.py restates torchaudio's _get_sinc_resample_kernel).  HBM-bound: each output
// sample is one K-tap dot product whose inputs are shared with its neighbours through L1.
#include "common.cuh"
#include <cstdint>

namespace p2p {
namespace {

constexpr int kTh = 256;
constexpr int kMaxBank = 28160;      // floats of filter bank staged in shared memory (110 KB: two CTAs per SM)

__global__ void __launch_bounds__(kTh) resample_kernel(const float* __restrict__ x, const float* __restrict__ bank,
                                                       float* __restrict__ out, long long T, long long T_out, int orig,
                                                       int nw, int width, int K, int bank_in_smem) {
    extern __shared__ float sbank[];
    if (bank_in_smem) {
        for (int i = threadIdx.x; i < nw * K; i += kTh) sbank[i] = __ldg(bank + i);
        __syncthreads();
    }
    const float* bk = bank_in_smem ? sbank : bank;
    const float* xb = x + (long long)blockIdx.y * T;
    float* ob = out + (long long)blockIdx.y * T_out;
    for (long long o = (long long)blockIdx.x * kTh + threadIdx.x; o < T_out; o += (long long)gridDim.x * kTh) {
        const long long i = o / nw;
        const int j = (int)(o - i * nw);
        const long long t0 = i * orig - width;
        const float* kj = bk + j * K;
        float acc = 0.f;
        for (int k = 0; k < K; ++k) {
            const long long t = t0 + k;
            if (t >= 0 && t < T) acc += kj[k] * __ldg(xb + t);
        }
        ob[o] = acc;
    }
}

}  // namespace
}  // namespace p2p

extern "C" int p2p_resample(const float* x, const float* bank, float* out, int B, long long T, long long T_out, int orig,
                            int new_freq, int width, void* stream) {
    using namespace p2p;
    if (B < 0 || T < 0 || T_out < 0 || orig < 1 || new_freq < 1 || width < 0)
        return fail(P2P_EINVAL, "bad resample geometry B=%d T=%lld T_out=%lld orig=%d new=%d width=%d", B, T, T_out, orig, new_freq, width);
    if (B == 0 || T_out == 0) return P2P_OK;
    if (!bank || !out || (!x && T > 0)) return fail(P2P_EINVAL, "null pointer argument");
    if (B > 65535) return fail(P2P_ENOTSUP, "B=%d exceeds 65535 rows per launch", B);
    const int K = 2 * width + orig;
    // every output index must exist in the conv1d the reference runs: i <= (T + 2*width + orig - K) / orig = T / orig
    if ((T_out - 1) / new_freq > T / orig) return fail(P2P_EINVAL, "T_out=%lld exceeds what %lld input samples produce", T_out, T);
    const long long words = (long long)new_freq * K;
    const int in_smem = words <= kMaxBank;
    // a CTA stages the whole bank: with a large bank (e.g. 44.1 -> 48 kHz: 160 phases x 161 taps = 103 KB) it has to
    // amortise that over many outputs (64 per thread), with a small one (48 <-> 8 kHz: 80-90 words) one per thread
    const long long per_cta = (long long)kTh * (words > 2048 ? 64 : 1);
    long long blocks = (T_out + per_cta - 1) / per_cta;
    if (blocks > 148 * 16) blocks = 148 * 16;
    dim3 grid((unsigned)blocks, B);
    if (in_smem && words * 4 > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(resample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(words * 4));
        if (e != cudaSuccess) return fail(P2P_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    resample_kernel<<<grid, kTh, in_smem ? words * 4 : 0, (cudaStream_t)stream>>>(x, bank, out, T, T_out, orig, new_freq, width, K, in_smem);
    return check_launch("resample_kernel");
}
