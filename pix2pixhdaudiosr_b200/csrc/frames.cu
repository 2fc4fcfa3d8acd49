// Time-domain side of the loss path (reference: models/pix2pixHD_model.py:251-258 to_frames, :314-320
// discriminate_time_D, :375-387 time-D losses, :408-415 frame-match loss; models/mdct.py:497-503 the windowed
// frames MDCT2 returns with return_ola).  All HBM-bound element-wise / reduction kernels:
//   frame_window:        audio [B,T] -> windowed frames [B,F,win]  (w[n] * x_pad[f*hop + n])
//   frames_db fwd / bwd: dB(|gain * w[n] * y[b,f,n]|) and its gradient (amplitude_to_DB(|.|, 20, min_value, 1))
//   frame_match:         lambda * mean((y[f, half+i] w[i] - y[f+1, i] w[half+i])^2), value and gradient
//   spectro_decode_bwd:  gradient of p2p_spectro_decode (explicit +/- encoding) back to the generator's image
#include "dtype.cuh"

namespace p2p {
namespace {

constexpr int kTh = 256;
constexpr float kLn10 = 2.302585092994046f;

inline int grid_for(long long n, int cap) {
    long long b = (n + kTh - 1) / kTh;
    if (b > cap) b = cap;
    return b < 1 ? 1 : (int)b;
}

__global__ void __launch_bounds__(kTh) frame_window_kernel(const float* __restrict__ audio, const float* __restrict__ window,
                                                            float* __restrict__ frames, long long T, int F, int hop,
                                                            int win, long long offset, long long total) {
    for (long long i = (long long)blockIdx.x * kTh + threadIdx.x; i < total; i += (long long)gridDim.x * kTh) {
        const int n = (int)(i % win);
        const long long bf = i / win;
        const int f = (int)(bf % F);
        const long long b = bf / F;
        const long long t = (long long)f * hop + n - offset;
        const float x = (t >= 0 && t < T) ? __ldg(audio + b * T + t) : 0.f;
        frames[i] = __ldg(window + n) * x;
    }
}

__global__ void __launch_bounds__(kTh) frames_db_fwd_kernel(const float* __restrict__ y, const float* __restrict__ window,
                                                             float gain, float* __restrict__ out, int N, float amin,
                                                             long long total) {
    for (long long i = (long long)blockIdx.x * kTh + threadIdx.x; i < total; i += (long long)gridDim.x * kTh) {
        const float g = window ? gain * __ldg(window + (int)(i % N)) : gain;
        out[i] = 20.f * log10f(fmaxf(fabsf(g * __ldg(y + i)), amin)) - 20.f;
    }
}

// d/dy of 20 log10(max(|g y|, amin)) = (20 / ln 10) * g * sign(g y) / |g y| where |g y| >= amin, 0 below the clamp
__global__ void __launch_bounds__(kTh) frames_db_bwd_kernel(const float* __restrict__ y, const float* __restrict__ window,
                                                             float gain, const float* __restrict__ dout,
                                                             float* __restrict__ dy, int N, float amin, long long total) {
    for (long long i = (long long)blockIdx.x * kTh + threadIdx.x; i < total; i += (long long)gridDim.x * kTh) {
        const float g = window ? gain * __ldg(window + (int)(i % N)) : gain;
        const float v = g * __ldg(y + i);
        const float a = fabsf(v);
        float d = 0.f;
        if (a >= amin && a > 0.f) d = __ldg(dout + i) * (20.f / kLn10) * g * (v > 0.f ? 1.f : -1.f) / a;
        dy[i] = d;
    }
}

// pairs (f, f+1), i < half:  d = y[b,f,half+i] w[i] - y[b,f+1,i] w[half+i];  loss = wgt * mean(d^2).
// Every element of dy receives at most one term: [f, n >= half] from the pair (f, f+1) (f < F-1), [f, n < half] from
// the pair (f-1, f) (f >= 1); the first half of frame 0 and the second half of frame F-1 get zero.
__global__ void __launch_bounds__(kTh) frame_match_kernel(const float* __restrict__ y, const float* __restrict__ window,
                                                           float wgt, float* __restrict__ loss, float* __restrict__ dy,
                                                           float gcoef, const float* __restrict__ gdev, int F, int N,
                                                           long long total /* B*F*N */, float inv_cnt) {
    __shared__ float scratch[33];
    const int half = N >> 1;
    if (gdev) gcoef *= __ldg(gdev);
    float acc = 0.f;
    for (long long i = (long long)blockIdx.x * kTh + threadIdx.x; i < total; i += (long long)gridDim.x * kTh) {
        const int n = (int)(i % N);
        const int f = (int)((i / N) % F);
        float g = 0.f;
        if (n >= half) {
            if (f < F - 1) {
                const int k = n - half;
                const float d = __ldg(y + i) * __ldg(window + k) - __ldg(y + i + N - half) * __ldg(window + half + k);
                acc += d * d;                                   // each pair is counted once, by its first member
                g = 2.f * gcoef * d * __ldg(window + k);
            }
        } else if (f >= 1) {
            const float d = __ldg(y + i - N + half) * __ldg(window + n) - __ldg(y + i) * __ldg(window + half + n);
            g = -2.f * gcoef * d * __ldg(window + half + n);
        }
        if (dy) dy[i] = g;
    }
    if (loss) {
        acc = block_sum(acc, scratch);
        if (threadIdx.x == 0) atomicAdd(loss, acc * wgt * inv_cnt);
    }
}

// grid (M/32, ceil(F/32), B), block (32, 8): dcoef[b, f, k] -> dx[b, k, f, 0..1] (channels-last fp32)
template <typename T>
__global__ void __launch_bounds__(256) spectro_decode_bwd_kernel(const T* __restrict__ x, long long sb, long long sc,
                                                                 long long sh, long long sw, const float* __restrict__ norm,
                                                                 const float* __restrict__ dcoef, float* __restrict__ dx,
                                                                 int F, int M, float alpha, float gain) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z, k0 = blockIdx.x * 32, f0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
    const float vmax = norm[0], vmin = norm[1];
    for (int r = ty; r < 32; r += 8) {
        const int f = f0 + r, k = k0 + tx;
        if (k < M && f < F) tile[r][tx] = __ldg(dcoef + ((long long)b * F + f) * M + k);
    }
    __syncthreads();
    // amp(x) = 10 * 10^(0.05 s) - min_value, s = |x| (max - min) + min  =>  d amp / dx = 0.05 ln10 (max - min) sign(x) * 10 * 10^(0.05 s)
    const float c = gain / (2.f * alpha - 1.f) * 0.05f * kLn10 * (vmax - vmin);
    for (int r = ty; r < 32; r += 8) {
        const int k = k0 + r, f = f0 + tx;
        if (k < M && f < F) {
            const T* p = x + (long long)b * sb + (long long)k * sh + (long long)f * sw;
            const float x0 = ldf(p), x1 = ldf(p + sc);
            const float e0 = 10.f * exp10f(0.05f * (fabsf(x0) * (vmax - vmin) + vmin));
            const float e1 = 10.f * exp10f(0.05f * (fabsf(x1) * (vmax - vmin) + vmin));
            const float g = tile[tx][r] * c;
            const float s0 = x0 > 0.f ? 1.f : (x0 < 0.f ? -1.f : 0.f), s1 = x1 > 0.f ? 1.f : (x1 < 0.f ? -1.f : 0.f);
            reinterpret_cast<float2*>(dx)[((long long)b * M + k) * F + f] = make_float2(g * e0 * s0, -g * e1 * s1);
        }
    }
}

}  // namespace
}  // namespace p2p

extern "C" int p2p_frame_window(const float* audio, const float* window, float* frames, int B, long long T, int F, int hop,
                                int win, long long offset, void* stream) {
    using namespace p2p;
    if (B < 0 || T < 0 || F < 0 || hop < 1 || win < 1) return fail(P2P_EINVAL, "bad frame geometry");
    const long long total = (long long)B * F * win;
    if (total == 0) return P2P_OK;
    if (!audio || !window || !frames) return fail(P2P_EINVAL, "null pointer argument");
    frame_window_kernel<<<grid_for(total, 148 * 8), kTh, 0, (cudaStream_t)stream>>>(audio, window, frames, T, F, hop, win, offset, total);
    return check_launch("frame_window_kernel");
}

extern "C" int p2p_frames_db_fwd(const float* frames, const float* window, float gain, float* out, long long rows, int N,
                                 float min_value, void* stream) {
    using namespace p2p;
    if (rows < 0 || N < 1) return fail(P2P_EINVAL, "bad frame geometry");
    const long long total = rows * N;
    if (total == 0) return P2P_OK;
    if (!frames || !out) return fail(P2P_EINVAL, "null pointer argument");
    frames_db_fwd_kernel<<<grid_for(total, 148 * 8), kTh, 0, (cudaStream_t)stream>>>(frames, window, gain, out, N, min_value, total);
    return check_launch("frames_db_fwd_kernel");
}

extern "C" int p2p_frames_db_bwd(const float* frames, const float* window, float gain, const float* dout, float* dframes,
                                 long long rows, int N, float min_value, void* stream) {
    using namespace p2p;
    if (rows < 0 || N < 1) return fail(P2P_EINVAL, "bad frame geometry");
    const long long total = rows * N;
    if (total == 0) return P2P_OK;
    if (!frames || !dout || !dframes) return fail(P2P_EINVAL, "null pointer argument");
    frames_db_bwd_kernel<<<grid_for(total, 148 * 8), kTh, 0, (cudaStream_t)stream>>>(frames, window, gain, dout, dframes, N, min_value, total);
    return check_launch("frames_db_bwd_kernel");
}

extern "C" int p2p_frame_match(const float* frames, const float* window, float weight, float* loss, float* dframes,
                               float gscale, const float* gscale_dev, int B, int F, int N, void* stream) {
    using namespace p2p;
    if (B < 1 || F < 2 || N < 2 || (N & 1)) return fail(P2P_EINVAL, "frame match needs B >= 1, F >= 2 and an even frame length (B=%d F=%d N=%d)", B, F, N);
    if (!frames || !window) return fail(P2P_EINVAL, "null pointer argument");
    const long long total = (long long)B * F * N;
    const float inv_cnt = 1.f / ((float)B * (float)(F - 1) * (float)(N / 2));
    frame_match_kernel<<<(loss && deterministic()) ? 1 : grid_for(total, 592), kTh, 0, (cudaStream_t)stream>>>(frames, window, weight, loss, dframes,
                                                                             gscale * weight * inv_cnt, gscale_dev, F, N, total, inv_cnt);
    return check_launch("frame_match_kernel");
}

extern "C" int p2p_spectro_decode_bwd(const void* x, long long sb, long long sc, long long sh, long long sw, const float* norm,
                                      const float* dcoef, float* dx, int B, int F, int M, float alpha, float gain, int dtype,
                                      void* stream) {
    using namespace p2p;
    if (!x || !norm || !dcoef || !dx) return fail(P2P_EINVAL, "null pointer argument");
    if (B < 1 || F < 1 || M < 1 || B > 65535) return fail(P2P_EINVAL, "bad spectrogram geometry");
    dim3 grid((M + 31) / 32, (F + 31) / 32, B), block(32, 8);
    cudaStream_t st = (cudaStream_t)stream;
    P2P_DISPATCH_DTYPE(dtype, T, spectro_decode_bwd_kernel<T><<<grid, block, 0, st>>>((const T*)x, sb, sc, sh, sw, norm, dcoef, dx, F, M, alpha, gain));
    return check_launch("spectro_decode_bwd_kernel");
}
