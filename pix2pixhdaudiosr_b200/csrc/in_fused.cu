// InstanceNorm + activation + residual fused with the PADDING of the next convolution (round 2).
//
// The reference chain  ReflectionPad2d -> Conv2d -> InstanceNorm2d -> ReLU  (models/networks.py:190-207, 223-247, 145-160,
// 340-361) ran here as conv -> in_stats -> in_finalize -> in_apply -> pad_planes.  With the statistics accumulated by the conv
// epilogue (p2p_conv_window_*_stats) ONE pass remains: read the fp32 conv output, normalise, activate, add the residual and
// write bf16 directly in the layout the NEXT convolution's TMA reads -- the padded tensor / stride-2 parity planes that
// p2p_pad_planes would have produced, reflection or zero halo included -- plus, if asked, the plain NHWC tensor.
//
// HBM-bound: algorithmic bytes per output element = sizeof(y) + 2 (+ 2 residual) (+ 2 plain copy); halo re-reads are
// (Hh*Wh*s*s)/(H*W) - 1 of that (1.5-6 %).
#include "dtype.cuh"

namespace p2p {
namespace {

__device__ __forceinline__ float act_fwd1(float x, int act, float slope) {
    if (act == 1) return x > 0.f ? x : 0.f;
    if (act == 2) return x > 0.f ? x : slope * x;
    return x;
}
__device__ __forceinline__ void ld8(const float* p, float (&v)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]));
        v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

struct PlanesGeom {
    int N, H, W, C8;            // source tensor [N,H,W,C], C = 8 * C8
    int pt, pl, Hh, Wh, s;      // plane (a,b)[n, yy, xx] = z[n, s*yy + a - pt, s*xx + b - pl]
    long long plane_rows;       // rows of one plane (>= N*Hh*Wh); the tail is zero-filled
    int reflect;
    int R;                      // pixel lanes per block: blockDim.x = C8 * R
    int chunk;                  // plane positions per block (per image)
};

// grid (chunks, N + 1): slice y < N handles image y; slice N zero-fills the plane tails
template <typename TY>
__global__ void __launch_bounds__(256) in_apply_planes_kernel(const TY* __restrict__ y, const float* __restrict__ bias,
                                                              const float* __restrict__ stats_in, int raw_mode,
                                                              float* __restrict__ stats_out, const __nv_bfloat16* __restrict__ res,
                                                              long long res_sn, long long res_sh, long long res_sw,
                                                              __nv_bfloat16* __restrict__ planes, __nv_bfloat16* __restrict__ z_out,
                                                              const PlanesGeom g, int norm, int act, float slope, float eps) {
    const int C8 = g.C8, tid = threadIdx.x;
    const int c8 = tid % C8, lane_r = tid / C8;
    pdl_trigger();
    pdl_wait();
    const int n = blockIdx.y;
    const int per_plane = g.Hh * g.Wh, npos = g.s * g.s * per_plane;
    uint4* dst = reinterpret_cast<uint4*>(planes);
    if (n == g.N) {             // zero tails: rows [N*Hh*Wh, plane_rows) of every plane
        const long long tail = g.plane_rows - (long long)g.N * per_plane;
        const long long total = tail * g.s * g.s * C8;
        for (long long i = (long long)blockIdx.x * blockDim.x + tid; i < total; i += (long long)gridDim.x * blockDim.x) {
            const long long pq = i / (tail * C8), r = i - pq * tail * C8;
            dst[(pq * g.plane_rows + (long long)g.N * per_plane) * C8 + r] = make_uint4(0, 0, 0, 0);
        }
        return;
    }
    float mean[8], rstd[8];
    if (norm) {
        const float* sp = stats_in + ((long long)n * C8 + c8) * 16;
        float st[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(sp) + i);
            st[4 * i] = t.x; st[4 * i + 1] = t.y; st[4 * i + 2] = t.z; st[4 * i + 3] = t.w;
        }
        if (raw_mode) {
            float b8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (bias) ld8(bias + c8 * 8, b8);
            const float inv = 1.f / (float)(g.H * g.W);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float m = st[2 * i] * inv;
                const float var = fmaxf(st[2 * i + 1] * inv - m * m, 0.f);      // biased variance (PyTorch instance_norm)
                mean[i] = m + b8[i];
                rstd[i] = rsqrtf(var + eps);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) { mean[i] = st[2 * i]; rstd[i] = st[2 * i + 1]; }
        }
    }
    const int p0 = blockIdx.x * g.chunk, p1 = min(npos, p0 + g.chunk);
#pragma unroll 2
    for (int pp = p0 + lane_r; pp < p1; pp += g.R) {
        const int pq = pp / per_plane, rem = pp - pq * per_plane;
        const int yy = rem / g.Wh, xx = rem - yy * g.Wh;
        int sy = g.s * yy + pq / g.s - g.pt, sx = g.s * xx + pq % g.s - g.pl;
        const bool interior = sy >= 0 && sy < g.H && sx >= 0 && sx < g.W;
        if (g.reflect) {
            sy = sy < 0 ? -sy : (sy >= g.H ? 2 * (g.H - 1) - sy : sy);
            sx = sx < 0 ? -sx : (sx >= g.W ? 2 * (g.W - 1) - sx : sx);
        }
        uint4 o = make_uint4(0, 0, 0, 0);
        if (sy >= 0 && sy < g.H && sx >= 0 && sx < g.W) {
            const long long src = (((long long)n * g.H + sy) * g.W + sx) * C8 + c8;
            float v[8];
            ld8(y + src * 8, v);
            if (norm) {
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = (v[i] - mean[i]) * rstd[i];
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = act_fwd1(v[i], act, slope);
            if (res) {
                float r[8];
                ld8(res + (long long)n * res_sn + (long long)sy * res_sh + (long long)sx * res_sw + c8 * 8, r);
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] += r[i];
            }
            o = pack8(v);
            if (interior) {
                if (z_out) reinterpret_cast<uint4*>(z_out)[src] = o;
                if (norm && stats_out && raw_mode && sy == 0 && sx == 0) {
                    float* so = stats_out + ((long long)n * C8 + c8) * 16;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        reinterpret_cast<float4*>(so)[i] = make_float4(mean[2 * i], rstd[2 * i], mean[2 * i + 1], rstd[2 * i + 1]);
                }
            }
        }
        dst[((long long)pq * g.plane_rows + (long long)n * per_plane + rem) * C8 + c8] = o;
    }
}

}  // namespace
}  // namespace p2p

using namespace p2p;

// z = act(InstanceNorm(y)) + res written as the next convolution's input planes (bf16): see the header comment of this file.
//   y [N,H,W,C] (y_dtype); bias [C] or NULL and stats_in [N,C,2]: raw_mode 1 = (sum, sum of squares) of (y - bias) as
//   accumulated by p2p_conv_window_*_stats (stats_out [N,C,2] then receives (mean, rstd) for the backward pass),
//   raw_mode 0 = (mean, rstd); res: bf16, element strides (res_sn, res_sh, res_sw) per image / row / pixel, channels
//   contiguous, or NULL; planes: s*s planes of plane_rows rows x C bf16, geometry as p2p_pad_planes; z_out: NULL or the plain
//   bf16 [N,H,W,C] tensor.  C % 8 == 0, C <= 2048.
extern "C" int p2p_instnorm_act_planes(const void* y, const float* bias, const float* stats_in, int raw_mode, float* stats_out,
                                       const void* res, long long res_sn, long long res_sh, long long res_sw, void* planes,
                                       void* z_out, int N, int H, int W, int C, int pt, int pl, int Hh, int Wh, int s,
                                       long long plane_rows, int reflect, int norm, int act, float slope, float eps, int y_dtype,
                                       void* stream) {
    if (!y || !planes || (norm && !stats_in)) return fail(P2P_EINVAL, "null pointer argument");
    if (N < 1 || H < 1 || W < 1 || Hh < 1 || Wh < 1 || s < 1 || s > 2 || pt < 0 || pl < 0) return fail(P2P_EINVAL, "bad plane geometry");
    if (C % 8 != 0 || C > 2048) return fail(P2P_ENOTSUP, "p2p_instnorm_act_planes needs C %% 8 == 0 and C <= 2048, got %d", C);
    if (plane_rows < (long long)N * Hh * Wh) return fail(P2P_EINVAL, "plane_rows smaller than N*Hh*Wh");
    if (reflect && (s != 1 || pt >= H || pl >= W || Hh - H - pt >= H || Wh - W - pl >= W))
        return fail(P2P_ENOTSUP, "reflection padding needs s == 1 and margins smaller than the input");
    if (N >= 65535) return fail(P2P_ENOTSUP, "N=%d too large", N);
    if ((long long)s * s * Hh * Wh > 2000000000LL) return fail(P2P_ENOTSUP, "plane too large");
    if (res && ((res_sn | res_sh | res_sw) % 8 != 0 || (((uintptr_t)res) & 15))) return fail(P2P_ENOTSUP, "residual strides must keep 16-byte alignment");
    if ((((uintptr_t)y) & 15) || (((uintptr_t)planes) & 15) || (((uintptr_t)z_out) & 15)) return fail(P2P_ENOTSUP, "unaligned pointer");
    PlanesGeom g;
    g.N = N; g.H = H; g.W = W; g.C8 = C / 8; g.pt = pt; g.pl = pl; g.Hh = Hh; g.Wh = Wh; g.s = s;
    g.plane_rows = plane_rows; g.reflect = reflect;
    g.R = 256 / g.C8;
    if (g.R < 1) return fail(P2P_ENOTSUP, "C too large");
    const int npos = s * s * Hh * Wh;
    // ~8 blocks per SM over the whole launch, at least 4 positions per pixel lane
    int chunks = (148 * 8 + N - 1) / N;
    int chunk = (npos + chunks - 1) / chunks;
    if (chunk < 4 * g.R) chunk = 4 * g.R;
    chunk = (chunk + g.R - 1) / g.R * g.R;
    chunks = (npos + chunk - 1) / chunk;
    g.chunk = chunk;
    const dim3 grid((unsigned)chunks, (unsigned)(N + 1));
    const int threads = g.C8 * g.R;
    cudaStream_t st = (cudaStream_t)stream;
    if (y_dtype == P2P_F32)
        launch_k(in_apply_planes_kernel<float>, grid, dim3(threads), 0, st, (const float*)y, bias, stats_in, raw_mode, stats_out,
                                                                (const __nv_bfloat16*)res, res_sn, res_sh, res_sw,
                                                                (__nv_bfloat16*)planes, (__nv_bfloat16*)z_out, g, norm, act, slope, eps);
    else if (y_dtype == P2P_BF16)
        launch_k(in_apply_planes_kernel<__nv_bfloat16>, grid, dim3(threads), 0, st, (const __nv_bfloat16*)y, bias, stats_in, raw_mode, stats_out,
                                                                        (const __nv_bfloat16*)res, res_sn, res_sh, res_sw,
                                                                        (__nv_bfloat16*)planes, (__nv_bfloat16*)z_out, g, norm, act, slope, eps);
    else
        return fail(P2P_EINVAL, "unknown dtype code %d", y_dtype);
    return check_launch("instnorm_act_planes");
}
