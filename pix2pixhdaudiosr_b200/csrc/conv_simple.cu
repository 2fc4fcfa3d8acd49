// Implicit-GEMM convolution passes, CUDA-core (FFMA) version: exact fp32 accumulation for every layer shape
// of models/networks.py (7x7 reflect, 3x3 s1 reflect, 3x3 s2 zero-pad, 4x4 s2/s1 pad 2, ConvTranspose 3x3 s2 as
// four parity passes) and their input / weight gradients.  This is the reference-accuracy path (dtype f32) and
// the fallback-free baseline that the tcgen05 kernels (conv_umma.cu) are checked against; see DESIGN.md §6.
#include "dtype.cuh"

namespace p2p {
namespace {

constexpr int BM = 64, BN = 64, BK = 16, kT = 256;

struct RowCoord { int n, y, x; bool ok; };

__device__ __forceinline__ RowCoord decode_row(const p2p_conv_pass& P, long long m, long long M) {
    RowCoord r;
    r.ok = m < M;
    const long long mm = r.ok ? m : 0;
    r.x = (int)(mm % P.gx);
    const long long q = mm / P.gx;
    r.y = (int)(q % P.gy);
    r.n = (int)(q / P.gy);
    return r;
}

// input pixel offset (in pixels, not elements) of tap t for row coordinate rc, or -1 when it reads zero padding
__device__ __forceinline__ long long gather_pixel(const p2p_conv_pass& P, const RowCoord& rc, int t) {
    int ih = rc.y * P.sy + P.dh[t], iw = rc.x * P.sx + P.dw[t];
    if (P.pad_mode == P2P_PAD_REFLECT) {
        ih = ih < 0 ? -ih : (ih >= P.H ? 2 * (P.H - 1) - ih : ih);
        iw = iw < 0 ? -iw : (iw >= P.W ? 2 * (P.W - 1) - iw : iw);
    }
    if (ih < 0 || ih >= P.H || iw < 0 || iw >= P.W) return -1;
    return ((long long)rc.n * P.H + ih) * P.W + iw;
}

template <typename T, typename TO>
__global__ void __launch_bounds__(kT)
conv_pass_simple(const __grid_constant__ p2p_conv_pass P, const T* __restrict__ in, const T* __restrict__ w,
                 const float* __restrict__ bias, TO* __restrict__ out) {
    __shared__ float As[BK][BM + 4], Bs[BK][BN + 4];
    const long long M = (long long)P.N * P.gy * P.gx;
    const int Ktot = P.ntaps * P.K;
    const long long m0 = (long long)blockIdx.x * BM;
    const int j0 = blockIdx.y * BN;
    const int tid = threadIdx.x, ty = tid / 16, tx = tid % 16;

    const int lr = tid / 4, lk = (tid % 4) * 4;           // A-load role: row lr, k offsets lk..lk+3
    const RowCoord rc = decode_row(P, m0 + lr, M);
    const int bk = tid / 16, bj = (tid % 16) * 4;         // B-load role: k row bk, columns bj..bj+3

    float acc[4][4] = {};
    for (int kb = 0; kb < Ktot; kb += BK) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int kk = kb + lk + i;
            float v = 0.f;
            if (rc.ok && kk < Ktot) {
                const int t = kk / P.K, c = kk - t * P.K;
                const long long px = gather_pixel(P, rc, t);
                if (px >= 0) v = ldf(in + px * P.K + c);
            }
            As[lk + i][lr] = v;
        }
        {
            const int kk = kb + bk;
            const int t = kk < Ktot ? kk / P.K : 0, c = kk - t * P.K;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int j = j0 + bj + i;
                float v = 0.f;
                if (kk < Ktot && j < P.J) v = ldf(w + (long long)P.wtap[t] * P.w_tap + (long long)c * P.w_k + (long long)j * P.w_j);
                Bs[bk][bj + i] = v;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(av[r], bv[c], acc[r][c]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const RowCoord oc = decode_row(P, m0 + ty * 4 + r, M);
        if (!oc.ok) continue;
        const long long opx = ((long long)oc.n * P.OH + (oc.y * P.oys + P.oyo)) * P.OW + (oc.x * P.oxs + P.oxo);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int j = j0 + tx * 4 + c;
            if (j < P.J) stf(out + opx * P.J + j, acc[r][c] + (bias ? bias[j] : 0.f));
        }
    }
}

// dW_t[k, j] += sum_m in[gather(m,t), k] * dout[orow(m), j].  grid (ceil(K/64), ceil(J/64), ntaps*S)
template <typename T>
__global__ void __launch_bounds__(kT)
conv_wgrad_simple(const __grid_constant__ p2p_conv_pass P, const T* __restrict__ in, const T* __restrict__ dout,
                  float* __restrict__ dw, int S, long long mchunk) {
    __shared__ float As[BK][BM + 4], Ds[BK][BN + 4];
    const long long M = (long long)P.N * P.gy * P.gx;
    const int t = blockIdx.z / S, s = blockIdx.z % S;
    const int k0 = blockIdx.x * BM, j0 = blockIdx.y * BN;
    const long long mbeg = (long long)s * mchunk, mend = min(M, mbeg + mchunk);
    const int tid = threadIdx.x, ty = tid / 16, tx = tid % 16;
    const int lm = tid / 16, lc = (tid % 16) * 4;          // load role: row lm of the 16-row m-slab, 4 channels

    float acc[4][4] = {};
    for (long long mb = mbeg; mb < mend; mb += BK) {
        const long long m = mb + lm;
        const RowCoord rc = decode_row(P, m, mend);
        long long px = -1, opx = 0;
        if (rc.ok) {
            px = gather_pixel(P, rc, t);
            opx = ((long long)rc.n * P.OH + (rc.y * P.oys + P.oyo)) * P.OW + (rc.x * P.oxs + P.oxo);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int k = k0 + lc + i, j = j0 + lc + i;
            As[lm][lc + i] = (px >= 0 && k < P.K) ? ldf(in + px * P.K + k) : 0.f;
            Ds[lm][lc + i] = (rc.ok && j < P.J) ? ldf(dout + opx * P.J + j) : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Ds[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(av[r], bv[c], acc[r][c]);
        }
        __syncthreads();
    }
    float* dwt = dw + (long long)P.wtap[t] * P.w_tap;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int k = k0 + ty * 4 + r;
        if (k >= P.K) continue;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int j = j0 + tx * 4 + c;
            if (j < P.J) atomicAdd(dwt + (long long)k * P.w_k + (long long)j * P.w_j, acc[r][c]);
        }
    }
}

}  // namespace

int check_pass(const p2p_conv_pass* P) {
    if (!P) return fail(P2P_EINVAL, "null pass descriptor");
    if (P->ntaps < 1 || P->ntaps > P2P_MAX_TAPS) return fail(P2P_EINVAL, "ntaps %d out of range", P->ntaps);
    if (P->N < 0 || P->H < 1 || P->W < 1 || P->K < 1 || P->J < 1 || P->gy < 0 || P->gx < 0 || P->OH < 1 || P->OW < 1)
        return fail(P2P_EINVAL, "bad conv pass geometry");
    if (P->pad_mode != P2P_PAD_ZERO && P->pad_mode != P2P_PAD_REFLECT) return fail(P2P_EINVAL, "bad pad mode");
    if (P->gy > 0 && ((P->gy - 1) * P->oys + P->oyo >= P->OH || P->oyo < 0)) return fail(P2P_EINVAL, "output rows exceed OH");
    if (P->gx > 0 && ((P->gx - 1) * P->oxs + P->oxo >= P->OW || P->oxo < 0)) return fail(P2P_EINVAL, "output cols exceed OW");
    return P2P_OK;
}

// entry points of the CUDA-core path; conv_dispatch.cu routes to these or to the tcgen05 kernels
int conv_forward_simple(const p2p_conv_pass* P, const void* in, const void* weight, const float* bias, void* out,
                        int dtype, int out_dtype, cudaStream_t st) {
    if (int rc = check_pass(P)) return rc;
    const long long M = (long long)P->N * P->gy * P->gx;
    if (M == 0) return P2P_OK;
    if (!in || !weight || !out) return fail(P2P_EINVAL, "null pointer argument");
    const long long gm = (M + BM - 1) / BM;
    if (gm > 2147483647LL) return fail(P2P_ENOTSUP, "too many rows");
    dim3 grid((unsigned)gm, (P->J + BN - 1) / BN);
    P2P_DISPATCH_DTYPE2(dtype, T, out_dtype, TO, conv_pass_simple<T, TO><<<grid, kT, 0, st>>>(*P, (const T*)in, (const T*)weight, bias, (TO*)out));
    return check_launch("conv_pass_simple");
}

int conv_wgrad_simple_launch(const p2p_conv_pass* P, const void* in, const void* dout, float* dw, int dtype, cudaStream_t st) {
    if (int rc = check_pass(P)) return rc;
    const long long M = (long long)P->N * P->gy * P->gx;
    if (M == 0) return P2P_OK;
    if (!in || !dout || !dw) return fail(P2P_EINVAL, "null pointer argument");
    const int gk = (P->K + BM - 1) / BM, gj = (P->J + BN - 1) / BN;
    long long S = (148LL * 4) / ((long long)gk * gj * P->ntaps);
    const long long maxS = (M + 255) / 256;
    if (S > maxS) S = maxS;
    if (S < 1 || deterministic()) S = 1;        // one CTA per (tap, k tile, j tile): its atomicAdd is the only write of the call
    if ((long long)P->ntaps * S > 65535) S = 65535 / P->ntaps;
    long long mchunk = (M + S - 1) / S;
    mchunk = (mchunk + BK - 1) / BK * BK;
    dim3 grid(gk, gj, (unsigned)(P->ntaps * S));
    P2P_DISPATCH_DTYPE(dtype, T, conv_wgrad_simple<T><<<grid, kT, 0, st>>>(*P, (const T*)in, (const T*)dout, dw, (int)S, mchunk));
    return check_launch("conv_wgrad_simple");
}

int conv_forward_umma(const p2p_conv_pass* P, const void* in, const void* weight, const float* bias, void* out, int dtype,
                      int out_dtype, cudaStream_t st);
int conv_wgrad_umma(const p2p_conv_pass* P, const void* in, const void* dout, float* dw, int dtype, cudaStream_t st);

}  // namespace p2p

extern "C" int p2p_conv_forward(const p2p_conv_pass* pass, const void* in, const void* weight, const float* bias,
                                void* out, int dtype, int out_dtype, void* stream) {
    if (int rc = p2p::check_pass(pass)) return rc;
    if (!in || !weight || !out) return p2p::fail(P2P_EINVAL, "null pointer argument");
    const int rc = p2p::conv_forward_umma(pass, in, weight, bias, out, dtype, out_dtype, (cudaStream_t)stream);
    if (rc <= 0) return rc;       // handled by the tcgen05 kernels (or failed there)
    return p2p::conv_forward_simple(pass, in, weight, bias, out, dtype, out_dtype, (cudaStream_t)stream);
}

extern "C" int p2p_conv_wgrad(const p2p_conv_pass* pass, const void* in, const void* dout, float* dweight, int dtype,
                              void* stream) {
    if (int rc = p2p::check_pass(pass)) return rc;
    if (!in || !dout || !dweight) return p2p::fail(P2P_EINVAL, "null pointer argument");
    const int rc = p2p::conv_wgrad_umma(pass, in, dout, dweight, dtype, (cudaStream_t)stream);
    if (rc <= 0) return rc;
    return p2p::conv_wgrad_simple_launch(pass, in, dout, dweight, dtype, (cudaStream_t)stream);
}
