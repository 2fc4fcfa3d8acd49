// Operand packing for the thin layers (2-4 channels on one side: G's 7x7 edge layers, D's first and logit layers):
// the kernel's columns are folded into the channel dimension so that (dw, c) fills <= 16 tensor-core channels and a
// KHxKW convolution becomes a KHx1 convolution (see ops.py "thin layers").  One thread per packed pixel (32 bytes).
#include "dtype.cuh"

namespace p2p {
namespace {

__device__ __forceinline__ int reflect_idx(int i, int n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

// xpk[n, y, ox, dw*C + c] = xpad[n, y, ox*s + dw, c]  (xpad = x padded by p, zeros or reflection); channels >= KW*C are zero
template <typename T>
__global__ void __launch_bounds__(256) pack_in_kernel(const T* __restrict__ x, T* __restrict__ xpk, int N, int H, int W, int C,
                                                      int KW, int s, int p, int reflect, int OW, unsigned total) {
    pdl_trigger();
    pdl_wait();
    const int Hp = H + 2 * p;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int ox = (int)(i % OW);
        unsigned r = i / OW;
        const int y = (int)(r % Hp);
        const int n = (int)(r / Hp);
        int sy = y - p;
        bool rowok = true;
        if (reflect) sy = reflect_idx(sy, H); else rowok = sy >= 0 && sy < H;
        T v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = T(0.f);
        const T* row = x + ((long long)n * H + sy) * W * C;
        for (int dw = 0; dw < KW; ++dw) {
            int sx = ox * s + dw - p;
            bool ok = rowok;
            if (reflect) sx = reflect_idx(sx, W); else ok = ok && sx >= 0 && sx < W;
            if (ok)
                for (int c = 0; c < C; ++c) v[dw * C + c] = row[(long long)sx * C + c];
        }
        T* o = xpk + (long long)i * 16;
#pragma unroll
        for (int k = 0; k < 16; ++k) o[k] = v[k];
    }
}

// y[n, oh, ow, co] = bias[co] + sum_dw ypk[n, oh, ow + dw, dw*Cout + co]
template <typename T>
__global__ void __launch_bounds__(256) fold_out_kernel(const float* __restrict__ ypk, const float* __restrict__ bias, T* __restrict__ y,
                                                       int OH, int Wp, int OW, int KW, int Cout, unsigned total) {
    pdl_trigger();
    pdl_wait();
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int co = (int)(i % Cout);
        unsigned r = i / Cout;
        const int ow = (int)(r % OW);
        r /= OW;       // = n * OH + oh
        const float* src = ypk + ((long long)r * Wp + ow) * 16 + co;
        float s = bias ? bias[co] : 0.f;
        for (int dw = 0; dw < KW; ++dw) s += src[dw * (16 + Cout)];
        stf(y + i, s);
    }
}

// row-packed form: y[n, oh, ow, co] = bias[co] + sum_dh ypk[n, oh + dh, ow, dh*Cout + co]     ypk: fp32 [N][Hp][Wy][16]
template <typename T>
__global__ void __launch_bounds__(256) fold_rows_kernel(const float* __restrict__ ypk, const float* __restrict__ bias, T* __restrict__ y,
                                                        int Hp, int Wy, int OH, int OW, int KH, int Cout, unsigned total) {
    pdl_trigger();
    pdl_wait();
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int co = (int)(i % Cout);
        unsigned r = i / Cout;
        const int ow = (int)(r % OW);
        r /= OW;
        const int oh = (int)(r % OH);
        const unsigned n = r / OH;
        const float* src = ypk + (((long long)n * Hp + oh) * Wy + ow) * 16 + co;
        float s = bias ? bias[co] : 0.f;
        for (int dh = 0; dh < KH; ++dh) s += src[(long long)dh * Wy * 16 + dh * Cout];
        stf(y + i, s);
    }
}

// dypk[n, y, x', dw*Cout + co] = dy[n, y, x' - dw, co] (zero outside), 16 channels per pixel
template <typename T>
__global__ void __launch_bounds__(256) pack_dy_kernel(const T* __restrict__ dy, T* __restrict__ dypk, int OW, int Cout, int KW, int Wp,
                                                      unsigned total) {
    pdl_trigger();
    pdl_wait();
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int xp = (int)(i % Wp);
        const unsigned r = i / Wp;     // = n * OH + y
        T v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = T(0.f);
        const T* row = dy + (long long)r * OW * Cout;
        for (int dw = 0; dw < KW; ++dw) {
            const int sx = xp - dw;
            if (sx >= 0 && sx < OW)
                for (int c = 0; c < Cout; ++c) v[dw * Cout + c] = row[sx * Cout + c];
        }
        T* o = dypk + (long long)i * 16;
#pragma unroll
        for (int k = 0; k < 16; ++k) o[k] = v[k];
    }
}


// ---- full im2col forms: the whole KHxKWxC patch (<= 128 values) becomes the channel vector of a 1x1 convolution -------
// xcol[q, (kh,kw,c)] = pad(x)[n, oy*s + kh, ox*s + kw, c], q = (n, oy, ox) < rows; rows .. rows_alloc and k >= KH*KW*C are zero.
// One thread per 8 consecutive k (16 bytes).
__global__ void __launch_bounds__(256) im2col_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ xcol, int N, int H,
                                                     int W, int C, int KH, int KW, int s, int p, int reflect, int OH, int OW, int Kc,
                                                     long long rows, unsigned total) {
    pdl_trigger();
    pdl_wait();
    const int G = Kc >> 3, KWC = KW * C, KK = KH * KWC;
    __shared__ int tab[128];            // k -> kh << 16 | kw << 8 | c
    for (int k = threadIdx.x; k < 128; k += blockDim.x) {
        const int kh = k / KWC, rem = k - kh * KWC, kw = rem / C;
        tab[k] = (kh << 16) | (kw << 8) | (rem - kw * C);
    }
    __syncthreads();
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int g = (int)(i % G);
        const unsigned q = i / G;
        __align__(16) __nv_bfloat16 v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __nv_bfloat16(0.f);
        if (q < rows) {
            const int ox = (int)(q % OW);
            const unsigned r = q / OW;
            const int oy = (int)(r % OH), n = (int)(r / OH);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int k = g * 8 + e;
                if (k < KK) {
                    const int t = tab[k], kh = t >> 16, kw = (t >> 8) & 255, c = t & 255;
                    int iy = oy * s + kh - p, ix = ox * s + kw - p;
                    bool ok = true;
                    if (reflect) { iy = reflect_idx(iy, H); ix = reflect_idx(ix, W); }
                    else ok = iy >= 0 && iy < H && ix >= 0 && ix < W;
                    if (ok) v[e] = x[(((long long)n * H + iy) * W + ix) * C + c];
                }
            }
        }
        *reinterpret_cast<uint4*>(xcol + (long long)i * 8) = *reinterpret_cast<const uint4*>(v);
    }
}

// adjoint of im2col with zero padding: dx[n, y, x, c] = sum over (kh, kw) with (y + p - kh, x + p - kw) = s * (oy, ox)
__global__ void __launch_bounds__(256) col2im_kernel(const __nv_bfloat16* __restrict__ dxcol, __nv_bfloat16* __restrict__ dx, int H, int W,
                                                     int C, int KH, int KW, int s, int p, int OH, int OW, int Kc, unsigned total) {
    pdl_trigger();
    pdl_wait();
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        unsigned r = i / C;
        const int x = (int)(r % W); r /= W;
        const int y = (int)(r % H);
        const long long n = r / H;
        float acc = 0.f;
        for (int kh = 0; kh < KH; ++kh) {
            const int ty = y + p - kh;
            if (ty < 0 || ty % s != 0 || ty / s >= OH) continue;
            for (int kw = 0; kw < KW; ++kw) {
                const int tx = x + p - kw;
                if (tx < 0 || tx % s != 0 || tx / s >= OW) continue;
                acc += __bfloat162float(dxcol[((n * OH + ty / s) * OW + tx / s) * Kc + (kh * KW + kw) * C + c]);
            }
        }
        dx[i] = __float2bfloat16_rn(acc);
    }
}

// dyfull[q, (kh,kw,co)] = dy[n, yp - kh, xp - kw, co] on the padded grid q = (n, yp, xp) (zero outside / beyond rows / k >= KH*KW*Cout)
__global__ void __launch_bounds__(256) pack_dy_full_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ out, int OH,
                                                           int OW, int Cout, int KH, int KW, int Hp, int Wp, int Kc, long long rows,
                                                           unsigned total) {
    pdl_trigger();
    pdl_wait();
    const int G = Kc >> 3, KWC = KW * Cout, KK = KH * KWC;
    __shared__ int tab[128];            // k -> kh << 16 | kw << 8 | co
    for (int k = threadIdx.x; k < 128; k += blockDim.x) {
        const int kh = k / KWC, rem = k - kh * KWC, kw = rem / Cout;
        tab[k] = (kh << 16) | (kw << 8) | (rem - kw * Cout);
    }
    __syncthreads();
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int g = (int)(i % G);
        const unsigned q = i / G;
        __align__(16) __nv_bfloat16 v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __nv_bfloat16(0.f);
        if (q < rows) {
            const int xp = (int)(q % Wp);
            const unsigned r = q / Wp;
            const int yp = (int)(r % Hp), n = (int)(r / Hp);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int k = g * 8 + e;
                if (k < KK) {
                    const int t = tab[k], kh = t >> 16, kw = (t >> 8) & 255, co = t & 255;
                    const int sy = yp - kh, sx = xp - kw;
                    if (sy >= 0 && sy < OH && sx >= 0 && sx < OW) v[e] = dy[(((long long)n * OH + sy) * OW + sx) * Cout + co];
                }
            }
        }
        *reinterpret_cast<uint4*>(out + (long long)i * 8) = *reinterpret_cast<const uint4*>(v);
    }
}

inline int grid_for(long long n) {
    long long b = (n + 255) / 256;
    return (int)(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}

}  // namespace
}  // namespace p2p

using namespace p2p;

extern "C" int p2p_pack_in(const void* x, void* xpk, int N, int H, int W, int C, int KW, int stride, int pad, int reflect,
                           int OW, int dtype, void* stream) {
    if (!x || !xpk) return fail(P2P_EINVAL, "null pointer argument");
    if (N < 1 || H < 1 || W < 1 || C < 1 || KW < 1 || KW * C > 16 || stride < 1 || pad < 0 || OW < 1)
        return fail(P2P_EINVAL, "bad packing geometry (KW*C must be <= 16)");
    if (reflect && (pad >= H || pad >= W)) return fail(P2P_EINVAL, "reflection padding must be smaller than the input");
    if ((OW - 1) * stride + KW > W + 2 * pad) return fail(P2P_EINVAL, "OW reaches past the padded row");
    const long long total = (long long)N * (H + 2 * pad) * OW;
    if (total > 4000000000LL) return fail(P2P_ENOTSUP, "tensor too large");
    P2P_DISPATCH_DTYPE(dtype, T, launch_k(pack_in_kernel<T>, dim3((unsigned)(grid_for(total))), dim3(256), 0, (cudaStream_t)stream, 
        (const T*)x, (T*)xpk, N, H, W, C, KW, stride, pad, reflect, OW, (unsigned)total));
    return check_launch("pack_in");
}

extern "C" int p2p_fold_out(const float* ypk, const float* bias, void* y, int N, int OH, int Wp, int OW, int KW, int Cout,
                            int out_dtype, void* stream) {
    if (!ypk || !y) return fail(P2P_EINVAL, "null pointer argument");
    if (N < 1 || OH < 1 || OW < 1 || KW < 1 || Cout < 1 || KW * Cout > 16 || OW + KW - 1 > Wp)
        return fail(P2P_EINVAL, "bad folding geometry (KW*Cout must be <= 16)");
    const long long total = (long long)N * OH * OW * Cout;
    if (total > 4000000000LL) return fail(P2P_ENOTSUP, "tensor too large");
    P2P_DISPATCH_DTYPE(out_dtype, T, launch_k(fold_out_kernel<T>, dim3((unsigned)(grid_for(total))), dim3(256), 0, (cudaStream_t)stream, 
        ypk, bias, (T*)y, OH, Wp, OW, KW, Cout, (unsigned)total));
    return check_launch("fold_out");
}

extern "C" int p2p_fold_out_rows(const float* ypk, const float* bias, void* y, int N, int Hp, int Wy, int OH, int OW, int KH, int Cout,
                                 int out_dtype, void* stream) {
    if (!ypk || !y) return fail(P2P_EINVAL, "null pointer argument");
    if (N < 1 || OH < 1 || OW < 1 || KH < 1 || Cout < 1 || KH * Cout > 16 || Wy < OW || Hp < OH + KH - 1)
        return fail(P2P_EINVAL, "bad packing geometry (KH*Cout must be <= 16)");
    const long long total = (long long)N * OH * OW * Cout;
    if (total > 4000000000LL) return fail(P2P_ENOTSUP, "tensor too large");
    P2P_DISPATCH_DTYPE(out_dtype, T, launch_k(fold_rows_kernel<T>, dim3((unsigned)(grid_for(total))), dim3(256), 0, (cudaStream_t)stream,
        ypk, bias, (T*)y, Hp, Wy, OH, OW, KH, Cout, (unsigned)total));
    return check_launch("fold_out_rows");
}

extern "C" int p2p_pack_dy(const void* dy, void* dypk, int N, int OH, int OW, int Cout, int KW, int Wp, int dtype, void* stream) {
    if (!dy || !dypk) return fail(P2P_EINVAL, "null pointer argument");
    if (N < 1 || OH < 1 || OW < 1 || KW < 1 || Cout < 1 || KW * Cout > 16 || Wp < OW)
        return fail(P2P_EINVAL, "bad packing geometry (KW*Cout must be <= 16)");
    const long long total = (long long)N * OH * Wp;
    if (total > 4000000000LL) return fail(P2P_ENOTSUP, "tensor too large");
    P2P_DISPATCH_DTYPE(dtype, T, launch_k(pack_dy_kernel<T>, dim3((unsigned)(grid_for(total))), dim3(256), 0, (cudaStream_t)stream, 
        (const T*)dy, (T*)dypk, OW, Cout, KW, Wp, (unsigned)total));
    return check_launch("pack_dy");
}

extern "C" int p2p_im2col(const void* x, void* xcol, int N, int H, int W, int C, int KH, int KW, int stride, int pad, int reflect,
                          int OH, int OW, int Kc, long long rows_alloc, void* stream) {
    if (!x || !xcol) return fail(P2P_EINVAL, "null pointer argument");
    if (N < 1 || H < 1 || W < 1 || C < 1 || KH < 1 || KW < 1 || stride < 1 || pad < 0 || OH < 1 || OW < 1 || Kc % 8 != 0 ||
        KH * KW * C > Kc || Kc > 128)
        return fail(P2P_EINVAL, "bad im2col geometry (KH*KW*C must fit Kc <= 128, Kc %% 8 == 0)");
    if ((OH - 1) * stride + KH > H + 2 * pad || (OW - 1) * stride + KW > W + 2 * pad) return fail(P2P_EINVAL, "output reaches past the padded input");
    if (reflect && (pad >= H || pad >= W)) return fail(P2P_EINVAL, "reflection padding must be smaller than the input");
    const long long rows = (long long)N * OH * OW;
    if (rows_alloc < rows) return fail(P2P_EINVAL, "rows_alloc smaller than N*OH*OW");
    const long long total = rows_alloc * (Kc / 8);
    if (total > 4000000000LL) return fail(P2P_ENOTSUP, "tensor too large");
    launch_k(im2col_kernel, dim3((unsigned)(grid_for(total))), dim3(256), 0, (cudaStream_t)stream, (const __nv_bfloat16*)x, (__nv_bfloat16*)xcol, N, H, W, C, KH, KW,
                                                                     stride, pad, reflect, OH, OW, Kc, rows, (unsigned)total);
    return check_launch("im2col");
}

extern "C" int p2p_col2im(const void* dxcol, void* dx, int N, int H, int W, int C, int KH, int KW, int stride, int pad, int OH,
                          int OW, int Kc, void* stream) {
    if (!dxcol || !dx) return fail(P2P_EINVAL, "null pointer argument");
    if (N < 1 || H < 1 || W < 1 || C < 1 || KH < 1 || KW < 1 || stride < 1 || pad < 0 || OH < 1 || OW < 1 || KH * KW * C > Kc)
        return fail(P2P_EINVAL, "bad col2im geometry");
    const long long total = (long long)N * H * W * C;
    if (total > 4000000000LL) return fail(P2P_ENOTSUP, "tensor too large");
    launch_k(col2im_kernel, dim3((unsigned)(grid_for(total))), dim3(256), 0, (cudaStream_t)stream, (const __nv_bfloat16*)dxcol, (__nv_bfloat16*)dx, H, W, C, KH, KW,
                                                                     stride, pad, OH, OW, Kc, (unsigned)total);
    return check_launch("col2im");
}

extern "C" int p2p_pack_dy_full(const void* dy, void* out, int N, int OH, int OW, int Cout, int KH, int KW, int Kc,
                                long long rows_alloc, void* stream) {
    if (!dy || !out) return fail(P2P_EINVAL, "null pointer argument");
    if (N < 1 || OH < 1 || OW < 1 || Cout < 1 || KH < 1 || KW < 1 || Kc % 8 != 0 || KH * KW * Cout > Kc || Kc > 128)
        return fail(P2P_EINVAL, "bad packing geometry (KH*KW*Cout must fit Kc <= 128, Kc %% 8 == 0)");
    const int Hp = OH + KH - 1, Wp = OW + KW - 1;
    const long long rows = (long long)N * Hp * Wp;
    if (rows_alloc < rows) return fail(P2P_EINVAL, "rows_alloc smaller than the padded grid");
    const long long total = rows_alloc * (Kc / 8);
    if (total > 4000000000LL) return fail(P2P_ENOTSUP, "tensor too large");
    launch_k(pack_dy_full_kernel, dim3((unsigned)(grid_for(total))), dim3(256), 0, (cudaStream_t)stream, (const __nv_bfloat16*)dy, (__nv_bfloat16*)out, OH, OW, Cout,
                                                                           KH, KW, Hp, Wp, Kc, rows, (unsigned)total);
    return check_launch("pack_dy_full");
}
