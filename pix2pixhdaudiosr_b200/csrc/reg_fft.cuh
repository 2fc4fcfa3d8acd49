// In-register radix-2 DIF FFT codelets (sizes 2..32) with compile-time twiddles.
// fft_reg<R>(a): forward FFT (e^{-2 pi j nk/R}) of a[0..R) in place; X[k] ends up in a[bitrev<R>(k)].
// Everything is resolved at compile time after unrolling, so `a` stays in registers.
#pragma once
#include <cuda_runtime.h>

namespace p2p {

__host__ __device__ constexpr int ilog2(int v) { return v <= 1 ? 0 : 1 + ilog2(v >> 1); }

template <int R>
__host__ __device__ constexpr int bitrev(int k) {
    int r = 0;
    for (int b = 0; b < ilog2(R); ++b) r |= ((k >> b) & 1) << (ilog2(R) - 1 - b);
    return r;
}

// cos(2 pi i / 32)
__host__ __device__ constexpr float cos32(int i) {
    i &= 31;
    if (i > 16) i = 32 - i;
    if (i > 8) return -cos32(16 - i);
    switch (i) {
        case 0: return 1.0f;
        case 1: return 0.98078528040323043f;
        case 2: return 0.92387953251128674f;
        case 3: return 0.83146961230254524f;
        case 4: return 0.70710678118654752f;
        case 5: return 0.55557023301960218f;
        case 6: return 0.38268343236508977f;
        case 7: return 0.19509032201612825f;
        default: return 0.0f;
    }
}
__host__ __device__ constexpr float sin32(int i) { return cos32(i - 8); }

// Packed fp32x2 arithmetic (Blackwell FADD2 / FMUL2 / FFMA2): one issue slot per complex add, two per
// complex multiply.  ptxas folds negation / broadcast / half-swap into operand modifiers.
#ifdef P2P_USE_F32X2   // measured on B200: no gain here (FADD2 is half rate, extra MOVs) -- kept as an option
__device__ __forceinline__ unsigned long long pk2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float2 upk2(unsigned long long r) {
    float2 c;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(c.x), "=f"(c.y) : "l"(r));
    return c;
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) {
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a.x, a.y)), "l"(pk2(b.x, b.y)));
    return upk2(r);
}
__device__ __forceinline__ float2 csub(float2 a, float2 b) {
    unsigned long long r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a.x, a.y)), "l"(pk2(b.x, b.y)));
    return upk2(r);
}
// (a.x + j a.y)(b.x + j b.y) = a.x*(b.x, b.y) + a.y*(-b.y, b.x)
__device__ __forceinline__ float2 cmulf(float2 a, float2 b) {
    unsigned long long t, r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(t) : "l"(pk2(a.y, a.y)), "l"(pk2(-b.y, b.x)));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pk2(a.x, a.x)), "l"(pk2(b.x, b.y)), "l"(t));
    return upk2(r);
}
// s * (a.x, a.y) for a real scalar pair (sx, sy)
__device__ __forceinline__ float2 cscale2(float2 a, float sx, float sy) {
    unsigned long long r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a.x, a.y)), "l"(pk2(sx, sy)));
    return upk2(r);
}
#else
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmulf(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cscale2(float2 a, float sx, float sy) { return make_float2(a.x * sx, a.y * sy); }
#endif

// d * e^{-2 pi j IDX/32}, IDX in [0,16), trivial factors special-cased
template <int IDX>
__device__ __forceinline__ float2 twmul32(float2 d) {
    if constexpr (IDX == 0) {
        return d;
    } else if constexpr (IDX == 8) {
        return make_float2(d.y, -d.x);
    } else if constexpr (IDX == 4) {
        constexpr float c = 0.70710678118654752f;          // (1 - j)/sqrt2
        return cscale2(cadd(d, make_float2(d.y, -d.x)), c, c);
    } else if constexpr (IDX == 12) {
        constexpr float c = 0.70710678118654752f;          // (-1 - j)/sqrt2
        return cscale2(cadd(make_float2(d.y, -d.x), make_float2(-d.x, -d.y)), c, c);
    } else {
        constexpr float wr = cos32(IDX), wi = -sin32(IDX);
        return cmulf(d, make_float2(wr, wi));
    }
}

template <int R, int S, int BLK, int I>
struct Bfly {
    __device__ __forceinline__ static void run(float2 (&a)[R]) {
        const float2 x = a[BLK + I], y = a[BLK + I + S / 2];
        a[BLK + I] = cadd(x, y);
        a[BLK + I + S / 2] = twmul32<I * (32 / S)>(csub(x, y));
        if constexpr (I + 1 < S / 2) Bfly<R, S, BLK, I + 1>::run(a);
        else if constexpr (BLK + S < R) Bfly<R, S, BLK + S, 0>::run(a);
    }
};

template <int R, int S>
struct Stage {
    __device__ __forceinline__ static void run(float2 (&a)[R]) {
        Bfly<R, S, 0, 0>::run(a);
        if constexpr (S > 2) Stage<R, S / 2>::run(a);
    }
};

template <int R>
__device__ __forceinline__ void fft_reg(float2 (&a)[R]) {
    static_assert(R >= 2 && R <= 32 && (R & (R - 1)) == 0, "fft_reg: R must be a power of two in [2,32]");
    Stage<R, R>::run(a);
}

}  // namespace p2p
