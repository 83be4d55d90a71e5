// sfe_fft.cuh -- register-level radix-R DFT butterflies (R = 2,4,8,16) used by the shared-memory
// FFT passes of the block convolver.
//
// dft_dif<R,SIGN>(v): in-place R-point DFT of the register array v with kernel exp(SIGN*2*pi*i*kq/R),
// computed as log2(R) radix-2 decimation-in-frequency stages whose twiddles are compile-time
// constants.  Input in natural order; the output for frequency q is left in slot brev<R>(q)
// (bit-reversed slot order), which the callers undo when they store.
#pragma once
#include "sfe_common.cuh"

namespace sfe {

template <int R> __host__ __device__ constexpr int brev(int x)
{
    int r = 0;
    for (int b = 1, c = R >> 1; b < R; b <<= 1, c >>= 1)
        if (x & b) r |= c;
    return r;
}

// cos/sin(2*pi*m/16), m = 0..15
__device__ constexpr float kCos16[16] = {1.0f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f,
                                         0.0f, -0.38268343236508977f, -0.70710678118654752f, -0.92387953251128674f,
                                         -1.0f, -0.92387953251128674f, -0.70710678118654752f, -0.38268343236508977f,
                                         0.0f, 0.38268343236508977f, 0.70710678118654752f, 0.92387953251128674f};
__device__ constexpr float kSin16[16] = {0.0f, 0.38268343236508977f, 0.70710678118654752f, 0.92387953251128674f,
                                         1.0f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f,
                                         0.0f, -0.38268343236508977f, -0.70710678118654752f, -0.92387953251128674f,
                                         -1.0f, -0.92387953251128674f, -0.70710678118654752f, -0.38268343236508977f};

// multiply by exp(SIGN * 2*pi*i * M16/16), M16 in [0,8): trivial cases folded at compile time
template <int M16, int SIGN> __device__ __forceinline__ float2 mul_w16(float2 a)
{
    if (M16 == 0) return a;
    if (M16 == 4) return SIGN > 0 ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
    if (M16 == 2) {
        const float h = 0.70710678118654752f;  // (1 + SIGN*i)/sqrt2
        return SIGN > 0 ? make_float2(h * (a.x - a.y), h * (a.x + a.y)) : make_float2(h * (a.x + a.y), h * (a.y - a.x));
    }
    if (M16 == 6) {
        const float h = 0.70710678118654752f;  // (-1 + SIGN*i)/sqrt2
        return SIGN > 0 ? make_float2(-h * (a.x + a.y), h * (a.x - a.y)) : make_float2(h * (a.y - a.x), -h * (a.x + a.y));
    }
    const float c = kCos16[M16], s = SIGN > 0 ? kSin16[M16] : -kSin16[M16];
    return make_float2(fmaf(a.x, c, -a.y * s), fmaf(a.x, s, a.y * c));
}

template <int R, int SIGN, int HALF, int G, int I> struct dif_bfly {
    static __device__ __forceinline__ void run(float2 (&v)[R])
    {
        const float2 a = v[G + I], b = v[G + I + HALF];
        v[G + I] = cadd(a, b);
        // twiddle exp(SIGN*2*pi*i*I/(2*HALF)) = w16^(I*8/HALF)
        v[G + I + HALF] = mul_w16<(I * 8) / HALF, SIGN>(csub(a, b));
    }
};

template <int R, int SIGN, int HALF, int G, int I> struct dif_inner {
    static __device__ __forceinline__ void run(float2 (&v)[R])
    {
        dif_bfly<R, SIGN, HALF, G, I>::run(v);
        if constexpr (I + 1 < HALF) dif_inner<R, SIGN, HALF, G, I + 1>::run(v);
    }
};

template <int R, int SIGN, int HALF, int G> struct dif_groups {
    static __device__ __forceinline__ void run(float2 (&v)[R])
    {
        dif_inner<R, SIGN, HALF, G, 0>::run(v);
        if constexpr (G + 2 * HALF < R) dif_groups<R, SIGN, HALF, G + 2 * HALF>::run(v);
    }
};

template <int R, int SIGN, int HALF> struct dif_stages {
    static __device__ __forceinline__ void run(float2 (&v)[R])
    {
        dif_groups<R, SIGN, HALF, 0>::run(v);
        if constexpr (HALF > 1) dif_stages<R, SIGN, HALF / 2>::run(v);
    }
};

template <int R, int SIGN> __device__ __forceinline__ void dft_dif(float2 (&v)[R])
{
    static_assert(R == 2 || R == 4 || R == 8 || R == 16, "radix");
    dif_stages<R, SIGN, R / 2>::run(v);
}

// ---- packed flavour: a complex value is one 64-bit register pair, so the butterfly's add and subtract are one
// FADD2 each (sm_100 f32x2); twiddles work on the unpacked halves (a register pair aliases its two floats).
typedef unsigned long long cpx2;
__device__ __forceinline__ cpx2 cpack(float2 a)
{
    cpx2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a.x), "f"(a.y));
    return r;
}
__device__ __forceinline__ float2 cunpack(cpx2 v)
{
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
__device__ __forceinline__ cpx2 cadd2(cpx2 a, cpx2 b)
{
    cpx2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ cpx2 csub2(cpx2 a, cpx2 b)
{
    cpx2 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// a * w with a packed, w = (wr, wi): (ar wr - ai wi, ai wr + ar wi) = a * (wr, wr) + (-ai, ar) * (wi, wi).  ptxas folds the
// half swap, the negation and the broadcasts into the operand selectors of FMUL2 / FFMA2 (tests/probes/f32x2_sel_probe.cu).
__device__ __forceinline__ cpx2 cmul2_pk(cpx2 a, float2 w)
{
    const float2 af = cunpack(a);
    cpx2 t, d, wr = cpack(make_float2(w.x, w.x)), wi = cpack(make_float2(w.y, w.y)), sw = cpack(make_float2(-af.y, af.x));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(t) : "l"(a), "l"(wr));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(sw), "l"(wi), "l"(t));
    return d;
}

template <int R, int SIGN, int HALF, int G, int I> struct dif2_inner {
    static __device__ __forceinline__ void run(cpx2 (&v)[R])
    {
        const cpx2 a = v[G + I], b = v[G + I + HALF];
        v[G + I] = cadd2(a, b);
        const cpx2 d = csub2(a, b);
        constexpr int M16 = (I * 8) / HALF;
        v[G + I + HALF] = M16 == 0 ? d : cpack(mul_w16<M16, SIGN>(cunpack(d)));
        if constexpr (I + 1 < HALF) dif2_inner<R, SIGN, HALF, G, I + 1>::run(v);
    }
};
template <int R, int SIGN, int HALF, int G> struct dif2_groups {
    static __device__ __forceinline__ void run(cpx2 (&v)[R])
    {
        dif2_inner<R, SIGN, HALF, G, 0>::run(v);
        if constexpr (G + 2 * HALF < R) dif2_groups<R, SIGN, HALF, G + 2 * HALF>::run(v);
    }
};
template <int R, int SIGN, int HALF> struct dif2_stages {
    static __device__ __forceinline__ void run(cpx2 (&v)[R])
    {
        dif2_groups<R, SIGN, HALF, 0>::run(v);
        if constexpr (HALF > 1) dif2_stages<R, SIGN, HALF / 2>::run(v);
    }
};
// same contract as dft_dif (natural order in, frequency q in slot brev<R>(q))
template <int R, int SIGN> __device__ __forceinline__ void dft_dif_packed(float2 (&v)[R])
{
    cpx2 pv[R];
#pragma unroll
    for (int k = 0; k < R; k++) pv[k] = cpack(v[k]);
    dif2_stages<R, SIGN, R / 2>::run(pv);
#pragma unroll
    for (int k = 0; k < R; k++) v[k] = cunpack(pv[k]);
}

}  // namespace sfe
