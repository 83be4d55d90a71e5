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
// ---- radix 16 with the 1/sqrt2 of the eighth-turn twiddles deferred --------------------------------------------------
// exp(+-2 pi i/8) = h (1 +- i) and exp(+-6 pi i/8) = h (-1 +- i), h = 1/sqrt2: the (+-1 +- i) factor is two additions, and the
// scale h stays PENDING on the value until the value meets one that carries no pending scale -- there the radix-2
// add/subtract pair a +- h b is two FFMA2 in place of two FADD2, so the scale costs nothing.  Two values that both carry h
// combine as they are.  Twelve FMUL fewer per butterfly (of ~250 FMA-pipe lane cycles), same slots, same order as
// dft_dif_packed; the results differ from it by rounding only.
#ifndef SFE_FFT_DEFER_SQRT2
#define SFE_FFT_DEFER_SQRT2 1
#endif
template <int SIGN> __device__ __forceinline__ cpx2 mul_p1(cpx2 d)    // (1 + SIGN i) d
{
    const float2 a = cunpack(d);
    return cpack(SIGN > 0 ? make_float2(a.x - a.y, a.x + a.y) : make_float2(a.x + a.y, a.y - a.x));
}
template <int SIGN> __device__ __forceinline__ cpx2 mul_m1(cpx2 d)    // (-1 + SIGN i) d
{
    const float2 a = cunpack(d);
    return cpack(SIGN > 0 ? make_float2(-(a.x + a.y), a.x - a.y) : make_float2(a.y - a.x, -(a.x + a.y)));
}
template <int SIGN> __device__ __forceinline__ cpx2 mul_i(cpx2 d)     // SIGN i d
{
    const float2 a = cunpack(d);
    return cpack(SIGN > 0 ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x));
}
__device__ __forceinline__ void bf_plain(cpx2 &a, cpx2 &b)             // (a, b) <- (a + b, a - b)
{
    const cpx2 s = cadd2(a, b);
    b = csub2(a, b);
    a = s;
}
__device__ __forceinline__ void bf_scaled(cpx2 &a, cpx2 &b)            // (a, b) <- (a + h b, a - h b)
{
    const cpx2 hp = cpack(make_float2(0.70710678118654752f, 0.70710678118654752f));
    const cpx2 hm = cpack(make_float2(-0.70710678118654752f, -0.70710678118654752f));
    cpx2 s, d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(s) : "l"(b), "l"(hp), "l"(a));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(b), "l"(hm), "l"(a));
    a = s;
    b = d;
}
// the four sixteenth-turn twiddles of stage 1: scalar (2 FMUL + 2 FFMA) or packed (FMUL2 + FFMA2, tuning switch)
#ifndef SFE_FFT_PACKED_W16
#define SFE_FFT_PACKED_W16 0
#endif
template <int M16, int SIGN> __device__ __forceinline__ cpx2 mul_w16_pk(cpx2 d)
{
    if (SFE_FFT_PACKED_W16) return cmul2_pk(d, make_float2(kCos16[M16], SIGN > 0 ? kSin16[M16] : -kSin16[M16]));
    return cpack(mul_w16<M16, SIGN>(cunpack(d)));
}
template <int SIGN> __device__ __forceinline__ void dft16_dif_packed_deferred(cpx2 (&v)[16])
{
    // stage 1: pairs (I, I + 8), twiddle w16^I; slots 10 and 14 leave with h pending
#pragma unroll
    for (int I = 0; I < 8; I++) bf_plain(v[I], v[I + 8]);
    v[9] = mul_w16_pk<1, SIGN>(v[9]);
    v[10] = mul_p1<SIGN>(v[10]);
    v[11] = mul_w16_pk<3, SIGN>(v[11]);
    v[12] = mul_i<SIGN>(v[12]);
    v[13] = mul_w16_pk<5, SIGN>(v[13]);
    v[14] = mul_m1<SIGN>(v[14]);
    v[15] = mul_w16_pk<7, SIGN>(v[15]);
    // stage 2: pairs (G + I, G + I + 4), twiddle w16^(2 I); (10, 14) both carry h; 5, 7, 13, 15 leave with h pending
#pragma unroll
    for (int G = 0; G < 16; G += 8) {
#pragma unroll
        for (int I = 0; I < 4; I++) bf_plain(v[G + I], v[G + I + 4]);
        v[G + 5] = mul_p1<SIGN>(v[G + 5]);
        v[G + 6] = mul_i<SIGN>(v[G + 6]);
        v[G + 7] = mul_m1<SIGN>(v[G + 7]);
    }
    // stage 3: pairs (G, G + 2) and (G + 1, G + 3), the latter with twiddle i; (8, 10) and (12, 14) resolve their h
#pragma unroll
    for (int G = 0; G < 16; G += 4) {
        if (G == 8 || G == 12) bf_scaled(v[G], v[G + 2]);
        else bf_plain(v[G], v[G + 2]);
        bf_plain(v[G + 1], v[G + 3]);          // (5, 7) and (13, 15): both carry h
        v[G + 3] = mul_i<SIGN>(v[G + 3]);
    }
    // stage 4: pairs (2m, 2m + 1); 5, 7, 13, 15 resolve their h
#pragma unroll
    for (int m = 0; m < 16; m += 2) {
        if (m == 4 || m == 6 || m == 12 || m == 14) bf_scaled(v[m], v[m + 1]);
        else bf_plain(v[m], v[m + 1]);
    }
}

// same contract as dft_dif (natural order in, frequency q in slot brev<R>(q))
template <int R, int SIGN> __device__ __forceinline__ void dft_dif_packed(float2 (&v)[R])
{
    cpx2 pv[R];
#pragma unroll
    for (int k = 0; k < R; k++) pv[k] = cpack(v[k]);
    if constexpr (R == 16 && SFE_FFT_DEFER_SQRT2) dft16_dif_packed_deferred<SIGN>(pv);
    else dif2_stages<R, SIGN, R / 2>::run(pv);
#pragma unroll
    for (int k = 0; k < R; k++) v[k] = cunpack(pv[k]);
}

}  // namespace sfe
