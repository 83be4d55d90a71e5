// polyphase_kernel.cu -- K2g / K2d: the polyphase FIR stages behind `decimate` and `resample`.
//
// Reference semantics (libdsp/decimate.cxx:102-125,132-140; libdsp/resample.cxx:100-148):
//   s(pos)  = sum_j taps[(pos mod L) + j*L] * x[floor(pos/L) - j]        (polyphase output on the
//                                                                          L-times upsampled grid)
//   out[k]  = s(pos_k) * (1 - mu_k) + mu_k * s(pos_k + 1)
// with (pos_k, mu_k) from the data-independent float32 time walk (schedule.cpp).  Both reference
// classes produce these numbers (bit-identically); `resample` merely evaluates all
// L phases for every input first.  The kernels below compute retained outputs only.
#include <algorithm>
#include <cstdlib>

#include <cmath>

#include "kernels.h"

namespace sfe {

namespace {

constexpr int POLY_NT = 128;                 // threads per CTA
constexpr int POLY_SPAN_BYTES = 32 * 1024;   // staged input span per CTA

// ---------------------------------------------------------------------------------------------
// K2g: any L, any rate, table or closed-form schedule, complex or real rows.  poly_point is the direct
// form (window read straight from global memory), used for tiles too sparse to stage.
// ---------------------------------------------------------------------------------------------
template <bool EXACT, bool REAL>
__device__ __forceinline__ float2 poly_point(const PolyParams &p, int s, long pos)
{
    // floor division (pos may be -1 for an output held over a call boundary)
    long n0 = pos / p.up;
    int ph = (int)(pos - n0 * p.up);
    if (ph < 0) { ph += p.up; n0 -= 1; }
    float2 acc = make_float2(0.f, 0.f);
    long n = n0;
    for (int m = ph; m < p.n_taps; m += p.up, n--) {
        const float t = __ldg(p.taps + m);
        float2 x;
        if (REAL) {
            float xr = 0.f;
            if (n < p.in.n) {
                if (n >= 0) xr = __ldg(reinterpret_cast<const float *>(p.in.cur) + (long)s * p.in.pitch + n);
                else if (p.in.H + n >= 0) xr = __ldg(reinterpret_cast<const float *>(p.in.hist) + (long)s * p.in.H + p.in.H + n);
            }
            x = make_float2(xr, 0.f);
        } else {
            x = load_sample(p.in, s, n);
        }
        if (EXACT) {  // accu += taps[m] * in[n-j]   as two IEEE operations (decimate.cxx:136)
            acc.x = __fadd_rn(acc.x, __fmul_rn(t, x.x));
            if (!REAL) acc.y = __fadd_rn(acc.y, __fmul_rn(t, x.y));
        } else {
            acc.x = fmaf(t, x.x, acc.x);
            if (!REAL) acc.y = fmaf(t, x.y, acc.y);
        }
    }
    return acc;
}

template <bool EXACT, bool REAL>
__device__ __forceinline__ void poly_store(const PolyParams &p, int s, long k, float2 s0, float2 s1, float mu, bool have1)
{
    float2 y;
    if (EXACT) {
        // get_sample(phase0,n0) * (1.0f-m_mu) + m_mu*get_sample(phase1,n1)   decimate.cxx:123
        const float w0 = __fsub_rn(1.0f, mu);
        y.x = __fadd_rn(__fmul_rn(s0.x, w0), __fmul_rn(mu, s1.x));
        y.y = __fadd_rn(__fmul_rn(s0.y, w0), __fmul_rn(mu, s1.y));
    } else if (have1) {
        const float w0 = 1.0f - mu;
        y.x = fmaf(mu, s1.x, s0.x * w0);
        y.y = fmaf(mu, s1.y, s0.y * w0);
    } else {
        y = s0;  // mu == 0: s0*1 + 0*s1
    }
    if (REAL) reinterpret_cast<float *>(p.out)[(long)s * p.out_pitch + k] = y.x;
    else reinterpret_cast<float2 *>(p.out)[(long)s * p.out_pitch + k] = y;
}

__device__ __forceinline__ void poly_sched(const PolyParams &p, long k, long &pos, float &mu)
{
    if (p.uniform) {
        pos = p.pos0 + k * p.step;
        mu = 0.f;
    } else {
        const SchedEntry e = p.sched[k];
        pos = e.pos;
        mu = e.mu;
    }
}

// grid positions of one call fit 32 bits (SchedEntry::pos is an int32): 32-bit division, not the 64-bit software routine
__device__ __forceinline__ long floordiv(long a, int b)
{
    const int ai = (int)a;
    if (b == 1) return ai;
    int q = ai / b;
    if (ai - q * b < 0) q -= 1;
    return q;
}

// Sparse outputs (the step is so large that a tile's span does not fit shared memory and the windows of neighbouring
// outputs hardly overlap): one thread per output straight from global memory.  (A warp-per-output flavour with
// coalesced window reads and a shuffle reduction was measured at 2x the time: ~500 instructions per output.)
template <bool EXACT, bool REAL>
__global__ void __launch_bounds__(POLY_NT) poly_direct_kernel(const PolyParams p)
{
    const long k = (long)blockIdx.x * POLY_NT + threadIdx.x;
    if (k >= p.n_out) return;
    long pos;
    float mu;
    poly_sched(p, k, pos, mu);
    const bool have1 = EXACT || mu != 0.f;
    for (int s = blockIdx.y; s < p.n_streams; s += gridDim.y) {
        const float2 s0 = poly_point<EXACT, REAL>(p, s, pos);
        const float2 s1 = have1 ? poly_point<EXACT, REAL>(p, s, pos + 1) : make_float2(0.f, 0.f);
        poly_store<EXACT, REAL>(p, s, k, s0, s1, mu, have1);
    }
}

// One tap, one phase: the stage is the linear interpolation out[k] = tap * ((1-mu) x[pos] + mu x[pos+1]) of its input
// (what remains of a fractional-rate decimator whose FIR was folded into the block convolver's taps, chain.cu).  Nothing
// to stage and nothing to share: one thread per (stream, output), the two neighbouring samples straight from global
// memory (a warp's reads cover one contiguous span of step * 32 samples), the schedule entry shared by all streams
// through L1/L2.  HBM bound: reads about min(2, step)/step of the input once, writes the output once.
__global__ void __launch_bounds__(256) poly_lerp_kernel(const PolyParams p, float tap)
{
    const long k = (long)blockIdx.x * 256 + threadIdx.x;
    if (k >= p.n_out) return;
    long pos;
    float mu;
    poly_sched(p, k, pos, mu);
    const float w0 = tap * (1.0f - mu), w1 = tap * mu;
    for (int s = blockIdx.y; s < p.n_streams; s += gridDim.y) {
        const float2 a = load_sample(p.in, s, pos), b = load_sample(p.in, s, pos + 1);
        reinterpret_cast<float2 *>(p.out)[(long)s * p.out_pitch + k] = make_float2(fmaf(w1, b.x, w0 * a.x), fmaf(w1, b.y, w0 * a.y));
    }
}

// One CTA = p.kt consecutive outputs of one stream.  The input span those outputs touch and the taps (phase
// major, tp[ph][j] = taps[ph + j*up], odd pitch) are staged in shared memory once; each thread then walks its
// outputs' windows there, sharing every sample between the two interpolation points s(pos) and s(pos+1):
// per tap 1 LDS.64 + 2 LDS.32 + 4 FMA instead of 4 global loads with bounds checks.  A tile whose span does not
// fit (very large rate: windows of neighbouring outputs do not overlap anyway) reads global memory directly.
template <bool EXACT, bool REAL>
__global__ void __launch_bounds__(POLY_NT) poly_generic_kernel(const PolyParams p)
{
    extern __shared__ __align__(16) unsigned char poly_sm[];
    float *tp = reinterpret_cast<float *>(poly_sm);                       // [up][jp]
    float2 *xs2 = reinterpret_cast<float2 *>(poly_sm + p.tap_bytes);      // staged samples (complex rows)
    float *xs1 = reinterpret_cast<float *>(poly_sm + p.tap_bytes);        // staged samples (real rows)

    const long k0 = (long)blockIdx.x * p.kt;
    const long k1 = min(k0 + (long)p.kt, p.n_out) - 1;
    long pos_lo, pos_hi;
    float mu_unused;
    poly_sched(p, k0, pos_lo, mu_unused);
    poly_sched(p, k1, pos_hi, mu_unused);
    const int J = p.jp;                                                    // taps per phase (padded to odd)
    const long n_lo = floordiv(pos_lo, p.up) - (J - 1);
    const long n_hi = floordiv(pos_hi + 1, p.up);
    const long span = n_hi - n_lo + 1;
    const bool tiled = p.tap_bytes > 0 && span <= p.span_cap;

    if (tiled)
        for (int i = threadIdx.x; i < p.up * J; i += POLY_NT) {
            const int ph = i / J, j = i - ph * J, m = ph + j * p.up;
            tp[i] = m < p.n_taps ? __ldg(p.taps + m) : 0.f;
        }

    for (int s = blockIdx.y; s < p.n_streams; s += gridDim.y) {
        if (!tiled) {
            for (long k = k0 + threadIdx.x; k <= k1; k += POLY_NT) {
                long pos;
                float mu;
                poly_sched(p, k, pos, mu);
                const float2 s0 = poly_point<EXACT, REAL>(p, s, pos);
                const bool have1 = EXACT || mu != 0.f;
                const float2 s1 = have1 ? poly_point<EXACT, REAL>(p, s, pos + 1) : make_float2(0.f, 0.f);
                poly_store<EXACT, REAL>(p, s, k, s0, s1, mu, have1);
            }
            continue;
        }
        __syncthreads();                                                   // previous stream's readers are done
        if (REAL) {
            for (int i = threadIdx.x; i < (int)span; i += POLY_NT) {
                const long n = n_lo + i;
                float xr = 0.f;
                if (n < p.in.n) {
                    if (n >= 0) xr = __ldg(reinterpret_cast<const float *>(p.in.cur) + (long)s * p.in.pitch + n);
                    else if (p.in.H + n >= 0) xr = __ldg(reinterpret_cast<const float *>(p.in.hist) + (long)s * p.in.H + p.in.H + n);
                }
                xs1[i] = xr;
            }
        } else {
            for (int i = threadIdx.x; i < (int)span; i += POLY_NT) xs2[i] = load_sample(p.in, s, n_lo + i);
        }
        __syncthreads();
        for (long k = k0 + threadIdx.x; k <= k1; k += POLY_NT) {
            long pos;
            float mu;
            poly_sched(p, k, pos, mu);
            const long n0 = floordiv(pos, p.up);
            const int ph0 = (int)(pos - n0 * p.up);
            int ph1 = ph0 + 1, d = 0;                                      // s(pos+1): next phase, or phase 0 one sample on
            if (ph1 == p.up) { ph1 = 0; d = 1; }
            const int i0 = (int)(n0 - n_lo);
            const float *t0 = tp + ph0 * J, *t1 = tp + ph1 * J;
            float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
            const bool have1 = EXACT || mu != 0.f;
            if (EXACT) {
                // reference order: ascending j, separate multiply and add, only the taps m < n_taps (decimate.cxx:132-140)
                const int c0 = (p.n_taps - ph0 + p.up - 1) / p.up, c1 = (p.n_taps - ph1 + p.up - 1) / p.up;
                for (int j = 0; j < c0; j++) {
                    const float2 x = REAL ? make_float2(xs1[i0 - j], 0.f) : xs2[i0 - j];
                    a0.x = __fadd_rn(a0.x, __fmul_rn(t0[j], x.x));
                    if (!REAL) a0.y = __fadd_rn(a0.y, __fmul_rn(t0[j], x.y));
                }
                for (int j = 0; j < c1; j++) {
                    const float2 x = REAL ? make_float2(xs1[i0 + d - j], 0.f) : xs2[i0 + d - j];
                    a1.x = __fadd_rn(a1.x, __fmul_rn(t1[j], x.x));
                    if (!REAL) a1.y = __fadd_rn(a1.y, __fmul_rn(t1[j], x.y));
                }
            } else if (!have1) {
#pragma unroll 4
                for (int j = 0; j < J; j++) {
                    const float2 x = REAL ? make_float2(xs1[i0 - j], 0.f) : xs2[i0 - j];
                    const float t = t0[j];
                    a0.x = fmaf(t, x.x, a0.x);
                    if (!REAL) a0.y = fmaf(t, x.y, a0.y);
                }
            } else {
                float2 xp = REAL ? make_float2(xs1[i0 + d], 0.f) : xs2[i0 + d];   // staged: n0 + d <= n_hi
#pragma unroll 4
                for (int j = 0; j < J; j++) {
                    const float2 x = REAL ? make_float2(xs1[i0 - j], 0.f) : xs2[i0 - j];
                    const float2 xb = d ? xp : x;                          // s(pos+1) reads one sample later when it wraps
                    const float ta = t0[j], tb = t1[j];
                    a0.x = fmaf(ta, x.x, a0.x);
                    a1.x = fmaf(tb, xb.x, a1.x);
                    if (!REAL) {
                        a0.y = fmaf(ta, x.y, a0.y);
                        a1.y = fmaf(tb, xb.y, a1.y);
                    }
                    xp = x;
                }
            }
            poly_store<EXACT, REAL>(p, s, k, a0, a1, mu, have1);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K2d: y[k] = sum_j g[j] x[n_first + D k - j], complex rows, CUDA cores.
// The CTA stages its input span in shared memory split into D polyphase planes
// (plane r holds x[n_first + D m - r]), so that per plane the work is a plain FIR whose
// neighbouring outputs share their window.  Each thread owns DR consecutive outputs and slides a
// register window over 8 taps at a time.  A complex sample is one 64-bit word and a tap is stored as the pair
// (g, g), so one packed FFMA2 advances the real and the imaginary sum: 64 FFMA2 per 15 LDS.64 + 4 LDS.128.
// Integer samples are converted exactly (exponent trick) while staging; their scale lives in the taps.
// ---------------------------------------------------------------------------------------------
constexpr int DR = 8;     // outputs per thread
constexpr int DNT = 128;  // threads per CTA
constexpr int DKT = DR * DNT;
constexpr int DSU = 16;   // staging: loads in flight per thread

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 d_pack2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ f32x2 d_ffma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 d_fadd2(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

template <int FMT> struct DRaw { typedef uint32_t type; };
template <> struct DRaw<SFE_IN_U8> { typedef uint16_t type; };
template <> struct DRaw<SFE_IN_F32> { typedef f32x2 type; };

// raw word -> (I, Q) as a packed pair of exact floats (unscaled integers for the integer formats)
__device__ __forceinline__ f32x2 d_cvt(uint32_t w)     // int16 I,Q: bits 0x4B40_0000 | (s ^ 0x8000) = 12582912 + 32768 + s
{
    const uint32_t a = w ^ 0x80008000u, K = 0x4B400000u;
    return d_fadd2(d_pack2(__uint_as_float(__byte_perm(a, K, 0x7610)), __uint_as_float(__byte_perm(a, K, 0x7632))),
                   d_pack2(-12615680.0f, -12615680.0f));
}
__device__ __forceinline__ f32x2 d_cvt(uint16_t w)     // u8 I,Q: bits 0x4B00_0000 | b = 8388608 + b; value b - 128
{
    const uint32_t K = 0x4B000000u;
    return d_fadd2(d_pack2(__uint_as_float(__byte_perm((uint32_t)w, K, 0x7650)), __uint_as_float(__byte_perm((uint32_t)w, K, 0x7651))),
                   d_pack2(-8388736.0f, -8388736.0f));
}
__device__ __forceinline__ f32x2 d_cvt(f32x2 w) { return w; }

template <int FMT> __device__ __forceinline__ f32x2 d_load_slow(const StreamView &v, int s, long i)
{
    typedef typename DRaw<FMT>::type RT;
    if (i >= v.n) return 0ull;
    if (i >= 0) return d_cvt(__ldg(reinterpret_cast<const RT *>(v.cur) + (long)s * v.pitch + i));
    const long h = (long)v.H + i;
    if (h < 0) return 0ull;
    return d_cvt(__ldg(reinterpret_cast<const RT *>(v.hist) + (long)s * v.H + h));
}

// Tile layout in shared memory: the span in natural order, sample e of the tile at word e + (e >> 6) (one pad word
// per 64 samples).  Staging is then a straight coalesced copy, and for D = 8 the windows of the 32 lanes (8 outputs =
// 64 samples apart) start 65 words apart: every LDS.64 of the inner loop is conflict free.  Plane r, entry m of the
// polyphase view is sample m*D + (D-1-r).
__device__ __forceinline__ int dphys(int e) { return e + (e >> 6); }

// The inner product of one thread: DR outputs, all planes and tap chunks; per chunk 15 window words and 8 tap pairs
// feed 64 FFMA2.  (Explicit double buffering of the chunk loads was measured slower: 128 registers, 3.55 vs 3.28 ms.)
template <int DT>
__device__ __forceinline__ void decim_load_chunk(const f32x2 *xs, const f32x2 *gs, int D, int QLP, int a0, int r, int qc,
                                                 f32x2 (&g)[8], f32x2 (&w)[DR + 7])
{
    const f32x2 *gr = gs + r * QLP + qc;
#pragma unroll
    for (int d = 0; d < 8; d += 2) {
        const ulonglong2 t = *reinterpret_cast<const ulonglong2 *>(gr + d);
        g[d] = t.x;
        g[d + 1] = t.y;
    }
    if (DT == 8) {   // (a + c)*8 + 7 - r: a*8 is a multiple of 64, so the pad term splits off at compile time
        const f32x2 *xr = xs + dphys((a0 - qc) * 8) + (7 - r);
#pragma unroll
        for (int c = 0; c < DR + 7; c++) w[c] = xr[c * 8 + (c >> 3)];
    } else {
        const int e = (a0 - qc) * D + (D - 1 - r);
#pragma unroll
        for (int c = 0; c < DR + 7; c++) w[c] = xs[dphys(e + c * D)];
    }
}
__device__ __forceinline__ void decim_fma_chunk(const f32x2 (&g)[8], const f32x2 (&w)[DR + 7], f32x2 (&acc)[DR])
{
#pragma unroll
    for (int d = 0; d < 8; d++)
#pragma unroll
        for (int i = 0; i < DR; i++) acc[i] = d_ffma2(g[d], w[7 - d + i], acc[i]);
}
template <int DT, int QT>
__device__ __forceinline__ void decim_compute(const f32x2 *xs, const f32x2 *gs, int D, int QLP, f32x2 (&acc)[DR])
{
#pragma unroll
    for (int i = 0; i < DR; i++) acc[i] = 0ull;
    // chunk qc of plane r reads entries a + c, c = 0..14, a = tid*DR + QLP - 8 - qc (a multiple of 8)
    const int a0 = threadIdx.x * DR + QLP - 8;
    if (DT && QT) {   // everything known at compile time: straight-line code, every shared-memory offset an immediate
#pragma unroll
        for (int ch = 0; ch < (DT ? DT : 1) * (QT / 8); ch++) {
            f32x2 g[8], w[DR + 7];
            decim_load_chunk<DT>(xs, gs, DT, QT, a0, ch / (QT ? QT / 8 : 1), (ch % (QT ? QT / 8 : 1)) * 8, g, w);
            decim_fma_chunk(g, w, acc);
        }
    } else {
        for (int r = 0; r < D; r++)
            for (int qc = 0; qc < QLP; qc += 8) {
                f32x2 g[8], w[DR + 7];
                decim_load_chunk<DT>(xs, gs, D, QLP, a0, r, qc, g, w);
                decim_fma_chunk(g, w, acc);
            }
    }
}

template <int FMT, int DT, int QT>     // DT, QT: decimation / taps per plane known at compile time (0 = from the parameters)
__global__ void __launch_bounds__(DNT) decim_fir_kernel(const DecimParams p)
{
    typedef typename DRaw<FMT>::type RT;
    extern __shared__ __align__(16) unsigned char smraw[];
    const int D = DT ? DT : p.D, QLP = QT ? QT : p.QLP;
    const int ML = DKT + QLP - 1;              // samples per plane
    f32x2 *gs = reinterpret_cast<f32x2 *>(smraw);                       // [D][QLP] pairs (g, g), scaled
    f32x2 *xs = gs + D * QLP;                                           // dphys(D * ML) words

    const float scale = FMT == SFE_IN_I16 ? 1.0f / 32767.0f : FMT == SFE_IN_U8 ? 1.0f / 127.0f : 1.0f;
    for (int i = threadIdx.x; i < D * QLP; i += DNT) {
        const float g = __ldg(p.gp + i) * scale;
        gs[i] = d_pack2(g, g);
    }

    const long k0 = (long)blockIdx.x * DKT;
    // tile sample e is x[xbase + e]; (plane r, entry mm) = sample mm*D + (D-1-r) = x[n_first + D*(k0 - QLP + 1 + mm) - r]
    const long xbase = p.n_first + (long)D * (k0 - QLP + 1) - (D - 1);
    const int total = D * ML;
    const bool interior = xbase >= 0 && xbase + total <= p.in.n;        // no history, no zero fill

    for (int s = blockIdx.y; s < p.n_streams; s += gridDim.y) {
        if (s != (int)blockIdx.y) __syncthreads();                      // the previous stream's readers are done
        if (interior) {
            const RT *src = reinterpret_cast<const RT *>(p.in.cur) + (long)s * p.in.pitch + xbase;
            for (int e0 = threadIdx.x; e0 < total; e0 += DNT * DSU) {
                RT raw[DSU];
#pragma unroll
                for (int u = 0; u < DSU; u++) raw[u] = e0 + u * DNT < total ? __ldg(src + e0 + u * DNT) : RT();
#pragma unroll
                for (int u = 0; u < DSU; u++)
                    if (e0 + u * DNT < total) xs[dphys(e0 + u * DNT)] = d_cvt(raw[u]);
            }
        } else {
            for (int e = threadIdx.x; e < total; e += DNT) xs[dphys(e)] = d_load_slow<FMT>(p.in, s, xbase + e);
        }
        __syncthreads();

        f32x2 acc[DR];
        decim_compute<DT, QT>(xs, gs, D, QLP, acc);
        const long kk = k0 + (long)threadIdx.x * DR;
        f32x2 *o = reinterpret_cast<f32x2 *>(p.out + (long)s * p.out_pitch + kk);
        if (kk + DR <= p.n_out && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
#pragma unroll
            for (int i = 0; i < DR; i += 2) *reinterpret_cast<ulonglong2 *>(o + i) = make_ulonglong2(acc[i], acc[i + 1]);
        } else {
#pragma unroll
            for (int i = 0; i < DR; i++)
                if (kk + i < p.n_out) o[i] = acc[i];
        }
    }
}

typedef void (*DecimKernel)(const DecimParams);
template <int FMT> DecimKernel decim_pick(int D, int QLP)
{
    static const bool no_qt = getenv("SFE_DECIM_NO_QT") != nullptr;   // tuning hook
    // the fully unrolled flavour is 8 % faster for the integer formats but measured 2.3x slower for complex64 rows
    if (D == 8) return QLP == 16 && !no_qt && FMT != SFE_IN_F32 ? decim_fir_kernel<FMT, 8, 16> : decim_fir_kernel<FMT, 8, 0>;
    return decim_fir_kernel<FMT, 0, 0>;
}

// ---------------------------------------------------------------------------------------------
// K2s: the generic polyphase stage when MANY streams share the filter (and therefore the schedule): lane = stream.
// One CTA = 32 streams x p.kt consecutive outputs.  The span is staged transposed, xs[sample][stream] with a row pitch
// of 33 words, so a warp reads one sample of its 32 streams as 256 contiguous bytes (conflict free whatever the rate;
// the thread-per-output kernel pays 5 wavefronts per LDS.64 for its 2.5-sample lane stride).  Per output the two
// interpolation points collapse into one tap set, warp uniform because mu is shared by the streams:
//   out = sum_i c_i x[n_top - i],  c_i = mu t1[i] + (1 - mu) t0[i - d],  n_top = floor((pos + 1) / L), d = n_top - n0
// computed once per output by the lanes, stored as (c, c) pairs and broadcast: 2 LDS.64 + 1 FFMA2 per tap and stream.
// Outputs are staged too and leave as contiguous runs per stream.  Not for SFE_FLAG_EXACT (different rounding order).
// ---------------------------------------------------------------------------------------------
constexpr int LS_WARPS = 4, LS_NT = 32 * LS_WARPS, LS_PITCH = 33;
constexpr int LS_KPW = 16, LS_KT = LS_WARPS * LS_KPW;   // outputs per warp and per CTA, at most
constexpr int LS_GUARD = 3;                              // zero rows in front of the staged span
constexpr int LS_G = 4;                                  // outputs computed together in the grouped flavour
static_assert(LS_G == 4 && LS_KPW % LS_G == 0, "the grouped loop is written for runs of four");

template <int FMT> __device__ __forceinline__ float2 ls_cvt(typename DRaw<FMT>::type w);
template <> __device__ __forceinline__ float2 ls_cvt<SFE_IN_I16>(uint32_t w) { return cvt_i16(w); }
template <> __device__ __forceinline__ float2 ls_cvt<SFE_IN_U8>(uint16_t w) { return cvt_u8(w); }
template <> __device__ __forceinline__ float2 ls_cvt<SFE_IN_F32>(f32x2 w)
{
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(w));
    return r;
}

// staging of complex64 rows: round IT copies samples 32 IT + lane of the warp's eight rows, every offset an immediate
constexpr int LS_ROUNDS = 8;                                                // 256 samples: the largest span a 64 KB tile holds
template <int IT, int RR, class RT>
__device__ __forceinline__ void ls_cp_rounds(uint32_t dst0, const RT *const (&src)[32 / LS_WARPS], int span, int lane)
{
    if constexpr (IT < LS_ROUNDS) {
        if (RR == 0 && IT * 32 >= span) return;                             // warp uniform: no later round has work
        if (IT * 32 + lane < span)
            asm volatile("cp.async.ca.shared.global [%0 + %2], [%1 + %3], 8;" ::"r"(dst0), "l"(src[RR]),
                         "n"(IT * 32 * LS_PITCH * 8 + LS_WARPS * RR * 8), "n"(IT * 32 * (int)sizeof(RT)) : "memory");
        if constexpr (RR + 1 < 32 / LS_WARPS) ls_cp_rounds<IT, RR + 1>(dst0, src, span, lane);
        else ls_cp_rounds<IT + 1, 0>(dst0, src, span, lane);
    }
}

template <int FMT, bool GRP>
__global__ void __launch_bounds__(LS_NT) poly_lanes_kernel(const PolyParams p)
{
    typedef typename DRaw<FMT>::type RT;
    extern __shared__ __align__(16) unsigned char ls_sm[];
    const int J = p.jp;                                                    // taps per phase (padded to odd)
    float *tp = reinterpret_cast<float *>(ls_sm);                          // [up][J] phase major
    const int JC = (J + 1 + 3) & ~3;                                       // combined taps per output, zero padded to fours
    f32x2 *cb = reinterpret_cast<f32x2 *>(ls_sm + p.tap_bytes);            // [LS_WARPS][JC] combined taps (c, c)
    // [LS_GUARD + span][LS_PITCH]; reused for the outputs.  The guard rows in front are zero: the padding taps read them.
    float2 *xs = reinterpret_cast<float2 *>(ls_sm + p.tap_bytes + LS_WARPS * p.cw_bytes) + LS_GUARD * LS_PITCH;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    const long k0 = (long)blockIdx.x * p.kt;
    const long k1 = min(k0 + (long)p.kt, p.n_out) - 1;
    const int s0 = blockIdx.y * 32;
    long pos_lo, pos_hi;
    float mu_unused;
    poly_sched(p, k0, pos_lo, mu_unused);
    poly_sched(p, k1, pos_hi, mu_unused);
    const long n_lo = floordiv(pos_lo, p.up) - (J - 1);
    const long n_hi = floordiv(pos_hi + 1, p.up);
    const int span = (int)(n_hi - n_lo + 1);
    if (span > p.span_cap) {   // cannot happen with the launcher's sizing unless the schedule is wildly non-uniform
        for (long k = k0 + warp; k <= k1; k += LS_WARPS) {
            long pos;
            float mu;
            poly_sched(p, k, pos, mu);
            if (s0 + lane < p.n_streams) {
                const float2 a = poly_point<false, false>(p, s0 + lane, pos);
                const float2 b = mu != 0.f ? poly_point<false, false>(p, s0 + lane, pos + 1) : make_float2(0.f, 0.f);
                poly_store<false, false>(p, s0 + lane, k, a, b, mu, mu != 0.f);
            }
        }
        return;
    }
    for (int i = threadIdx.x; i < p.up * J; i += LS_NT) {
        const int ph = i / J, j = i - ph * J, m = ph + j * p.up;
        tp[i] = m < p.n_taps ? __ldg(p.taps + m) : 0.f;
    }
    for (int i = threadIdx.x; i < LS_GUARD * LS_PITCH; i += LS_NT) xs[i - LS_GUARD * LS_PITCH] = make_float2(0.f, 0.f);
    // the warp's schedule entries, one per lane, fetched now so that their latency hides behind the staging loads
    // register slot q of warp w holds output k0 + kl(q): outputs w, w + 4, ... one at a time, or (GRP) runs of LS_G
    // consecutive outputs that are computed together
    auto kl_of = [&](int q) { return GRP ? LS_G * (warp + LS_WARPS * (q / LS_G)) + (q % LS_G) : warp + LS_WARPS * q; };
    long pos_l = 0;
    float mu_l = 0.f;
    if (lane < LS_KPW && k0 + kl_of(lane) <= k1) poly_sched(p, k0 + kl_of(lane), pos_l, mu_l);
    const int pos_li = (int)pos_l;                                         // positions of one call fit 32 bits (SchedEntry)
    // ---- stage the span transposed: warp w takes rows w, w + 4, ...; lanes run along the samples ----
    const bool interior = n_lo >= 0 && n_hi < p.in.n;
    if (interior) {
        // all 8 rows of the warp x 2 sample chunks per round: 16 independent loads in flight per lane
        const RT *src[32 / LS_WARPS];
#pragma unroll
        for (int rr = 0; rr < 32 / LS_WARPS; rr++) {
            const int s = min(s0 + warp + LS_WARPS * rr, p.n_streams - 1);   // rows past the end repeat the last stream (never stored)
            src[rr] = reinterpret_cast<const RT *>(p.in.cur) + (long)s * p.in.pitch + n_lo;
        }
        if (FMT == SFE_IN_F32) {
            // complex64 rows need no conversion: 8-byte asynchronous copies straight into the transposed tile, the whole
            // span of the CTA in flight at once instead of 16 loads per lane and round
            // (rounds unrolled with every offset an immediate of one shared-memory address and eight row pointers: the
            // per-copy address arithmetic was 16 % of the kernel's instructions)
            const uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(xs + lane * LS_PITCH + warp);
#pragma unroll
            for (int rr = 0; rr < 32 / LS_WARPS; rr++) src[rr] += lane;
            if (span <= 32 * LS_ROUNDS) {
                ls_cp_rounds<0, 0>(dst0, src, span, lane);
            } else {
                for (int i0 = lane; i0 < span; i0 += 32) {
#pragma unroll
                    for (int rr = 0; rr < 32 / LS_WARPS; rr++) {
                        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(xs + i0 * LS_PITCH + warp + LS_WARPS * rr);
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src[rr] + i0 - lane) : "memory");
                    }
                }
            }
            asm volatile("cp.async.commit_group;\n cp.async.wait_group 0;" ::: "memory");
        } else
        for (int i0 = lane; i0 < span; i0 += 64) {
            RT raw[2][32 / LS_WARPS];
#pragma unroll
            for (int u = 0; u < 2; u++)
#pragma unroll
                for (int rr = 0; rr < 32 / LS_WARPS; rr++) raw[u][rr] = i0 + 32 * u < span ? __ldg(src[rr] + i0 + 32 * u) : RT();
#pragma unroll
            for (int u = 0; u < 2; u++)
#pragma unroll
                for (int rr = 0; rr < 32 / LS_WARPS; rr++)
                    if (i0 + 32 * u < span) xs[(i0 + 32 * u) * LS_PITCH + warp + LS_WARPS * rr] = ls_cvt<FMT>(raw[u][rr]);
        }
    } else {
        for (int r = warp; r < 32; r += LS_WARPS) {
            const int s = s0 + r;
            for (int i = lane; i < span; i += 32)
                xs[i * LS_PITCH + r] = s < p.n_streams ? load_sample(p.in, s, n_lo + i) : make_float2(0.f, 0.f);
        }
    }
    __syncthreads();
    // ---- compute: warp w takes outputs k0 + w, k0 + w + 4, ...; every lane its own stream ----
    f32x2 res[LS_KPW];
    if (GRP) {
        // LS_G consecutive outputs share most of their samples (the launcher picks this flavour when the windows overlap
        // by half or more): their combined taps are laid out on one sample axis, m = newest sample of the run - sample,
        // as rows of LS_G (c, c) pairs, zero outside each output's window.  Per sample one LDS.64 of the lane's own
        // stream and two broadcast LDS.128 feed LS_G FFMA2 -- 4 shared-memory wavefronts instead of 10.
        float4 *ct = reinterpret_cast<float4 *>(ls_sm + p.tap_bytes + warp * p.cw_bytes);   // [rows] of LS_G combined taps
        const int g = lane & (LS_G - 1);
        // the window of every output of the warp, derived once by the lane that holds its schedule entry (one division per
        // lane instead of one per lane and run): row of x[n_top] in the staged span, and ph0 | d << 16
        int row_l = 0, phd_l = 0;
        {
            const int n0 = (int)floordiv(pos_li, p.up);
            const int ph0 = pos_li - n0 * p.up;
            const int d = ph0 + 1 == p.up ? 1 : 0;
            row_l = (int)(n0 + d - n_lo);
            phd_l = ph0 | (d << 16);
        }
#pragma unroll
        for (int qq = 0; qq < LS_KPW / LS_G; qq++) {
#pragma unroll
            for (int u = 0; u < LS_G; u++) res[qq * LS_G + u] = 0ull;
            const long kg = k0 + kl_of(qq * LS_G);
            if (kg > k1) continue;                                         // warp uniform
            // every lane derives the window of output g = lane % LS_G of the run
            const int row = __shfl_sync(0xffffffffu, row_l, qq * LS_G + g);
            const int phd = __shfl_sync(0xffffffffu, phd_l, qq * LS_G + g);
            const float mu = __shfl_sync(0xffffffffu, mu_l, qq * LS_G + g);
            const bool valid = kg + g <= k1;
            const int ph0 = phd & 0xffff, d = phd >> 16;
            const int ph1 = d ? 0 : ph0 + 1;
            const int row_hi = __reduce_max_sync(0xffffffffu, valid ? row : 0);
            const int off = row_hi - row;
            int M = __reduce_max_sync(0xffffffffu, valid ? off + J + d : 0);
            // two samples per round.  Consecutive outputs of a float32 time walk are at most ceil(step) + 1 grid positions
            // apart (decimate.cxx:124, also across a call seam and a held-over output), and the launcher sized the table
            // for (LS_G - 1) such gaps plus one window, so the clamp never cuts a live tap.
            M = min((M + 1) & ~1, p.grp_rows);
            const float *t0 = tp + ph0 * J, *t1 = tp + ph1 * J;
            const float w0 = 1.0f - mu;
            __syncwarp();                                                  // the previous run's readers are done
            float *cf = reinterpret_cast<float *>(ct);
            for (int m = lane / LS_G; m < M; m += 32 / LS_G) {
                const int i = m - off;
                const float a = (valid && i >= 0 && i < J) ? t1[i] : 0.f;
                const float b = (valid && i - d >= 0 && i - d < J) ? t0[i - d] : 0.f;
                cf[m * LS_G + g] = mu != 0.f ? fmaf(mu, a, w0 * b) : b;
            }
            __syncwarp();
            const f32x2 *xq = reinterpret_cast<const f32x2 *>(xs + row_hi * LS_PITCH + lane);
            const float4 *cq = ct;
            f32x2 a0 = 0ull, a1 = 0ull, a2 = 0ull, a3 = 0ull;
            // the taps are scalars: ptxas folds the (c, c) broadcast into the .F32 operand selector of FFMA2, so one
            // broadcast LDS.128 carries the four taps of a sample row (they were (c, c) pairs: two loads per row)
            for (int m = 0; m < M; m += 2, xq -= 2 * LS_PITCH, cq += 2) {
                const f32x2 x0 = xq[0], x1 = xq[-LS_PITCH];
                const float4 c = cq[0], e = cq[1];
                a0 = d_ffma2(d_pack2(c.x, c.x), x0, a0);
                a1 = d_ffma2(d_pack2(c.y, c.y), x0, a1);
                a2 = d_ffma2(d_pack2(c.z, c.z), x0, a2);
                a3 = d_ffma2(d_pack2(c.w, c.w), x0, a3);
                a0 = d_ffma2(d_pack2(e.x, e.x), x1, a0);
                a1 = d_ffma2(d_pack2(e.y, e.y), x1, a1);
                a2 = d_ffma2(d_pack2(e.z, e.z), x1, a2);
                a3 = d_ffma2(d_pack2(e.w, e.w), x1, a3);
            }
            res[qq * LS_G + 0] = a0;
            res[qq * LS_G + 1] = a1;
            res[qq * LS_G + 2] = a2;
            res[qq * LS_G + 3] = a3;
        }
    } else {
        f32x2 *cw = cb + warp * JC;
#pragma unroll
        for (int q = 0; q < LS_KPW; q++) {
            res[q] = 0ull;
            const long k = k0 + warp + LS_WARPS * q;
            if (k > k1) continue;                                              // warp uniform
            const long pos = __shfl_sync(0xffffffffu, pos_li, q);
            const float mu = __shfl_sync(0xffffffffu, mu_l, q);
            const long n0 = floordiv(pos, p.up);
            const int ph0 = (int)(pos - n0 * p.up);
            int ph1 = ph0 + 1, d = 0;
            if (ph1 == p.up) { ph1 = 0; d = 1; }
            const float *t0 = tp + ph0 * J, *t1 = tp + ph1 * J;
            const float w0 = 1.0f - mu;
            __syncwarp();                                                      // the previous output's readers are done
            for (int i = lane; i < JC; i += 32) {
                const float a = i < J ? t1[i] : 0.f, b = (i - d >= 0 && i - d < J) ? t0[i - d] : 0.f;
                const float c = mu != 0.f ? fmaf(mu, a, w0 * b) : b;           // mu == 0: exactly the taps of s(pos)
                cw[i] = d_pack2(c, c);                                         // i >= J + d: zero padding
            }
            __syncwarp();
            // x[n_top] of this lane's stream; four taps per round: two 128-bit tap loads, four sample loads at constant
            // offsets from one moving pointer.  The padding taps meet rows of the span or of the zero guard in front of it.
            const f32x2 *xq = reinterpret_cast<const f32x2 *>(xs + (int)(n0 + d - n_lo) * LS_PITCH + lane);
            const ulonglong2 *cq = reinterpret_cast<const ulonglong2 *>(cw);
            f32x2 acc = 0ull, acc2 = 0ull;
            for (int i = 0; i < JC; i += 4, xq -= 4 * LS_PITCH, cq += 2) {
                const ulonglong2 c01 = cq[0], c23 = cq[1];
                const f32x2 x0 = xq[0], x1 = xq[-LS_PITCH], x2 = xq[-2 * LS_PITCH], x3 = xq[-3 * LS_PITCH];
                acc = d_ffma2(c01.x, x0, acc);
                acc2 = d_ffma2(c01.y, x1, acc2);
                acc = d_ffma2(c23.x, x2, acc);
                acc2 = d_ffma2(c23.y, x3, acc2);
            }
            res[q] = d_fadd2(acc, acc2);
        }
    }
    __syncthreads();                                                       // all reads of xs are done: reuse it
    f32x2 *ob = reinterpret_cast<f32x2 *>(xs);                             // [32 streams][kt + 1]
    const int opitch = p.kt + 1;
    const int nk = (int)(k1 - k0 + 1);
    {
        // slot q of warp w is output kl_of(q) = w_first + (a constant of q): one pointer, immediate offsets, 32-bit bounds
        const int w_first = GRP ? LS_G * warp : warp;
        f32x2 *obl = ob + lane * opitch + w_first;
        const int left = nk - w_first;
#pragma unroll
        for (int q = 0; q < LS_KPW; q++) {
            const int rel = GRP ? LS_G * LS_WARPS * (q / LS_G) + (q % LS_G) : LS_WARPS * q;
            if (rel < left) obl[rel] = res[q];
        }
    }
    __syncthreads();
    f32x2 *o0 = reinterpret_cast<f32x2 *>(p.out) + (long)(s0 + warp) * p.out_pitch + k0;
    if (nk <= 64 && (nk & 1) == 0 && (p.out_pitch & 1) == 0 && (reinterpret_cast<uintptr_t>(o0) & 15) == 0) {
        // whole tiles (the common case): a lane takes two consecutive outputs of a row, all 8 rows of the warp unrolled --
        // one 128-bit store per row and lane instead of two 64-bit ones under a loop with its own index arithmetic
        // (that loop was 14 % of the kernel's instructions, more than the filter itself: profiles/r02m_cfg4_k2s_summary.json)
        const bool mine = 2 * lane < nk;
        const f32x2 *ol = ob + warp * opitch + 2 * lane;
#pragma unroll
        for (int rr = 0; rr < 32 / LS_WARPS; rr++) {
            if (mine && s0 + warp + LS_WARPS * rr < p.n_streams) {
                const f32x2 a = ol[rr * LS_WARPS * opitch], b = ol[rr * LS_WARPS * opitch + 1];
                *reinterpret_cast<ulonglong2 *>(o0 + (long)rr * LS_WARPS * p.out_pitch + 2 * lane) = make_ulonglong2(a, b);
            }
        }
    } else {
        for (int r = warp; r < 32; r += LS_WARPS) {
            const int s = s0 + r;
            if (s >= p.n_streams) continue;
            f32x2 *o = reinterpret_cast<f32x2 *>(p.out) + (long)s * p.out_pitch + k0;
            for (int kl = lane; kl < nk; kl += 32) o[kl] = ob[r * opitch + kl];
        }
    }
}

}  // namespace

int launch_poly_generic(const PolyParams &p_in, cudaStream_t st)
{
    PolyParams p = p_in;
    if (p.n_out <= 0 || p.n_streams <= 0) return SFE_OK;
    if (p.n_taps == 1 && p.up == 1 && !p.exact && !p.in_real && p.lerp_tap != 0.f) {   // pure interpolation stage
        dim3 grid((unsigned)((p.n_out + 255) / 256), (unsigned)(p.n_streams < 32768 ? p.n_streams : 32768));
        poly_lerp_kernel<<<grid, 256, 0, st>>>(p, p.lerp_tap);
        count_launch();
        return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
    }
    // tile: as many outputs per CTA as keep the staged span under POLY_SPAN_BYTES (the span grows with the step)
    const int eb = p.in_real ? 4 : 8;
    const int J = (p.n_taps + p.up - 1) / p.up;
    p.jp = J | 1;
    const size_t tap_bytes = ((size_t)p.up * p.jp * sizeof(float) + 15) & ~(size_t)15;
    double step = p.uniform ? (double)p.step : 0.0;                        // grid positions per output
    if (!p.uniform) step = (double)p.rate * p.up;
    const double per_out = step / p.up;                                    // input samples per output
    // many streams sharing the schedule: lane-per-stream flavour (K2s)
    static const bool no_lanes = getenv("SFE_POLY_NO_LANES") != nullptr;   // tuning hook
    if (!p.exact && !p.in_real && p.n_streams >= 16 && tap_bytes <= 32 * 1024 && !no_lanes) {
        // tile: 40 KB when that still holds whole runs for every warp (more CTAs per SM: 1.26 instead of 1.32 ms on cfg 4
        // stage a; 24 KB: 1.41 ms), else 64 KB (large steps: few outputs fit, and below 8 the tiled generic kernel takes over)
        static const long tile_kb = getenv("SFE_POLY_LANES_TILE_KB") ? atol(getenv("SFE_POLY_LANES_TILE_KB")) : 0;   // tuning hook
        long cap_s = ((tile_kb > 0 ? tile_kb : 40) * 1024) / (LS_PITCH * 8);   // samples of span per CTA
        long kl = (long)((double)(cap_s - p.jp - 8) / (per_out * 1.01 + 1e-9));
        if (tile_kb <= 0 && kl < LS_WARPS * LS_G) {
            cap_s = (64 * 1024) / (LS_PITCH * 8);
            kl = (long)((double)(cap_s - p.jp - 8) / (per_out * 1.01 + 1e-9));
        }
        if (kl > LS_KT) kl = LS_KT;
        kl = kl >= LS_WARPS * LS_G ? kl / (LS_WARPS * LS_G) * (LS_WARPS * LS_G) : kl / LS_WARPS * LS_WARPS;   // whole runs of LS_G per warp
        if (kl >= 8) {
            p.kt = (int)kl;
            p.tap_bytes = (int)tap_bytes;
            p.span_cap = (int)std::min<long>(cap_s, (long)((double)kl * per_out * 1.01) + p.jp + 8);
            const size_t body = std::max<size_t>((size_t)p.span_cap * LS_PITCH * 8, (size_t)32 * (kl + 1) * 8);
            // runs of LS_G outputs are computed together when their windows overlap by half or more (and the table stays small)
            static const bool no_grp = getenv("SFE_POLY_NO_GROUP") != nullptr;   // tuning hook
            const int grp_rows = ((int)std::ceil((LS_G - 1) * (per_out + 1.0)) + p.jp + 1 + 4 + 1) & ~1;
            const bool grp = !no_grp && (LS_G - 1) * per_out <= p.jp && grp_rows <= 128 && kl % (LS_WARPS * LS_G) == 0;
            p.grp_rows = grp ? grp_rows : 0;
            p.cw_bytes = grp ? grp_rows * LS_G * 8 : ((p.jp + 1 + 3) & ~3) * 8;
            const size_t smem_l = tap_bytes + (size_t)LS_WARPS * p.cw_bytes + (size_t)LS_GUARD * LS_PITCH * 8 + body;
            typedef void (*LanesKernel)(const PolyParams);
            LanesKernel kl_kern = grp ? (p.in.fmt == SFE_IN_I16 ? poly_lanes_kernel<SFE_IN_I16, true> : p.in.fmt == SFE_IN_U8 ? poly_lanes_kernel<SFE_IN_U8, true> : poly_lanes_kernel<SFE_IN_F32, true>)
                                      : (p.in.fmt == SFE_IN_I16 ? poly_lanes_kernel<SFE_IN_I16, false> : p.in.fmt == SFE_IN_U8 ? poly_lanes_kernel<SFE_IN_U8, false> : poly_lanes_kernel<SFE_IN_F32, false>);
            if (smem_l > 48 * 1024 && !SFE_CUDA(cudaFuncSetAttribute(kl_kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_l)))
                return SFE_ERR_CUDA;
            dim3 grid((unsigned)((p.n_out + kl - 1) / kl), (unsigned)((p.n_streams + 31) / 32));
            kl_kern<<<grid, LS_NT, smem_l, st>>>(p);
            count_launch();
            return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
        }
    }
    const long cap = POLY_SPAN_BYTES / eb;
    long kt = (long)((double)(cap - p.jp - 8) / (per_out * 1.01 + 1e-9));
    kt = kt / POLY_NT * POLY_NT;
    if (kt > 8 * POLY_NT) kt = 8 * POLY_NT;
    while (kt > POLY_NT && (p.n_out + kt - 1) / kt * p.n_streams < 2 * 148) kt -= POLY_NT;   // small jobs: more CTAs
    p.tap_bytes = (int)tap_bytes;
    p.span_cap = (int)std::min<long>(cap, (long)((double)kt * per_out * 1.01) + p.jp + 8);
    if (kt < POLY_NT || tap_bytes > 48 * 1024) {                           // no useful tile: direct global reads
        kt = POLY_NT;
        p.tap_bytes = 0;
        p.span_cap = 0;
    }
    p.kt = (int)kt;
    if (!p.tap_bytes) {                                                    // sparse outputs: one thread per output
        dim3 grid((unsigned)((p.n_out + POLY_NT - 1) / POLY_NT), (unsigned)(p.n_streams < 32768 ? p.n_streams : 32768));
        auto kd = p.exact ? (p.in_real ? poly_direct_kernel<true, true> : poly_direct_kernel<true, false>)
                          : (p.in_real ? poly_direct_kernel<false, true> : poly_direct_kernel<false, false>);
        kd<<<grid, POLY_NT, 0, st>>>(p);
        count_launch();
        return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
    }
    const size_t smem = p.tap_bytes ? tap_bytes + (size_t)p.span_cap * eb : 0;
    dim3 grid((unsigned)((p.n_out + kt - 1) / kt), (unsigned)(p.n_streams < 32768 ? p.n_streams : 32768));
    auto kern = p.exact ? (p.in_real ? poly_generic_kernel<true, true> : poly_generic_kernel<true, false>)
                        : (p.in_real ? poly_generic_kernel<false, true> : poly_generic_kernel<false, false>);
    if (smem > 48 * 1024 && !SFE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)))
        return SFE_ERR_CUDA;
    kern<<<grid, POLY_NT, smem, st>>>(p);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

int launch_decim_fir(const DecimParams &p, cudaStream_t st)
{
    if (p.n_out <= 0 || p.n_streams <= 0) return SFE_OK;
    const int ML = DKT + p.QLP - 1;
    const int words = p.D * ML + (p.D * ML >> 6) + 2;
    const size_t smem = sizeof(float2) * (size_t)p.D * p.QLP + sizeof(float2) * (size_t)words;
    if (smem > 220 * 1024) {
        set_error("decim_fir: tile does not fit shared memory (D * taps too large)");
        return SFE_ERR_ARG;
    }
    DecimKernel kern = p.in.fmt == SFE_IN_I16 ? decim_pick<SFE_IN_I16>(p.D, p.QLP) : p.in.fmt == SFE_IN_U8 ? decim_pick<SFE_IN_U8>(p.D, p.QLP) : decim_pick<SFE_IN_F32>(p.D, p.QLP);
    if (smem > 48 * 1024 && !SFE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)))
        return SFE_ERR_CUDA;
    dim3 grid((unsigned)((p.n_out + DKT - 1) / DKT), (unsigned)(p.n_streams < 32768 ? p.n_streams : 32768));
    kern<<<grid, DNT, smem, st>>>(p);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

}  // namespace sfe
