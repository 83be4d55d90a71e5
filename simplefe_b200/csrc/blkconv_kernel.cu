// blkconv_kernel.cu -- K1: FFT block convolver for batches of IQ streams (sm_100a).
//
// Replaces blkconv::process (libdsp/blkconv.cxx:77-110) plus the caller-side sample conversion.
// The reference does overlap-ADD on one real stream with FFTW r2c/c2r.  Here one complex FFT
// carries I and Q of a stream at once (real taps => (I*h) + j(Q*h)), and blocks use overlap-SAVE
// so that every FFT block is independent: block b of a stream reads fft_len samples starting
// n_taps-1 before its first output and discards the first n_taps-1 results.  The hop equals the
// reference's block size fft_len+1-n_taps (blkconv.cxx:47).  Both formulations compute the same
// causal linear convolution from zero state; they differ by float rounding only.
//
// One FFT = N/16 threads, 16 points per thread and pass:
//   forward  : radix-16 DIF passes (last pass radix 2/4/8 when log2 N is not a multiple of 4),
//              data exchanged through padded shared memory between passes
//   multiply : by the tap spectrum, stored in the same digit-reversed order and pre-scaled by 1/N
//              (blkconv.cxx:92-101), fused between the last forward and first inverse butterfly
//   inverse  : the mirrored DIT passes with conjugated twiddles, then overlap discard + store.
// No reordering pass is ever needed: forward leaves digit-reversed order, inverse consumes it.
#include <cmath>
#include <cstdlib>
#include <vector>

// resident threads per SM the register allocation must allow (CTAs of <= 256 threads): 1024 = four 4096-point FFTs per SM
// at 64 registers (no spills since the twiddle products replaced fifteen live table values by four)
#ifndef SFE_K1_THREADS
#define SFE_K1_THREADS 1024
#endif
#ifndef SFE_FFT_PACKED
#define SFE_FFT_PACKED 1
#endif
#include "sfe_fft.cuh"
#include "sfe_fft_plan.h"
#include "kernels.h"

namespace sfe {

namespace {

using fftplan::BetaPerm;
using fftplan::ilog2;
using fftplan::phi;       // bank-conflict padding
using fftplan::Plan;

// padded index of element base + q*S: when S is a multiple of 16 the padding of the stride is a constant, so the compiler
// keeps ONE address per thread and pass and folds q * (S + S/16) into the immediate offset of every LDS / STS
template <int N, int S> __device__ __forceinline__ int phi_at(int pbase, int base, int q)
{
    if (Plan<N>::ROWPAD) return pbase + q * (S + S / (N / 16));   // base + q S stays in the row of `base` unless S = N/16 (whole rows)
    return S % 16 == 0 ? pbase + q * (S + S / 16) : phi<N>(base + q * S);
}
// input flavours the kernel is compiled for (one instantiation each: no per-sample format dispatch)
enum { KF_I16 = 0, KF_U8 = 1, KF_C64 = 2, KF_REAL = 3 };

struct Ctx {
    const BlkconvParams &p;
    int stream;     // stream row of this FFT
    long blk0;      // first output sample of this block
    bool active;
    bool in_interior;    // all N input samples of the block lie in the current rows: no history, no zero fill
    bool out_interior;   // all N - ov outputs of the block are valid
    const void *in_base; // sample blk0 - ov of this stream's current row (interior blocks)
    void *out_base;      // output sample blk0 of this stream
    // Pass 0 lets thread t work on column (t + rot) mod S instead of column t, so that every warp's 32 consecutive samples
    // start on a 128-byte line of the row: a block starts wherever hop * block index puts it (3842 samples), and a warp
    // access that straddles one more line than it has to costs the LSU data pipe -- this kernel's busiest unit -- one more
    // wavefront per load and per store (3 instead of 2 for complex64, 2 instead of 1 for int16).
    int rot_in, rot_out;
};

template <int KF> struct RawOf { typedef uint32_t type; };
template <> struct RawOf<KF_U8> { typedef uint16_t type; };
template <> struct RawOf<KF_C64> { typedef float2 type; };
template <> struct RawOf<KF_REAL> { typedef float type; };

__device__ __forceinline__ float2 cvt_raw(uint32_t w) { return cvt_i16(w); }
__device__ __forceinline__ float2 cvt_raw(uint16_t w) { return cvt_u8(w); }
__device__ __forceinline__ float2 cvt_raw(float2 w) { return w; }
__device__ __forceinline__ float2 cvt_raw(float w) { return make_float2(w, 0.f); }

template <int KF> __device__ __forceinline__ typename RawOf<KF>::type load_raw_at(const void *base, long i)
{
    return __ldg(reinterpret_cast<const typename RawOf<KF>::type *>(base) + i);
}
template <int KF> __device__ __forceinline__ float2 load_at(const void *base, long i) { return cvt_raw(load_raw_at<KF>(base, i)); }

// source of pass 0 (forward): global memory, with format conversion, history and zero fill
template <int KF> __device__ __forceinline__ float2 load_input(const Ctx &c, int i)
{
    if (!c.active) return make_float2(0.f, 0.f);
    const long n = c.blk0 - c.p.ov + i;
    if (n >= c.p.in.n) return make_float2(0.f, 0.f);
    if (n >= 0) return load_at<KF>(c.p.in.cur, (long)c.stream * c.p.in.pitch + n);
    const long h = (long)c.p.in.H + n;
    return h < 0 ? make_float2(0.f, 0.f) : load_at<KF>(c.p.in.hist, (long)c.stream * c.p.in.H + h);
}

template <int KF> __device__ __forceinline__ void store_output(const Ctx &c, int i, float2 v)
{
    if (i < c.p.ov) return;                                 // overlap discard
    if (!c.out_interior && (!c.active || c.blk0 + (i - c.p.ov) >= c.p.in.n)) return;
    if (KF == KF_REAL) reinterpret_cast<float *>(c.out_base)[i - c.p.ov] = v.x;
    else reinterpret_cast<float2 *>(c.out_base)[i - c.p.ov] = v;
}

// ---- twiddles of one radix-16 pass ------------------------------------------------------------------------------
// w[q] = exp(-2 pi i j q / M), q = 1..15.  Only the powers 1, 2, 4, 8 come from the table (one coalesced row piece each);
// the other eleven are products of those, at most three multiplications deep (relative error < 3e-7, two orders inside
// the parity budget).  With fifteen loads per thread and pass the twiddle rows were 18 % of this kernel's LSU
// wavefronts, and the LSU data pipe is what bounds it (profiles/r02_ncu_k1_c64_before.json: 79.5 % of peak).
// data x twiddle as FMUL2 + FFMA2 instead of four scalar instructions: slower while the shared-memory bank conflicts kept the
// LSU pipe as the bound (1.36 vs 1.23 ms), faster since the row layout removed them (cfg 1 int16 2.267 -> 2.163 ms)
#ifndef SFE_FFT_PACKED_TWMUL
#define SFE_FFT_PACKED_TWMUL 1
#endif
#ifndef SFE_FFT_PACKED_HMUL
#define SFE_FFT_PACKED_HMUL 0      // the same for the tap-spectrum multiply (tuning switch)
#endif
#ifndef SFE_FFT_TW_PRODUCTS
#define SFE_FFT_TW_PRODUCTS 1
#endif
__device__ __forceinline__ float2 cmul_pk(float2 a, float2 w)   // a * w as FMUL2 + FFMA2 (operand swizzles fold into the selectors)
{
    const cpx2 t = cmul2_pk(cpack(a), w);
    return cunpack(t);
}
template <int S> __device__ __forceinline__ void load_twiddles(const float2 *row0, int j, float2 (&w)[16])
{
    // row0 = table row of q = 1 for this pass: row q-1 starts at row0 + (q-1) S
#if SFE_FFT_TW_PRODUCTS
    w[1] = __ldg(row0 + j);
    w[2] = __ldg(row0 + S + j);
    w[4] = __ldg(row0 + 3 * S + j);
    w[8] = __ldg(row0 + 7 * S + j);
    w[3] = cmul_pk(w[1], w[2]);
    w[5] = cmul_pk(w[4], w[1]);
    w[6] = cmul_pk(w[4], w[2]);
    w[7] = cmul_pk(w[4], w[3]);
    w[9] = cmul_pk(w[8], w[1]);
    w[10] = cmul_pk(w[8], w[2]);
    w[11] = cmul_pk(w[8], w[3]);
    w[12] = cmul_pk(w[8], w[4]);
    w[13] = cmul_pk(w[8], w[5]);
    w[14] = cmul_pk(w[8], w[6]);
    w[15] = cmul_pk(w[8], w[7]);
#else
#pragma unroll
    for (int q = 1; q < 16; q++) w[q] = __ldg(row0 + (q - 1) * S + j);
#endif
}

// ---- one forward (DIF) pass -------------------------------------------------------------------
// P = pass index. FIRST: read from global. The last pass leaves its result in registers (vout).
template <int N, int P, bool KEEP, int KF>
__device__ __forceinline__ void fwd_pass(float2 *X, int t, const Ctx &c, float2 (&keep)[16])
{
    using PL = Plan<N>;
    constexpr int R = PL::radix(P), M = PL::size(P), S = M / R, NB = 16 / R;
#pragma unroll
    for (int u = 0; u < NB; u++) {
        const int beta = BetaPerm<N, P>::apply(t + u * PL::TPF);
        const int b = (beta / S) * M, j = P == 0 ? ((beta + c.rot_in) & (S - 1)) : beta % S;
        const int pb = phi<N>(b + j);
        float2 v[R];
        if (P == 0) {
            if (c.in_interior) {   // all loads first, conversions after: R independent requests in flight per thread
                typename RawOf<KF>::type raw[R];
#pragma unroll
                for (int k = 0; k < R; k++) raw[k] = load_raw_at<KF>(c.in_base, b + j + k * S);
#pragma unroll
                for (int k = 0; k < R; k++) v[k] = cvt_raw(raw[k]);
            } else {
#pragma unroll
                for (int k = 0; k < R; k++) v[k] = load_input<KF>(c, b + j + k * S);
            }
        } else {
#pragma unroll
            for (int k = 0; k < R; k++) v[k] = X[phi_at<N, S>(pb, b + j, k)];
        }
        float2 w[16];
        if constexpr (S > 1) load_twiddles<S>(c.p.tw + PL::tw_off(P), j, w);     // (S > 1 only in radix-16 passes)
        if (SFE_FFT_PACKED) dft_dif_packed<R, -1>(v);
        else dft_dif<R, -1>(v);
#pragma unroll
        for (int sl = 0; sl < R; sl++) {
            const int q = brev<R>(sl);
            float2 y = v[sl];
            if (S > 1 && q != 0) y = SFE_FFT_PACKED_TWMUL ? cmul_pk(y, w[q & 15]) : cmul(y, w[q & 15]);
            if (KEEP) keep[u * R + q] = y;
            else X[phi_at<N, S>(pb, b + j, q)] = y;
        }
    }
}

// ---- one inverse (DIT) pass: exact mirror of fwd_pass<P> ----------------------------------------
template <int N, int P, bool FROMREG, int KF>
__device__ __forceinline__ void inv_pass(float2 *X, int t, const Ctx &c, float2 (&keep)[16])
{
    using PL = Plan<N>;
    constexpr int R = PL::radix(P), M = PL::size(P), S = M / R, NB = 16 / R;
#pragma unroll
    for (int u = 0; u < NB; u++) {
        const int beta = BetaPerm<N, P>::apply(t + u * PL::TPF);
        const int b = (beta / S) * M, j = P == 0 ? ((beta + c.rot_out) & (S - 1)) : beta % S;
        const int pb = phi<N>(b + j);
        float2 v[R];
        float2 w[16];
        if constexpr (S > 1) load_twiddles<S>(c.p.tw + PL::tw_off(P), j, w);
#pragma unroll
        for (int q = 0; q < R; q++) {
            float2 y = FROMREG ? keep[u * R + q] : X[phi_at<N, S>(pb, b + j, q)];
            if (S > 1 && q != 0) y = SFE_FFT_PACKED_TWMUL ? cmul_pk(y, make_float2(w[q & 15].x, -w[q & 15].y)) : cmul_conj(y, w[q & 15]);
            v[q] = y;
        }
        if (SFE_FFT_PACKED) dft_dif_packed<R, +1>(v);
        else dft_dif<R, +1>(v);
#pragma unroll
        for (int sl = 0; sl < R; sl++) {
            const int k = brev<R>(sl);
            // (interpolating epilogue: the time-domain block goes back into the slots this thread has just read)
            if (P == 0 && c.p.lerp_sched == nullptr) store_output<KF>(c, b + j + k * S, v[sl]);
            else X[phi_at<N, S>(pb, b + j, k)] = v[sl];
        }
    }
}

template <int N, int P, int KF> struct FwdChain {
    static __device__ __forceinline__ void run(float2 *X, int t, const Ctx &c, float2 (&keep)[16])
    {
        constexpr bool last = (P == Plan<N>::NP - 1);
        fwd_pass<N, P, last, KF>(X, t, c, keep);
        if constexpr (!last) {
            __syncthreads();
            FwdChain<N, P + 1, KF>::run(X, t, c, keep);
        }
    }
};

template <int N, int P, int KF> struct InvChain {
    static __device__ __forceinline__ void run(float2 *X, int t, const Ctx &c, float2 (&keep)[16])
    {
        constexpr bool first = (P == Plan<N>::NP - 1);
        inv_pass<N, P, first, KF>(X, t, c, keep);
        if constexpr (P > 0) {
            __syncthreads();
            InvChain<N, P - 1, KF>::run(X, t, c, keep);
        }
    }
};

template <int N, int FPB, int MINB, int KF>
__global__ void __launch_bounds__(Plan<N>::TPF *FPB, MINB) blkconv_fft_kernel(const BlkconvParams p)
{
    using PL = Plan<N>;
    extern __shared__ float2 smem[];
    const int f = threadIdx.x / PL::TPF, t = threadIdx.x % PL::TPF;
    const long id = (long)blockIdx.x * FPB + f;
    const long total = (long)p.n_streams * p.n_blocks;
    Ctx c{p, 0, 0, id < total, false, false, nullptr, nullptr, 0, 0};
    if (c.active) {
        c.stream = (int)(id / p.n_blocks);
        c.blk0 = p.blk_start + (id % p.n_blocks) * (long)p.hop;
        const long first = c.blk0 - p.ov;
        c.in_interior = first >= 0 && first + N <= p.in.n;
        c.out_interior = c.blk0 + (N - p.ov) <= p.in.n;
        constexpr int EB = KF == KF_I16 ? 4 : KF == KF_U8 ? 2 : KF == KF_C64 ? 8 : 4;
        c.in_base = static_cast<const char *>(p.in.cur) + ((long)c.stream * p.in.pitch + first) * EB;
        c.out_base = static_cast<char *>(p.out) + ((long)c.stream * p.out_pitch + c.blk0) * (KF == KF_REAL ? 4 : 8);
        constexpr int OB = KF == KF_REAL ? 4 : 8;
        c.rot_in = (int)(((128u - (unsigned)(reinterpret_cast<uintptr_t>(c.in_base) & 127u)) & 127u) / EB);
        c.rot_out = (int)(((128u - (unsigned)((reinterpret_cast<uintptr_t>(c.out_base) - (uintptr_t)p.ov * OB) & 127u)) & 127u) / OB);
    }
    float2 *X = smem + f * PL::PADN;
    float2 keep[16];

    FwdChain<N, 0, KF>::run(X, t, c, keep);

    // thread-private spectrum slice: last pass has S = 1, element index = R*beta + q.  The tap spectrum is stored
    // slot-major, hp[(u*R + q) * TPF + t], so that the threads of a warp read consecutive words.
    constexpr int RL = PL::radix(PL::NP - 1), NB = 16 / RL;
    if (p.mode == BLKCONV_MODE_SPECTRUM) {
        // tap spectrum in the kernel's own digit-reversed order, pre-scaled (blkconv.cxx:50,60-69)
        if (c.active) {
#pragma unroll
            for (int u = 0; u < NB; u++)
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    const float2 v = keep[u * RL + q];
                    reinterpret_cast<float2 *>(p.out)[(u * RL + q) * PL::TPF + t] = make_float2(v.x * p.scale, v.y * p.scale);
                }
        }
        return;
    }
#pragma unroll
    for (int u = 0; u < NB; u++)
#pragma unroll
        for (int q = 0; q < RL; q++)
            keep[u * RL + q] = SFE_FFT_PACKED_HMUL ? cmul_pk(keep[u * RL + q], __ldg(p.hp + (u * RL + q) * PL::TPF + t))
                                                   : cmul(keep[u * RL + q], __ldg(p.hp + (u * RL + q) * PL::TPF + t));

    InvChain<N, PL::NP - 1, KF>::run(X, t, c, keep);

    if (p.lerp_sched != nullptr) {
        // X[phi(i)] now holds s[blk0 - ov + i]; valid for i >= ov.  The schedule is the same for every stream.
        __syncthreads();
        if (c.active) {
            const long blk = id % p.n_blocks;
            const int k_lo = __ldg(p.lerp_kidx + blk), k_hi = __ldg(p.lerp_kidx + blk + 1);
            const long org = c.blk0 - p.ov;
            float2 *orow = reinterpret_cast<float2 *>(p.out) + (long)c.stream * p.out_pitch;
            for (int k = k_lo + t; k < k_hi; k += PL::TPF) {
                const int2 e = __ldg(reinterpret_cast<const int2 *>(p.lerp_sched) + k);   // {pos, mu}
                const int i = (int)((long)e.x - org);
                const float2 a = X[phi<N>(i)], b2 = X[phi<N>(i + 1)];
                const float w1 = __int_as_float(e.y), w0 = 1.0f - w1;
                orow[k] = make_float2(fmaf(w1, b2.x, w0 * a.x), fmaf(w1, b2.y, w0 * a.y));
            }
        }
    }
}

template <int N> int launch_n(const BlkconvParams &p, cudaStream_t st)
{
    using PL = Plan<N>;
    constexpr int FPB = PL::TPF >= 128 ? 1 : 128 / PL::TPF;
    const long total = (long)p.n_streams * p.n_blocks;
    if (total <= 0) return SFE_OK;
    const size_t smem = sizeof(float2) * PL::PADN * FPB;
    // resident CTAs per SM the register allocation must allow (SFE_K1_THREADS)
    // (1024 threads only where it was measured to pay and fits 64 registers without spills: N >= 2048; the two-FFT CTAs of the
    // smaller sizes keep 768)
    constexpr int MINB = PL::TPF * FPB <= 256 ? (N >= 2048 ? SFE_K1_THREADS : 768) / (PL::TPF * FPB) : 1;
    void (*kern)(const BlkconvParams) =
        p.in_real ? blkconv_fft_kernel<N, FPB, MINB, KF_REAL>
                  : p.in.fmt == SFE_IN_I16 ? blkconv_fft_kernel<N, FPB, MINB, KF_I16>
                                           : p.in.fmt == SFE_IN_U8 ? blkconv_fft_kernel<N, FPB, MINB, KF_U8> : blkconv_fft_kernel<N, FPB, MINB, KF_C64>;
    if (smem > 48 * 1024 &&
        !SFE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)))
        return SFE_ERR_CUDA;
    const long grid = (total + FPB - 1) / FPB;
    kern<<<(unsigned)grid, PL::TPF * FPB, smem, st>>>(p);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

template <int N> void tw_table_n(std::vector<float2> &h)
{
    using PL = Plan<N>;
    h.assign((size_t)PL::tw_off(PL::NP) + 1, make_float2(1.f, 0.f));
    for (int p = 0; p < PL::NP; p++) {
        const int R = PL::radix(p), M = PL::size(p), S = M / R;
        if (S <= 1) continue;
        for (int q = 1; q < R; q++)
            for (int j = 0; j < S; j++) {
                const double a = -2.0 * M_PI * (double)j * (double)q / (double)M;
                h[(size_t)PL::tw_off(p) + (size_t)(q - 1) * S + j] = make_float2((float)cos(a), (float)sin(a));
            }
    }
}

}  // namespace

// host image of the per-pass twiddle rows the kernel for fft_len expects (Plan<N>::tw_off)
bool blkconv_twiddle_table(int fft_len, std::vector<float2> &h)
{
    switch (fft_len) {
        case 32: tw_table_n<32>(h); return true;
        case 64: tw_table_n<64>(h); return true;
        case 128: tw_table_n<128>(h); return true;
        case 256: tw_table_n<256>(h); return true;
        case 512: tw_table_n<512>(h); return true;
        case 1024: tw_table_n<1024>(h); return true;
        case 2048: tw_table_n<2048>(h); return true;
        case 4096: tw_table_n<4096>(h); return true;
        case 8192: tw_table_n<8192>(h); return true;
        default: return false;
    }
}

int launch_blkconv(int fft_len, const BlkconvParams &p, cudaStream_t st)
{
    switch (fft_len) {
        case 32: return launch_n<32>(p, st);
        case 64: return launch_n<64>(p, st);
        case 128: return launch_n<128>(p, st);
        case 256: return launch_n<256>(p, st);
        case 512: return launch_n<512>(p, st);
        case 1024: return launch_n<1024>(p, st);
        case 2048: return launch_n<2048>(p, st);
        case 4096: return launch_n<4096>(p, st);
        case 8192: return launch_n<8192>(p, st);
        default:
            set_error("blkconv: fft_len must be a power of two in [32, 8192]");
            return SFE_ERR_ARG;
    }
}

}  // namespace sfe
