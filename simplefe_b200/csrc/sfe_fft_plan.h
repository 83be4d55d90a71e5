// sfe_fft_plan.h -- compile-time plan of the block convolver's shared-memory FFT (blkconv_kernel.cu): pass radices,
// twiddle-table offsets, the padded shared-memory layout and the per-pass assignment of sub-transforms to threads.
// Plain constexpr C++: tests/cpp/fft_layout_check.cxx includes it on the host and checks every half-warp access of every
// pass for bank conflicts.
#pragma once
#ifndef __CUDACC__
#ifndef __host__
#define __host__
#endif
#ifndef __device__
#define __device__
#endif
#ifndef __forceinline__
#define __forceinline__ inline
#endif
#endif
#ifndef SFE_K1_ROWPAD
#define SFE_K1_ROWPAD 1
#endif

namespace sfe {
namespace fftplan {

__host__ __device__ constexpr int ilog2(int n) { return n <= 1 ? 0 : 1 + ilog2(n >> 1); }

template <int N> struct Plan {
    static constexpr int LOGN = ilog2(N);
    static constexpr int FULL = LOGN / 4;                  // radix-16 passes
    static constexpr int LASTR = 1 << (LOGN % 4);          // 1 = none
    static constexpr int NP = FULL + (LASTR > 1 ? 1 : 0);  // passes
    static constexpr int TPF = N / 16;                     // threads per FFT
    // Shared-memory layout.  Small transforms: one element of padding per 16 (stride-16 accesses of the last pass hit 16
    // banks).  N >= 1024 (SFE_K1_ROWPAD): 16 rows of N/16 elements with ONE element of padding per row, so the bank of
    // element e is (e[3:0] + e[s0+3:s0]) mod 16 with s0 = log2 N - 4: low four bits plus low four ROW bits.  Pass 0 lets
    // thread t work on column (t + rot) mod N/16 (Ctx::rot_in) -- with padding inside the rows a half-warp whose 16
    // columns do not start on a multiple of 16 straddles a padding slot, its first and last lane share a bank, and every
    // pass-0 exchange costs two wavefronts instead of one (35.8 M conflicts on 143 M ideal wavefronts,
    // profiles/r02k_k1_c64_summary.json); with padding only between rows any 16 consecutive columns are 16 banks.
    // The later passes keep every half-warp on 16 banks by choosing WHICH sub-transform a thread takes (BetaPerm below).
    static constexpr bool ROWPAD = SFE_K1_ROWPAD && N >= 1024;
    static constexpr int PADN = ROWPAD ? N + 16 : N + N / 16;   // padded shared-memory length
    static __host__ __device__ constexpr int radix(int p) { return p < FULL ? 16 : LASTR; }
    static __host__ __device__ constexpr int size(int p)   // sub-transform length entering pass p
    {
        int m = N;
        for (int i = 0; i < p; i++) m /= radix(i);
        return m;
    }
    // Twiddles are stored per pass as rows q = 1..R-1 of S = M/R entries, tw[q-1][j] = exp(-2 pi i j q / M): the
    // threads of a warp hold consecutive j, so every twiddle load is one contiguous 128/256-byte piece of L1
    // (the flat exp(-2 pi i k / N) table indexed by (N/M) j q cost up to 16 L1 wavefronts per load).
    static __host__ __device__ constexpr int tw_off(int p)
    {
        int o = 0;
        for (int i = 0; i < p; i++) o += (radix(i) - 1) * (size(i) / radix(i));
        return o;
    }
};

// Which sub-transform a thread takes in pass P >= 1 (row-padded layout).  Pass P works on the elements
// e = b + j + k S (k = 0..R-1 in registers): beta = (b / M) S + j numbers the sub-transforms, and element bits
// [sigma, sigma + rho) (sigma = log2 S, rho = log2 R) are the register index k.  The 16 lanes of a half-warp must differ
// in (e[3:0] + e[s0+3:s0]) mod 16: lane bit i goes to element bit i where that bit is free, and to ROW bit s0 + i where
// bit i is one of the k bits -- the two groups occupy disjoint bit positions of the sum, so the 16 lanes hit 16 banks
// whatever k and the other bits are.  The thread's remaining index bits fill the remaining positions in order.  For
// N = 4096 this is: passes 0 and 1 as they were, the last pass with the two nibbles of t exchanged (a transpose).
template <int N, int P> struct BetaPerm {
    using PL = Plan<N>;
    static constexpr int R = PL::radix(P), S = PL::size(P) / R;
    static constexpr int SIG = ilog2(S), RHO = ilog2(R), S0B = PL::LOGN - 4, NBITS = PL::LOGN - RHO;
    static __host__ __device__ constexpr int dst_of_lane(int i)
    {
        const int T = (i >= SIG && i < SIG + RHO) ? S0B + i : i;   // element bit
        return T < SIG ? T : T - RHO;                              // bit of beta (the k bits are not part of beta)
    }
    static __host__ __device__ constexpr bool lane_dst(int d)
    {
        return dst_of_lane(0) == d || dst_of_lane(1) == d || dst_of_lane(2) == d || dst_of_lane(3) == d;
    }
    static __host__ __device__ constexpr int src_of(int d)
    {
        for (int i = 0; i < 4; i++)
            if (dst_of_lane(i) == d) return i;
        int s = 4;
        for (int x = 0; x < d; x++)
            if (!lane_dst(x)) s++;
        return s;
    }
    static __host__ __device__ constexpr int run_len(int d)
    {
        int len = 1;
        while (d + len < NBITS && src_of(d + len) == src_of(d) + len) len++;
        return len;
    }
    template <int D> static __host__ __device__ __forceinline__ int from(int beta)
    {
        if constexpr (D >= NBITS) return 0;
        else {
            constexpr int s = src_of(D), len = run_len(D);
            return (((beta >> s) & ((1 << len) - 1)) << D) | from<D + len>(beta);
        }
    }
    static __host__ __device__ __forceinline__ int apply(int beta)
    {
        if constexpr (!PL::ROWPAD || P == 0) return beta;
        else return from<0>(beta);
    }
};

// padded shared-memory index of element a
template <int N> __host__ __device__ __forceinline__ int phi(int a)
{
    return Plan<N>::ROWPAD ? a + (a >> (Plan<N>::LOGN - 4)) : a + (a >> 4);
}

}  // namespace fftplan
}  // namespace sfe
