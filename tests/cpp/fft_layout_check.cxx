// fft_layout_check.cxx -- host check of the block convolver's shared-memory layout (simplefe_b200/csrc/sfe_fft_plan.h):
// for every transform size with the row-padded layout, every pass, every register index k and every half-warp, the 16
// lanes must touch 16 different banks (8-byte elements: bank = padded index mod 16), pass 0 for EVERY column rotation;
// and the assignment of sub-transforms to threads must be a bijection.  Test infrastructure (tests/test_abi.py).
#include <cstdio>
#include <set>

#include "sfe_fft_plan.h"

using namespace sfe::fftplan;

template <int N, int P> int check_pass()
{
    using PL = Plan<N>;
    constexpr int R = PL::radix(P), M = PL::size(P), S = M / R, NB = 16 / R;
    int bad = 0;
    std::set<int> seen;
    const int rots = P == 0 ? 32 : 1;
    for (int rot = 0; rot < rots; rot++)
        for (int u = 0; u < NB; u++)
            for (int t0 = 0; t0 < PL::TPF; t0 += 16)
                for (int k = 0; k < R; k++) {
                    std::set<int> banks;
                    for (int l = 0; l < 16; l++) {
                        const int beta = BetaPerm<N, P>::apply(t0 + l + u * PL::TPF);
                        if (k == 0 && rot == 0) seen.insert(beta);
                        const int b = (beta / S) * M, j = P == 0 ? ((beta + rot) & (S - 1)) : beta % S;
                        banks.insert(phi<N>(b + j + k * S) & 15);
                    }
                    if (banks.size() != 16) bad++;
                }
    const bool bij = seen.size() == (size_t)PL::TPF * NB && *seen.begin() == 0 && *seen.rbegin() == PL::TPF * NB - 1;
    printf("N=%d pass %d (radix %d, stride %d): %s, %d conflicted half-warp accesses\n", N, P, R, S, bij ? "bijection" : "NOT a bijection", bad);
    return bad + (bij ? 0 : 1);
}

template <int N, int P = 0> int check_size()
{
    int bad = check_pass<N, P>();
    if constexpr (P + 1 < Plan<N>::NP) bad += check_size<N, P + 1>();
    return bad;
}

int main()
{
    static_assert(Plan<4096>::ROWPAD && Plan<1024>::ROWPAD && !Plan<512>::ROWPAD, "row layout from 1024 points on");
    const int bad = check_size<1024>() + check_size<2048>() + check_size<4096>() + check_size<8192>();
    printf(bad ? "FAIL\n" : "OK: every half-warp access is conflict free\n");
    return bad ? 1 : 0;
}
