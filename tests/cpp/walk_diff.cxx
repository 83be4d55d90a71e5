// walk_diff.cxx -- the schedule walk of csrc/schedule.cpp (closed-form runs inside a binade) against the plain
// sequential float32 recurrence of the reference (decimate.cxx:73,96-127), bit for bit, over random parameters.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>

#include "schedule.h"

// the reference's walk, one float addition per output (the shape schedule.cpp had before the closed-form runs)
static int walk_seq(sfe::WalkState &st, int up, int n_in, int out_len, float rate, int64_t base_in, sfe::EntryVec &out)
{
    int n_out = 0;
    const int64_t base = base_in * (int64_t)up;
    float t = (float)st.pos + st.mu;
    const float step = rate * (float)up;
    if (st.held) {
        out.push_back({(int32_t)(base - 1), st.mu});
        n_out++;
        st.held = false;
        t = t + step;
    }
    for (;;) {
        const int pos = (int)floorf(t);
        st.pos = pos;
        st.mu = t - (float)pos;
        const int n0 = pos / up, n1 = (pos + 1) / up;            // decimate.cxx:109-113 (C division)
        if (n0 >= n_in || n_out >= out_len) break;
        if (n1 >= n_in) {
            st.held = true;
            break;
        }
        out.push_back({(int32_t)(base + pos), st.mu});
        n_out++;
        t = t + step;
    }
    st.pos -= n_in * up;
    return n_out;
}

int main(int argc, char **argv)
{
    std::mt19937 rng(argc > 1 ? atoi(argv[1]) : 1);
    const int iters = argc > 2 ? atoi(argv[2]) : 3000;
    long cases = 0, entries = 0, run_total = 0;
    for (int it = 0; it < iters; it++) {
        const int up = 1 + rng() % 24;
        const int blk = 16 + rng() % 8192;
        float rate;
        switch (rng() % 6) {
            case 0: rate = 1.0f / up + (rng() % 100000) * 1e-5f * 8.0f; break;
            case 1: rate = (float)(1 + rng() % 700) + (rng() % 1000) * 1e-3f; break;
            case 2: rate = (float)(rng() % 64 + 1) / (float)(rng() % 7 + 1); break;
            case 3: rate = 2.5003f; break;
            case 4: rate = (float)(rng() % 2048 + 1) / 1024.0f; break;   // few mantissa bits: ties and exact additions
            default: rate = 1.0f + (rng() % 4096) / 4096.0f * (rng() % 40); break;
        }
        if (rate * up < 1.0f) rate = 1.0f / up;
        sfe::WalkState a, b, r;
        sfe::RunVec runs;
        int64_t kr = 0;
        const int out_len = (int)floorf((float)blk / rate) + 2;
        sfe::EntryVec ea, eb;
        long off = 0;
        for (int call = 0; call < 12; call++) {
            const int n = (rng() % 4 == 0) ? 1 + (int)(rng() % blk) : blk;
            const int na = sfe::walk_call(a, up, n, out_len, rate, off, ea);
            const int nb = walk_seq(b, up, n, out_len, rate, off, eb);
            const int nr = sfe::walk_call_runs(r, up, n, out_len, rate, off, runs, &kr);   // the run-recording flavour
            if (nr != nb || r.pos != b.pos || memcmp(&r.mu, &b.mu, 4) || r.held != b.held || kr != (int64_t)eb.size()) {
                printf("RUN MISMATCH (state) up %d blk %d rate %.9g call %d n %d: %d vs %d outputs\n", up, blk, rate, call, n, nr, nb);
                return 1;
            }
            if (na != nb || a.pos != b.pos || memcmp(&a.mu, &b.mu, 4) || a.held != b.held || ea.size() != eb.size() ||
                memcmp(ea.data(), eb.data(), ea.size() * sizeof(sfe::SchedEntry))) {
                printf("MISMATCH up %d blk %d rate %.9g call %d n %d: %d vs %d outputs\n", up, blk, rate, call, n, na, nb);
                return 1;
            }
            off += n;
            cases++;
        }
        sfe::EntryVec er;
        sfe::expand_runs(runs, er);
        if (er.size() != eb.size() || memcmp(er.data(), eb.data(), er.size() * sizeof(sfe::SchedEntry))) {
            printf("RUN MISMATCH (entries) up %d blk %d rate %.9g\n", up, blk, rate);
            return 1;
        }
        run_total += (long)runs.size();
        entries += (long)ea.size();
    }
    printf("ok: %ld calls, %ld entries identical (walk_call, and %ld runs of walk_call_runs expanded)\n", cases, entries, run_total);
    return 0;
}
