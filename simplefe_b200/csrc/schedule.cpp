// schedule.cpp -- see schedule.h.  Pure host code, IEEE single precision, no FMA contraction
// (compiled with -ffp-contract=off by simplefe_b200/build.py).
#include "schedule.h"
#include <cmath>

namespace sfe {

int walk_call(WalkState &st, int up, int n_in, int out_len, float rate, int64_t base_in,
              std::vector<SchedEntry> &out)
{
    int n_out = 0;
    const int64_t base = base_in * (int64_t)up;
    float t = (float)st.pos + st.mu;          // decimate.cxx:73  (float t = m_pos + m_mu)
    const float step = rate * (float)up;               // rate * m_up_ratio, rounded to float
    if (st.held) {                                     // decimate.cxx:96-100
        // held-over output: left point is the last grid position of the previous call (-1),
        // right point is (phase 0, n 0); weight is the mu saved with it
        out.push_back({(int32_t)(base - 1), st.mu});
        n_out++;
        st.held = false;
        t = t + step;
    }
    for (;;) {
        const float tf = t;
        st.pos = (int)floorf(tf);                      // decimate.cxx:106
        st.mu = tf - (float)st.pos;                    // decimate.cxx:107
        const int n0 = st.pos / up;                    // C truncating division, as the reference
        const int n1 = (st.pos + 1) / up;
        if (n0 >= n_in || n_out >= out_len) break;     // decimate.cxx:115
        if (n1 >= n_in) {                              // decimate.cxx:118-121
            st.held = true;
            break;
        }
        out.push_back({(int32_t)(base + st.pos), st.mu});
        n_out++;
        t = tf + step;                                 // decimate.cxx:124
    }
    st.pos -= n_in * up;                               // decimate.cxx:127
    return n_out;
}

void walk_batch(WalkState &st, int up, int blk, float rate, int64_t n_in, Schedule &s)
{
    s.e.clear();
    const int out_len = (int)floorf((float)blk / rate) + 2;
    for (int64_t off = 0; off < n_in; off += blk) {
        const int n = (int)((n_in - off) < blk ? (n_in - off) : blk);
        walk_call(st, up, n, out_len, rate, off, s.e);
    }
    s.uniform = false;
    s.pos0 = s.step = 0;
    const size_t m = s.e.size();
    if (m == 0) return;
    bool uni = true;
    const int64_t p0 = s.e[0].pos;
    const int64_t d = m > 1 ? (int64_t)s.e[1].pos - p0 : (int64_t)llroundf(rate * (float)up);
    for (size_t k = 0; k < m && uni; k++)
        uni = (s.e[k].mu == 0.0f) && ((int64_t)s.e[k].pos == p0 + (int64_t)k * d);
    if (uni && d > 0) {
        s.uniform = true;
        s.pos0 = p0;
        s.step = d;
    }
}

}  // namespace sfe
