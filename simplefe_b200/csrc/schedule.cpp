// schedule.cpp -- see schedule.h.  Pure host code, IEEE single precision, no FMA contraction
// (compiled with -ffp-contract=off by simplefe_b200/build.py).
#include "schedule.h"
#include <cmath>
#include <cstring>

namespace sfe {

int walk_call(WalkState &st, int up, int n_in, int out_len, float rate, int64_t base_in,
              EntryVec &out)
{
    int n_out = 0;
    const int64_t base = base_in * (int64_t)up;
    float t = (float)st.pos + st.mu;          // decimate.cxx:73  (float t = m_pos + m_mu)
    const float step = rate * (float)up;               // rate * m_up_ratio, rounded to float
    const int64_t lim = (int64_t)n_in * up;
    if (st.held) {                                     // decimate.cxx:96-100
        // held-over output: left point is the last grid position of the previous call (-1),
        // right point is (phase 0, n 0); weight is the mu saved with it
        out.push_back({(int32_t)(base - 1), st.mu});
        n_out++;
        st.held = false;
        t = t + step;
    }
    // the loop writes through a raw pointer into space reserved up front: at most out_len outputs per call, and never
    // more than one per grid position (step >= 1 grid unit), whatever capacity the caller claims to have
    const size_t at = out.size();
    int64_t room = out_len > n_out ? (int64_t)(out_len - n_out) : 0;
    const int64_t most = lim + 2 + (st.pos < 0 ? -(int64_t)st.pos : 0);
    if (room > most) room = most;
    out_len = n_out + (int)room;                       // the loop below never writes past the reservation
    out.resize(at + (size_t)room);
    SchedEntry *dst = out.data() + at;
    const int32_t base32 = (int32_t)base;
    // state lives in locals inside the loop (the only loop-carried dependency is t = tf + step); written back after it
    int pos = st.pos;
    float mu = st.mu;
    bool held = st.held;
    for (;;) {
        const float tf = t;
        // Closed-form run.  Inside one binade [2^e, 2^(e+1)) every time value is a multiple of u = 2^(e-23), and
        // RN(t + step) - t is the same multiple INC of u for every t whose sum stays representable at that ulp (step
        // rounded to the ulp; a tie only makes the first addition differ, hence the three probes).  So the next values
        // are X*u with X += INC in integers: no floating-point dependency chain (the reference's recurrence costs one
        // add latency per output), identical bits.  Anything else -- t < 1, the addition that leaves the binade, the
        // end-of-call checks -- goes through the reference's own sequence below.
        if (tf >= 1.0f && tf < 4194304.0f) {
            uint32_t b0, b1, b2, b3;
            const float t1 = tf + step, t2 = t1 + step, t3 = t2 + step;
            memcpy(&b0, &tf, 4);
            memcpy(&b1, &t1, 4);
            memcpy(&b2, &t2, 4);
            memcpy(&b3, &t3, 4);
            if ((b3 >> 23) == (b0 >> 23)) {            // all four in the binade of tf
                const int e = (int)(b0 >> 23) - 127;
                const uint32_t X0 = (b0 & 0x7FFFFFu) | 0x800000u, X1 = (b1 & 0x7FFFFFu) | 0x800000u;
                const uint32_t X2 = (b2 & 0x7FFFFFu) | 0x800000u, X3 = (b3 & 0x7FFFFFu) | 0x800000u;
                const uint32_t INC = X1 - X0;
                if (INC > 0 && X2 - X1 == INC && X3 - X2 == INC) {
                    const int sh = 23 - e;
                    const uint32_t fm = (1u << sh) - 1u;
                    uint32_t ub = (uint32_t)(e - 23 + 127) << 23;
                    float u;
                    memcpy(&u, &ub, 4);
                    // length of the run: (a) the sum from X must still round at this ulp, X + INC + 1 <= 2^24;
                    // (b) the reference's end-of-call checks stay false, (X >> sh) + 1 < lim; (c) room for outputs
                    int64_t n_run = (int64_t)((0xFFFFFFu - X0) / INC);
                    const int64_t L = lim > 1 ? ((lim - 1) << sh) : 0;          // X < L  <=>  (X >> sh) + 1 < lim
                    const int64_t n_b = L > (int64_t)X0 ? (L - (int64_t)X0 + INC - 1) / INC : 0;
                    if (n_b < n_run) n_run = n_b;
                    if (out_len - n_out < n_run) n_run = out_len - n_out;
                    if (n_run > 0) {
                        const int nr = (int)n_run;
                        for (int j = 0; j < nr; j++) {                           // no dependency between iterations
                            const uint32_t X = X0 + (uint32_t)j * INC;
                            dst[j] = SchedEntry{base32 + (int32_t)(X >> sh), (float)(X & fm) * u};
                        }
                        dst += nr;
                        n_out += nr;
                        t = (float)(X0 + (uint32_t)nr * INC) * u;                // exact: at most 2^24
                        continue;
                    }
                }
            }
        }
        pos = (int)tf;                                 // floorf(t) without the libm call: truncate, then fix t < 0
        if ((float)pos > tf) pos--;                    // decimate.cxx:106
        mu = tf - (float)pos;                          // decimate.cxx:107
        // n0 = pos / up and n1 = (pos + 1) / up (C truncating division, decimate.cxx:109-113) are only compared
        // with n_in: pos / up >= n_in  <=>  pos >= n_in * up for any sign of pos, so the walk needs no division
        if (pos >= lim || n_out >= out_len) break;     // decimate.cxx:115
        if (pos + 1 >= lim) {                          // decimate.cxx:118-121
            held = true;
            break;
        }
        *dst++ = SchedEntry{base32 + pos, mu};
        n_out++;
        t = tf + step;                                 // decimate.cxx:124
    }
    st.pos = pos;
    st.mu = mu;
    st.held = held;
    out.resize((size_t)(dst - out.data()));
    st.pos -= n_in * up;                               // decimate.cxx:127
    return n_out;
}

// The same walk, recorded run by run (see SchedRun).  Every decision is the one walk_call takes -- the closed-form test,
// the run length, the sequential fall-back -- only the per-output loop is gone.
int walk_call_runs(WalkState &st, int up, int n_in, int out_len, float rate, int64_t base_in, RunVec &runs, int64_t *k)
{
    int n_out = 0;
    const int64_t base = base_in * (int64_t)up;
    float t = (float)st.pos + st.mu;
    const float step = rate * (float)up;
    const int64_t lim = (int64_t)n_in * up;
    auto single = [&](int32_t pos, float mu) {
        uint32_t mb;
        memcpy(&mb, &mu, 4);
        runs.push_back(SchedRun{*k, 1, pos, 0u, 0u, -1, mb});
        (*k)++;
    };
    if (st.held) {
        single((int32_t)(base - 1), st.mu);
        n_out++;
        st.held = false;
        t = t + step;
    }
    {
        int64_t room = out_len > n_out ? (int64_t)(out_len - n_out) : 0;
        const int64_t most = lim + 2 + (st.pos < 0 ? -(int64_t)st.pos : 0);
        if (room > most) room = most;
        out_len = n_out + (int)room;
    }
    const int32_t base32 = (int32_t)base;
    int pos = st.pos;
    float mu = st.mu;
    bool held = st.held;
    for (;;) {
        const float tf = t;
        if (tf >= 1.0f && tf < 4194304.0f) {
            uint32_t b0, b1, b2, b3;
            const float t1 = tf + step, t2 = t1 + step, t3 = t2 + step;
            memcpy(&b0, &tf, 4);
            memcpy(&b1, &t1, 4);
            memcpy(&b2, &t2, 4);
            memcpy(&b3, &t3, 4);
            if ((b3 >> 23) == (b0 >> 23)) {
                const int e = (int)(b0 >> 23) - 127;
                const uint32_t X0 = (b0 & 0x7FFFFFu) | 0x800000u, X1 = (b1 & 0x7FFFFFu) | 0x800000u;
                const uint32_t X2 = (b2 & 0x7FFFFFu) | 0x800000u, X3 = (b3 & 0x7FFFFFu) | 0x800000u;
                const uint32_t INC = X1 - X0;
                if (INC > 0 && X2 - X1 == INC && X3 - X2 == INC) {
                    const int sh = 23 - e;
                    const uint32_t ub = (uint32_t)(e - 23 + 127) << 23;
                    float u;
                    memcpy(&u, &ub, 4);
                    int64_t n_run = (int64_t)((0xFFFFFFu - X0) / INC);
                    const int64_t L = lim > 1 ? ((lim - 1) << sh) : 0;
                    const int64_t n_b = L > (int64_t)X0 ? (L - (int64_t)X0 + INC - 1) / INC : 0;
                    if (n_b < n_run) n_run = n_b;
                    if (out_len - n_out < n_run) n_run = out_len - n_out;
                    if (n_run > 0) {
                        const int nr = (int)n_run;
                        runs.push_back(SchedRun{*k, nr, base32, X0, INC, sh, ub});
                        *k += nr;
                        n_out += nr;
                        t = (float)(X0 + (uint32_t)nr * INC) * u;
                        continue;
                    }
                }
            }
        }
        pos = (int)tf;
        if ((float)pos > tf) pos--;
        mu = tf - (float)pos;
        if (pos >= lim || n_out >= out_len) break;
        if (pos + 1 >= lim) {
            held = true;
            break;
        }
        single(base32 + pos, mu);
        n_out++;
        t = tf + step;
    }
    st.pos = pos;
    st.mu = mu;
    st.held = held;
    st.pos -= n_in * up;
    return n_out;
}

int64_t walk_batch_runs(WalkState &st, int up, int blk, float rate, int64_t n_in, RunVec &runs)
{
    runs.clear();
    const int out_len = (int)floorf((float)blk / rate) + 2;
    int64_t k = 0;
    for (int64_t off = 0; off < n_in; off += blk) {
        const int n = (int)((n_in - off) < blk ? (n_in - off) : blk);
        walk_call_runs(st, up, n, out_len, rate, off, runs, &k);
    }
    return k;
}

void expand_runs(const RunVec &runs, EntryVec &out)
{
    for (const SchedRun &r : runs) {
        if (r.sh < 0) {
            float mu;
            memcpy(&mu, &r.ubits, 4);
            out.push_back(SchedEntry{r.base, mu});
            continue;
        }
        float u;
        memcpy(&u, &r.ubits, 4);
        const uint32_t fm = (1u << r.sh) - 1u;
        for (int j = 0; j < r.n; j++) {
            const uint32_t X = r.X0 + (uint32_t)j * r.INC;
            out.push_back(SchedEntry{r.base + (int32_t)(X >> r.sh), (float)(X & fm) * u});
        }
    }
}

void walk_batch(WalkState &st, int up, int blk, float rate, int64_t n_in, Schedule &s, bool want_entries)
{
    s.e.clear();
    s.count = 0;
    s.uniform = false;
    s.pos0 = s.step = 0;
    const int out_len = (int)floorf((float)blk / rate) + 2;
    if (s.memo_up != up || s.memo_out_len != out_len || s.memo_rate != rate || s.memo.size() > 4096) {
        s.memo.clear();   // other parameters, or a fractional rate whose states never repeat
        s.memo_up = up;
        s.memo_out_len = out_len;
        s.memo_rate = rate;
    }
    const float stepf = rate * (float)up;
    if (stepf != floorf(stepf)) {
        // fractional step: (pos, mu, held) never repeats, so memoising calls only costs allocations; walk straight
        // into the entry list.  Such a schedule is never closed-form (mu != 0 somewhere).
        EntryVec scratch;
        EntryVec &dst = want_entries ? s.e : scratch;
        if (want_entries) dst.reserve((size_t)((double)n_in / (double)rate) + (size_t)(n_in / blk) * 2 + 8);
        for (int64_t off = 0; off < n_in; off += blk) {
            const int n = (int)((n_in - off) < blk ? (n_in - off) : blk);
            if (!want_entries) scratch.clear();
            s.count += walk_call(st, up, n, out_len, rate, off, dst);
        }
        return;
    }
    bool uni = true, have_prev = false;
    int64_t d = 0, prev = 0, first = 0;
    for (int64_t off = 0; off < n_in; off += blk) {
        const int n = (int)((n_in - off) < blk ? (n_in - off) : blk);
        uint32_t mu_bits;
        memcpy(&mu_bits, &st.mu, 4);
        const auto key = std::make_tuple(st.pos, mu_bits, st.held, n);
        auto it = s.memo.find(key);
        if (it == s.memo.end()) {
            CallMemo m;
            WalkState w = st;
            walk_call(w, up, n, out_len, rate, 0, m.e);
            m.after = w;
            for (size_t k = 0; k < m.e.size(); k++) {
                m.mu0 = m.mu0 && m.e[k].mu == 0.0f;
                if (k == 1) m.d = (int64_t)m.e[1].pos - m.e[0].pos;
                if (k >= 1) m.arith = m.arith && ((int64_t)m.e[k].pos - m.e[k - 1].pos == m.d);
            }
            it = s.memo.emplace(key, std::move(m)).first;
        }
        const CallMemo &m = it->second;
        st = m.after;
        const size_t cnt = m.e.size();
        if (cnt == 0) continue;
        const int64_t base = off * (int64_t)up;
        if (want_entries) {
            const size_t at = s.e.size();
            s.e.resize(at + cnt);
            for (size_t k = 0; k < cnt; k++) s.e[at + k] = SchedEntry{(int32_t)(base + m.e[k].pos), m.e[k].mu};
        }
        // uniformity of the batch from the calls' summaries
        uni = uni && m.mu0 && m.arith;
        const int64_t f = base + m.e.front().pos, l = base + m.e.back().pos;
        if (cnt >= 2) {
            if (d == 0) d = m.d;
            else uni = uni && d == m.d;
        }
        if (have_prev) {
            const int64_t gap = f - prev;
            if (d == 0) d = gap;
            else uni = uni && gap == d;
        } else {
            first = f;
        }
        have_prev = true;
        prev = l;
        s.count += (int64_t)cnt;
    }
    if (s.count == 0) return;
    if (s.count == 1) d = (int64_t)llroundf(rate * (float)up);
    if (uni && d > 0) {
        s.uniform = true;
        s.pos0 = first;
        s.step = d;
    }
}

}  // namespace sfe
