// schedule.h -- host-side output-time schedule of the polyphase stages.
//
// The reference walks a *float32* time variable over the L-times upsampled grid
// (libdsp/decimate.cxx:73,96-127; libdsp/resample.cxx:89,119-150).  Which grid positions are
// produced, the interpolation weight mu and the left-over rule depend only on (rate, upsample,
// n_in, out_len) and on three scalars carried between calls -- never on the samples.  The GPU
// kernels therefore take the schedule as input; this file generates it with the reference's
// exact single-precision recurrence so that output counts and sample indices are bit-exact.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <tuple>
#include <utility>
#include <vector>

namespace sfe {

struct WalkState {        // decimate.h:55-59 / resample.h:56-59
    int pos = 0;          // m_pos: next output position on the upsampled grid, relative to the
                          //        first sample of the next call
    float mu = 0.0f;      // m_mu
    bool held = false;    // m_is_leftover: an output straddles the call boundary
};

struct SchedEntry {
    int32_t pos;          // grid position of the left interpolation point, relative to the first
                          // sample of the *batch* (may be -1 for a held-over output)
    float mu;             // out = s(pos)*(1-mu) + mu*s(pos+1)
};

// entry lists are filled by raw stores right after a resize(): an allocator whose default construction is a no-op keeps
// resize() from zero-filling memory that is about to be overwritten
template <class T> struct DefaultInitAlloc : std::allocator<T> {
    template <class U> struct rebind { typedef DefaultInitAlloc<U> other; };
    template <class U> void construct(U *p) noexcept { ::new (static_cast<void *>(p)) U; }
    template <class U, class... A> void construct(U *p, A &&...a) { ::new (static_cast<void *>(p)) U(std::forward<A>(a)...); }
};
typedef std::vector<SchedEntry, DefaultInitAlloc<SchedEntry>> EntryVec;

// what one reference process() call does from a given state: memoised, because for integral steps the
// state (pos, mu, held) cycles through a handful of values and a batch is thousands of identical calls
struct CallMemo {
    EntryVec e;                 // entries relative to the call's first sample (base_in = 0)
    WalkState after;
    bool mu0 = true;            // all mu == 0
    bool arith = true;          // positions form an arithmetic progression
    int64_t d = 0;              // its difference (n >= 2)
};

struct Schedule {
    EntryVec e;                 // one per output, in output order (empty in summary mode)
    int64_t count = 0;          // number of outputs of the batch
    bool uniform = false;       // all mu == 0 and pos = pos0 + k*step
    int64_t pos0 = 0, step = 0;
    // key: (pos, mu bits, held, n_in of the call); parameters (up, out_len, rate) are fixed per owner
    std::map<std::tuple<int, uint32_t, bool, int>, CallMemo> memo;
    int memo_up = 0, memo_out_len = 0;
    float memo_rate = 0.f;
};

// A stretch of the walk in closed form (schedule.cpp): entry j < n of the run is
//     X = X0 + j * INC;  pos = base + (X >> sh);  mu = (float)(X & ((1u << sh) - 1)) * u        (u = float with bits ubits)
// i.e. the integer-mantissa arithmetic of one binade; sh < 0 marks a single entry {pos = base, mu = float(ubits)} that
// went through the reference's own sequence (t < 1, the addition that leaves a binade, the end of a call).  A fractional-
// rate schedule of 10^6 samples is a few thousand runs instead of 4 * 10^5 entries: the host walks run by run and the
// device expands them (expand_sched_kernel), so neither the per-output host loop nor the 3 MB upload remain.
struct SchedRun {
    int64_t k0;           // index of the run's first output in the batch
    int32_t n;            // outputs in the run
    int32_t base;         // grid position offset (call offset * up), or the position itself for a single entry
    uint32_t X0, INC;
    int32_t sh;
    uint32_t ubits;
};
typedef std::vector<SchedRun> RunVec;

// walk_call with the outputs as runs: appends to `runs`, *k is the running output index of the batch
int walk_call_runs(WalkState &st, int up, int n_in, int out_len, float rate, int64_t base_in, RunVec &runs, int64_t *k);
// walk_batch for fractional steps, runs instead of entries; returns the number of outputs
int64_t walk_batch_runs(WalkState &st, int up, int blk, float rate, int64_t n_in, RunVec &runs);
// host expansion of a run list (tests; the device does the same in expand_sched_kernel)
void expand_runs(const RunVec &runs, EntryVec &out);

// One reference process() call: appends the outputs of a call with n_in samples whose first
// sample sits at batch offset `base_in`; updates st.  Returns the number of outputs appended.
int walk_call(WalkState &st, int up, int n_in, int out_len, float rate, int64_t base_in,
              EntryVec &out);

// A batch of n_in samples fed as ceil(n_in/blk) calls of blk samples, out_len = floor(blk/rate)+2.
// want_entries = false (summary mode): only count / uniform / pos0 / step are produced -- all that the
// closed-form kernels need -- and a batch costs O(calls) host time instead of O(outputs).
void walk_batch(WalkState &st, int up, int blk, float rate, int64_t n_in, Schedule &s, bool want_entries = true);

}  // namespace sfe
