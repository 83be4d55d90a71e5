// chain.cu -- batched chain object (sfe_chain_*): S independent IQ streams through
// [blkconv] -> [polyphase stage a] -> [polyphase stage b], all streams sharing one filter set.
//
// Kernel plan chosen at create time:
//   * a polyphase stage whose float32 step rate*L is an integer multiple of L produces a uniform,
//     phase-0 schedule for ever (mu == 0, exact in float32 while blk*L < 2^24): it runs on the
//     decimating-FIR kernel (K2d);
//   * blkconv followed by such a stage is FUSED: the two FIRs are convolved into one tap set (in
//     double) and only the retained outputs are ever computed -- the FFT convolver's full-rate
//     intermediate never exists;
//   * everything else runs stage by stage: K1 (FFT blkconv) and K2g (generic polyphase with the
//     host-generated float32 schedule).
// Stream state (input history per stage, resampler phase) is carried across calls so a stream may
// be fed in arbitrary pieces.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "kernels.h"

namespace sfe {
const float2 *chain_twiddles(int dev, int N);
float2 *chain_spectrum(int dev, const float *taps, int n_taps, int N, cudaStream_t st);
bool chain_device_ready(int *dev);
int chain_blkconv_fft(int n_taps, int fft_len);
}  // namespace sfe

using namespace sfe;

namespace {

enum OpKind { OP_BLKCONV, OP_DECIM, OP_POLY, OP_TOEP, OP_BANK };

struct Op {
    OpKind kind;
    int in_fmt;                    // format of this op's input rows
    int H = 0;                     // history samples kept of the input
    void *hist[2] = {nullptr, nullptr};
    // blkconv
    int N = 0, bc_ntaps = 0, hop = 0, ov = 0;
    const float2 *tw = nullptr;
    float2 *hp = nullptr;
    // polyphase
    int up = 1, blk = 0, T = 0;
    float rate = 1.f;
    float *d_taps = nullptr;
    WalkState ws;
    Schedule sched;
    SchedEntry *d_sched = nullptr;
    size_t d_sched_cap = 0;
    SchedEntry *h_sched = nullptr;   // pinned staging copy of the schedule
    size_t h_sched_cap = 0;
    cudaEvent_t h_sched_done = nullptr;
    void *reg_ptr = nullptr;         // storage of sched.e while it is registered as pinned memory (copied from in place)
    // fractional steps: the schedule as closed-form runs, expanded on the device
    bool use_runs = false;
    RunVec runs;
    SchedRun *h_runs = nullptr, *d_runs = nullptr;   // pinned staging copy / device copy
    size_t h_runs_cap = 0, d_runs_cap = 0;
    // decimating FIR
    int D = 0, QLP = 0;
    float *d_gp = nullptr;
    // tensor-core Toeplitz FIR
    ToepHost toep;
    int8_t *d_bmat = nullptr;
    // fused bank (K3): optional pre-stage (blkconv and/or integral decimator) folded into the bank stage's taps
    bool has_pre = false;
    int pre_up = 1, pre_blk = 0;
    float pre_rate = 1.f;
    WalkState pre_ws;
    Schedule pre_sched;
    std::vector<double> g;         // pre-stage FIR at the raw rate (double)
    std::vector<float> r;          // bank stage taps as the reference ctor pads them
    std::map<int, float *> d_tables;  // bank tables by alignment key (bank_table)
    int row_len = 0, n_rows = 0, rows_per_phase = 1, nc = 0, n_live = 1;
    long n_first = 0;
    bool bad_sched = false;
    bool lerp = false;             // OP_BLKCONV whose kernel also does the interpolation of the one-phase stage behind it
    int *d_kidx = nullptr;         //   first schedule entry of every FFT block (launch_block_index)
    size_t d_kidx_cap = 0;
    float lerp_tap = 0.f;          // OP_POLY with one tap and one phase: interpolation only (poly_lerp_kernel)
    bool direct_fir = false;       // OP_POLY standing in for blkconv: every sample is an output (pos = k, mu = 0)
    // per call
    long n_in = 0, n_out = 0;
    std::string name;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev;  // profiling: (start, stop) per launch
};

struct Workspace {
    std::vector<void *> mid;       // intermediate row buffers between ops (complex64)
    std::vector<size_t> mid_bytes;
    void *d_in = nullptr, *d_out = nullptr;  // staging for process_host
    size_t in_bytes = 0, out_bytes = 0;
    void *d_align = nullptr;                 // realigned copy of caller rows that TMA cannot address
    size_t align_bytes = 0;
    cudaStream_t st = nullptr;
};

}  // namespace

struct sfe_chain {
    int dev = 0, S = 0, in_fmt = 0, flags = 0;
    std::vector<Op> ops;
    int cur = 0;                   // which history buffer holds the live state
    Workspace ws[4];               // [0] also serves process_dev; process_host rotates over the first n_ws
    int n_ws = 2;
    std::string plan;
    bool profile = false;          // bracket the main kernels of THIS call with events
    int profile_every = 0;         // 0 = off, n = every n-th process call
    long profile_calls = 0;
};

namespace {

bool integral_step(const sfe_stage_desc &s, long *step)
{
    const float f = s.rate * (float)s.up;
    if (!(f >= 1.0f) || f != floorf(f)) return false;
    if ((double)s.blk * s.up + f >= 16777216.0) return false;  // float32 time stays exact
    *step = (long)f;
    return true;
}

std::vector<float> padded_taps(const sfe_stage_desc &s)
{
    // decimate.cxx:42-51 pads an even count with one zero tap; zero padding never changes results
    const int T = (s.klass == SFE_KLASS_DECIMATE && s.n_taps % 2 == 0) ? s.n_taps + 1 : s.n_taps;
    std::vector<float> t((size_t)T, 0.f);
    memcpy(t.data(), s.taps, sizeof(float) * s.n_taps);
    return t;
}

// Filter tables are uploaded at create time only.  A cudaMemcpy from pageable memory may return once the data is
// staged, before the DMA has landed, and the chain's streams are non-blocking (they do not order against the legacy
// stream): sfe_chain_create ends with one cudaDeviceSynchronize that covers every upload made through here.
bool upload(float **d, const std::vector<float> &h)
{
    return SFE_CUDA(cudaMalloc(d, sizeof(float) * h.size())) &&
           SFE_CUDA(cudaMemcpy(*d, h.data(), sizeof(float) * h.size(), cudaMemcpyHostToDevice));
}

bool make_decim(Op &op, const std::vector<double> &g, int D)
{
    op.kind = OP_DECIM;
    op.D = D;
    op.T = (int)g.size();
    int QL = (op.T + D - 1) / D;
    op.QLP = (QL + 7) / 8 * 8;
    std::vector<float> gp((size_t)D * op.QLP, 0.f);
    for (int j = 0; j < op.T; j++) gp[(size_t)(j % D) * op.QLP + j / D] = (float)g[j];
    op.H = op.T + D;
    return upload(&op.d_gp, gp);
}

bool make_toep(Op &op, const std::vector<double> &g, int in_fmt)
{
    if (!toeplitz_prepare(g, op.toep, in_fmt)) return false;
    op.kind = OP_TOEP;
    op.D = 8;
    op.T = (int)g.size();
    op.H = op.T + op.D;
    return SFE_CUDA(cudaMalloc(&op.d_bmat, op.toep.bmat.size())) &&
           SFE_CUDA(cudaMemcpy(op.d_bmat, op.toep.bmat.data(), op.toep.bmat.size(), cudaMemcpyHostToDevice));
}

// K3 bank tables for every chunk alignment a call can present (n_first mod spc), or the one table that holds all
// alignments as rows: built and uploaded once, so that no process call ever copies from pageable memory.
bool make_bank_tables(Op &op)
{
    const int spc = 16 / fmt_bytes(op.in_fmt);
    // the sample scale of the format (SURVEY.md D2 / source_c_impl.cc:126) lives in the taps
    const double scale = op.in_fmt == SFE_IN_U8 ? (double)(1.0f / 127.0f) : (double)(1.0f / 32767.0f);
    const int k0 = op.rows_per_phase > 1 ? -1 : 0, k1 = op.rows_per_phase > 1 ? -1 : spc - 1;
    for (int key = k0; key <= k1; key++) {
        std::vector<float> rows;
        int rl = 0, nr = 0, tc = 0;
        bank_table(op.g, op.D, op.r, op.up, scale, key, spc, rows, &rl, &nr, &tc);
        float *d = nullptr;
        if (!upload(&d, rows)) return false;
        op.d_tables.emplace(key, d);
    }
    return true;
}

bool alloc_hist(Op &op, int S)
{
    const size_t bytes = (size_t)S * op.H * fmt_bytes(op.in_fmt);
    for (int i = 0; i < 2; i++)
        // zero state: offset-binary bytes encode 0.0 as 128 (source_c_impl.cc:126-129)
        if (!SFE_CUDA(cudaMalloc(&op.hist[i], bytes ? bytes : 1)) ||
            !SFE_CUDA(cudaMemset(op.hist[i], op.in_fmt == SFE_IN_U8 ? 0x80 : 0, bytes)))
            return false;
    return true;
}

long stage_max_out(const Op &op, long n)
{
    if ((op.kind == OP_BLKCONV && !op.lerp) || op.direct_fir) return n;
    if (op.kind == OP_BANK) {
        if (op.has_pre) {
            const long calls = (n + op.pre_blk - 1) / op.pre_blk;
            n = (long)((double)n / op.D) + 2 * calls + 4;
        }
        const long calls = (n + op.blk - 1) / op.blk;
        return (long)((double)n / (double)op.rate) + 2 * calls + 4;
    }
    const double r = (op.kind == OP_DECIM || op.kind == OP_TOEP) ? (double)op.D : (double)op.rate;
    const long calls = op.blk > 0 ? (n + op.blk - 1) / op.blk : 1;
    return (long)((double)n / r) + 2 * calls + 4;
}

bool ensure(void **p, size_t *have, size_t need)
{
    if (*have >= need) return true;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *have = 0;
    if (!SFE_CUDA(cudaMalloc(p, need))) return false;
    *have = need;
    return true;
}

// host part of a call: advance every op's schedule for n_in new samples per stream
void plan_call(sfe_chain *c, long n_in)
{
    c->profile = c->profile_every > 0 && (c->profile_calls++ % c->profile_every) == 0;
    long n = n_in;
    for (Op &op : c->ops) {
        op.n_in = n;
        if (op.kind == OP_BLKCONV && op.lerp) {
            op.use_runs = true;
            op.n_out = (long)walk_batch_runs(op.ws, 1, op.blk, op.rate, n, op.runs);
            op.sched.uniform = false;
            op.sched.count = op.n_out;
        } else if (op.kind == OP_BLKCONV) {
            op.n_out = n;
        } else if (op.kind == OP_BANK) {
            long na = n;
            op.n_first = 0;
            op.bad_sched = false;
            if (op.has_pre) {
                walk_batch(op.pre_ws, op.pre_up, op.pre_blk, op.pre_rate, n, op.pre_sched, false);
                na = (long)op.pre_sched.count;
                if (na > 0) {
                    op.bad_sched = !op.pre_sched.uniform || op.pre_sched.step != (int64_t)op.D * op.pre_up;
                    op.n_first = op.pre_sched.pos0 >= 0 ? op.pre_sched.pos0 / op.pre_up : -1;
                }
            }
            walk_batch(op.ws, op.up, op.blk, op.rate, na, op.sched, false);
            op.n_out = (long)op.sched.count;
            if (op.n_out > 0 && !op.sched.uniform) op.bad_sched = true;
        } else if (op.direct_fir) {
            op.n_out = n;
            op.sched.uniform = true;
            op.sched.pos0 = 0;
            op.sched.step = 1;
            op.sched.count = n;
        } else {
            const float stepf = op.rate * (float)op.up;
            op.use_runs = op.kind == OP_POLY && stepf != floorf(stepf);
            if (op.use_runs) {
                // fractional step: a few closed-form runs per reference call; the device expands them (upload_schedules)
                op.n_out = (long)walk_batch_runs(op.ws, op.up, op.blk, op.rate, n, op.runs);
                op.sched.uniform = false;
                op.sched.count = op.n_out;
                n = op.n_out;
                continue;
            }
            // the closed-form kernels only need the summary; the generic kernel wants every (pos, mu)
            if (op.kind == OP_POLY) {
                // The entry list is copied to the device from where the walk writes it (its storage is registered as pinned
                // memory, upload_schedules): wait for the previous copy, and grow the storage here, unregistered -- never
                // inside the walk.
                if (op.h_sched_done) cudaEventSynchronize(op.h_sched_done);
                const size_t calls = (size_t)(n / (op.blk > 0 ? op.blk : 1)) + 2;
                const size_t bound = (size_t)((double)n / (double)op.rate * 1.01) + 4 * calls + (size_t)((double)op.blk / (double)op.rate) + 64;
                if (op.sched.e.capacity() < bound) {
                    if (op.reg_ptr) cudaHostUnregister(op.reg_ptr);
                    op.reg_ptr = nullptr;
                    op.sched.e.reserve(2 * bound);
                }
            }
            walk_batch(op.ws, op.up, op.blk, op.rate, n, op.sched, op.kind == OP_POLY);
            op.n_out = (long)op.sched.count;
        }
        n = op.n_out;
    }
}

// The schedule state of every stage, saved before plan_call so that a call that fails (capacity, arguments) leaves the
// stream exactly where it was: the histories are only swapped by a successful call, and the walk states come back
// from here.
struct WalkSnap {
    std::vector<WalkState> ws, pre;
    long profile_calls = 0;
};
WalkSnap snapshot(const sfe_chain *c)
{
    WalkSnap s;
    for (const Op &op : c->ops) {
        s.ws.push_back(op.ws);
        s.pre.push_back(op.pre_ws);
    }
    s.profile_calls = c->profile_calls;
    return s;
}
void restore(sfe_chain *c, const WalkSnap &s)
{
    for (size_t i = 0; i < c->ops.size(); i++) {
        c->ops[i].ws = s.ws[i];
        c->ops[i].pre_ws = s.pre[i];
    }
    c->profile_calls = s.profile_calls;
}

// Argument guards that need no state: schedule entries and kernel indices carry 32-bit positions on the L-times grid
// of one call (SchedEntry.pos), so a call whose grid does not fit is refused instead of wrapping silently.
int check_call(const sfe_chain *c, long n_in, const char *who)
{
    if (n_in > (1L << 30)) {
        set_error(std::string(who) + ": at most 2^30 samples per stream and call");
        return SFE_ERR_ARG;
    }
    long n = n_in;
    for (const Op &op : c->ops) {
        if (op.kind == OP_POLY && !op.direct_fir) {
            const double grid = (double)n * op.up + (double)op.rate * op.up + op.up;
            if (grid >= 2147483648.0) {
                set_error(std::string(who) + ": n_in * upsample of stage '" + op.name + "' reaches 2^31 (" + std::to_string(n) + " * " +
                          std::to_string(op.up) + "): feed the stream in smaller calls");
                return SFE_ERR_ARG;
            }
        }
        n = stage_max_out(op, n);
    }
    return SFE_OK;
}

// device part: run all ops for stream rows [s0, s0+ns) on workspace w
int run_ops(sfe_chain *c, Workspace &w, int s0, int ns, const void *d_in, long in_pitch, void *d_out, long out_pitch,
            cudaStream_t st)
{
    const void *src = d_in;
    long src_pitch = in_pitch;
    const size_t nops = c->ops.size();
    if (w.mid.size() < nops) {
        w.mid.resize(nops, nullptr);
        w.mid_bytes.resize(nops, 0);
    }
    for (size_t i = 0; i < nops; i++) {
        Op &op = c->ops[i];
        const bool last = i + 1 == nops;
        void *dst = d_out;
        long dst_pitch = out_pitch;
        if (!last) {
            dst_pitch = op.n_out > 0 ? op.n_out : 1;
            if (!ensure(&w.mid[i], &w.mid_bytes[i], sizeof(float2) * (size_t)ns * dst_pitch)) return SFE_ERR_CUDA;
            dst = w.mid[i];
        }
        const NvtxRange range(op.name.c_str());      // one range per op: its main kernel and its state carry
        const int eb = fmt_bytes(op.in_fmt);
        const char *hist_old = static_cast<const char *>(op.hist[c->cur]) + (size_t)s0 * op.H * eb;
        char *hist_new = static_cast<char *>(op.hist[c->cur ^ 1]) + (size_t)s0 * op.H * eb;
        StreamView in{src, hist_old, src_pitch, op.H, op.n_in, op.in_fmt};
        int rc = SFE_OK;
        bool carried = false;                    // the op's kernel wrote the new history itself
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (c->profile && op.ev.size() < 256) {
            if (!SFE_CUDA(cudaEventCreate(&e0)) || !SFE_CUDA(cudaEventCreate(&e1)) || !SFE_CUDA(cudaEventRecord(e0, st)))
                return SFE_ERR_CUDA;
        }
        if (op.kind == OP_BLKCONV) {
            BlkconvParams p{};
            p.in = in;
            p.in_real = 0;
            p.out = dst;
            p.out_pitch = dst_pitch;
            p.n_streams = ns;
            p.n_blocks = op.lerp ? (op.n_out > 0 ? (op.n_in - 1) / op.hop + 1 : 0) : (op.n_in + op.hop - 1) / op.hop;
            if (op.lerp) {
                p.blk_start = -1;
                p.lerp_sched = op.d_sched;
                p.lerp_kidx = op.d_kidx;
            }
            p.hop = op.hop;
            p.ov = op.ov;
            p.tw = op.tw;
            p.hp = op.hp;
            p.mode = BLKCONV_MODE_CONV;
            rc = launch_blkconv(op.N, p, st);
        } else if ((op.kind == OP_DECIM || (op.kind == OP_TOEP && op.d_gp && op.n_out < 4 * 1024)) && op.n_out > 0) {
            DecimParams p{};
            p.in = in;
            p.out = static_cast<float2 *>(dst);
            p.out_pitch = dst_pitch;
            p.n_streams = ns;
            p.n_out = op.n_out;
            p.gp = op.d_gp;
            p.D = op.D;
            p.QLP = op.QLP;
            // uniform phase-0 schedule: positions are multiples of `up` (or -1 when up == 1)
            p.n_first = op.sched.pos0 >= 0 ? op.sched.pos0 / op.up : -1;
            rc = launch_decim_fir(p, st);
        } else if (op.kind == OP_TOEP && op.n_out > 0) {
            // the kernel stages the rows by TMA, which needs 16-byte aligned rows: realign caller buffers that are not
            // (u8 rows of odd length too: the tensor map's elements are two samples wide, so its last word reaches one sample
            // past the row's end -- harmless inside a buffer of our own, not something to read from the caller's last row)
            if (((reinterpret_cast<uintptr_t>(src) & 15) || ((src_pitch * eb) & 15) || (op.in_fmt == SFE_IN_U8 && (op.n_in & 1))) &&
                op.n_in >= 128) {
                const long ap = (op.n_in + 7) & ~7L;
                if (!ensure(&w.d_align, &w.align_bytes, (size_t)ns * ap * eb)) return SFE_ERR_CUDA;
                if (!SFE_CUDA(cudaMemcpy2DAsync(w.d_align, (size_t)ap * eb, src, (size_t)src_pitch * eb, (size_t)op.n_in * eb, ns,
                                                cudaMemcpyDeviceToDevice, st)))
                    return SFE_ERR_CUDA;
                in.cur = w.d_align;
                in.pitch = ap;
            }
            ToepParams p{};
            p.in = in;
            p.out = static_cast<float2 *>(dst);
            p.out_pitch = dst_pitch;
            p.n_streams = ns;
            p.n_out = op.n_out;
            p.n_first = op.sched.pos0 >= 0 ? op.sched.pos0 / op.up : -1;
            p.KB = op.toep.KB;
            p.bmat = op.d_bmat;
            for (int t = 0; t < 3; t++) {
                p.w[t] = op.toep.w[t];
                p.corr[t] = op.toep.corr[t];
            }
            rc = launch_toeplitz(p, st);
        } else if (op.kind == OP_BANK && op.n_out > 0) {
            if (op.bad_sched) {
                set_error("fused bank: the output schedule of this call is not uniform (float32 time walk lost exactness)");
                return SFE_ERR_ARG;
            }
            // TMA needs 16-byte aligned rows: realign caller buffers that are not
            if ((reinterpret_cast<uintptr_t>(src) & 15) || ((src_pitch * eb) & 15)) {
                const long ap = (op.n_in + 7) & ~7L;                     // 16-byte rows in either format
                if (!ensure(&w.d_align, &w.align_bytes, (size_t)ns * ap * eb)) return SFE_ERR_CUDA;
                if (!SFE_CUDA(cudaMemcpy2DAsync(w.d_align, (size_t)ap * eb, src, (size_t)src_pitch * eb, (size_t)op.n_in * eb, ns,
                                                cudaMemcpyDeviceToDevice, st)))
                    return SFE_ERR_CUDA;
                in.cur = w.d_align;
                in.pitch = ap;
            }
            const int spc = 16 / eb;                                     // samples per 16-byte chunk
            const int key = op.rows_per_phase > 1 ? -1 : (int)(((op.n_first % spc) + spc) % spc);
            auto it = op.d_tables.find(key);
            if (it == op.d_tables.end()) {   // every key is built by sfe_chain_create (make_bank_tables)
                set_error("fused bank: no table for alignment key " + std::to_string(key));
                return SFE_ERR_ARG;
            }
            BankParams p{};
            p.in = in;
            p.out = static_cast<float2 *>(dst);
            p.out_pitch = dst_pitch;
            p.n_streams = ns;
            p.n_out = op.n_out;
            p.pos0 = op.sched.pos0;
            p.step = op.sched.step;
            p.L = op.up;
            p.D = op.D;
            p.n_first = op.n_first;
            p.H = op.H;
            p.table = it->second;
            p.row_len = op.row_len;
            p.nc = op.nc;
            p.n_live = op.n_live;
            p.n_rows = op.n_rows;
            p.rows_per_phase = op.rows_per_phase;
            p.hist_new = hist_new;               // the kernel carries the state itself: one launch per call
            rc = launch_bank(p, st);
            carried = rc == SFE_OK;
        } else if (op.kind == OP_POLY && op.n_out > 0) {
            PolyParams p{};
            p.in = in;
            p.in_real = 0;
            p.out = dst;
            p.out_pitch = dst_pitch;
            p.n_streams = ns;
            p.n_out = op.n_out;
            p.taps = op.d_taps;
            p.n_taps = op.T;
            p.up = op.up;
            p.uniform = op.sched.uniform ? 1 : 0;
            p.pos0 = op.sched.pos0;
            p.step = op.sched.step;
            p.sched = op.d_sched;
            p.exact = (c->flags & SFE_FLAG_EXACT) ? 1 : 0;
            p.rate = op.rate;
            p.lerp_tap = op.lerp_tap;
            rc = launch_poly_generic(p, st);
        }
        if (rc != SFE_OK) return rc;
        if (e0) {
            if (!SFE_CUDA(cudaEventRecord(e1, st))) return SFE_ERR_CUDA;
            op.ev.emplace_back(e0, e1);
        }
        if (!carried) {
            rc = launch_save_tail(hist_new, hist_old, src, src_pitch, op.n_in, op.H, eb, ns, st);
            if (rc != SFE_OK) return rc;
        }
        src = dst;
        src_pitch = dst_pitch;
    }
    return SFE_OK;
}

// schedules that are not closed-form must be on the device before the kernels run
int upload_schedules(sfe_chain *c, cudaStream_t st)
{
    for (Op &op : c->ops) {
        if (!(op.kind == OP_POLY || (op.kind == OP_BLKCONV && op.lerp)) || op.sched.uniform || op.n_out == 0) continue;
        const size_t need = sizeof(SchedEntry) * (size_t)op.n_out;
        if (op.d_sched_cap < need) {
            if (op.d_sched) cudaFree(op.d_sched);
            op.d_sched = nullptr;
            op.d_sched_cap = 0;
            if (!SFE_CUDA(cudaMalloc(&op.d_sched, need * 2))) return SFE_ERR_CUDA;
            op.d_sched_cap = need * 2;
        }
        if (!op.h_sched_done && !SFE_CUDA(cudaEventCreateWithFlags(&op.h_sched_done, cudaEventDisableTiming))) return SFE_ERR_CUDA;
        if (op.use_runs) {
            // run list -> pinned staging -> device, then one small kernel writes the n_out entries where the kernels read them
            const size_t nr = op.runs.size(), rb = nr * sizeof(SchedRun);
            if (op.h_runs_cap < rb) {
                if (op.h_runs) {
                    cudaEventSynchronize(op.h_sched_done);
                    cudaFreeHost(op.h_runs);
                }
                cudaFree(op.d_runs);
                op.h_runs = nullptr;
                op.d_runs = nullptr;
                op.h_runs_cap = op.d_runs_cap = 0;
                if (!SFE_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&op.h_runs), rb * 2, cudaHostAllocDefault)) ||
                    !SFE_CUDA(cudaMalloc(&op.d_runs, rb * 2)))
                    return SFE_ERR_CUDA;
                op.h_runs_cap = op.d_runs_cap = rb * 2;
            }
            if (!SFE_CUDA(cudaEventSynchronize(op.h_sched_done))) return SFE_ERR_CUDA;   // the previous call's copy has left the staging buffer
            memcpy(op.h_runs, op.runs.data(), rb);
            if (!SFE_CUDA(cudaMemcpyAsync(op.d_runs, op.h_runs, rb, cudaMemcpyHostToDevice, st)) ||
                !SFE_CUDA(cudaEventRecord(op.h_sched_done, st)))
                return SFE_ERR_CUDA;
            int rc = launch_expand_sched(op.d_runs, (int)nr, op.d_sched, st);
            if (rc != SFE_OK) return rc;
            if (op.lerp) {                       // which entries fall into which FFT block of the convolver
                const long nb = (op.n_in - 1) / op.hop + 1;
                if (op.d_kidx_cap < (size_t)nb + 1) {
                    cudaFree(op.d_kidx);
                    op.d_kidx = nullptr;
                    op.d_kidx_cap = 0;
                    if (!SFE_CUDA(cudaMalloc(&op.d_kidx, sizeof(int) * 2 * ((size_t)nb + 1)))) return SFE_ERR_CUDA;
                    op.d_kidx_cap = 2 * ((size_t)nb + 1);
                }
                rc = launch_block_index(op.d_sched, op.n_out, -1, op.hop, nb, op.d_kidx, st);
                if (rc != SFE_OK) return rc;
            }
            continue;
        }
        // straight from the vector the walk filled, once its storage is registered (plan_call keeps it from moving)
        void *hp = op.sched.e.data();
        if (op.reg_ptr != hp) {
            if (op.reg_ptr) cudaHostUnregister(op.reg_ptr);
            op.reg_ptr = nullptr;
            if (cudaHostRegister(hp, op.sched.e.capacity() * sizeof(SchedEntry), cudaHostRegisterDefault) == cudaSuccess) op.reg_ptr = hp;
            else cudaGetLastError();   // refused: the staging copy below still works
        }
        if (op.reg_ptr == hp) {
            if (!SFE_CUDA(cudaMemcpyAsync(op.d_sched, hp, need, cudaMemcpyHostToDevice, st)) ||
                !SFE_CUDA(cudaEventRecord(op.h_sched_done, st)))
                return SFE_ERR_CUDA;
            continue;
        }
        // through a pinned staging buffer: the copy is a real asynchronous DMA instead of a staged pageable one.  The
        // buffer may only be rewritten once its previous copy has completed.
        if (op.h_sched_cap < need) {
            if (op.h_sched) {
                cudaEventSynchronize(op.h_sched_done);
                cudaFreeHost(op.h_sched);
            }
            op.h_sched = nullptr;
            op.h_sched_cap = 0;
            if (!SFE_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&op.h_sched), need * 2, cudaHostAllocDefault))) return SFE_ERR_CUDA;
            op.h_sched_cap = need * 2;
        }
        if (!op.h_sched_done && !SFE_CUDA(cudaEventCreateWithFlags(&op.h_sched_done, cudaEventDisableTiming))) return SFE_ERR_CUDA;
        if (!SFE_CUDA(cudaEventSynchronize(op.h_sched_done))) return SFE_ERR_CUDA;
        memcpy(op.h_sched, op.sched.e.data(), need);
        if (!SFE_CUDA(cudaMemcpyAsync(op.d_sched, op.h_sched, need, cudaMemcpyHostToDevice, st)) ||
            !SFE_CUDA(cudaEventRecord(op.h_sched_done, st)))
            return SFE_ERR_CUDA;
    }
    return SFE_OK;
}

}  // namespace

static sfe_chain * sfe_chain_create_impl(const sfe_chain_desc *d)
{
    if (!d || d->n_streams < 1 || d->in_fmt < SFE_IN_I16 || d->in_fmt > SFE_IN_F32) {
        set_error("chain: bad descriptor (n_streams >= 1, in_fmt in {I16,U8,F32})");
        return nullptr;
    }
    const bool has_bc = d->bc_ntaps > 0, has_a = d->a.klass != SFE_KLASS_NONE, has_b = d->b.klass != SFE_KLASS_NONE;
    if (!has_bc && !has_a && !has_b) {
        set_error("chain: no stage present");
        return nullptr;
    }
    if (has_b && !has_a) {
        set_error("chain: stage b without stage a");
        return nullptr;
    }
    if (has_bc && (!d->bc_taps || d->bc_ntaps > d->bc_fft || d->bc_fft > (1 << 28))) {
        set_error("chain: blkconv needs taps and 1 <= n_taps <= fft_len <= 2^28");
        return nullptr;
    }
    for (const sfe_stage_desc *s : {&d->a, &d->b}) {
        if (s->klass == SFE_KLASS_NONE) continue;
        const bool rate_ok = s->klass == SFE_KLASS_DECIMATE ? s->rate >= 1.0f : s->rate >= 1.0f / (float)s->up;
        if (!s->taps || s->n_taps < 1 || s->up < 1 || s->blk < 1 || !rate_ok) {
            set_error("chain: bad polyphase stage (taps, upsample >= 1, blk >= 1, rate >= 1 for decimate, >= 1/upsample for resample)");
            return nullptr;
        }
    }
    int dev = 0;
    if (!chain_device_ready(&dev)) return nullptr;

    // ---- blkconv followed by a one-phase (upsample 1) stage with a FRACTIONAL rate ------------------------------------
    // With L = 1 the polyphase sum of decimate.cxx:132-140 is an ordinary causal FIR, s = a * y, and the stage's output
    // is the linear interpolation (1-mu) s[pos] + mu s[pos+1] of it (decimate.cxx:123).  blkconv is a causal FIR too, so
    // the FFT convolver applies g = h * a in one go -- at the cost it had for h alone -- and the stage that follows keeps
    // only the interpolation: a polyphase stage with the single tap 1.  The schedule (rate, blk, held-over outputs) is
    // untouched.  cfg 4: the 31-tap stage a was the most expensive kernel of the chain (1.27 of 2.7 ms).
    sfe_chain_desc fused_desc;
    std::vector<float> fused_taps;
    static const float unit_tap = 1.0f;
    bool lerp_fused = false;
    {
        long st = 0;
        if (has_bc && has_a && d->a.up == 1 && !(d->flags & (SFE_FLAG_EXACT | SFE_FLAG_NO_FUSE)) && !integral_step(d->a, &st) &&
            d->a.rate * (float)d->a.up == d->a.rate) {
            std::vector<double> g((size_t)d->bc_ntaps + d->a.n_taps - 1, 0.0);
            for (int i = 0; i < d->a.n_taps; i++)
                for (int j = 0; j < d->bc_ntaps; j++) g[(size_t)i + j] += (double)d->a.taps[i] * (double)d->bc_taps[j];
            while (g.size() > 1 && g.back() == 0.0) g.pop_back();
            fused_taps.assign(g.begin(), g.end());
            fused_desc = *d;
            fused_desc.bc_taps = fused_taps.data();
            fused_desc.bc_ntaps = (int)fused_taps.size();
            fused_desc.a.taps = &unit_tap;
            fused_desc.a.n_taps = 1;
            d = &fused_desc;
            lerp_fused = true;
        }
    }

    sfe_chain *c = new sfe_chain;
    c->dev = dev;
    c->S = d->n_streams;
    c->in_fmt = d->in_fmt;
    c->flags = d->flags;
    bool ok = true;
    int fmt = d->in_fmt;
    long step_a = 0;
    const bool a_decim = has_a && integral_step(d->a, &step_a) && step_a % d->a.up == 0 && !(d->flags & SFE_FLAG_EXACT);
    const bool fuse = has_bc && a_decim && !(d->flags & SFE_FLAG_NO_FUSE);

    // ---- plan K3: the whole chain as one polyphase bank over the raw samples -------------------------------
    // needs: int16 input, a last stage with an integral multi-phase step, and before it nothing but FIRs that
    // keep every D-th sample (blkconv and/or an integral decimator).  Everything else uses the staged plans.
    bool bank = false;
    if (!(d->flags & (SFE_FLAG_EXACT | SFE_FLAG_NO_BANK)) && (d->in_fmt == SFE_IN_I16 || d->in_fmt == SFE_IN_U8) && has_a && (!has_b || a_decim)) {
        const sfe_stage_desc &bs = has_b ? d->b : d->a;
        long step_b = 0;
        if (integral_step(bs, &step_b) && step_b % bs.up != 0) {
            Op op{};
            op.kind = OP_BANK;
            op.in_fmt = fmt;
            op.up = bs.up;
            op.blk = bs.blk;
            op.rate = bs.rate;
            op.r = padded_taps(bs);
            op.D = 1;
            op.g.assign(1, 1.0);
            if (has_b) {
                op.has_pre = true;
                op.pre_up = d->a.up;
                op.pre_blk = d->a.blk;
                op.pre_rate = d->a.rate;
                op.D = (int)(step_a / d->a.up);
                const std::vector<float> ta = padded_taps(d->a);
                op.g.clear();
                for (size_t q = 0; q * d->a.up < ta.size(); q++) op.g.push_back((double)ta[q * d->a.up]);
                while (op.g.size() > 1 && op.g.back() == 0.0) op.g.pop_back();
            }
            if (has_bc) {
                std::vector<double> f(op.g.size() + d->bc_ntaps - 1, 0.0);
                for (size_t i = 0; i < op.g.size(); i++)
                    for (int j = 0; j < d->bc_ntaps; j++) f[i + j] += op.g[i] * (double)d->bc_taps[j];
                op.g.swap(f);
            }
            const int J = ((int)op.r.size() + op.up - 1) / op.up;
            const int Tc = op.D * (J - 1) + (int)op.g.size();
            const int spc = 16 / fmt_bytes(fmt), box = 8 * spc;          // samples per 16-byte chunk / per 128-byte box row
            const int NC = bank_nc(Tc, spc);
            op.nc = NC;
            op.row_len = bank_row_len(Tc, spc);
            op.rows_per_phase = op.D % spc == 0 ? 1 : spc;
            op.n_rows = op.up * op.rows_per_phase;
            const long spacing = (long)op.D * (step_b / op.up);          // samples between consecutive outputs, at least
            // windows open and close on box boundaries: at most one may open per box, and a box is overlapped by
            // the windows whose first chunk lies in a range of NC + 7 chunks
            const long live = spacing >= 2 * box ? (NC + 6) / (spacing / spc) + 1 : 1 << 30;
            if (live <= bank_max_live() && bank_smem_bytes(op.n_rows, op.row_len, op.in_fmt) <= 227 * 1024) {
                op.n_live = (int)live;
                op.H = (spc * NC + 2 * op.D + spc + box - 1) / box * box;
                op.name = std::string("bank_fused(") + (has_bc ? "blkconv*" : "") + (has_b ? "decim*" : "") + "polyphase)/L" +
                          std::to_string(op.up) + "D" + std::to_string(op.D) + "T" + std::to_string(Tc);
                c->plan = op.name;
                c->ops.push_back(op);
                bank = true;
            }
        }
    }
    const bool staged = !bank;

    if (staged && has_bc && !fuse) {
        // The reference's block size fft_len+1-n_taps (blkconv.cxx:47) only shapes its own calls; the stream it emits
        // is the causal convolution whatever fft_len is, so the kernels run overlap-save with an internal power-of-two
        // transform (the caller's fft_len when it is one the radix-16 kernel is compiled for).
        Op op{};
        op.in_fmt = fmt;
        op.bc_ntaps = d->bc_ntaps;
        op.ov = d->bc_ntaps - 1;               // blkconv.cxx:48
        op.N = chain_blkconv_fft(d->bc_ntaps, d->bc_fft);
        if (op.N) {
            op.kind = OP_BLKCONV;
            op.hop = op.N - op.ov;
            op.H = op.ov;
            op.tw = chain_twiddles(dev, op.N);
            op.hp = chain_spectrum(dev, d->bc_taps, d->bc_ntaps, op.N, nullptr);
            ok = ok && op.tw && op.hp;
            op.name = "blkconv_fft" + std::to_string(op.N) + (lerp_fused ? "(blkconv*stage_a_fir)" : "");
            static const bool no_epi = getenv("SFE_NO_LERP_EPILOGUE") != nullptr;   // tuning hook: separate interpolation kernel
            if (lerp_fused && !no_epi) {
                // the interpolation that is left of stage a runs in the convolver's epilogue, from its shared-memory block:
                // the full-rate convolved stream is never written (blocks overlap by one more sample, for s[pos + 1], and
                // start at -1, for an output held over from the previous call)
                op.lerp = true;
                op.hop = op.N - op.ov - 1;
                op.H = op.ov + 1;
                op.up = 1;
                op.blk = d->a.blk;
                op.rate = d->a.rate;
                op.name += "+lerp_epilogue";
            }
        } else {
            // taps longer than any compiled transform: direct FIR = one-phase polyphase stage with step 1
            op.kind = OP_POLY;
            op.direct_fir = true;
            op.T = d->bc_ntaps;
            op.up = 1;
            op.rate = 1.f;
            op.H = op.T;
            ok = ok && upload(&op.d_taps, std::vector<float>(d->bc_taps, d->bc_taps + d->bc_ntaps));
            op.name = "blkconv_direct_fir/T" + std::to_string(op.T);
        }
        c->ops.push_back(op);
        c->plan += op.name;
        fmt = SFE_IN_F32;
    }
    for (int which = 0; which < 2 && ok && staged; which++) {
        const sfe_stage_desc &s = which == 0 ? d->a : d->b;
        if (s.klass == SFE_KLASS_NONE) continue;
        if (which == 0 && !c->ops.empty() && c->ops.back().lerp) continue;   // stage a lives in the convolver's epilogue
        Op op{};
        op.in_fmt = fmt;
        op.up = s.up;
        op.blk = s.blk;
        op.rate = s.rate;
        const std::vector<float> t = padded_taps(s);
        long step = 0;
        const bool decim = integral_step(s, &step) && step % s.up == 0 && !(d->flags & SFE_FLAG_EXACT);
        if (!c->plan.empty()) c->plan += "+";
        const size_t plan_mark = c->plan.size();
        if (decim) {
            // phase-0 taps g[q] = taps[q*up]; optionally convolved with the blkconv taps (fusion)
            std::vector<double> g;
            for (size_t q = 0; q * s.up < t.size(); q++) g.push_back((double)t[q * s.up]);
            while (g.size() > 1 && g.back() == 0.0) g.pop_back();
            if (which == 0 && fuse) {
                std::vector<double> f(g.size() + d->bc_ntaps - 1, 0.0);
                for (size_t i = 0; i < g.size(); i++)
                    for (int j = 0; j < d->bc_ntaps; j++) f[i + j] += g[i] * (double)d->bc_taps[j];
                g.swap(f);
                c->plan += "fused(blkconv*decim_fir)";
            } else {
                c->plan += "decim_fir";
            }
            // tensor cores: many streams sharing the filter, or a FEW long streams (each row is time-split over the tile rows;
            // short calls of fewer than 8 streams fall back to the CUDA-core FIR at run time, so that op carries both tap images)
            // (u8 rows -- one signed byte per component -- run on the same kernel from 8 streams on; fewer keep the CUDA-core FIR)
            const bool tensor = !(d->flags & SFE_FLAG_NO_TENSOR) && (fmt == SFE_IN_I16 || (fmt == SFE_IN_U8 && d->n_streams >= 8)) &&
                                step / s.up == 8 && (int)g.size() <= toeplitz_max_taps();
            const bool both = tensor && d->n_streams < 8;
            if (both) ok = ok && make_decim(op, g, (int)(step / s.up));
            if (tensor && make_toep(op, g, fmt)) {
                c->plan += both ? "[toeplitz_i8_tcgen05 time-split | decim_fir for short calls]" : "[toeplitz_i8_tcgen05]";
            } else if (!both) {
                ok = ok && make_decim(op, g, (int)(step / s.up));
            }
            c->plan += "/D" + std::to_string(op.D) + "T" + std::to_string(op.T);
        } else {
            op.kind = OP_POLY;
            op.T = (int)t.size();
            op.H = (op.T - 1) / s.up + 1;
            ok = ok && upload(&op.d_taps, t);
            if (op.T == 1 && s.up == 1 && fmt == SFE_IN_F32 && !(d->flags & SFE_FLAG_EXACT)) {
                op.lerp_tap = t[0];
                op.H = 2;
                c->plan += "lerp";
            } else {
                c->plan += "polyphase_generic/L" + std::to_string(s.up) + "T" + std::to_string(op.T);
            }
        }
        op.name = c->plan.substr(plan_mark);
        c->ops.push_back(op);
        fmt = SFE_IN_F32;
    }
    for (Op &op : c->ops) ok = ok && alloc_hist(op, c->S) && (op.kind != OP_BANK || make_bank_tables(op));
    static const int n_ws_env = getenv("SFE_HOST_WORKSPACES") ? atoi(getenv("SFE_HOST_WORKSPACES")) : 0;   // tuning hook
    c->n_ws = n_ws_env >= 2 && n_ws_env <= 4 ? n_ws_env : 2;
    for (int i = 0; i < 4 && ok; i++) ok = SFE_CUDA(cudaStreamCreateWithFlags(&c->ws[i].st, cudaStreamNonBlocking));
    // every table, history fill and spectrum above is on the device before the first process call may read it
    ok = ok && SFE_CUDA(cudaDeviceSynchronize());
    if (!ok) {
        sfe_chain_destroy(c);
        return nullptr;
    }
    return c;
}

extern "C" sfe_chain * sfe_chain_create(const sfe_chain_desc *d)
{
    SFE_GUARDED(nullptr, return sfe_chain_create_impl(d);)
}

extern "C" void sfe_chain_destroy(sfe_chain *c)
{
    if (!c) return;
    cudaSetDevice(c->dev);
    cudaDeviceSynchronize();
    for (Op &op : c->ops) {
        cudaFree(op.hist[0]);
        cudaFree(op.hist[1]);
        cudaFree(op.hp);
        cudaFree(op.d_taps);
        cudaFree(op.d_gp);
        cudaFree(op.d_bmat);
        cudaFree(op.d_sched);
        cudaFree(op.d_runs);
        cudaFree(op.d_kidx);
        if (op.h_runs) cudaFreeHost(op.h_runs);
        if (op.reg_ptr) cudaHostUnregister(op.reg_ptr);
        if (op.h_sched) cudaFreeHost(op.h_sched);
        if (op.h_sched_done) cudaEventDestroy(op.h_sched_done);
        for (auto &kv : op.d_tables) cudaFree(kv.second);
    }
    for (Workspace &w : c->ws) {
        cudaFree(w.d_align);
        for (void *m : w.mid) cudaFree(m);
        cudaFree(w.d_in);
        cudaFree(w.d_out);
        if (w.st) cudaStreamDestroy(w.st);
    }
    delete c;
}

extern "C" int sfe_chain_reset(sfe_chain *c)
{
    if (!c) return SFE_ERR_ARG;
    if (!SFE_CUDA(cudaSetDevice(c->dev)) || !SFE_CUDA(cudaDeviceSynchronize())) return SFE_ERR_CUDA;
    for (Op &op : c->ops) {
        op.ws = WalkState{};
        op.pre_ws = WalkState{};
        const size_t bytes = (size_t)c->S * op.H * fmt_bytes(op.in_fmt);
        for (int i = 0; i < 2; i++)
            if (!SFE_CUDA(cudaMemset(op.hist[i], op.in_fmt == SFE_IN_U8 ? 0x80 : 0, bytes))) return SFE_ERR_CUDA;
    }
    return SFE_OK;
}

extern "C" long sfe_chain_max_out(sfe_chain *c, long n_in)
{
    if (!c) return SFE_ERR_ARG;
    long n = n_in;
    for (const Op &op : c->ops) n = stage_max_out(op, n);
    return n;
}

extern "C" const char *sfe_chain_plan(sfe_chain *c) { return c ? c->plan.c_str() : ""; }
extern "C" int sfe_chain_n_ops(sfe_chain *c) { return c ? (int)c->ops.size() : SFE_ERR_ARG; }
extern "C" const char *sfe_chain_op_name(sfe_chain *c, int op)
{
    return (c && op >= 0 && op < (int)c->ops.size()) ? c->ops[op].name.c_str() : "";
}
extern "C" int sfe_chain_profile_enable(sfe_chain *c, int on)
{
    if (!c) return SFE_ERR_ARG;
    c->profile_every = on > 0 ? on : 0;
    c->profile_calls = 0;
    c->profile = false;
    return SFE_OK;
}
extern "C" int sfe_chain_profile_read(sfe_chain *c, int op, float *mean_ms, int *n_samples)
{
    if (!c || op < 0 || op >= (int)c->ops.size() || !mean_ms || !n_samples) return SFE_ERR_ARG;
    Op &o = c->ops[op];
    double sum = 0;
    int n = 0;
    for (auto &pr : o.ev) {
        float ms = 0;
        if (!SFE_CUDA(cudaEventSynchronize(pr.second)) || !SFE_CUDA(cudaEventElapsedTime(&ms, pr.first, pr.second))) return SFE_ERR_CUDA;
        sum += ms;
        n++;
        cudaEventDestroy(pr.first);
        cudaEventDestroy(pr.second);
    }
    o.ev.clear();
    *mean_ms = n ? (float)(sum / n) : 0.f;
    *n_samples = n;
    return SFE_OK;
}

static int sfe_chain_process_dev_impl(sfe_chain *c, const void *d_in, long n_in, long in_pitch, void *d_out,
                                     long out_pitch, long *n_out, void *stream)
{
    if (!c || !d_in || !d_out || n_in < 0 || in_pitch < n_in) {
        set_error("chain_process_dev: bad arguments");
        return SFE_ERR_ARG;
    }
    int rc = check_call(c, n_in, "chain_process_dev");
    if (rc != SFE_OK) return rc;
    if (!SFE_CUDA(cudaSetDevice(c->dev))) return SFE_ERR_CUDA;
    static const bool trace = getenv("SFE_TRACE_HOST") != nullptr;   // host time of the three parts of a call, on stderr
    const auto t0 = std::chrono::steady_clock::now();
    const WalkSnap snap = snapshot(c);
    {
        const NvtxRange range("sfe schedule walk");
        plan_call(c, n_in);
    }
    const auto t1 = std::chrono::steady_clock::now();
    const long no = c->ops.back().n_out;
    if (n_out) *n_out = no;
    if (out_pitch < no) {
        restore(c, snap);    // nothing was launched and no history moved: the stream is where it was
        set_error("chain_process_dev: out_pitch smaller than the output count " + std::to_string(no));
        return SFE_ERR_CAP;
    }
    cudaStream_t st = (cudaStream_t)stream;
    rc = upload_schedules(c, st);
    if (rc == SFE_OK) rc = run_ops(c, c->ws[0], 0, c->S, d_in, in_pitch, d_out, out_pitch, st);
    const auto t2 = std::chrono::steady_clock::now();
    if (rc != SFE_OK) {
        restore(c, snap);    // histories are swapped only below; kernels already queued wrote the spare buffers
        return rc;
    }
    if (trace) {
        auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
            return std::chrono::duration<double, std::micro>(b - a).count();
        };
        fprintf(stderr, "sfe host trace: schedule %.0f us, upload + launches %.0f us\n", us(t0, t1), us(t1, t2));
    }
    c->cur ^= 1;
    return SFE_OK;
}

extern "C" int sfe_chain_process_dev(sfe_chain *c, const void *d_in, long n_in, long in_pitch, void *d_out,
                                     long out_pitch, long *n_out, void *stream)
{
    SFE_GUARDED(SFE_ERR_HOST, return sfe_chain_process_dev_impl(c, d_in, n_in, in_pitch, d_out, out_pitch, n_out, stream);)
}

static int process_host_groups(sfe_chain *c, const void *h_in, long n_in, long in_pitch, void *h_out, long out_pitch, long no);

static int sfe_chain_process_host_impl(sfe_chain *c, const void *h_in, long n_in, long in_pitch, void *h_out,
                                      long out_pitch, long *n_out)
{
    if (!c || !h_in || !h_out || n_in < 0 || in_pitch < n_in) {
        set_error("chain_process_host: bad arguments");
        return SFE_ERR_ARG;
    }
    int rc = check_call(c, n_in, "chain_process_host");
    if (rc != SFE_OK) return rc;
    if (!SFE_CUDA(cudaSetDevice(c->dev))) return SFE_ERR_CUDA;
    const WalkSnap snap = snapshot(c);
    plan_call(c, n_in);
    const long no = c->ops.back().n_out;
    if (n_out) *n_out = no;
    if (out_pitch < no) {
        restore(c, snap);
        set_error("chain_process_host: out_pitch smaller than the output count " + std::to_string(no));
        return SFE_ERR_CAP;
    }
    rc = process_host_groups(c, h_in, n_in, in_pitch, h_out, out_pitch, no);
    if (rc != SFE_OK) {
        restore(c, snap);
        for (Workspace &w : c->ws) cudaStreamSynchronize(w.st);
        return rc;
    }
    c->cur ^= 1;
    return SFE_OK;
}

extern "C" int sfe_chain_process_host(sfe_chain *c, const void *h_in, long n_in, long in_pitch, void *h_out,
                                      long out_pitch, long *n_out)
{
    SFE_GUARDED(SFE_ERR_HOST, return sfe_chain_process_host_impl(c, h_in, n_in, in_pitch, h_out, out_pitch, n_out);)
}

static int process_host_groups(sfe_chain *c, const void *h_in, long n_in, long in_pitch, void *h_out, long out_pitch, long no)
{
    const int eb = fmt_bytes(c->in_fmt);
    // stream groups of ~64 MiB input: H2D(g+1) overlaps kernels(g) overlaps D2H(g-1) on n_ws CUDA streams with a staging
    // buffer each (two; three and four were measured within the run-to-run spread of the box: profiles/r02_host_workspaces.json)
    const size_t row_in = (size_t)n_in * eb;
    const long dpitch = (n_in + 7) & ~7L;   // device staging rows start 16-byte aligned in every format (TMA)
    static const long group_mb = getenv("SFE_HOST_GROUP_MB") ? atol(getenv("SFE_HOST_GROUP_MB")) : 64;   // tuning hook
    int gs = (int)(((size_t)(group_mb > 0 ? group_mb : 64) << 20) / (row_in ? row_in : 1));
    if (gs < 1) gs = 1;
    if (gs > c->S) gs = c->S;
    if (c->S >= 4 && gs > (c->S + 3) / 4) gs = (c->S + 3) / 4;  // at least 4 groups so that copies overlap
    if (c->S >= 2 * c->n_ws && gs > (c->S + 2 * c->n_ws - 1) / (2 * c->n_ws)) gs = (c->S + 2 * c->n_ws - 1) / (2 * c->n_ws);
    int rc = upload_schedules(c, c->ws[0].st);
    if (rc != SFE_OK) return rc;
    if (!SFE_CUDA(cudaStreamSynchronize(c->ws[0].st))) return SFE_ERR_CUDA;
    const long opitch = no > 0 ? no : 1;
    int g = 0;
    for (int s0 = 0; s0 < c->S; s0 += gs, g++) {
        Workspace &w = c->ws[g % c->n_ws];
        const int ns = c->S - s0 < gs ? c->S - s0 : gs;
        if (!ensure(&w.d_in, &w.in_bytes, (size_t)gs * dpitch * eb) ||
            !ensure(&w.d_out, &w.out_bytes, sizeof(float2) * (size_t)gs * opitch))
            return SFE_ERR_CUDA;
        const char *src = static_cast<const char *>(h_in) + (size_t)s0 * in_pitch * eb;
        if (!SFE_CUDA(cudaMemcpy2DAsync(w.d_in, (size_t)dpitch * eb, src, (size_t)in_pitch * eb, row_in, ns, cudaMemcpyHostToDevice, w.st)))
            return SFE_ERR_CUDA;
        rc = run_ops(c, w, s0, ns, w.d_in, dpitch, w.d_out, opitch, w.st);
        if (rc != SFE_OK) return rc;
        if (no > 0) {
            char *dst = static_cast<char *>(h_out) + sizeof(float2) * (size_t)s0 * out_pitch;
            if (!SFE_CUDA(cudaMemcpy2DAsync(dst, sizeof(float2) * out_pitch, w.d_out, sizeof(float2) * opitch,
                                            sizeof(float2) * no, ns, cudaMemcpyDeviceToHost, w.st)))
                return SFE_ERR_CUDA;
        }
    }
    for (Workspace &w : c->ws)
        if (!SFE_CUDA(cudaStreamSynchronize(w.st))) return SFE_ERR_CUDA;
    return SFE_OK;
}
