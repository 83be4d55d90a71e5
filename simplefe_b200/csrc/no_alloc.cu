// no_alloc.cu -- sfe_polyphase_process_noalloc: one decimate/resample process() call that allocates nothing
// (include/sfe_gpu.h section 7).  No counterpart in the reference; added for
// callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// It is the one-thread-per-output flavour of K2g (polyphase_kernel.cu: poly_direct_kernel, real rows) with the piece
// of scratch memory that kernel needs -- the (pos, mu) schedule table of the call, which sfe_polyphase_process keeps in
// a cudaMalloc'ed, grown-on-demand buffer (capi.cu) -- in a file-scope __device__ array of SFE_NOALLOC_MAX_OUT entries
// (2 MB, held by every GPU that loads the library until the process exits).  Taps, history, input and output are the
// caller's device buffers; the three scalars a reference object carries between calls (decimate.h:55-59) travel in a
// caller-owned sfe_walk_state.  Nothing here calls cudaMalloc, cudaFree or a library routine that does.
//
// The array is shared by every call on a device and keeps the last call's entries.  Nothing in it accumulates: a call
// overwrites entries [0, n_out) on its stream before its kernel reads exactly those, so there is nothing to clear --
// but two calls that overlap on the GPU (different streams, or different host threads) race on it.
#include <mutex>

#include "kernels.h"

using namespace sfe;

namespace {

constexpr int NA_MAX_OUT = SFE_NOALLOC_MAX_OUT;
constexpr int NA_NT = 128;
constexpr int NA_MAX_DEV = 64;

static __device__ __align__(256) SchedEntry g_na_sched[NA_MAX_OUT];

// In host code `g_na_sched` is only a handle for the runtime's symbol table, not a device address: the address is
// looked up once per device, with that device current, and the kernel gets the pointer.
std::mutex g_na_mutex;
SchedEntry *g_na_ptr[NA_MAX_DEV];

SchedEntry *scratch_of_current_device()
{
    int dev = -1;
    if (!SFE_CUDA(cudaGetDevice(&dev))) return nullptr;
    if (dev < 0 || dev >= NA_MAX_DEV) {
        set_error("polyphase_noalloc: device ordinal out of range");
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(g_na_mutex);
    if (!g_na_ptr[dev]) {
        void *ptr = nullptr;
        if (!SFE_CUDA(cudaGetSymbolAddress(&ptr, g_na_sched))) return nullptr;
        g_na_ptr[dev] = static_cast<SchedEntry *>(ptr);
    }
    return g_na_ptr[dev];
}

struct NaParams {
    const float *in, *hist, *taps;
    float *out;
    const SchedEntry *sched;      // the scratch table: device address from cudaGetSymbolAddress
    int n_in, n_hist, n_taps, up, n_out;
};

// s(pos) = sum_j taps[(pos mod L) + j L] * x[floor(pos / L) - j]      (decimate.cxx:132-140); x[n < 0] comes from the
// caller's history (the last n_hist samples before this call), zeros before that (decimate.cxx:54-57)
template <bool EXACT> __device__ __forceinline__ float na_point(const NaParams &p, int pos)
{
    int n = pos / p.up;
    int ph = pos - n * p.up;
    if (ph < 0) { ph += p.up; n -= 1; }            // pos = -1: an output held over the call boundary
    float acc = 0.f;
    for (int m = ph; m < p.n_taps; m += p.up, n--) {
        float x = 0.f;
        if (n < p.n_in) {
            if (n >= 0) x = __ldg(p.in + n);
            else if (p.n_hist + n >= 0) x = __ldg(p.hist + p.n_hist + n);
        }
        const float t = __ldg(p.taps + m);
        acc = EXACT ? __fadd_rn(acc, __fmul_rn(t, x)) : fmaf(t, x, acc);
    }
    return acc;
}

template <bool EXACT> __global__ void __launch_bounds__(NA_NT) poly_noalloc_kernel(const NaParams p)
{
    const int k = blockIdx.x * NA_NT + threadIdx.x;
    if (k >= p.n_out) return;
    const SchedEntry e = p.sched[k];
    const float s0 = na_point<EXACT>(p, e.pos);
    float y;
    if (EXACT) {                                   // s0 * (1.0f - mu) + mu * s1, four IEEE operations (decimate.cxx:123)
        const float s1 = na_point<EXACT>(p, e.pos + 1);
        y = __fadd_rn(__fmul_rn(s0, __fsub_rn(1.0f, e.mu)), __fmul_rn(e.mu, s1));
    } else if (e.mu != 0.f) {
        y = fmaf(e.mu, na_point<EXACT>(p, e.pos + 1), s0 * (1.0f - e.mu));
    } else {
        y = s0;
    }
    p.out[k] = y;
}

int process_noalloc(int klass, const float *d_taps, int n_taps, int up, sfe_walk_state *state, const float *d_hist, int n_hist,
                    const float *d_in, int n_in, float *d_out, int out_len, float rate, int exact, cudaStream_t st)
{
    if ((klass != SFE_KLASS_DECIMATE && klass != SFE_KLASS_RESAMPLE) || !d_taps || n_taps < 1 || up < 1 || !state || !d_in || !d_out ||
        n_hist < 0 || (n_hist > 0 && !d_hist)) {
        set_error("polyphase_noalloc: bad arguments");
        return SFE_ERR_ARG;
    }
    // the reference's argument checks, messages and return value (as sfe_polyphase_process, minus the blksize it has not)
    if (klass == SFE_KLASS_DECIMATE) {
        if (rate < 1.0) { printf("rate should be larger than 1.0\n"); return 0; }                      // decimate.cxx:75-78
    } else if (rate < 1.0 / up) {                                                                       // resample.cxx:91-94
        printf("input parameter is wrong, rate <= 1/upsample, n_in <= blksize\n");
        return 0;
    }
    if (out_len < floorf(n_in * 1.0f / rate)) { printf("output buffer is not large enough"); return 0; }   // decimate.cxx:84-87
    if (n_in <= 0) return 0;
    if ((double)n_in * up + (double)rate * up + up >= 2147483648.0) {
        set_error("polyphase_noalloc: n_in * upsample reaches 2^31; feed the stream in smaller calls");
        return SFE_ERR_ARG;
    }
    if ((double)n_in / rate + 2.0 > (double)NA_MAX_OUT) {                                               // before any work
        set_error("polyphase_noalloc: about " + std::to_string((long)((double)n_in / rate)) + " outputs in one call, the static scratch holds " +
                  std::to_string(NA_MAX_OUT) + " (SFE_NOALLOC_MAX_OUT): feed the stream in smaller calls");
        return SFE_ERR_CAP;
    }
    SchedEntry *d_sched = scratch_of_current_device();
    if (!d_sched) return SFE_ERR_CUDA;

    // host side: the walk of this call into a per-thread list whose room is reserved once (host memory only)
    static thread_local EntryVec sched;
    if (sched.capacity() < (size_t)NA_MAX_OUT + 8) sched.reserve((size_t)NA_MAX_OUT + 8);
    sched.clear();
    WalkState ws;
    ws.pos = state->pos;
    ws.mu = state->mu;
    ws.held = state->held != 0;
    const int n_out = walk_call(ws, up, n_in, out_len, rate, 0, sched);
    if (n_out > NA_MAX_OUT) {                      // the caller's state is untouched
        set_error("polyphase_noalloc: " + std::to_string(n_out) + " outputs in one call, the static scratch holds " + std::to_string(NA_MAX_OUT));
        return SFE_ERR_CAP;
    }
    if (n_out > 0) {
        // pageable source: the runtime has copied it to its own staging memory when this returns, `sched` may be reused
        if (!SFE_CUDA(cudaMemcpyAsync(d_sched, sched.data(), sizeof(SchedEntry) * (size_t)n_out, cudaMemcpyHostToDevice, st))) return SFE_ERR_CUDA;
        // decimate pads an even tap count with one zero tap, resample pads to a multiple of upsample (decimate.cxx:42-51,
        // resample.cxx:58-62): zero taps change no result, so the caller's n_taps are used as they are
        NaParams p{d_in, n_hist > 0 ? d_hist : d_in, d_taps, d_out, d_sched, n_in, n_hist, n_taps, up, n_out};
        const unsigned grid = (unsigned)((n_out + NA_NT - 1) / NA_NT);
        if (exact) poly_noalloc_kernel<true><<<grid, NA_NT, 0, st>>>(p);
        else poly_noalloc_kernel<false><<<grid, NA_NT, 0, st>>>(p);
        count_launch();
        if (!SFE_CUDA(cudaGetLastError())) return SFE_ERR_CUDA;
    }
    state->pos = ws.pos;
    state->mu = ws.mu;
    state->held = ws.held ? 1 : 0;
    return n_out;
}

}  // namespace

extern "C" int sfe_polyphase_process_noalloc(int klass, const float *d_taps, int n_taps, int upsample, sfe_walk_state *state,
                                             const float *d_hist, int n_hist, const float *d_in, int n_in, float *d_out, int out_len,
                                             float rate, int exact, void *stream)
{
    SFE_GUARDED(SFE_ERR_HOST, return process_noalloc(klass, d_taps, n_taps, upsample, state, d_hist, n_hist, d_in, n_in, d_out, out_len,
                                                     rate, exact, (cudaStream_t)stream);)
}
