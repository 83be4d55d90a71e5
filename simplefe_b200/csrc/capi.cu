// capi.cu -- the C ABI of libsfegpu.so (include/sfe_gpu.h): object lifetimes, stream state,
// kernel selection.  Host-side C++; all arithmetic on samples happens in the kernels.
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "kernels.h"

namespace sfe {

NvtxRange::NvtxRange(const char *name) { nvtxRangePushA(name); }
NvtxRange::~NvtxRange() { nvtxRangePop(); }

// ---- error / bookkeeping -----------------------------------------------------------------------
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};

void set_error(const std::string &msg) { g_err = msg; }
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool cuda_ok(cudaError_t e, const char *what, const char *file, int line)
{
    if (e == cudaSuccess) return true;
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %s (%s) at %s:%d: %s", cudaGetErrorName(e), cudaGetErrorString(e), file,
             line, what);
    g_err = buf;
    fprintf(stderr, "libsfegpu: %s\n", buf);
    return false;
}

static thread_local int g_device = 0;

static bool device_ready()
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        set_error(std::string("no CUDA device available (libsfegpu has no CPU fallback): ") +
                  (e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e)));
        return false;
    }
    if (!SFE_CUDA(cudaSetDevice(g_device))) return false;
    return true;
}

// ---- twiddle tables, one per (device, N) ---------------------------------------------------------
static std::mutex g_tw_mutex;
static std::map<std::pair<int, int>, float2 *> g_tw;

static const float2 *twiddles(int dev, int N)
{
    std::lock_guard<std::mutex> lk(g_tw_mutex);
    auto it = g_tw.find({dev, N});
    if (it != g_tw.end()) return it->second;
    std::vector<float2> h;
    if (!blkconv_twiddle_table(N, h)) {
        set_error("blkconv: fft_len must be a power of two in [32, 8192]");
        return nullptr;
    }
    float2 *d = nullptr;
    if (!SFE_CUDA(cudaMalloc(&d, sizeof(float2) * h.size()))) return nullptr;
    // pageable source: the copy may return before the DMA has landed, and the kernels that read the table run on
    // non-blocking streams -- wait for the device once, here (one-off per (device, N))
    if (!SFE_CUDA(cudaMemcpy(d, h.data(), sizeof(float2) * h.size(), cudaMemcpyHostToDevice)) || !SFE_CUDA(cudaDeviceSynchronize()))
        return nullptr;
    g_tw[{dev, N}] = d;
    return d;
}

static bool valid_fft(int n) { return n >= 32 && n <= 8192 && (n & (n - 1)) == 0; }

// The FFT size the kernels run for a blkconv(taps, n_taps, fft_len).  The reference hands ANY fft_len to FFTW
// (blkconv.cxx:60,72-73) and its result -- the causal convolution, emitted in blocks of fft_len+1-n_taps -- does not
// depend on it, so the size of the internal overlap-save transform is free: a power of two the radix-16
// kernel is compiled for.  0 = taps too long for any compiled size: the convolution then runs as a direct FIR.
int blkconv_internal_fft(int n_taps, int fft_len)
{
    if (valid_fft(fft_len) && n_taps <= fft_len) return fft_len;
    for (int n = 32; n <= 8192; n *= 2)
        if (n >= 4 * n_taps) return n;          // at least 3/4 of every transform is new samples
    return n_taps <= 4097 ? 8192 : 0;           // at least half
}

// tap spectrum in kernel order, computed by the forward half of the convolution kernel itself
static float2 *make_spectrum(int dev, const float *taps, int n_taps, int N, cudaStream_t st)
{
    const float2 *tw = twiddles(dev, N);
    if (!tw) return nullptr;
    float *d_t = nullptr;
    float2 *d_h = nullptr;
    if (!SFE_CUDA(cudaMalloc(&d_t, sizeof(float) * n_taps))) return nullptr;
    if (!SFE_CUDA(cudaMalloc(&d_h, sizeof(float2) * N))) return nullptr;
    bool ok = SFE_CUDA(cudaMemcpyAsync(d_t, taps, sizeof(float) * n_taps, cudaMemcpyHostToDevice, st));
    BlkconvParams p{};
    p.in = StreamView{d_t, d_t, n_taps, 0, n_taps, SFE_IN_F32};
    p.in_real = 1;
    p.out = d_h;
    p.out_pitch = N;
    p.n_streams = 1;
    p.n_blocks = 1;
    p.hop = N;
    p.ov = 0;
    p.tw = tw;
    p.hp = nullptr;
    p.mode = BLKCONV_MODE_SPECTRUM;
    p.scale = 1.0f / (float)N;  // blkconv.cxx:50
    ok = ok && launch_blkconv(N, p, st) == SFE_OK && SFE_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_t);
    if (!ok) {
        cudaFree(d_h);
        return nullptr;
    }
    return d_h;
}

// entry points for chain.cu
int chain_blkconv_fft(int n_taps, int fft_len) { return blkconv_internal_fft(n_taps, fft_len); }
const float2 *chain_twiddles(int dev, int N) { return twiddles(dev, N); }
float2 *chain_spectrum(int dev, const float *taps, int n_taps, int N, cudaStream_t st)
{
    return make_spectrum(dev, taps, n_taps, N, st);
}
bool chain_device_ready(int *dev)
{
    if (!device_ready()) return false;
    *dev = g_device;
    return true;
}

}  // namespace sfe

using namespace sfe;

// =================================================================================================
// library
// =================================================================================================
extern "C" const char *sfe_version(void) { return "simplefe_b200 0.1 (sm_100a)"; }
extern "C" const char *sfe_last_error(void) { return g_err.c_str(); }
extern "C" int64_t sfe_kernel_launches(void) { return g_launches.load(); }

extern "C" int sfe_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        set_error(std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
        return SFE_ERR_NODEV;
    }
    return n;
}

extern "C" int sfe_set_device(int device)
{
    if (!SFE_CUDA(cudaSetDevice(device))) return SFE_ERR_CUDA;
    g_device = device;
    return SFE_OK;
}

extern "C" void *sfe_pinned_alloc(size_t bytes)
{
    void *p = nullptr;
    if (!device_ready()) return nullptr;
    if (!SFE_CUDA(cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable))) return nullptr;
    return p;
}
extern "C" void sfe_pinned_free(void *p) { if (p) cudaFreeHost(p); }
extern "C" void *sfe_dev_alloc(size_t bytes)
{
    void *p = nullptr;
    if (!device_ready()) return nullptr;
    if (!SFE_CUDA(cudaMalloc(&p, bytes ? bytes : 1))) return nullptr;
    return p;
}
extern "C" void sfe_dev_free(void *p) { if (p) cudaFree(p); }
extern "C" int sfe_memcpy_h2d(void *d, const void *h, size_t n, void *st)
{
    return SFE_CUDA(cudaMemcpyAsync(d, h, n, cudaMemcpyHostToDevice, (cudaStream_t)st)) ? SFE_OK : SFE_ERR_CUDA;
}
extern "C" int sfe_memcpy_d2h(void *h, const void *d, size_t n, void *st)
{
    return SFE_CUDA(cudaMemcpyAsync(h, d, n, cudaMemcpyDeviceToHost, (cudaStream_t)st)) ? SFE_OK : SFE_ERR_CUDA;
}
extern "C" int sfe_stream_sync(void *st)
{
    return SFE_CUDA(cudaStreamSynchronize((cudaStream_t)st)) ? SFE_OK : SFE_ERR_CUDA;
}

// =================================================================================================
// (1a) blkconv object -- libdsp/blkconv.h:35-48
// =================================================================================================
struct sfe_blkconv {
    int dev, N, n_taps, blk, ov;    // N: internal transform size (0 = direct FIR), blk: fft_len + 1 - n_taps
    float *d_taps = nullptr;        // direct FIR only
    float *h_buf = nullptr;       // pinned, N+2 floats (blkconv.cxx:44)
    float *d_in = nullptr, *d_out = nullptr;
    float *d_hist[2] = {nullptr, nullptr};
    int cur = 0;
    float2 *d_hp = nullptr;
    const float2 *d_tw = nullptr;
    cudaStream_t st = nullptr;
};

static sfe_blkconv * sfe_blkconv_create_impl(const float *taps, int n_taps, int fft_len)
{
    if (!taps || n_taps < 1 || n_taps > fft_len || fft_len > (1 << 28)) {
        set_error("blkconv: need taps and 1 <= n_taps <= fft_len <= 2^28");
        return nullptr;
    }
    if (!device_ready()) return nullptr;
    sfe_blkconv *b = new sfe_blkconv;
    b->dev = g_device;
    b->N = blkconv_internal_fft(n_taps, fft_len);
    b->n_taps = n_taps;
    b->blk = fft_len + 1 - n_taps;   // blkconv.cxx:47
    b->ov = n_taps - 1;              // blkconv.cxx:48
    bool ok = SFE_CUDA(cudaStreamCreateWithFlags(&b->st, cudaStreamNonBlocking)) &&
              SFE_CUDA(cudaHostAlloc((void **)&b->h_buf, sizeof(float) * (fft_len + 2), cudaHostAllocDefault)) &&
              SFE_CUDA(cudaMalloc(&b->d_in, sizeof(float) * b->blk)) &&
              SFE_CUDA(cudaMalloc(&b->d_out, sizeof(float) * b->blk)) &&
              SFE_CUDA(cudaMalloc(&b->d_hist[0], sizeof(float) * (b->ov + 1))) &&
              SFE_CUDA(cudaMalloc(&b->d_hist[1], sizeof(float) * (b->ov + 1))) &&
              SFE_CUDA(cudaMemsetAsync(b->d_hist[0], 0, sizeof(float) * (b->ov + 1), b->st)) &&  // blkconv.cxx:53-55
              SFE_CUDA(cudaMemsetAsync(b->d_hist[1], 0, sizeof(float) * (b->ov + 1), b->st));
    if (ok) {
        memset(b->h_buf, 0, sizeof(float) * (fft_len + 2));
        if (b->N) {
            b->d_tw = twiddles(b->dev, b->N);
            b->d_hp = make_spectrum(b->dev, taps, n_taps, b->N, b->st);
            ok = b->d_tw && b->d_hp;
        } else {
            ok = SFE_CUDA(cudaMalloc(&b->d_taps, sizeof(float) * n_taps)) &&
                 SFE_CUDA(cudaMemcpyAsync(b->d_taps, taps, sizeof(float) * n_taps, cudaMemcpyHostToDevice, b->st)) &&
                 SFE_CUDA(cudaStreamSynchronize(b->st));
        }
    }
    if (!ok) {
        sfe_blkconv_destroy(b);
        return nullptr;
    }
    return b;
}

extern "C" sfe_blkconv * sfe_blkconv_create(const float *taps, int n_taps, int fft_len)
{
    SFE_GUARDED(nullptr, return sfe_blkconv_create_impl(taps, n_taps, fft_len);)
}

extern "C" int sfe_blkconv_blksize(sfe_blkconv *b) { return b ? b->blk : 0; }
extern "C" float *sfe_blkconv_buf(sfe_blkconv *b) { return b ? b->h_buf : nullptr; }

extern "C" int sfe_blkconv_process(sfe_blkconv *b)
{
    if (!b) return SFE_ERR_ARG;
    if (!SFE_CUDA(cudaSetDevice(b->dev))) return SFE_ERR_CUDA;
    if (!SFE_CUDA(cudaMemcpyAsync(b->d_in, b->h_buf, sizeof(float) * b->blk, cudaMemcpyHostToDevice, b->st)))
        return SFE_ERR_CUDA;
    int rc;
    if (b->N) {
        // overlap-save with the internal transform: ceil(blk / hop) FFT blocks cover the call's blk samples
        BlkconvParams p{};
        p.in = StreamView{b->d_in, b->d_hist[b->cur], b->blk, b->ov, b->blk, SFE_IN_F32};
        p.in_real = 1;
        p.out = b->d_out;
        p.out_pitch = b->blk;
        p.n_streams = 1;
        p.hop = b->N - b->ov;
        p.n_blocks = (b->blk + p.hop - 1) / p.hop;
        p.ov = b->ov;
        p.tw = b->d_tw;
        p.hp = b->d_hp;
        p.mode = BLKCONV_MODE_CONV;
        rc = launch_blkconv(b->N, p, b->st);
    } else {
        // taps longer than any compiled transform: y[n] = sum_j h[j] x[n-j] as a one-phase polyphase stage with step 1
        PolyParams p{};
        p.in = StreamView{b->d_in, b->d_hist[b->cur], b->blk, b->ov, b->blk, SFE_IN_F32};
        p.in_real = 1;
        p.out = b->d_out;
        p.out_pitch = b->blk;
        p.n_streams = 1;
        p.n_out = b->blk;
        p.taps = b->d_taps;
        p.n_taps = b->n_taps;
        p.up = 1;
        p.uniform = 1;
        p.pos0 = 0;
        p.step = 1;
        p.rate = 1.f;
        rc = launch_poly_generic(p, b->st);
    }
    if (rc != SFE_OK) return rc;
    rc = launch_save_tail(b->d_hist[b->cur ^ 1], b->d_hist[b->cur], b->d_in, b->blk, b->blk, b->ov, 4, 1, b->st);
    if (rc != SFE_OK) return rc;
    b->cur ^= 1;
    if (!SFE_CUDA(cudaMemcpyAsync(b->h_buf, b->d_out, sizeof(float) * b->blk, cudaMemcpyDeviceToHost, b->st)))
        return SFE_ERR_CUDA;
    return SFE_CUDA(cudaStreamSynchronize(b->st)) ? SFE_OK : SFE_ERR_CUDA;
}

extern "C" void sfe_blkconv_destroy(sfe_blkconv *b)
{
    if (!b) return;
    cudaSetDevice(b->dev);
    if (b->st) cudaStreamSynchronize(b->st);
    cudaFreeHost(b->h_buf);
    cudaFree(b->d_in);
    cudaFree(b->d_out);
    cudaFree(b->d_hist[0]);
    cudaFree(b->d_hist[1]);
    cudaFree(b->d_hp);
    cudaFree(b->d_taps);
    if (b->st) cudaStreamDestroy(b->st);
    delete b;
}

// =================================================================================================
// (1b) decimate / resample objects -- libdsp/decimate.h:33-46, resample.h:33-47
// =================================================================================================
struct sfe_polyphase {
    int dev, klass, up, blksize, T, H;
    int exact = 0;
    WalkState ws;
    float *d_taps = nullptr, *d_in = nullptr, *d_out = nullptr;
    float *d_hist[2] = {nullptr, nullptr};
    int cur = 0;
    int out_cap = 0;
    SchedEntry *d_sched = nullptr;
    int sched_cap = 0;
    EntryVec sched;
    cudaStream_t st = nullptr;
};

static sfe_polyphase *poly_create(int klass, const float *taps, int n_taps, int up, int blksize)
{
    if (!taps || n_taps < 1 || up < 1 || blksize < 1) {
        set_error("decimate/resample: need taps, n_taps >= 1, upsample >= 1, blksize >= 1");
        return nullptr;
    }
    if (!device_ready()) return nullptr;
    sfe_polyphase *h = new sfe_polyphase;
    h->dev = g_device;
    h->klass = klass;
    h->up = up;
    h->blksize = blksize;
    // decimate pads an even tap count with one zero tap (decimate.cxx:42-51); resample zero-pads
    // to a multiple of `upsample` (resample.cxx:43,58-62).  Zero taps do not change results.
    h->T = (klass == SFE_KLASS_DECIMATE && n_taps % 2 == 0) ? n_taps + 1 : n_taps;
    h->H = (h->T - 1) / up + 1;
    std::vector<float> t((size_t)h->T, 0.0f);
    memcpy(t.data(), taps, sizeof(float) * n_taps);
    bool ok = SFE_CUDA(cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking)) &&
              SFE_CUDA(cudaMalloc(&h->d_taps, sizeof(float) * h->T)) &&
              SFE_CUDA(cudaMemcpyAsync(h->d_taps, t.data(), sizeof(float) * h->T, cudaMemcpyHostToDevice, h->st)) &&
              SFE_CUDA(cudaMalloc(&h->d_in, sizeof(float) * blksize)) &&
              SFE_CUDA(cudaMalloc(&h->d_hist[0], sizeof(float) * h->H)) &&
              SFE_CUDA(cudaMalloc(&h->d_hist[1], sizeof(float) * h->H)) &&
              SFE_CUDA(cudaMemsetAsync(h->d_hist[0], 0, sizeof(float) * h->H, h->st)) &&  // decimate.cxx:54-57
              SFE_CUDA(cudaMemsetAsync(h->d_hist[1], 0, sizeof(float) * h->H, h->st)) &&
              SFE_CUDA(cudaStreamSynchronize(h->st));
    if (!ok) {
        sfe_polyphase_destroy(h);
        return nullptr;
    }
    return h;
}

extern "C" sfe_polyphase *sfe_decimate_create(const float *taps, int n_taps, int upsample, int blksize)
{
    return poly_create(SFE_KLASS_DECIMATE, taps, n_taps, upsample, blksize);
}
extern "C" sfe_polyphase *sfe_resample_create(const float *taps, int n_taps, int upsample, int blksize)
{
    return poly_create(SFE_KLASS_RESAMPLE, taps, n_taps, upsample, blksize);
}
extern "C" int sfe_polyphase_set_exact(sfe_polyphase *h, int exact)
{
    if (!h) return SFE_ERR_ARG;
    h->exact = exact ? 1 : 0;
    return SFE_OK;
}

static int sfe_polyphase_process_impl(sfe_polyphase *h, const float *in, int n_in, float *out, int out_len, float rate)
{
    if (!h || !in || !out) return SFE_ERR_ARG;
    // the reference's argument checks, messages and return value
    if (h->klass == SFE_KLASS_DECIMATE) {
        if (rate < 1.0) { printf("rate should be larger than 1.0\n"); return 0; }                   // decimate.cxx:75-78
        if (n_in > h->blksize) { printf("number of samples should be less than blksize\n"); return 0; } // :79-82
    } else if (n_in > h->blksize || rate < 1.0 / h->up) {                                           // resample.cxx:91-94
        printf("input parameter is wrong, rate <= 1/upsample, n_in <= blksize\n");
        return 0;
    }
    if (out_len < floorf(n_in * 1.0f / rate)) { printf("output buffer is not large enough"); return 0; } // decimate.cxx:84-87
    if (n_in <= 0) return 0;
    if ((double)n_in * h->up + (double)rate * h->up + h->up >= 2147483648.0) {   // 32-bit grid positions (SchedEntry.pos)
        set_error("decimate/resample: n_in * upsample reaches 2^31; feed the stream in smaller calls");
        return SFE_ERR_ARG;
    }
    if (!SFE_CUDA(cudaSetDevice(h->dev))) return SFE_ERR_CUDA;

    h->sched.clear();
    const int n_out = walk_call(h->ws, h->up, n_in, out_len, rate, 0, h->sched);
    if (n_out > h->sched_cap) {
        cudaFree(h->d_sched);
        cudaFree(h->d_out);
        h->d_sched = nullptr;
        h->d_out = nullptr;
        h->sched_cap = n_out + 64;
        if (!SFE_CUDA(cudaMalloc(&h->d_sched, sizeof(SchedEntry) * h->sched_cap)) ||
            !SFE_CUDA(cudaMalloc(&h->d_out, sizeof(float) * h->sched_cap)))
            return SFE_ERR_CUDA;
    }
    if (!SFE_CUDA(cudaMemcpyAsync(h->d_in, in, sizeof(float) * n_in, cudaMemcpyHostToDevice, h->st))) return SFE_ERR_CUDA;
    if (n_out > 0) {
        if (!SFE_CUDA(cudaMemcpyAsync(h->d_sched, h->sched.data(), sizeof(SchedEntry) * n_out, cudaMemcpyHostToDevice, h->st)))
            return SFE_ERR_CUDA;
        PolyParams p{};
        p.in = StreamView{h->d_in, h->d_hist[h->cur], n_in, h->H, n_in, SFE_IN_F32};
        p.in_real = 1;
        p.out = h->d_out;
        p.out_pitch = n_out;
        p.n_streams = 1;
        p.n_out = n_out;
        p.taps = h->d_taps;
        p.n_taps = h->T;
        p.up = h->up;
        p.uniform = 0;
        p.sched = h->d_sched;
        p.exact = h->exact;
        p.rate = rate;
        int rc = launch_poly_generic(p, h->st);
        if (rc != SFE_OK) return rc;
        if (!SFE_CUDA(cudaMemcpyAsync(out, h->d_out, sizeof(float) * n_out, cudaMemcpyDeviceToHost, h->st))) return SFE_ERR_CUDA;
    }
    int rc = launch_save_tail(h->d_hist[h->cur ^ 1], h->d_hist[h->cur], h->d_in, n_in, n_in, h->H, 4, 1, h->st);
    if (rc != SFE_OK) return rc;
    h->cur ^= 1;
    if (!SFE_CUDA(cudaStreamSynchronize(h->st))) return SFE_ERR_CUDA;
    return n_out;
}

extern "C" int sfe_polyphase_process(sfe_polyphase *h, const float *in, int n_in, float *out, int out_len, float rate)
{
    SFE_GUARDED(SFE_ERR_HOST, return sfe_polyphase_process_impl(h, in, n_in, out, out_len, rate);)
}

extern "C" void sfe_polyphase_destroy(sfe_polyphase *h)
{
    if (!h) return;
    cudaSetDevice(h->dev);
    if (h->st) cudaStreamSynchronize(h->st);
    cudaFree(h->d_taps);
    cudaFree(h->d_in);
    cudaFree(h->d_out);
    cudaFree(h->d_hist[0]);
    cudaFree(h->d_hist[1]);
    cudaFree(h->d_sched);
    if (h->st) cudaStreamDestroy(h->st);
    delete h;
}

extern "C" int sfe_pack_dac10_dev(const float *d_in, long n, uint8_t *d_out, void *stream)
{
    if (!d_in || !d_out || n < 0 || (n & 3)) {
        set_error("pack_dac10: n must be a non-negative multiple of 4");
        return SFE_ERR_ARG;
    }
    return launch_pack_dac10(d_in, n, d_out, (cudaStream_t)stream);
}
