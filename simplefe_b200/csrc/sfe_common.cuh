// sfe_common.cuh -- shared device/host helpers for libsfegpu (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <exception>
#include <string>

#include "../../include/sfe_gpu.h"

namespace sfe {

// ---- error plumbing (no CPU fallback: errors surface, they are never swallowed) -------------
void set_error(const std::string &msg);
bool cuda_ok(cudaError_t e, const char *what, const char *file, int line);
void count_launch(int n = 1);
#define SFE_CUDA(call) ::sfe::cuda_ok((call), #call, __FILE__, __LINE__)
// extern "C" entry points that allocate on the host run their body under this guard: no exception crosses the C ABI
#define SFE_GUARDED(ret_on_throw, ...)                                                        \
    try {                                                                                     \
        __VA_ARGS__                                                                           \
    } catch (const std::exception &e_) {                                                      \
        ::sfe::set_error(std::string("host-side failure: ") + e_.what());                     \
        return ret_on_throw;                                                                  \
    } catch (...) {                                                                           \
        ::sfe::set_error("host-side failure: unknown exception");                             \
        return ret_on_throw;                                                                  \
    }

// ---- NVTX ranges around the launches of one op (header-only NVTX 3: a no-op unless a profiler is attached) -----------
struct NvtxRange {
    explicit NvtxRange(const char *name);
    ~NvtxRange();
};

// ---- complex helpers -------------------------------------------------------------------------
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cmul_conj(float2 a, float2 b)  // a * conj(b)
{
    return make_float2(fmaf(a.x, b.x, a.y * b.y), fmaf(a.y, b.x, -a.x * b.y));
}

// ---- sample formats ---------------------------------------------------------------------------
// int16: f = (float)s * (1.0f/32767.0f): one IEEE multiply, never contracted into an FMA, so the
// conversion is bit-exact against the CPU oracle.
// u8:    f = (b-128) * (1.0f/127.0f)   gr-simplefe/lib/source_c_impl.cc:121-132
__device__ __forceinline__ float2 cvt_i16(uint32_t w)
{
    const float q = 1.0f / 32767.0f;
    const int lo = (int)(short)(w & 0xffffu), hi = (int)(short)(w >> 16);
    return make_float2(__fmul_rn((float)lo, q), __fmul_rn((float)hi, q));
}
__device__ __forceinline__ float2 cvt_u8(uint16_t w)
{
    const float q = 1.0f / 127.0f;
    const int lo = (int)(w & 0xffu) - 128, hi = (int)(w >> 8) - 128;
    return make_float2(__fmul_rn((float)lo, q), __fmul_rn((float)hi, q));
}
__host__ __device__ __forceinline__ int fmt_bytes(int fmt) { return fmt == SFE_IN_I16 ? 4 : fmt == SFE_IN_U8 ? 2 : 8; }

// A stream view: `cur` holds samples [0, n) of this call, `hist` the H samples before them
// (hist[H-1] is sample -1); samples >= n read as zero.
struct StreamView {
    const void *cur;   // row 0 of the batch
    const void *hist;  // row 0 of the history, H samples per row, same format
    long pitch;        // row pitch of cur in complex samples
    int H;             // history length (row pitch of hist)
    long n;            // valid samples per row
    int fmt;           // SFE_IN_*
};

__device__ __forceinline__ float2 load_raw(const void *p, long idx, int fmt)
{
    if (fmt == SFE_IN_I16) return cvt_i16(__ldg(reinterpret_cast<const uint32_t *>(p) + idx));
    if (fmt == SFE_IN_U8) return cvt_u8(__ldg(reinterpret_cast<const uint16_t *>(p) + idx));
    return __ldg(reinterpret_cast<const float2 *>(p) + idx);
}

// sample i (may be negative or >= n) of stream row s
__device__ __forceinline__ float2 load_sample(const StreamView &v, int s, long i)
{
    if (i >= v.n) return make_float2(0.f, 0.f);
    if (i >= 0) return load_raw(v.cur, (long)s * v.pitch + i, v.fmt);
    const long h = (long)v.H + i;
    if (h < 0) return make_float2(0.f, 0.f);
    return load_raw(v.hist, (long)s * v.H + h, v.fmt);
}

}  // namespace sfe
