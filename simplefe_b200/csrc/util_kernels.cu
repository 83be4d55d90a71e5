// util_kernels.cu -- state carry and wire-format helpers.
#include "kernels.h"

#include <cstdlib>

namespace sfe {

namespace {

// hist_new[s][i] = (hist_old[s] ++ cur[s][0..n))[n + i], i < H.  Works on 16-bit units so that
// every sample format (2, 4, 8 bytes) shares one kernel.
__global__ void save_tail_kernel(uint16_t *hn, const uint16_t *ho, const uint16_t *cur, long pitch_u, long n_u,
                                 int H_u, int n_streams)
{
    // the next call's main kernel may be queued behind this grid as a programmatic dependent: let its CTAs start their
    // prologue as soon as there is room (it waits for this grid -- and through it for this call's main kernel -- before
    // it touches anything either of them reads or writes)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    // one flat grid-stride loop over all rows: a few hundred CTAs whatever the stream count, all resident at once
    const long total = (long)n_streams * H_u;
    for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
        const long s = e / H_u, i = e - s * H_u;
        const long src = n_u + i - H_u;  // index into cur when >= 0, else into the old history
        hn[e] = src >= 0 ? cur[s * pitch_u + src] : ho[s * H_u + (H_u + src)];
    }
    // Launched as a programmatic dependent of the main kernel (which executes griddepcontrol.launch_dependents on
    // entry), this grid runs BESIDE it: the copies above need nothing the main kernel writes.  The wait at the end
    // keeps the stream order well defined -- this grid completes only after the primary has completed and flushed, so
    // whatever follows on the stream (a D2H copy of the outputs, the caller's own kernels) sees the main kernel's
    // results.  Without the launch attribute the instruction is a no-op.
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// Device-side expansion of a fractional-rate schedule: the host walks the reference's float32 recurrence run by run
// (schedule.cpp, SchedRun: integer mantissa arithmetic inside one binade), the device turns every run into its entries.
// Bit for bit the reference's sequence (tests/cpp/walk_diff.cxx expands the same runs on the host).
__global__ void expand_sched_kernel(const SchedRun *runs, int n_runs, SchedEntry *out)
{
    for (int r = blockIdx.x; r < n_runs; r += gridDim.x) {
        const SchedRun run = runs[r];
        if (run.sh < 0) {
            if (threadIdx.x == 0) out[run.k0] = SchedEntry{run.base, __uint_as_float(run.ubits)};
            continue;
        }
        const float u = __uint_as_float(run.ubits);
        const uint32_t fm = (1u << run.sh) - 1u;
        for (int j = threadIdx.x; j < run.n; j += blockDim.x) {
            const uint32_t X = run.X0 + (uint32_t)j * run.INC;
            out[run.k0 + j] = SchedEntry{run.base + (int32_t)(X >> run.sh), __fmul_rn((float)(X & fm), u)};
        }
    }
}

// examples/bpsk/bpsk.cxx:76-101 / gr-simplefe/lib/sink_c_impl.cc:118-144:
//   u = ((short)(x*511) + 512) & 0x3FF ; 4 words -> 5 bytes, the four 2-bit tops first
__global__ void pack_dac10_kernel(const float *in, long n4, uint8_t *out)
{
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n4) return;
    const float4 x = reinterpret_cast<const float4 *>(in)[g];
    const float v[4] = {x.x, x.y, x.z, x.w};
    unsigned u[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const short sh = (short)(int)__fmul_rn(v[k], 511.0f);  // C float->short: truncation toward zero
        u[k] = ((unsigned)(sh + 512)) & 0x3FFu;
    }
    uint8_t *o = out + g * 5;
    o[0] = (uint8_t)((u[0] >> 8) | ((u[1] >> 8) << 2) | ((u[2] >> 8) << 4) | ((u[3] >> 8) << 6));
    o[1] = (uint8_t)(u[0] & 0xFF);
    o[2] = (uint8_t)(u[1] & 0xFF);
    o[3] = (uint8_t)(u[2] & 0xFF);
    o[4] = (uint8_t)(u[3] & 0xFF);
}

}  // namespace

int launch_save_tail(void *hist_new, const void *hist_old, const void *cur, long pitch, long n, int H, int elem,
                     int n_streams, cudaStream_t st, bool overlap_prev)
{
    if (H <= 0 || n_streams <= 0) return SFE_OK;
    static const bool no_pdl = getenv("SFE_NO_PDL") != nullptr;
    const int u = elem / 2;
    const int H_u = H * u;
    long ctas = ((long)n_streams * H_u + 255) / 256;
    if (ctas > 592) ctas = 592;          // 4 per SM
    dim3 grid((unsigned)ctas);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = overlap_prev && !no_pdl ? 1 : 0;
    if (!SFE_CUDA(cudaLaunchKernelEx(&cfg, save_tail_kernel, static_cast<uint16_t *>(hist_new), static_cast<const uint16_t *>(hist_old),
                                     static_cast<const uint16_t *>(cur), pitch * u, n * u, H_u, n_streams)))
        return SFE_ERR_CUDA;
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

__global__ void block_index_kernel(const SchedEntry *sched, long n_out, long blk_start, int hop, long n_blocks, int *kidx)
{
    const long b = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b > n_blocks) return;
    const long target = blk_start + b * hop;
    long lo = 0, hi = n_out;                  // lower bound: positions increase strictly along the schedule
    while (lo < hi) {
        const long mid = (lo + hi) >> 1;
        if ((long)sched[mid].pos < target) lo = mid + 1;
        else hi = mid;
    }
    kidx[b] = (int)lo;
}

int launch_block_index(const SchedEntry *d_sched, long n_out, long blk_start, int hop, long n_blocks, int *d_kidx, cudaStream_t st)
{
    if (n_blocks < 0) return SFE_OK;
    block_index_kernel<<<(unsigned)((n_blocks + 1 + 127) / 128), 128, 0, st>>>(d_sched, n_out, blk_start, hop, n_blocks, d_kidx);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

int launch_expand_sched(const SchedRun *d_runs, int n_runs, SchedEntry *d_out, cudaStream_t st)
{
    if (n_runs <= 0) return SFE_OK;
    expand_sched_kernel<<<n_runs < 2368 ? n_runs : 2368, 128, 0, st>>>(d_runs, n_runs, d_out);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

int launch_pack_dac10(const float *in, long n, uint8_t *out, cudaStream_t st)
{
    if (n <= 0) return SFE_OK;
    const long n4 = n / 4;
    pack_dac10_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(in, n4, out);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

}  // namespace sfe
