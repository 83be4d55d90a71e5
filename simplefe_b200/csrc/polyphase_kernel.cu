// polyphase_kernel.cu -- K2g / K2d: the polyphase FIR stages behind `decimate` and `resample`.
//
// Reference semantics (libdsp/decimate.cxx:102-125,132-140; libdsp/resample.cxx:100-148):
//   s(pos)  = sum_j taps[(pos mod L) + j*L] * x[floor(pos/L) - j]        (polyphase output on the
//                                                                          L-times upsampled grid)
//   out[k]  = s(pos_k) * (1 - mu_k) + mu_k * s(pos_k + 1)
// with (pos_k, mu_k) from the data-independent float32 time walk (schedule.cpp).  Both reference
// classes produce these numbers (bit-identically); `resample` merely evaluates all
// L phases for every input first.  The kernels below compute retained outputs only.
#include "kernels.h"

namespace sfe {

namespace {

// ---------------------------------------------------------------------------------------------
// K2g: any L, any rate, table or closed-form schedule, complex or real rows.  One thread per
// output; window read straight from global memory (L1/L2 serve the overlap between neighbours).
// ---------------------------------------------------------------------------------------------
template <bool EXACT, bool REAL>
__device__ __forceinline__ float2 poly_point(const PolyParams &p, int s, long pos)
{
    // floor division (pos may be -1 for an output held over a call boundary)
    long n0 = pos / p.up;
    int ph = (int)(pos - n0 * p.up);
    if (ph < 0) { ph += p.up; n0 -= 1; }
    float2 acc = make_float2(0.f, 0.f);
    long n = n0;
    for (int m = ph; m < p.n_taps; m += p.up, n--) {
        const float t = __ldg(p.taps + m);
        float2 x;
        if (REAL) {
            float xr = 0.f;
            if (n < p.in.n) {
                if (n >= 0) xr = __ldg(reinterpret_cast<const float *>(p.in.cur) + (long)s * p.in.pitch + n);
                else if (p.in.H + n >= 0) xr = __ldg(reinterpret_cast<const float *>(p.in.hist) + (long)s * p.in.H + p.in.H + n);
            }
            x = make_float2(xr, 0.f);
        } else {
            x = load_sample(p.in, s, n);
        }
        if (EXACT) {  // accu += taps[m] * in[n-j]   as two IEEE operations (decimate.cxx:136)
            acc.x = __fadd_rn(acc.x, __fmul_rn(t, x.x));
            if (!REAL) acc.y = __fadd_rn(acc.y, __fmul_rn(t, x.y));
        } else {
            acc.x = fmaf(t, x.x, acc.x);
            if (!REAL) acc.y = fmaf(t, x.y, acc.y);
        }
    }
    return acc;
}

template <bool EXACT, bool REAL>
__global__ void __launch_bounds__(128) poly_generic_kernel(const PolyParams p)
{
    const long k = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= p.n_out) return;
    long pos;
    float mu;
    if (p.uniform) {
        pos = p.pos0 + k * p.step;
        mu = 0.f;
    } else {
        const SchedEntry e = p.sched[k];
        pos = e.pos;
        mu = e.mu;
    }
    for (int s = blockIdx.y; s < p.n_streams; s += gridDim.y) {
        const float2 s0 = poly_point<EXACT, REAL>(p, s, pos);
        float2 y;
        if (EXACT) {
            // get_sample(phase0,n0) * (1.0f-m_mu) + m_mu*get_sample(phase1,n1)   decimate.cxx:123
            const float2 s1 = poly_point<EXACT, REAL>(p, s, pos + 1);
            const float w0 = __fsub_rn(1.0f, mu);
            y.x = __fadd_rn(__fmul_rn(s0.x, w0), __fmul_rn(mu, s1.x));
            y.y = __fadd_rn(__fmul_rn(s0.y, w0), __fmul_rn(mu, s1.y));
        } else if (mu != 0.f) {
            const float2 s1 = poly_point<EXACT, REAL>(p, s, pos + 1);
            const float w0 = 1.0f - mu;
            y.x = fmaf(mu, s1.x, s0.x * w0);
            y.y = fmaf(mu, s1.y, s0.y * w0);
        } else {
            y = s0;  // mu == 0: s0*1 + 0*s1
        }
        if (REAL) reinterpret_cast<float *>(p.out)[(long)s * p.out_pitch + k] = y.x;
        else reinterpret_cast<float2 *>(p.out)[(long)s * p.out_pitch + k] = y;
    }
}

// ---------------------------------------------------------------------------------------------
// K2d: y[k] = sum_j g[j] x[n_first + D k - j], complex rows, CUDA cores.
// The CTA stages its input span in shared memory split into D polyphase planes
// (plane r holds x[n_first + D m - r]), so that per plane the work is a plain FIR whose
// neighbouring outputs share their window.  Each thread owns R consecutive outputs and slides a
// register window over 8 taps at a time: 8*R*2 FMAs per (R+7) 64-bit shared loads.
// ---------------------------------------------------------------------------------------------
constexpr int DR = 8;     // outputs per thread
constexpr int DNT = 128;  // threads per CTA
constexpr int DKT = DR * DNT;

__device__ __forceinline__ int dphi(int a) { return a + a / DR; }

__global__ void __launch_bounds__(DNT) decim_fir_kernel(const DecimParams p)
{
    extern __shared__ __align__(16) unsigned char smraw[];
    const int D = p.D, QLP = p.QLP;
    const int ML = DKT + QLP - 1;              // samples per plane
    const int PITCH = dphi(ML) + 1;
    float *gs = reinterpret_cast<float *>(smraw);                       // [D][QLP]
    float2 *xs = reinterpret_cast<float2 *>(smraw + sizeof(float) * D * QLP);  // [D][PITCH]

    for (int i = threadIdx.x; i < D * QLP; i += DNT) gs[i] = __ldg(p.gp + i);

    const long k0 = (long)blockIdx.x * DKT;
    const int s = blockIdx.y;
    // x index of (plane r, entry mm): n_first + D*(k0 - QLP + 1 + mm) - r ; enumerate e = mm*D + (D-1-r)
    const long xbase = p.n_first + (long)D * (k0 - QLP + 1) - (D - 1);
    const int total = D * ML;
    for (int e = threadIdx.x; e < total; e += DNT) {
        const int mm = e / D, r = D - 1 - (e - mm * D);
        xs[r * PITCH + dphi(mm)] = load_sample(p.in, s, xbase + e);
    }
    __syncthreads();

    float2 acc[DR];
#pragma unroll
    for (int i = 0; i < DR; i++) acc[i] = make_float2(0.f, 0.f);

    const int a0 = threadIdx.x * DR + QLP - 1 - 7;  // window start for chunk qc = 0
    for (int r = 0; r < D; r++) {
        const float2 *xr = xs + r * PITCH;
        const float *gr = gs + r * QLP;
        for (int qc = 0; qc < QLP; qc += 8) {
            const float4 ga = *reinterpret_cast<const float4 *>(gr + qc);
            const float4 gb = *reinterpret_cast<const float4 *>(gr + qc + 4);
            const float g[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
            float2 w[DR + 7];
            const int a = a0 - qc;
#pragma unroll
            for (int c = 0; c < DR + 7; c++) w[c] = xr[dphi(a + c)];
#pragma unroll
            for (int d = 0; d < 8; d++)
#pragma unroll
                for (int i = 0; i < DR; i++) {
                    acc[i].x = fmaf(g[d], w[7 - d + i].x, acc[i].x);
                    acc[i].y = fmaf(g[d], w[7 - d + i].y, acc[i].y);
                }
        }
    }
    const long kk = k0 + (long)threadIdx.x * DR;
    float2 *o = p.out + (long)s * p.out_pitch + kk;
#pragma unroll
    for (int i = 0; i < DR; i++)
        if (kk + i < p.n_out) o[i] = acc[i];
}

}  // namespace

int launch_poly_generic(const PolyParams &p, cudaStream_t st)
{
    if (p.n_out <= 0 || p.n_streams <= 0) return SFE_OK;
    dim3 grid((unsigned)((p.n_out + 127) / 128), (unsigned)(p.n_streams < 32768 ? p.n_streams : 32768));
    if (p.exact) {
        if (p.in_real) poly_generic_kernel<true, true><<<grid, 128, 0, st>>>(p);
        else poly_generic_kernel<true, false><<<grid, 128, 0, st>>>(p);
    } else {
        if (p.in_real) poly_generic_kernel<false, true><<<grid, 128, 0, st>>>(p);
        else poly_generic_kernel<false, false><<<grid, 128, 0, st>>>(p);
    }
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

int launch_decim_fir(const DecimParams &p, cudaStream_t st)
{
    if (p.n_out <= 0 || p.n_streams <= 0) return SFE_OK;
    const int ML = DKT + p.QLP - 1;
    const int PITCH = ML + ML / DR + 1;
    const size_t smem = sizeof(float) * p.D * p.QLP + sizeof(float2) * (size_t)p.D * PITCH;
    if (smem > 220 * 1024 || p.n_streams > 65535) {
        set_error("decim_fir: tile does not fit shared memory (D * taps too large)");
        return SFE_ERR_ARG;
    }
    if (smem > 48 * 1024 &&
        !SFE_CUDA(cudaFuncSetAttribute(decim_fir_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)))
        return SFE_ERR_CUDA;
    dim3 grid((unsigned)((p.n_out + DKT - 1) / DKT), (unsigned)p.n_streams);
    decim_fir_kernel<<<grid, DNT, smem, st>>>(p);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

}  // namespace sfe
