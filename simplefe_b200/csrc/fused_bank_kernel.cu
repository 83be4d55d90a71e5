// fused_bank_kernel.cu -- K3: the whole blkconv -> decimate -> resample chain of many int16 IQ streams as ONE
// polyphase filter bank that reads the raw samples once and writes only the final outputs.
//
// The three reference stages are linear and, for the integral-step configurations, periodically time
// invariant (blkconv.cxx:77-110 is a causal FIR h; decimate.cxx:102-140 with rate D keeps every D-th sample of
// a FIR a; resample.cxx:100-148 evaluates phase (pos mod L) of a polyphase FIR r at decimated index pos/L).
// With g = h * a (chain.cu, in double) the output at grid position pos = pos0 + k*step is
//     y[k] = sum_j r[ph + L j] * d[M - j],   d[m] = sum_i g[i] x[n_first + D m - i],   M = floor(pos/L), ph = pos mod L
//          = sum_u c_ph[u] * x[R_k - u],     c_ph[u] = sum_j r[ph + L j] g[u - D j],   R_k = n_first + D M
// i.e. L tap sets of length Tc = D (J-1) + len(g) applied at the retained output instants only.  For the
// BASELINE chain (255-tap blkconv, 128-tap /8, 512-tap L=24 at step 625) this is 24 x 550 taps and 5.3 MAC per
// input sample instead of 95.5: the kernel is bound by the single HBM read of the int16 samples.
//
// Mapping (streams share the filter, so the output schedule and the taps are warp-uniform):
//   lane   = one stream: 32 streams per warp walk forward in time together
//   warp   = one work item (group of 32 streams, contiguous range of outputs); 8 independent warps per CTA
//   input  = TMA 2-D tiles (32 samples x 32 streams, SWIZZLE_128B) into a per-warp shared-memory ring,
//            completion on mbarriers; the swizzle makes the row-per-lane 128-bit reads conflict free
//   taps   = the bank table in shared memory, time-reversed and chunk aligned; every lane reads the same
//            128-bit word (broadcast), one load feeds 8 FMAs
//   samples are converted int16 -> float once (I2F), the 1/32767 of SURVEY.md D2 is folded into the taps
//   up to four outputs are live per lane (their windows overlap); accumulators rotate as outputs finish.
#include <cuda.h>

#include <map>
#include <mutex>

#include "kernels.h"

namespace sfe {

namespace {

constexpr int BK_WARPS = 8;          // work items per CTA
constexpr int BK_NST = 2;            // ring stages per warp
constexpr int BK_BPS = 2;            // TMA boxes per stage
constexpr int BOX_SAMPLES = 32, BOX_ROWS = 32;
constexpr int BOX_BYTES = BOX_SAMPLES * BOX_ROWS * 4;
constexpr int STAGE_BYTES = BK_BPS * BOX_BYTES;
constexpr int A_MAX = 4;             // live outputs per lane

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_box(uint32_t dst, const CUtensorMap *map, int x, int y, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar) : "memory");
}

// the two int16 halves of a word as int32: PRMT with sign replication (selector nibble bit 3), one ALU op each
__device__ __forceinline__ int sext_lo(uint32_t w)
{
    int r;
    asm("prmt.b32 %0, %1, %1, 0x9910;" : "=r"(r) : "r"(w));
    return r;
}
__device__ __forceinline__ int sext_hi(uint32_t w)
{
    int r;
    asm("prmt.b32 %0, %1, %1, 0xBB32;" : "=r"(r) : "r"(w));
    return r;
}

struct OutInfo {
    int E;     // chunk (4 samples, counted from sample -H) that holds the newest sample of the window
    int row;   // row of the bank table
};

__device__ __forceinline__ OutInfo out_info(const BankParams &p, long k)
{
    const long pos = p.pos0 + k * p.step;
    long M = pos / p.L;
    int ph = (int)(pos - M * p.L);
    if (ph < 0) { ph += p.L; M -= 1; }                    // floor division: pos = -1 is a held-over output
    const long rh = p.n_first + (long)p.D * M + p.H;      // newest sample, counted from sample -H
    OutInfo o;
    o.E = (int)(rh >> 2);
    o.row = p.rows_per_phase == 4 ? ph * 4 + (int)(rh & 3) : ph;
    return o;
}

// NA live windows over chunks [c0, c1) of the current ring stage (chunk = 4 samples = one 128-bit word per lane)
template <int NA>
__device__ __forceinline__ void run_chunks(const uint8_t *sp, int c0, int c1, uint32_t lane_x, float2 (&acc)[A_MAX],
                                           const float4 *(&tp)[A_MAX])
{
#pragma unroll 2
    for (int c = c0; c < c1; c++) {
        // box (c >> 3) of the stage, 16-byte slot (c & 7) ^ (row & 7) of this lane's 128-byte row (SWIZZLE_128B)
        const uint4 w = *reinterpret_cast<const uint4 *>(sp + (c >> 3) * BOX_BYTES + ((((uint32_t)c & 7u) ^ lane_x) << 4));
        float xi[4], xq[4];
        xi[0] = (float)sext_lo(w.x); xq[0] = (float)sext_hi(w.x);
        xi[1] = (float)sext_lo(w.y); xq[1] = (float)sext_hi(w.y);
        xi[2] = (float)sext_lo(w.z); xq[2] = (float)sext_hi(w.z);
        xi[3] = (float)sext_lo(w.w); xq[3] = (float)sext_hi(w.w);
#pragma unroll
        for (int j = 0; j < NA; j++) {
            const float4 t = *tp[j];
            tp[j]++;
            acc[j].x = fmaf(t.x, xi[0], acc[j].x); acc[j].y = fmaf(t.x, xq[0], acc[j].y);
            acc[j].x = fmaf(t.y, xi[1], acc[j].x); acc[j].y = fmaf(t.y, xq[1], acc[j].y);
            acc[j].x = fmaf(t.z, xi[2], acc[j].x); acc[j].y = fmaf(t.z, xq[2], acc[j].y);
            acc[j].x = fmaf(t.w, xi[3], acc[j].x); acc[j].y = fmaf(t.w, xq[3], acc[j].y);
        }
    }
}

__global__ void __launch_bounds__(BK_WARPS * 32, 1)
fused_bank_kernel(const __grid_constant__ CUtensorMap map_cur, const __grid_constant__ CUtensorMap map_hist, const BankParams p)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[BK_WARPS * BK_NST];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *sbase = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);   // SWIZZLE_128B boxes: 1024 B alignment
    uint8_t *ring = sbase + (size_t)warp * BK_NST * STAGE_BYTES;
    float *table = reinterpret_cast<float *>(sbase + (size_t)BK_WARPS * BK_NST * STAGE_BYTES);

    // ---- one-time: bank table -> shared memory, barriers ------------------------------------------------
    {
        const int n16 = p.n_rows * p.row_len / 4;
        const float4 *src = reinterpret_cast<const float4 *>(p.table);
        float4 *dst = reinterpret_cast<float4 *>(table);
        for (int i = threadIdx.x; i < n16; i += BK_WARPS * 32) dst[i] = __ldg(src + i);
    }
    if (lane == 0) {
        for (int s = 0; s < BK_NST; s++) mbar_init(smem_u32(&bars[warp * BK_NST + s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    // ---- this warp's work item ---------------------------------------------------------------------------
    const long item = (long)blockIdx.x * BK_WARPS + warp;
    if (item >= (long)p.n_groups * p.n_seg) return;
    const int g = (int)(item % p.n_groups), seg = (int)(item / p.n_groups);
    const long k_lo = p.n_out * seg / p.n_seg, k_hi = p.n_out * (seg + 1) / p.n_seg;
    if (k_lo >= k_hi) return;
    const int NC = p.row_len >> 2;

    const OutInfo first = out_info(p, k_lo), last = out_info(p, k_hi - 1);
    const int q_begin = first.E - NC + 1, q_end = last.E;          // chunks this item reads (inclusive)
    const int box0 = q_begin >> 3, box1 = q_end >> 3;              // 8 chunks per box
    const int n_batches = (box1 - box0) / BK_BPS + 1;
    const int hist_boxes = p.H >> 5;
    const uint32_t bar0 = smem_u32(&bars[warp * BK_NST]);
    const uint32_t ring_a = smem_u32(ring);

    auto issue = [&](int batch) {                                  // lane 0 only
        const int st = batch % BK_NST;
        const uint32_t bar = bar0 + 8 * st;
        mbar_expect_tx(bar, STAGE_BYTES);
#pragma unroll
        for (int b = 0; b < BK_BPS; b++) {
            const int box = box0 + batch * BK_BPS + b;             // samples [32 box - H, 32 box - H + 32)
            const uint32_t dst = ring_a + st * STAGE_BYTES + b * BOX_BYTES;
            if (box < hist_boxes) tma_box(dst, &map_hist, box * BOX_SAMPLES, g * BOX_ROWS, bar);
            else tma_box(dst, &map_cur, (box - hist_boxes) * BOX_SAMPLES, g * BOX_ROWS, bar);
        }
    };
    if (lane == 0)
        for (int b = 0; b < BK_NST && b < n_batches; b++) issue(b);

    // ---- live outputs (slot 0 finishes first) ---------------------------------------------------------
    float2 acc[A_MAX];
    const float4 *tp[A_MAX];
    int Es[A_MAX];
#pragma unroll
    for (int j = 0; j < A_MAX; j++) {
        acc[j] = make_float2(0.f, 0.f);
        tp[j] = reinterpret_cast<const float4 *>(table);
        Es[j] = 0x3fffffff;
    }
    int nact = 0;
    long k_next = k_lo, k_head = k_lo;
    OutInfo nxt = first;
    int S_next = q_begin;
    const int s_row = g * BOX_ROWS + lane;
    float2 *orow = p.out + (long)s_row * p.out_pitch;
    const bool s_ok = s_row < p.n_streams;
    const uint32_t lane_off = (uint32_t)lane * 128u, lane_x = (uint32_t)(lane & 7);

    for (int batch = 0; batch < n_batches; batch++) {
        const int st = batch % BK_NST;
        mbar_wait(bar0 + 8 * st, (uint32_t)((batch / BK_NST) & 1));
        const int qb = (box0 + batch * BK_BPS) << 3;               // first chunk of this stage
        const uint8_t *sp = ring + st * STAGE_BYTES + lane_off;
        int q = qb > q_begin ? qb : q_begin;
        const int q_stop = (qb + 8 * BK_BPS) <= q_end + 1 ? (qb + 8 * BK_BPS) : q_end + 1;
        while (q < q_stop) {
            if (q == S_next) {                                     // a window opens (at most one per chunk)
#pragma unroll
                for (int j = 0; j < A_MAX; j++)
                    if (j == nact) {
                        acc[j] = make_float2(0.f, 0.f);
                        tp[j] = reinterpret_cast<const float4 *>(table + (size_t)nxt.row * p.row_len);
                        Es[j] = nxt.E;
                    }
                nact++;
                k_next++;
                if (k_next < k_hi) {
                    nxt = out_info(p, k_next);
                    S_next = nxt.E - NC + 1;
                } else {
                    S_next = 0x3fffffff;
                }
            }
            // chunks [q, run_end): no window opens, the oldest one may close with the last chunk
            int run_end = q_stop < S_next ? q_stop : S_next;
            if (Es[0] + 1 < run_end) run_end = Es[0] + 1;
            switch (nact) {
                case 1: run_chunks<1>(sp, q - qb, run_end - qb, lane_x, acc, tp); break;
                case 2: run_chunks<2>(sp, q - qb, run_end - qb, lane_x, acc, tp); break;
                case 3: run_chunks<3>(sp, q - qb, run_end - qb, lane_x, acc, tp); break;
                case 4: run_chunks<4>(sp, q - qb, run_end - qb, lane_x, acc, tp); break;
                default: break;
            }
            q = run_end;
            if (q - 1 == Es[0]) {                                  // the oldest window closes
                if (s_ok) orow[k_head] = acc[0];
                k_head++;
#pragma unroll
                for (int j = 0; j + 1 < A_MAX; j++) {
                    acc[j] = acc[j + 1];
                    tp[j] = tp[j + 1];
                    Es[j] = Es[j + 1];
                }
                Es[A_MAX - 1] = 0x3fffffff;
                nact--;
            }
        }
        __syncwarp();
        if (lane == 0 && batch + BK_NST < n_batches) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // our reads before the next TMA write
            issue(batch + BK_NST);
        }
    }
}

// ---- cuTensorMapEncodeTiled through the runtime (no link against libcuda) ---------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(f);
    });
    return fn;
}

bool encode_rows(CUtensorMap *m, const void *base, long n_cols, long n_rows, long pitch)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) {
        set_error("fused bank: cuTensorMapEncodeTiled is not available from this driver");
        return false;
    }
    const cuuint64_t dims[2] = {(cuuint64_t)n_cols, (cuuint64_t)n_rows};
    const cuuint64_t strides[1] = {(cuuint64_t)pitch * 4};
    const cuuint32_t box[2] = {BOX_SAMPLES, BOX_ROWS};
    const cuuint32_t es[2] = {1, 1};
    const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void *>(base), dims, strides, box, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("fused bank: cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
        return false;
    }
    return true;
}

}  // namespace

size_t bank_smem_bytes(int n_rows, int row_len)
{
    return (size_t)BK_WARPS * BK_NST * STAGE_BYTES + (size_t)n_rows * row_len * sizeof(float) + 1024;
}

int bank_max_live() { return A_MAX; }

int launch_bank(const BankParams &p_in, cudaStream_t st)
{
    BankParams p = p_in;
    if (p.n_out <= 0 || p.n_streams <= 0) return SFE_OK;
    if ((reinterpret_cast<uintptr_t>(p.in.cur) & 15) || (p.in.pitch & 3) || (p.H & 31) || p.in.fmt != SFE_IN_I16) {
        set_error("fused bank: input rows must be int16 IQ, 16-byte aligned with a pitch that is a multiple of 4 samples");
        return SFE_ERR_ARG;
    }
    CUtensorMap mc, mh;
    if (!encode_rows(&mc, p.in.cur, p.in.n, p.n_streams, p.in.pitch) || !encode_rows(&mh, p.in.hist, p.H, p.n_streams, p.H))
        return SFE_ERR_CUDA;
    int dev = 0, sms = 0;
    if (!SFE_CUDA(cudaGetDevice(&dev)) || !SFE_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev))) return SFE_ERR_CUDA;
    p.n_groups = (p.n_streams + BOX_ROWS - 1) / BOX_ROWS;
    // one work item per warp slot of the machine, but never fewer than 8 outputs per item
    long n_seg = ((long)sms * BK_WARPS + p.n_groups - 1) / p.n_groups;
    if (n_seg > p.n_out / 8) n_seg = p.n_out / 8;
    if (n_seg < 1) n_seg = 1;
    p.n_seg = (int)n_seg;
    const long items = (long)p.n_groups * p.n_seg;
    const size_t smem = bank_smem_bytes(p.n_rows, p.row_len);
    if (!SFE_CUDA(cudaFuncSetAttribute(fused_bank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))) return SFE_ERR_CUDA;
    fused_bank_kernel<<<(unsigned)((items + BK_WARPS - 1) / BK_WARPS), BK_WARPS * 32, smem, st>>>(mc, mh, p);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

// Host side: the bank table.  c_ph[u] = sum_j r[ph + L j] g[u - D j] in double, scaled, rounded to float, stored
// time-reversed and chunk aligned: row (ph, e) holds, for chunk-relative sample i, the tap u = 4 NC - 4 + e - i
// where e = (newest sample + H) mod 4.
void bank_table(const std::vector<double> &g, int D, const std::vector<float> &r, int L, double scale, int e_only,
                std::vector<float> &rows, int *row_len, int *n_rows, int *Tc_out)
{
    const int J = ((int)r.size() + L - 1) / L;
    const int Tc = D * (J - 1) + (int)g.size();
    const int NC = (Tc + 3 + 3) / 4;
    const int RL = 4 * NC;
    const int per = e_only >= 0 ? 1 : 4;
    rows.assign((size_t)L * per * RL, 0.f);
    std::vector<double> c((size_t)Tc);
    for (int ph = 0; ph < L; ph++) {
        std::fill(c.begin(), c.end(), 0.0);
        for (int j = 0; ph + L * j < (int)r.size(); j++) {
            const double rv = (double)r[ph + L * j];
            if (rv == 0.0) continue;
            for (size_t i = 0; i < g.size(); i++) c[(size_t)D * j + i] += rv * g[i];
        }
        for (int v = 0; v < per; v++) {
            const int e = e_only >= 0 ? e_only : v;
            float *row = rows.data() + ((size_t)ph * per + v) * RL;
            for (int i = 0; i < RL; i++) {
                const int u = RL - 4 + e - i;
                if (u >= 0 && u < Tc) row[i] = (float)(c[u] * scale);
            }
        }
    }
    *row_len = RL;
    *n_rows = L * per;
    *Tc_out = Tc;
}

}  // namespace sfe
