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

#include <algorithm>
#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>

#include "kernels.h"

namespace sfe {

namespace {

constexpr int BOX_SAMPLES = 32, BOX_ROWS = 32;          // a box row is 32 words = 128 bytes: 32 int16 or 64 u8 complex samples
constexpr int BOX_BYTES = BOX_SAMPLES * BOX_ROWS * 4;   // one TMA box = 8 chunks of 16 bytes per lane
constexpr int SLOT_BOXES = 2;         // boxes (time blocks of 128 B per row) fetched by one TMA instruction
constexpr int SLOT_BYTES = SLOT_BOXES * BOX_BYTES;
constexpr int A_MAX = 4;             // window slots per lane, at most
constexpr int PAD_CHUNKS = 8;        // zero chunks in front of and behind the taps of every table row

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

__device__ __forceinline__ void tma_box3(uint32_t dst, const CUtensorMap *map, int tb, int y, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(map), "r"(0), "r"(tb), "r"(y), "r"(bar) : "memory");
}

// ---- packed single precision (sm_100: FFMA2 / FADD2 work on an aligned register pair) -------------------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float2 unpack2(f32x2 v)
{
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 fadd2(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// shared-memory loads by 32-bit address.  volatile: the ring slots are rewritten by the TMA (async proxy) behind
// the compiler's back, so these loads must neither be merged across laps nor moved over the mbarrier wait.
__device__ __forceinline__ uint4 lds_u4(uint32_t addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
template <int OFF>
__device__ __forceinline__ void lds_taps(uint32_t addr, f32x2 &t01, f32x2 &t23)
{
    asm volatile("ld.shared.v2.u64 {%0, %1}, [%2+%3];" : "=l"(t01), "=l"(t23) : "r"(addr), "n"(OFF));
}

// ---- int16 I/Q -> float, four complex samples (one 128-bit word) into (even, odd) pairs ------------------
// CONV 0: PRMT sign extension + I2FP (ALU pipe);  1: I2F.S16 on the packed halves (XU pipe);
// 2: I through the XU, Q through the ALU;  3: exponent trick, exact: bits 0x4B40_0000 | (s ^ 0x8000) are the float
//    12582912 + 32768 + s, one packed FADD2 removes the offset for two values.
__device__ __forceinline__ float cvt_lo_alu(uint32_t w)
{
    int r;
    asm("prmt.b32 %0, %1, %1, 0x9910;" : "=r"(r) : "r"(w));   // bytes (b0, b1, sign, sign)
    return (float)r;
}
__device__ __forceinline__ float cvt_hi_alu(uint32_t w)
{
    int r;
    asm("prmt.b32 %0, %1, %1, 0xBB32;" : "=r"(r) : "r"(w));
    return (float)r;
}
__device__ __forceinline__ float cvt_lo_xu(uint32_t w)
{
    float f;
    asm("{\n .reg .b16 l, h;\n mov.b32 {l, h}, %1;\n cvt.rn.f32.s16 %0, l;\n}\n" : "=f"(f) : "r"(w));
    return f;
}
__device__ __forceinline__ float cvt_hi_xu(uint32_t w)
{
    float f;
    asm("{\n .reg .b16 l, h;\n mov.b32 {l, h}, %1;\n cvt.rn.f32.s16 %0, h;\n}\n" : "=f"(f) : "r"(w));
    return f;
}
template <int CONV>
__device__ __forceinline__ void convert4(const uint4 &w, f32x2 &i01, f32x2 &i23, f32x2 &q01, f32x2 &q23)
{
    if (CONV == 3) {
        const uint32_t K = 0x4B400000u;
        const f32x2 off = pack2(-12615680.0f, -12615680.0f);
        const uint32_t a = w.x ^ 0x80008000u, b = w.y ^ 0x80008000u, c = w.z ^ 0x80008000u, d = w.w ^ 0x80008000u;
        i01 = fadd2(pack2(__uint_as_float(__byte_perm(a, K, 0x7610)), __uint_as_float(__byte_perm(b, K, 0x7610))), off);
        q01 = fadd2(pack2(__uint_as_float(__byte_perm(a, K, 0x7632)), __uint_as_float(__byte_perm(b, K, 0x7632))), off);
        i23 = fadd2(pack2(__uint_as_float(__byte_perm(c, K, 0x7610)), __uint_as_float(__byte_perm(d, K, 0x7610))), off);
        q23 = fadd2(pack2(__uint_as_float(__byte_perm(c, K, 0x7632)), __uint_as_float(__byte_perm(d, K, 0x7632))), off);
    } else {
        const bool ixu = CONV == 1 || CONV == 2, qxu = CONV == 1;
        i01 = pack2(ixu ? cvt_lo_xu(w.x) : cvt_lo_alu(w.x), ixu ? cvt_lo_xu(w.y) : cvt_lo_alu(w.y));
        i23 = pack2(ixu ? cvt_lo_xu(w.z) : cvt_lo_alu(w.z), ixu ? cvt_lo_xu(w.w) : cvt_lo_alu(w.w));
        q01 = pack2(qxu ? cvt_hi_xu(w.x) : cvt_hi_alu(w.x), qxu ? cvt_hi_xu(w.y) : cvt_hi_alu(w.y));
        q23 = pack2(qxu ? cvt_hi_xu(w.z) : cvt_hi_alu(w.z), qxu ? cvt_hi_xu(w.w) : cvt_hi_alu(w.w));
    }
}

// ---- u8 offset-binary I/Q -> float, eight complex samples per 128-bit word (bytes I0 Q0 I1 Q1 per 32 bits) ----
// The value is b - 128 (gr-simplefe/lib/source_c_impl.cc:126-129; the 1/127 lives in the taps), exact in float.
// CONV 10: exponent trick: bits 0x4B00_0000 | b are the float 8388608 + b, a packed FADD2 subtracts 8388736;
// 11: xor 0x80 makes the byte a signed b - 128, I2F.S8 with a byte selector converts it (XU pipe);
// 12: I through the XU, Q through the exponent trick.
__device__ __forceinline__ float u8_magic(uint32_t w, int byte)
{
    const uint32_t K = 0x4B000000u;
    return __uint_as_float(__byte_perm(w, K, 0x7650 + byte));      // (w.byte, K.b1, K.b2, K.b3)
}
__device__ __forceinline__ float s8_xu(uint32_t wx, int byte)     // wx = w ^ 0x80808080
{
    return (float)(signed char)((wx >> (8 * byte)) & 0xffu);
}
template <int CONV>
__device__ __forceinline__ void convert_u8_pair(uint32_t wa, uint32_t wb, f32x2 &iA, f32x2 &qA, f32x2 &iB, f32x2 &qB)
{
    // wa holds samples (0,1), wb samples (2,3) of a group of four
    const f32x2 off = pack2(-8388736.0f, -8388736.0f);
    const bool ixu = CONV == 11 || CONV == 12, qxu = CONV == 11;
    const uint32_t xa = wa ^ 0x80808080u, xb = wb ^ 0x80808080u;
    iA = ixu ? pack2(s8_xu(xa, 0), s8_xu(xa, 2)) : fadd2(pack2(u8_magic(wa, 0), u8_magic(wa, 2)), off);
    iB = ixu ? pack2(s8_xu(xb, 0), s8_xu(xb, 2)) : fadd2(pack2(u8_magic(wb, 0), u8_magic(wb, 2)), off);
    qA = qxu ? pack2(s8_xu(xa, 1), s8_xu(xa, 3)) : fadd2(pack2(u8_magic(wa, 1), u8_magic(wa, 3)), off);
    qB = qxu ? pack2(s8_xu(xb, 1), s8_xu(xb, 3)) : fadd2(pack2(u8_magic(wb, 1), u8_magic(wb, 3)), off);
}

__host__ __device__ constexpr int conv_spc(int conv) { return conv >= 10 ? 8 : 4; }   // samples per 16-byte chunk

// one chunk (16 bytes of this lane's stream: 4 int16 or 8 u8 complex samples) into the NA window slots.
// Accumulators hold (even, odd) partial sums: a pair of consecutive samples meets the pair of consecutive taps
// in one FFMA2.  TOFF = byte offset of this chunk's taps from the slot's tap pointer.
template <int NA, int CONV, int TOFF>
__device__ __forceinline__ void chunk(uint32_t addr, f32x2 (&accI)[NA], f32x2 (&accQ)[NA], const uint32_t (&tp)[NA])
{
    const uint4 w = lds_u4(addr);
    if (conv_spc(CONV) == 4) {
        f32x2 i01, i23, q01, q23;
        convert4<CONV>(w, i01, i23, q01, q23);
#pragma unroll
        for (int j = 0; j < NA; j++) {
            f32x2 t01, t23;
            lds_taps<TOFF>(tp[j], t01, t23);
            accI[j] = ffma2(i01, t01, accI[j]);
            accQ[j] = ffma2(q01, t01, accQ[j]);
            accI[j] = ffma2(i23, t23, accI[j]);
            accQ[j] = ffma2(q23, t23, accQ[j]);
        }
    } else {
        f32x2 i01, i23, i45, i67, q01, q23, q45, q67;
        convert_u8_pair<CONV>(w.x, w.y, i01, q01, i23, q23);
        convert_u8_pair<CONV>(w.z, w.w, i45, q45, i67, q67);
#pragma unroll
        for (int j = 0; j < NA; j++) {
            f32x2 t01, t23, t45, t67;
            lds_taps<TOFF>(tp[j], t01, t23);
            lds_taps<TOFF + 16>(tp[j], t45, t67);
            accI[j] = ffma2(i01, t01, accI[j]);
            accQ[j] = ffma2(q01, t01, accQ[j]);
            accI[j] = ffma2(i23, t23, accI[j]);
            accQ[j] = ffma2(q23, t23, accQ[j]);
            accI[j] = ffma2(i45, t45, accI[j]);
            accQ[j] = ffma2(q45, t45, accQ[j]);
            accI[j] = ffma2(i67, t67, accI[j]);
            accQ[j] = ffma2(q67, t67, accQ[j]);
        }
    }
}

// one whole box (8 chunks) of this lane's 128-byte line.  ax carries the SWIZZLE_128B term (line & 7) << 4 in
// bits 4-6: chunk c sits at ax ^ (c << 4).  Every window slot advances by 8 chunks of taps.
template <int NA, int CONV>
__device__ __forceinline__ void run_box(uint32_t ax, f32x2 (&accI)[NA], f32x2 (&accQ)[NA], uint32_t (&tp)[NA])
{
    constexpr int TB = conv_spc(CONV) * 4;     // tap bytes per chunk
    chunk<NA, CONV, 0 * TB>(ax, accI, accQ, tp);
    chunk<NA, CONV, 1 * TB>(ax ^ 0x10u, accI, accQ, tp);
    chunk<NA, CONV, 2 * TB>(ax ^ 0x20u, accI, accQ, tp);
    chunk<NA, CONV, 3 * TB>(ax ^ 0x30u, accI, accQ, tp);
    chunk<NA, CONV, 4 * TB>(ax ^ 0x40u, accI, accQ, tp);
    chunk<NA, CONV, 5 * TB>(ax ^ 0x50u, accI, accQ, tp);
    chunk<NA, CONV, 6 * TB>(ax ^ 0x60u, accI, accQ, tp);
    chunk<NA, CONV, 7 * TB>(ax ^ 0x70u, accI, accQ, tp);
#pragma unroll
    for (int j = 0; j < NA; j++) tp[j] += 8 * TB;
}

struct OutInfo {
    int E;     // chunk (16 bytes, counted from sample -H) that holds the newest sample of the window
    int row;   // row of the bank table
};

// output at decimated index M, phase ph
template <int SPC>
__device__ __forceinline__ OutInfo out_info(const BankParams &p, long M, int ph)
{
    const long rh = p.n_first + (long)p.D * M + p.H;      // newest sample, counted from sample -H
    OutInfo o;
    o.E = (int)(rh / SPC);
    o.row = p.rows_per_phase > 1 ? ph * SPC + (int)(rh % SPC) : ph;
    return o;
}
__device__ __forceinline__ void split_pos(const BankParams &p, long k, long &M, int &ph)
{
    const long pos = p.pos0 + k * p.step;
    M = pos / p.L;
    ph = (int)(pos - M * p.L);
    if (ph < 0) { ph += p.L; M -= 1; }                    // floor division: pos = -1 is a held-over output
}

// ceil(a / b) for b > 0 and any sign of a
template <class T> __host__ __device__ __forceinline__ T cdiv_t(T a, T b) { return a >= 0 ? (a + b - 1) / b : -((-a) / b); }

// first output index k in [0, n_out] whose newest sample, counted from sample -H, is >= X
// (rh_k = n_first + D floor(pos_k / L) + H with pos_k = pos0 + k step: monotonic in k).  T = int when the host has
// checked that every position of the call fits 31 bits (p.fits32): 64-bit division is a software routine of ~80
// instructions, and every warp does seven of them before it can issue its first load.
template <class T> __host__ __device__ __forceinline__ long bank_first_k(const BankParams &p, long X)
{
    const T Mmin = cdiv_t<T>((T)(X - p.n_first - p.H), (T)p.D);
    T k = cdiv_t<T>((T)p.L * Mmin - (T)p.pos0, (T)p.step);
    if (k < 0) k = 0;
    if (k > (T)p.n_out) k = (T)p.n_out;
    return (long)k;
}

// ring slots [lo, hi) (SB boxes each, absolute box grid from sample -H) that hold every window of the call
struct BankSpan {
    int lo, hi;
};
template <int SPC, int SB, class T> __host__ __device__ __forceinline__ BankSpan bank_span(const BankParams &p)
{
    T pos = (T)p.pos0, M = pos / (T)p.L;
    if (pos - M * (T)p.L < 0) M -= 1;
    const T rh0 = (T)p.n_first + (T)p.D * M + (T)p.H;
    pos = (T)p.pos0 + (T)(p.n_out - 1) * (T)p.step;
    M = pos / (T)p.L;
    if (pos - M * (T)p.L < 0) M -= 1;
    const T rh1 = (T)p.n_first + (T)p.D * M + (T)p.H;
    const T q0 = rh0 / SPC - p.nc + 1, q1 = rh1 / SPC;             // first and last chunk read
    BankSpan s;
    s.lo = (int)((q0 >> 3) / SB);
    s.hi = (int)((q1 >> 3) / SB) + 1;
    return s;
}

template <int WARPS, int NST, int CONV, int NLIVE, int SB = SLOT_BOXES>   // SB: boxes per ring slot (2: one 3-D box, 1: one 2-D box)
__global__ void __launch_bounds__(WARPS * 32, 1)
fused_bank_kernel(const __grid_constant__ CUtensorMap map_cur, const __grid_constant__ CUtensorMap map_hist,
                  const __grid_constant__ CUtensorMap map_cur3, const __grid_constant__ CUtensorMap map_hist3, const BankParams p)
{
    constexpr int SLOT_B = SB * BOX_BYTES;                           // bytes per ring slot
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[WARPS * NST];
    __shared__ __align__(8) uint64_t table_bar;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t *sbase = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);   // SWIZZLE_128B boxes: 1024 B alignment
    float *table = reinterpret_cast<float *>(sbase + (size_t)WARPS * NST * SLOT_B);
    const uint32_t ring_a = smem_u32(sbase) + (uint32_t)warp * NST * SLOT_B;
    const uint32_t table_a = smem_u32(table);

    // ---- one-time: barriers; bank table -> shared memory by bulk async copies (one thread issues them, their completion
    // is awaited only before the first box is processed, so the copy overlaps the first input boxes' flight) -----------
    const uint32_t tbar = smem_u32(&table_bar);
    const uint32_t table_bytes = (uint32_t)(p.n_rows * p.row_len) * 4u;      // rows are whole 16-byte chunks
    if (threadIdx.x == 0) {
        for (int s = 0; s < WARPS * NST; s++) mbar_init(smem_u32(&bars[s]), 1);
        mbar_init(tbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(tbar, table_bytes);
        for (uint32_t off = 0; off < table_bytes; off += 32768u) {
            const uint32_t n = table_bytes - off < 32768u ? table_bytes - off : 32768u;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(table_a + off), "l"(reinterpret_cast<const char *>(p.table) + off), "r"(n), "r"(tbar) : "memory");
        }
    }
    __syncthreads();

    // ---- state carry, folded into this kernel: hist_new[s] <- the last H samples of (hist_old[s] ++ cur[s][0..n)) ----
    // A few elements per thread, spread over the whole grid; it needs nothing this kernel computes and writes a buffer
    // this kernel never reads, so a call is ONE launch (the separate save_tail kernel remains for calls without outputs).
    // Every thread runs it once: behind its warp's first TMA issues (the copies' latency then overlaps the boxes' flight),
    // or on its way out when the warp has no work item.
    auto carry = [&]() {
        asm volatile("griddepcontrol.wait;" ::: "memory");           // (a no-op once passed) the target was the previous call's hist_old
        if (p.hist_new == nullptr) return;
        const unsigned total = (unsigned)p.n_streams * (unsigned)p.H, stride = gridDim.x * (WARPS * 32), H = (unsigned)p.H;   // (checked on the host: fits)
        const long n_h = p.in.n - p.H;
        if (conv_spc(CONV) == 4) {
            const uint32_t *cur = static_cast<const uint32_t *>(p.in.cur), *ho = static_cast<const uint32_t *>(p.in.hist);
            uint32_t *hn = static_cast<uint32_t *>(p.hist_new);
            for (unsigned e = blockIdx.x * (WARPS * 32) + threadIdx.x; e < total; e += stride) {
                const unsigned srow = e / H, i = e - srow * H;
                const long src = n_h + i;
                hn[e] = src >= 0 ? __ldg(cur + (long)srow * p.in.pitch + src) : __ldg(ho + (long)srow * p.H + (p.H + src));
            }
        } else {
            const uint16_t *cur = static_cast<const uint16_t *>(p.in.cur), *ho = static_cast<const uint16_t *>(p.in.hist);
            uint16_t *hn = static_cast<uint16_t *>(p.hist_new);
            for (unsigned e = blockIdx.x * (WARPS * 32) + threadIdx.x; e < total; e += stride) {
                const unsigned srow = e / H, i = e - srow * H;
                const long src = n_h + i;
                hn[e] = src >= 0 ? __ldg(cur + (long)srow * p.in.pitch + src) : __ldg(ho + (long)srow * p.H + (p.H + src));
            }
        }
    };

    // ---- this warp's work item ---------------------------------------------------------------------------
    // An item is a group of 32 streams and a contiguous range of BOXES (32-sample time blocks), not of outputs: every
    // box of the call is read by exactly one warp.  A window that straddles the seam between two items is accumulated
    // in two parts -- the older samples by the item before the seam, the newer ones by the item behind it -- and
    // whichever warp finishes second adds the two (older + newer, always in that order) and writes the output.  With
    // items cut by outputs every item re-read a window length of lead-in: 2 % of the traffic at 1024 streams, but
    // 15 % at 128 (the per-GPU shard of the 8-GPU split), where an item is only 17 outputs long.
    // (a warp without work still waits for the table copy: a CTA must not retire with a bulk copy into its shared memory in flight)
    const long item = (long)blockIdx.x * WARPS + warp;
    if (item >= (long)p.n_groups * p.n_seg) { carry(); mbar_wait(tbar, 0); return; }
    const int g = (int)(item % p.n_groups), seg = (int)(item / p.n_groups);
    constexpr int SPC = conv_spc(CONV);             // samples per 16-byte chunk
    const int NC = p.nc;
    const int row_bytes = p.row_len * 4;

    const BankSpan sp = p.fits32 ? bank_span<SPC, SB, int>(p) : bank_span<SPC, SB, long>(p);   // slots [sp.lo, sp.hi) hold every window of the call
    const int sA = sp.lo + (int)((long)(sp.hi - sp.lo) * seg / p.n_seg), sB = sp.lo + (int)((long)(sp.hi - sp.lo) * (seg + 1) / p.n_seg);
    if (sA >= sB) { carry(); mbar_wait(tbar, 0); return; }
    const int box0 = sA * SB, box_end = sB * SB, n_boxes = box_end - box0;
    // outputs this item touches: [k_lo, k_end) -- the first window that ends at or behind box0, up to the first window that
    // opens at or behind box_end; the first n_inh of them opened before box0 (inherited from the item before)
    const long k_lo = p.fits32 ? bank_first_k<int>(p, (long)SPC * 8 * box0) : bank_first_k<long>(p, (long)SPC * 8 * box0);
    const long k_end = p.fits32 ? bank_first_k<int>(p, (long)SPC * (8L * box_end + NC - 1)) : bank_first_k<long>(p, (long)SPC * (8L * box_end + NC - 1));
    if (k_lo >= k_end) { carry(); mbar_wait(tbar, 0); return; }   // no window meets these boxes (cannot happen between two outputs of a call)

    long M_next;
    int ph_next;
    split_pos(p, k_lo, M_next, ph_next);
    const int slot0 = sA, n_slots = sB - sA;
    const int hist_boxes = p.H / (8 * SPC);                        // a box = 128 bytes = 8 chunks per row
    const int full_tb = (int)(p.in.n / (8 * SPC));                 // whole 128-byte time blocks of the current rows
    const uint32_t bar0 = smem_u32(&bars[warp * NST]);

    // A slot whose two boxes are whole blocks of the same buffer comes in as ONE 3-D box (32 samples, 2 blocks, 32 rows):
    // each row contributes 256 contiguous bytes, which is what lets HBM run at full rate (128-byte pieces cap it near
    // 5.3 TB/s, tests/probes/tma_box_bw.cu).  Edge slots (history/current seam, ragged end) use two 2-D boxes, whose
    // out-of-range samples the TMA zero-fills.
    auto slot_3d = [&](int a) -> bool {
        const int b0 = 2 * a, b1 = 2 * a + 1;
        return SB == 2 && (b1 < hist_boxes || (b0 >= hist_boxes && b1 - hist_boxes < full_tb));
    };
    auto issue = [&](int a, int rs) {                              // lane 0 only; a = absolute slot, rs = ring slot
        const uint32_t bar = bar0 + 8 * rs;
        mbar_expect_tx(bar, SLOT_B);
        const uint32_t dst = ring_a + rs * SLOT_B;
        const int b0 = SB * a;                                     // box b holds samples [32 b - H, 32 b - H + 32)
        if (slot_3d(a)) {
            if (b0 < hist_boxes) tma_box3(dst, &map_hist3, b0, g * BOX_ROWS, bar);
            else tma_box3(dst, &map_cur3, b0 - hist_boxes, g * BOX_ROWS, bar);
        } else {
#pragma unroll
            for (int t = 0; t < SB; t++) {
                const int box = b0 + t;
                if (box < hist_boxes) tma_box(dst + t * BOX_BYTES, &map_hist, box * BOX_SAMPLES, g * BOX_ROWS, bar);
                else tma_box(dst + t * BOX_BYTES, &map_cur, (box - hist_boxes) * BOX_SAMPLES, g * BOX_ROWS, bar);
            }
        }
    };
    // (This grid may have been launched as a programmatic dependent of the previous call's kernel, launch_bank_rows: the
    // barriers and the table copy above then ran beside that call's last CTAs; griddepcontrol.wait above restored the
    // stream order before anything earlier work produced is read -- history, samples -- or anything it may still use is
    // written -- outputs, seam buffers, the new history.)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (lane == 0)
        for (int b = 0; b < NST && b < n_slots; b++) issue(slot0 + b, b);
    carry();
    // The next call's grid may start its prologue as soon as every CTA of this one has got here (it then waits, as this
    // one just did, for this grid to complete before it reads or writes anything this grid touches).
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    // ---- window slots (slot 0 closes first) ------------------------------------------------------------
    // Every table row carries PAD_CHUNKS zero chunks in front of and behind its taps, so a window may start at
    // the beginning of the box that holds its first sample and run to the end of the box that holds its last
    // one: all bookkeeping happens between boxes and every box is the same straight-line code over NLIVE slots.
    // A slot without a window rests on the zero pad of row 0.
    f32x2 accI[NLIVE], accQ[NLIVE];
    uint32_t tp[NLIVE];
    int Eb[NLIVE];                                                 // box in which the slot's window ends
#pragma unroll
    for (int j = 0; j < NLIVE; j++) {
        accI[j] = accQ[j] = 0ull;
        tp[j] = table_a;
        Eb[j] = 0x3fffffff;
    }
    int nact = 0;
    long k_next = k_lo, k_head = k_lo;
    OutInfo nxt = out_info<SPC>(p, M_next, ph_next);
    int open_box = (nxt.E - NC + 1) >> 3;
    const int sM = (int)(p.step / p.L), sP = (int)(p.step % p.L);   // schedule stepping without division
    auto advance = [&]() {                                           // nxt <- the output after k_next
        k_next++;
        if (k_next < k_end) {
            ph_next += sP;
            M_next += sM;
            if (ph_next >= p.L) { ph_next -= p.L; M_next += 1; }
            nxt = out_info<SPC>(p, M_next, ph_next);
            open_box = (nxt.E - NC + 1) >> 3;
        } else {
            open_box = 0x3fffffff;
        }
    };
    // windows inherited from the item before: they opened in an earlier box, so their taps start (box0 - opening box)
    // boxes into the row; the item before has accumulated the samples up to box0
    int n_inh = 0;
    while (open_box < box0) {
        const int S = nxt.E - NC + 1;
        const uint32_t t0 = table_a + (uint32_t)(nxt.row * row_bytes) + (uint32_t)(PAD_CHUNKS - (S & 7)) * (uint32_t)(SPC * 4) +
                            (uint32_t)(box0 - open_box) * (uint32_t)(8 * SPC * 4);
#pragma unroll
        for (int j = 0; j < NLIVE; j++)
            if (j == nact) {
                tp[j] = t0;
                Eb[j] = nxt.E >> 3;
            }
        nact++;
        n_inh++;
        advance();
    }
    const int n_inh0 = n_inh;
    const int s_row = g * BOX_ROWS + lane;
    float2 *orow = p.out + (long)s_row * p.out_pitch;
    const bool s_ok = s_row < p.n_streams;
    // seam exchange: parts[(seam, side, slot, lane)], seam = g * n_seg + (index of the item behind the seam)
    auto part_at = [&](int seam_seg, int side, int j) -> float2 * {
        return p.parts + ((((long)g * p.n_seg + seam_seg) * 2 + side) * A_MAX + j) * 32 + lane;
    };
    // both parts of the n outputs from k0 are (or are about to be) in memory: the second warp to get here adds them
    auto seam_arrive = [&](int seam_seg, int n, long k0) {
        __threadfence();
        __syncwarp();
        int old = 0;
        if (lane == 0) old = atomicAdd(p.seam_cnt + (long)g * p.n_seg + seam_seg, 1);
        old = __shfl_sync(0xffffffffu, old, 0);
        if (old == 0) return;
        __threadfence();
        for (int j = 0; j < n; j++) {
            const float2 a = __ldcg(part_at(seam_seg, 0, j)), b = __ldcg(part_at(seam_seg, 1, j));
            if (s_ok) orow[k0 + j] = make_float2(a.x + b.x, a.y + b.y);      // older samples + newer samples
        }
        if (lane == 0) p.seam_cnt[(long)g * p.n_seg + seam_seg] = 0;          // ready for the next launch
    };

    mbar_wait(tbar, 0);                                              // the bank table has landed
    int slot = 0;
    uint32_t phase = 0;
    for (int bi = 0; bi < n_boxes; bi++) {
        const int B = box0 + bi, a = B / SB, tb = B % SB;
        if (tb == 0) mbar_wait(bar0 + 8 * slot, phase);              // first box of a slot (items start on slot boundaries)
        if (B == open_box) {                                         // a window opens in this box (at most one)
            const int S = nxt.E - NC + 1;                            // its first chunk
            const uint32_t t0 = table_a + (uint32_t)(nxt.row * row_bytes) + (uint32_t)(PAD_CHUNKS - (S & 7)) * (uint32_t)(SPC * 4);
#pragma unroll
            for (int j = 0; j < NLIVE; j++)
                if (j == nact) {
                    accI[j] = accQ[j] = 0ull;
                    tp[j] = t0;
                    Eb[j] = nxt.E >> 3;
                }
            nact++;
            advance();
        }
        // this lane's 128-byte line of the box: 3-D slot: line 2 lane + tb; 2-D boxes: box tb, line lane
        const uint32_t line = slot_3d(a) ? (uint32_t)(2 * lane + tb) : (uint32_t)(tb * BOX_ROWS + lane);
        const uint32_t ax = ring_a + slot * SLOT_B + line * 128u + ((line & 7u) << 4);
        run_box<NLIVE, CONV>(ax, accI, accQ, tp);
        if (B == Eb[0]) {                                            // the oldest window is complete
            const float2 yi = unpack2(accI[0]), yq = unpack2(accQ[0]);
            const float2 y = make_float2(yi.x + yi.y, yq.x + yq.y);
            if (n_inh > 0) {                                         // ... its newer part, that is: leave it at the seam
                *part_at(seg, 1, n_inh0 - n_inh) = y;
                if (--n_inh == 0) seam_arrive(seg, n_inh0, k_lo);
            } else if (s_ok) {
                orow[k_head] = y;
            }
            k_head++;
#pragma unroll
            for (int j = 0; j + 1 < NLIVE; j++) {
                accI[j] = accI[j + 1];
                accQ[j] = accQ[j + 1];
                tp[j] = tp[j + 1];
                Eb[j] = Eb[j + 1];
            }
            Eb[NLIVE - 1] = 0x3fffffff;
            nact--;
        }
#pragma unroll
        for (int j = 0; j < NLIVE; j++)
            if (j >= nact) tp[j] = table_a;                          // idle slots stay on the zero pad
        if (tb == SB - 1) {                                          // slot consumed: refill it, move on
            __syncwarp();
            if (lane == 0 && (a - slot0) + NST < n_slots) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // our reads before the next TMA write
                issue(a + NST, slot);
            }
            if (++slot == NST) { slot = 0; phase ^= 1; }
        }
    }
    // windows still open continue in the next item: leave their older parts at the seam behind this item
    if (nact > 0) {
#pragma unroll
        for (int j = 0; j < NLIVE; j++)
            if (j < nact) {
                const float2 yi = unpack2(accI[j]), yq = unpack2(accQ[j]);
                *part_at(seg + 1, 0, j) = make_float2(yi.x + yi.y, yq.x + yq.y);
            }
        seam_arrive(seg + 1, nact, k_head);
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

// 3-D view of the same rows: (32 samples, time blocks of 128 B, rows); box = (32, SLOT_BOXES, 32 rows)
bool encode_rows3(CUtensorMap *m, const void *base, long n_blocks, long n_rows, long pitch)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[3] = {BOX_SAMPLES, (cuuint64_t)(n_blocks > 0 ? n_blocks : 1), (cuuint64_t)n_rows};
    const cuuint64_t strides[2] = {BOX_SAMPLES * 4, (cuuint64_t)pitch * 4};
    const cuuint32_t box[3] = {BOX_SAMPLES, SLOT_BOXES, BOX_ROWS};
    const cuuint32_t es[3] = {1, 1, 1};
    const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<void *>(base), dims, strides, box, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("fused bank: cuTensorMapEncodeTiled (3-D) failed with CUresult " + std::to_string((int)r));
        return false;
    }
    return true;
}

// ---- launch variants: warps per CTA, ring slots per warp, conversion flavour ----------------------------------
struct BankVariant {
    int warps, nst, conv, sb;      // conv: int16 flavour (0-3) or u8 flavour (10-12); sb: boxes per ring slot
};
typedef void (*BankKernel)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const BankParams);

// int16 rows: 8 warps x 2 slots of one 3-D box (256 contiguous bytes per row keep HBM at full rate; the kernel is HBM bound).
// u8 rows carry twice the samples per byte, so the kernel is bound by its two warps per scheduler instead: 16 warps with
// one-box slots (the same 128 KB of ring) give the schedulers four.
BankVariant bank_variant(int fmt)
{
    static const BankVariant v16 = [] {
        BankVariant d{8, 2, 2, 2};
        if (const char *e = getenv("SFE_BANK_VARIANT")) {   // tuning hook: "warps,slots,conv"
            int w = 0, n = 0, c = 0;
            if (sscanf(e, "%d,%d,%d", &w, &n, &c) >= 3) d = BankVariant{w, n, c, 2};
        }
        return d;
    }();
    static const BankVariant v8 = [] {
        BankVariant d{16, 2, 12, 1};
        if (const char *e = getenv("SFE_BANK_VARIANT_U8")) {   // tuning hook: "warps,conv,boxes per slot"
            int w = 0, c = 0, b = 0;
            if (sscanf(e, "%d,%d,%d", &w, &c, &b) >= 3) d = BankVariant{w, 2, c, b};
        }
        return d;
    }();
    return fmt == SFE_IN_U8 ? v8 : v16;
}

template <int W, int N, int C, int SB = SLOT_BOXES>
BankKernel bank_pick_live(int live)
{
    switch (live) {
        case 1: return fused_bank_kernel<W, N, C, 1, SB>;
        case 2: return fused_bank_kernel<W, N, C, 2, SB>;
        case 3: return fused_bank_kernel<W, N, C, 3, SB>;
        case 4: return fused_bank_kernel<W, N, C, 4, SB>;
        default: return nullptr;
    }
}

// The shipped library carries the two flavours that won the sweeps (profiles/r01_bank_v3_variant_sweep.txt,
// r01_bank_u8_variant_sweep.txt); the others exist only in a -DSFE_BANK_SWEEP build (SFE_BANK_SWEEP=1 python -m
// simplefe_b200.build, scripts/bank_sweep.sh), where SFE_BANK_VARIANT* picks one.
BankKernel bank_pick(const BankVariant &v, int live, int fmt)
{
    if (fmt == SFE_IN_U8) {
        if (v.nst != 2) return nullptr;
        switch (v.warps * 1000 + v.conv * 10 + v.sb) {
            case 16121: return bank_pick_live<16, 2, 12, 1>(live);
#ifdef SFE_BANK_SWEEP
            case 8102: return bank_pick_live<8, 2, 10>(live);
            case 8122: return bank_pick_live<8, 2, 12>(live);
            case 12121: return bank_pick_live<12, 2, 12, 1>(live);
            case 16101: return bank_pick_live<16, 2, 10, 1>(live);
#endif
            default: return nullptr;
        }
    }
    if (v.sb != 2) return nullptr;
    const int key = v.warps * 100 + v.nst * 10 + v.conv;
    switch (key) {
        case 822: return bank_pick_live<8, 2, 2>(live);
#ifdef SFE_BANK_SWEEP
        case 820: return bank_pick_live<8, 2, 0>(live);
        case 1020: return bank_pick_live<10, 2, 0>(live);
        case 1021: return bank_pick_live<10, 2, 1>(live);
        case 1022: return bank_pick_live<10, 2, 2>(live);
        case 1023: return bank_pick_live<10, 2, 3>(live);
#endif
        default: return nullptr;
    }
}

}  // namespace

size_t bank_smem_bytes(int n_rows, int row_len, int fmt)
{
    const BankVariant v = bank_variant(fmt);
    return (size_t)v.warps * v.nst * v.sb * BOX_BYTES + (size_t)n_rows * row_len * sizeof(float) + 1024;
}

int bank_max_live() { return A_MAX; }
int bank_nc(int Tc, int spc) { return (Tc + 2 * spc - 2) / spc; }
int bank_row_len(int Tc, int spc) { return spc * (bank_nc(Tc, spc) + 2 * PAD_CHUNKS); }

// A tensor map depends only on (base, extents, pitch): streaming callers present the same few buffers call after call,
// and an encode costs about as much as a small launch, so the descriptors are memoised (bounded, per process).
struct MapKey {
    const void *base;
    long a, rows, pitch;
    int dims;
    bool operator<(const MapKey &o) const
    {
        return std::tie(base, a, rows, pitch, dims) < std::tie(o.base, o.a, o.rows, o.pitch, o.dims);
    }
};
static bool cached_map(CUtensorMap *m, int dims, const void *base, long a, long rows, long pitch)
{
    static std::mutex mu;
    static std::map<MapKey, CUtensorMap> cache;
    const MapKey k{base, a, rows, pitch, dims};
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(k);
        if (it != cache.end()) {
            *m = it->second;
            return true;
        }
    }
    if (!(dims == 2 ? encode_rows(m, base, a, rows, pitch) : encode_rows3(m, base, a, rows, pitch))) return false;
    std::lock_guard<std::mutex> lk(mu);
    if (cache.size() > 256) cache.clear();
    cache[k] = *m;
    return true;
}

// Seam exchange buffers (parts of the outputs whose window straddles two work items, arrival counters), one set per
// (device, stream): launches on one stream are serialised, sfe_chain_process_host's two streams run side by side.
// Grown on demand, kept for the life of the process (2 KB per work item: ~2.5 MB for a full machine).
static bool seam_scratch(int dev, cudaStream_t st, size_t items, float2 **parts, int **cnt)
{
    struct Buf {
        float2 *parts = nullptr;
        int *cnt = nullptr;
        size_t items = 0;
    };
    static std::mutex mu;
    static std::map<std::pair<int, cudaStream_t>, Buf> bufs;
    std::lock_guard<std::mutex> lk(mu);
    Buf &b = bufs[{dev, st}];
    if (b.items < items) {
        const size_t want = std::max<size_t>(items, 2048);
        cudaFree(b.parts);        // synchronises with the kernels that may still use the old buffers
        cudaFree(b.cnt);
        b = Buf{};
        if (!SFE_CUDA(cudaMalloc(&b.parts, want * 2 * A_MAX * 32 * sizeof(float2))) || !SFE_CUDA(cudaMalloc(&b.cnt, want * sizeof(int))) ||
            !SFE_CUDA(cudaMemset(b.cnt, 0, want * sizeof(int))) || !SFE_CUDA(cudaDeviceSynchronize()))
            return false;
        b.items = want;
    }
    *parts = b.parts;
    *cnt = b.cnt;
    return true;
}

}  // namespace sfe

// Plain (unswizzled) 2-D map of rows of 32-bit words, box = box_cols x box_rows, out-of-range elements read as zero:
// what the Toeplitz kernel stages its raw int16 IQ rows with.  Memoised like the bank kernel's maps.  `out` is a CUtensorMap.
bool sfe_plain_map_2d(void *out, const void *base, long n_cols, long n_rows, long pitch_words, int box_cols, int box_rows)
{
    using namespace sfe;
    static std::mutex mu;
    static std::map<std::tuple<const void *, long, long, long, int, int>, CUtensorMap> cache;
    const auto key = std::make_tuple(base, n_cols, n_rows, pitch_words, box_cols, box_rows);
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) {
            *static_cast<CUtensorMap *>(out) = it->second;
            return true;
        }
    }
    EncodeTiledFn fn = encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from this driver");
        return false;
    }
    CUtensorMap m;
    const cuuint64_t dims[2] = {(cuuint64_t)n_cols, (cuuint64_t)n_rows};
    const cuuint64_t strides[1] = {(cuuint64_t)pitch_words * 4};
    const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    const cuuint32_t es[2] = {1, 1};
    const CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void *>(base), dims, strides, box, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (plain 2-D) failed with CUresult " + std::to_string((int)r));
        return false;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (cache.size() > 256) cache.clear();
    cache[key] = m;
    *static_cast<CUtensorMap *>(out) = m;
    return true;
}

namespace sfe {

static int launch_bank_rows(const BankParams &p_in, cudaStream_t st)
{
    BankParams p = p_in;
    if (p.n_out <= 0 || p.n_streams <= 0) return SFE_OK;
    const int eb = fmt_bytes(p.in.fmt);                      // bytes per complex sample: 4 (int16) or 2 (u8)
    const int spc = 16 / eb;
    if ((reinterpret_cast<uintptr_t>(p.in.cur) & 15) || ((p.in.pitch * eb) & 15) || (p.H % (8 * spc)) ||
        (p.in.fmt != SFE_IN_I16 && p.in.fmt != SFE_IN_U8)) {
        set_error("fused bank: input rows must be int16 or u8 IQ, 16-byte aligned with a pitch that is a multiple of 16 bytes");
        return SFE_ERR_ARG;
    }
    const BankVariant v = bank_variant(p.in.fmt);
    BankKernel kern = bank_pick(v, p.n_live, p.in.fmt);
    if (!kern) {
        set_error("fused bank: SFE_BANK_VARIANT names a variant that is not compiled in (or n_live is out of range)");
        return SFE_ERR_ARG;
    }
    CUtensorMap mc, mh, mc3, mh3;
    // tensor maps count 32-bit words: one int16 sample or two u8 samples
    const long n_words = (p.in.n * eb + 3) / 4, pitch_words = p.in.pitch * eb / 4, hist_words = (long)p.H * eb / 4;
    if (!cached_map(&mc, 2, p.in.cur, n_words, p.n_streams, pitch_words) ||
        !cached_map(&mc3, 3, p.in.cur, p.in.n * eb / (BOX_SAMPLES * 4), p.n_streams, pitch_words))
        return SFE_ERR_CUDA;
    if (p.H == 0) {                          // no history rows (time-split launches): the maps are never used
        mh = mc;
        mh3 = mc3;
    } else if (!cached_map(&mh, 2, p.in.hist, hist_words, p.n_streams, hist_words) ||
               !cached_map(&mh3, 3, p.in.hist, hist_words / BOX_SAMPLES, p.n_streams, hist_words)) {
        return SFE_ERR_CUDA;
    }
    int dev = 0;
    if (!SFE_CUDA(cudaGetDevice(&dev))) return SFE_ERR_CUDA;
    static std::mutex cache_mutex;
    static std::map<int, int> sm_count;                                   // per device
    static std::map<std::pair<int, const void *>, size_t> smem_set;      // (device, kernel) -> opted-in bytes
    int sms = 0;
    {
        std::lock_guard<std::mutex> lk(cache_mutex);
        auto it = sm_count.find(dev);
        if (it == sm_count.end()) {
            if (!SFE_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev))) return SFE_ERR_CUDA;
            sm_count[dev] = sms;
        } else {
            sms = it->second;
        }
    }
    p.n_groups = (p.n_streams + BOX_ROWS - 1) / BOX_ROWS;
    // One work item per warp slot of the machine (rounded down: one item too many would cost a whole second wave).  Items
    // are ranges of ring slots; a window may straddle ONE seam between items, so an item is at least a window long.
    const BankSpan sp = spc == 4 ? (v.sb == 2 ? bank_span<4, 2, long>(p) : bank_span<4, 1, long>(p)) : (v.sb == 2 ? bank_span<8, 2, long>(p) : bank_span<8, 1, long>(p));
    const long span_slots = sp.hi - sp.lo;
    const long win_boxes = ((p.nc + 7) >> 3) + 1;
    const long min_slots = (win_boxes + v.sb - 1) / v.sb + 1;
    long n_seg = ((long)sms * v.warps) / p.n_groups;
    const long cap = span_slots / min_slots;
    if (n_seg > cap) n_seg = cap;
    if (n_seg < 1) n_seg = 1;
    p.n_seg = (int)n_seg;
    const long items = (long)p.n_groups * p.n_seg;
    if (!seam_scratch(dev, st, (size_t)items, &p.parts, &p.seam_cnt)) return SFE_ERR_CUDA;
    const size_t smem = bank_smem_bytes(p.n_rows, p.row_len, p.in.fmt);
    {
        std::lock_guard<std::mutex> lk(cache_mutex);
        size_t &have = smem_set[{dev, reinterpret_cast<const void *>(kern)}];
        if (have < smem) {
            if (!SFE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))) return SFE_ERR_CUDA;
            have = smem;
        }
    }
    // Programmatic dependent launch: when the kernel before this one on the stream is the previous call's state carry
    // (save_tail, which triggers its dependents on entry), this grid's CTAs start as that call's CTAs retire and run
    // their prologue early; griddepcontrol.wait in the kernel restores the stream order before any dependent access.
    // Behind anything else (a copy, a kernel that never triggers) the launch is an ordinary one.
    static const bool no_pdl = getenv("SFE_NO_PDL") != nullptr;
    {   // every position of the call in 31 bits: the kernel's index arithmetic then uses 32-bit division
        const double top = (double)p.pos0 + (double)p.n_out * (double)p.step + (double)p.L;
        const double rh = (double)p.n_first + (double)p.D * (top / p.L + 1) + p.H + 64.0 * p.nc;
        p.fits32 = (top < 2.0e9 && rh * 8 < 2.0e9 && (double)p.L * (rh / p.D + 2) < 2.0e9 && p.pos0 > -2.0e9) ? 1 : 0;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((items + v.warps - 1) / v.warps));
    cfg.blockDim = dim3(v.warps * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = no_pdl ? 0 : 1;
    if (!SFE_CUDA(cudaLaunchKernelEx(&cfg, kern, mc, mh, mc3, mh3, p))) return SFE_ERR_CUDA;
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

// A single stream (or a few) leaves 31 of the 32 lanes of every warp idle.  The bank is periodically time invariant, so
// a long row can be cut into P pieces of Kp outputs each (Kp a multiple of the phase period L / gcd(step, L)) that all
// follow the schedule of piece 0 shifted by Q = D Kp step / L samples: pieces 1..P-1 become the rows of a second launch
// whose "current" samples start H samples before the piece (their history is simply the preceding part of the row),
// and the outputs behind the last whole piece form a third, one-row launch.  Results are those of the unsplit launch
// bit for bit (same windows, same taps, same summation order).
static long gcd_l(long a, long b) { return b ? gcd_l(b, a % b) : a; }

int launch_bank(const BankParams &p, cudaStream_t st)
{
    const int eb = fmt_bytes(p.in.fmt), spc = 16 / (eb > 0 ? eb : 4);
    static const bool no_split = getenv("SFE_BANK_NO_SPLIT") != nullptr;
    if (p.n_streams != 1 || p.n_out < 2048 || no_split || p.step <= 0) return launch_bank_rows(p, st);
    const long sl = ((p.step % p.L) + p.L) % p.L;
    const long period = p.L / gcd_l(sl ? sl : p.L, p.L);
    // outputs per piece: whole periods, at least 16 outputs, and a piece length that keeps rows 16-byte aligned
    long Kp = 0, Q = 0;
    for (long m = (16 + period - 1) / period; m <= (16 + period - 1) / period + 4 * spc; m++) {
        const long k = m * period;
        const long q = (long)p.D * (k * p.step / p.L);
        if ((k * p.step) % p.L == 0 && q % spc == 0 && q >= p.H) { Kp = k; Q = q; break; }
    }
    if (!Kp) return launch_bank_rows(p, st);
    // samples a piece may touch behind its own Q: the offset of output 0 plus one window step
    long M0 = p.pos0 / p.L;
    if (p.pos0 - M0 * p.L < 0) M0 -= 1;
    const long extra = std::max<long>(0, p.n_first + (long)p.D * M0) + (long)p.D * (p.step / p.L + 2) + 8 * spc;
    long P = p.n_out / Kp;
    while (P > 1 && P * Q + extra > p.in.n) P--;
    if (P < 64) return launch_bank_rows(p, st);

    BankParams a = p;                        // piece 0: the real history
    a.n_out = Kp;
    int rc = launch_bank_rows(a, st);
    if (rc != SFE_OK) return rc;
    BankParams b = p;                        // pieces 1..P-1 as rows
    b.hist_new = nullptr;                    // (the state carry rides on piece 0, which sees the call's real row and history)
    b.in.cur = static_cast<const char *>(p.in.cur) + (Q - p.H) * eb;
    b.in.hist = b.in.cur;
    b.in.pitch = Q;
    b.in.H = 0;
    b.in.n = Q + p.H + extra;
    b.H = 0;
    b.n_first = p.n_first + p.H;
    b.n_streams = (int)(P - 1);
    b.n_out = Kp;
    b.out = p.out + Kp;
    b.out_pitch = Kp;
    rc = launch_bank_rows(b, st);
    if (rc != SFE_OK) return rc;
    if (p.n_out > P * Kp) {                  // the ragged end: one row from the end of the last whole piece
        BankParams c = b;
        c.in.cur = static_cast<const char *>(p.in.cur) + (P * Q - p.H) * eb;
        c.in.hist = c.in.cur;
        c.in.pitch = p.in.n - P * Q + p.H;
        c.in.n = p.in.n - P * Q + p.H;
        c.n_streams = 1;
        c.n_out = p.n_out - P * Kp;
        c.out = p.out + P * Kp;
        c.out_pitch = c.n_out;
        // a one-row tensor map still wants a 16-byte pitch
        c.in.pitch = (c.in.pitch + spc - 1) / spc * spc;
        rc = launch_bank_rows(c, st);
    }
    return rc;
}

// Host side: the bank table.  c_ph[u] = sum_j r[ph + L j] g[u - D j] in double, scaled, rounded to float, stored
// time-reversed and chunk aligned (spc samples per 16-byte chunk): row (ph, e) holds, behind PAD_CHUNKS zero
// chunks, for chunk-relative sample i the tap u = spc NC - spc + e - i, where e = (newest sample + H) mod spc.
void bank_table(const std::vector<double> &g, int D, const std::vector<float> &r, int L, double scale, int e_only, int spc,
                std::vector<float> &rows, int *row_len, int *n_rows, int *Tc_out)
{
    const int J = ((int)r.size() + L - 1) / L;
    const int Tc = D * (J - 1) + (int)g.size();
    const int NC = bank_nc(Tc, spc);
    const int RL = bank_row_len(Tc, spc);         // zero pads: windows open and close on box boundaries (kernel)
    const int per = e_only >= 0 ? 1 : spc;
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
            for (int i = 0; i < spc * NC; i++) {
                const int u = spc * NC - spc + e - i;
                if (u >= 0 && u < Tc) row[spc * PAD_CHUNKS + i] = (float)(c[u] * scale);
            }
        }
    }
    *row_len = RL;
    *n_rows = L * per;
    *Tc_out = Tc;
}

}  // namespace sfe
