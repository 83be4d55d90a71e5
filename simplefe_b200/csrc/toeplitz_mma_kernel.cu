// toeplitz_mma_kernel.cu -- K2t: decimate-by-8 FIR over many int16 IQ streams that share one filter,
// as a Toeplitz-batched dense contraction on the 5th-generation tensor cores (tcgen05, int8).
//
//   y[s][k] = sum_j g[j] * x[s][n_first + 8k - j]           (g: up to 641 taps; x: int16 I/Q)
//
// This is the reference's blkconv -> decimate pair for the many-streams case (blkconv.cxx:77-110 and
// decimate.cxx:102-140 with rate 8, upsample 1), with the two FIRs pre-convolved into g (chain.cu).
//
// Exact integer formulation.  Every int16 sample is the byte pair (lo unsigned, hi signed): s = 256*hi + lo.
// The hi byte is biased to unsigned (hi ^ 0x80 = hi + 128) so that all four byte planes of an IQ stream
// (I_lo, I_hi, Q_lo, Q_hi) are u8 rows of the A operand.  The taps are quantised once to 24-bit fixed
// point and split into three signed base-256 digits d0, d1, d2 (s8), which form the B operand:
//   B[kappa][(tau, o)] = d_tau[128*(KB-1) + 8*o - kappa]       (a banded Toeplitz block, 16 outputs o per step)
// One tile step covers 128 input samples of 32 streams: M = 128 rows (32 streams x 4 byte planes),
// N = 48 columns (3 digits x 16 outputs), K = 128*KB window bytes, accumulated in int32 in TMEM -- the
// arithmetic is exact; the only approximation is the 2^-23 (relative to the largest tap) tap quantisation.
// The epilogue recombines  y = sum_tau w_tau * (256*(A_hi - 128*sum d_tau) + A_lo)  in fp32.
//
// u8 rows (the device's wire format, offset-binary I,Q bytes): sample b is the int16 value b - 128, i.e. lo plane
// b ^ 0x80 and hi plane its sign extension; the producers build those two planes straight from the bytes (4 PRMT + 4 LOP
// per 4 samples), the staged K-block is half as large, and everything behind the operand tiles is the int16 kernel with
// the scale 1/127 in place of 1/32767 (toeplitz_prepare).
//
// Persistent, warp-specialised CTA (two per SM):
//   warps 0-3  epilogue   tcgen05.ld of the int32 accumulators, digit/byte recombination, float2 stores
//   warp  4    MMA issue  one thread; tcgen05.mma.cta_group::1.kind::i8, commits to mbarriers
//   warp  5    TMA issue  one thread; the raw int16 IQ rows of a K-block (32 streams x 128 samples = 16 KB) as ONE
//                         cp.async.bulk.tensor box into a staging ring, several K-blocks ahead
//   warps 6-9  producers  (eight stream rows of every K-block each) 128-bit shared-memory loads of the staged rows, byte-plane split with PRMT, stores into the
//                         K-major SWIZZLE_128B operand tiles of a second ring (no im2col copy: consecutive steps
//                         reuse KB-1 of their KB K-block tiles).  Round 1 loaded the rows with LDG into registers: the
//                         producers then spent 70 % of their 1650 warp instructions per step on 64-bit row addresses,
//                         bounds checks and rematerialised swizzle offsets (72-register cap), and a third of all stall
//                         samples on those loads' scoreboard (profiles/r02_k2t_summary.json).  K-blocks that reach
//                         into the history rows (the first of a call) still take that path.
#include <cuda.h>

#include <cstdlib>
#include <cstring>

#include "kernels.h"

bool sfe_plain_map_2d(void *out, const void *base, long n_cols, long n_rows, long pitch_words, int box_cols, int box_rows);

namespace sfe {

namespace {

#ifndef SFE_TOEP_QUAD_STORE
#define SFE_TOEP_QUAD_STORE 1
#endif
#ifndef SFE_TOEP_EARLY_RELEASE
#define SFE_TOEP_EARLY_RELEASE 0
#endif
#ifndef SFE_TOEP_WAIT_NS
#define SFE_TOEP_WAIT_NS 0           // suspend-time hint of mbarrier.try_wait: a waiting warp sleeps instead of spinning through issue slots
#endif
constexpr int TD = 8;              // decimation
constexpr int TNO = 128 / TD;      // outputs per tile step and row
constexpr int TNT = 3;             // tap digits
constexpr int TNCOL = TNT * TNO;   // 48 accumulator columns
constexpr int TILE_BYTES = 128 * 128;
#ifndef SFE_TOEP_PROD_WARPS
#define SFE_TOEP_PROD_WARPS 4        // 4 warps of 8 stream rows each; 8 x 4 (round 1) and 2 x 16 measured slower
#endif
constexpr int N_EPI_WARPS = 4, N_PROD_WARPS = SFE_TOEP_PROD_WARPS;
constexpr int RPW = 32 / N_PROD_WARPS;                    // stream rows of a K-block per producer warp
constexpr uint32_t ROW_STEP = N_PROD_WARPS * 512u;        // bytes between a warp's consecutive rows: staged rows are 512 B, and
                                                          // their four plane rows 4 x 128 B of the operand tile (multiple of 1024: same swizzle)
static_assert(N_PROD_WARPS == 2 || N_PROD_WARPS == 4 || N_PROD_WARPS == 8, "producer warps");
constexpr int W_MMA = N_EPI_WARPS, W_TMA = N_EPI_WARPS + 1, W_PROD0 = N_EPI_WARPS + 2;
constexpr int N_THREADS = 32 * (N_EPI_WARPS + 2 + N_PROD_WARPS);
constexpr int MAX_RING = 12, MAX_STAGE = 6;
constexpr int ACC_STRIDE = 64;     // TMEM columns between the two accumulators

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t done = 0;
    while (!done)
#if SFE_TOEP_WAIT_NS
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n selp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity), "r"(SFE_TOEP_WAIT_NS) : "memory");
#else
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
#endif
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_rows(uint32_t dst, const CUtensorMap *map, int x, int y, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Operand tiles are K-major SWIZZLE_128B: rows 128 B apart, 8-row groups 1024 B apart (SBO), descriptor version 1;
// the MMA issuer builds the 64-bit descriptors from a constant high word and a 14-bit address field (see there).
__device__ __forceinline__ uint32_t swz(int row, int byte)  // byte offset inside a tile
{
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((((byte >> 4) ^ (row & 7)) << 4) | (byte & 15)));
}

struct Piece {
    int g;
    long j0, j1;
};

// the flat (group, step) range of this CTA, cut at group boundaries
struct PieceIter {
    long f, f1, J;
    __device__ PieceIter(const ToepParams &p)
    {
        J = p.n_steps;
        const long F = (long)p.n_groups * J;
        const long W = (F + gridDim.x - 1) / gridDim.x;
        f = (long)blockIdx.x * W;
        f1 = f + W < F ? f + W : F;
    }
    __device__ bool next(Piece &pc)
    {
        if (f >= f1) return false;
        pc.g = (int)(f / J);
        pc.j0 = f - (long)pc.g * J;
        const long left = f1 - f;
        pc.j1 = pc.j0 + left < J ? pc.j0 + left : J;
        f += pc.j1 - pc.j0;
        return true;
    }
};

__device__ __forceinline__ uint32_t load_i16(const ToepParams &p, int s, long n)
{
    if (s >= p.n_streams || n >= p.in.n) return 0u;
    if (n >= 0) return __ldg(reinterpret_cast<const uint32_t *>(p.in.cur) + (long)s * p.in.pitch + n);
    const long h = (long)p.in.H + n;
    return h < 0 ? 0u : __ldg(reinterpret_cast<const uint32_t *>(p.in.hist) + (long)s * p.in.H + h);
}

// prmt.b32 in its default mode (__byte_perm only defines the low three bits of a selector nibble)
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// u8 sample as the int16 pair word the int16 path loads: (b - 128) sign extended, I in the low half
__device__ __forceinline__ uint32_t load_u8_as_i16(const ToepParams &p, int s, long n)
{
    if (s >= p.n_streams || n >= p.in.n) return 0u;
    uint32_t w;
    if (n >= 0) w = __ldg(reinterpret_cast<const uint16_t *>(p.in.cur) + (long)s * p.in.pitch + n);
    else {
        const long h = (long)p.in.H + n;
        if (h < 0) return 0u;
        w = __ldg(reinterpret_cast<const uint16_t *>(p.in.hist) + (long)s * p.in.H + h);
    }
    const int i = (int)(w & 0xFF) - 128, q = (int)(w >> 8) - 128;
    return ((uint32_t)i & 0xFFFFu) | ((uint32_t)q << 16);
}

template <int FMT>
__global__ void __launch_bounds__(N_THREADS, 2) toeplitz_i8_kernel(const __grid_constant__ CUtensorMap map_cur, const ToepParams p)
{
    constexpr bool U8 = FMT == SFE_IN_U8;
    constexpr int ST_ROW = U8 ? 256 : 512;                 // bytes of one stream's K-block in the staging ring
    constexpr int ST_BYTES = 32 * ST_ROW;
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar_full[MAX_RING], bar_empty[MAX_RING], bar_acc_full[2], bar_acc_empty[2];
    __shared__ __align__(8) uint64_t bar_st_full[MAX_STAGE], bar_st_empty[MAX_STAGE];
    __shared__ uint32_t tmem_base_s;

    const int KB = p.KB, R = p.ring, Q = p.n_stage;
    uint8_t *sbase = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);   // SWIZZLE_128B tiles need 1024 B alignment
    uint8_t *sB = sbase;                                  // KB tiles of TNCOL rows x 128 B (swizzled image)
    uint8_t *ring = sbase + ((KB * TNCOL * 128 + 1023) & ~1023);  // R tiles of 128 rows x 128 B
    uint8_t *stage = ring + (size_t)R * TILE_BYTES;       // Q tiles of 32 rows x 512 B: raw int16 IQ of one K-block
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    // ---- one-time setup ---------------------------------------------------------------------------
    for (int i = tid; i < KB * TNCOL * 128 / 16; i += N_THREADS)
        reinterpret_cast<uint4 *>(sB)[i] = __ldg(reinterpret_cast<const uint4 *>(p.bmat) + i);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (tid == 0) {
        for (int i = 0; i < R; i++) {
            mbar_init(&bar_full[i], 32 * N_PROD_WARPS);
            mbar_init(&bar_empty[i], 1);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&bar_acc_full[i], 1);
            mbar_init(&bar_acc_empty[i], 32 * N_EPI_WARPS);
        }
        for (int i = 0; i < Q; i++) {
            mbar_init(&bar_st_full[i], 1);
            mbar_init(&bar_st_empty[i], 32 * N_PROD_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == W_MMA) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(128));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_base_s;

    const long X0 = p.n_first - 128L * (KB - 1);          // first sample of K-block 0

    // a K-block is staged by TMA when it lies in the current rows (the TMA zero-fills what is behind their end and the
    // rows of a ragged last group); K-blocks that reach into the history rows are loaded by the producers themselves
    const bool use_tma = p.use_tma != 0;
    auto staged = [&](long m) -> bool { return use_tma && X0 + 128 * m >= 0; };

    if (warp == W_TMA) {
        // =================================== TMA issuer =============================================
        if (lane == 0) {
            int q_slot = 0;
            uint32_t q_par = 0;
            PieceIter it(p);
            Piece pc;
            while (it.next(pc)) {
                const long m_end = pc.j1 + KB - 1;
                for (long m = pc.j0; m < m_end; m++) {
                    if (!staged(m)) continue;
                    mbar_wait(&bar_st_empty[q_slot], q_par ^ 1);                 // every producer has read the slot
                    mbar_expect_tx(&bar_st_full[q_slot], ST_BYTES);
                    // (the map's elements are 32-bit words: one int16 sample, or two u8 samples)
                    tma_rows(smem_u32(stage) + (uint32_t)q_slot * TILE_BYTES, &map_cur, (int)((X0 + 128 * m) >> (U8 ? 1 : 0)), pc.g * 32,
                             &bar_st_full[q_slot]);
                    if (++q_slot == Q) { q_slot = 0; q_par ^= 1; }
                }
            }
        }
    } else if (warp >= W_PROD0) {
        // =================================== producers ==============================================
        const int pw = warp - W_PROD0;
        int p_slot = 0, q_slot = 0;
        uint32_t p_par = 0, q_par = 0;
        const uint32_t ring_sa = smem_u32(ring), stage_sa = smem_u32(stage);
        uint32_t soff[4];                                  // swizzled byte offsets of this thread's four plane rows (q = 0)
#pragma unroll
        for (int k = 0; k < 4; k++) soff[k] = swz(4 * pw + k, 4 * lane);
        const uint32_t st_off = (uint32_t)pw * ST_ROW + (uint32_t)lane * (U8 ? 8u : 16u);   // this lane's 4 samples of staged row pw (+ N_PROD_WARPS q rows)
        PieceIter it(p);
        Piece pc;
        while (it.next(pc)) {
            const int s_base = pc.g * 32;
            const long m_end = pc.j1 + KB - 1;
            for (long m = pc.j0; m < m_end; m++) {
                uint4 v[RPW];
                if (staged(m)) {
                    mbar_wait(&bar_st_full[q_slot], q_par);
                    const uint32_t sa = stage_sa + (uint32_t)q_slot * TILE_BYTES + st_off;
#pragma unroll
                    for (int q = 0; q < RPW; q++) {
                        if (U8) {
                            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v[q].x), "=r"(v[q].y) : "r"(sa + q * (N_PROD_WARPS * ST_ROW)));
                            v[q].z = v[q].w = 0u;
                        } else {
                            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                                         : "=r"(v[q].x), "=r"(v[q].y), "=r"(v[q].z), "=r"(v[q].w) : "r"(sa + q * ROW_STEP));
                        }
                    }
                    // release the slot BEFORE waiting for room in the operand ring, so that the TMA runs ahead of the MMA
                    // instead of in step with it (the arrive has release semantics: this thread's loads above are ordered
                    // before it, and the issuer puts a proxy fence between its acquire and the refill)
#if SFE_TOEP_EARLY_RELEASE
                    mbar_arrive(&bar_st_empty[q_slot]);
                    if (++q_slot == Q) { q_slot = 0; q_par ^= 1; }
#endif
                } else {
                    const long n0 = X0 + 128 * m + 4 * lane;
#pragma unroll
                    for (int q = 0; q < RPW; q++) {
                        const int s = s_base + pw + N_PROD_WARPS * q;
                        // (u8 rows: the four samples as int16 pair words -- this path only serves the K-blocks that reach
                        // into the history and rows TMA cannot address, so the int16 split below is reused as it is)
                        v[q].x = U8 ? load_u8_as_i16(p, s, n0) : load_i16(p, s, n0);
                        v[q].y = U8 ? load_u8_as_i16(p, s, n0 + 1) : load_i16(p, s, n0 + 1);
                        v[q].z = U8 ? load_u8_as_i16(p, s, n0 + 2) : load_i16(p, s, n0 + 2);
                        v[q].w = U8 ? load_u8_as_i16(p, s, n0 + 3) : load_i16(p, s, n0 + 3);
                    }
                }
                const int slot = p_slot;                   // ring position and lap parity advance without divisions
                const uint32_t par = p_par;
                if (++p_slot == R) { p_slot = 0; p_par ^= 1; }
                mbar_wait(&bar_empty[slot], par ^ 1);      // slot free (first pass falls through)
                const uint32_t tile = ring_sa + (uint32_t)slot * TILE_BYTES;
#pragma unroll
                for (int q = 0; q < RPW; q++) {
                    uint32_t il, ih, ql, qh;
                    if (U8 && staged(m)) {
                        // x = (I0 Q0 I1 Q1), y = (I2 Q2 I3 Q3) offset binary: lo plane b ^ 0x80, biased hi plane 0x80 for
                        // b >= 128 and 0x7F below (sign-replicating PRMT: 0xFF where the byte's top bit is set)
                        il = prmt(v[q].x, v[q].y, 0x6420) ^ 0x80808080u;
                        ql = prmt(v[q].x, v[q].y, 0x7531) ^ 0x80808080u;
                        ih = prmt(v[q].x, v[q].y, 0xECA8) ^ 0x7F7F7F7Fu;      // selector bit 3: the byte's sign, replicated
                        qh = prmt(v[q].x, v[q].y, 0xFDB9) ^ 0x7F7F7F7Fu;
                    } else {
                        const uint32_t t01 = __byte_perm(v[q].x, v[q].y, 0x5140), t23 = __byte_perm(v[q].z, v[q].w, 0x5140);
                        const uint32_t u01 = __byte_perm(v[q].x, v[q].y, 0x7362), u23 = __byte_perm(v[q].z, v[q].w, 0x7362);
                        il = __byte_perm(t01, t23, 0x5410), ih = __byte_perm(t01, t23, 0x7632) ^ 0x80808080u;
                        ql = __byte_perm(u01, u23, 0x5410), qh = __byte_perm(u01, u23, 0x7632) ^ 0x80808080u;
                    }
                    // rows 4*(pw + N_PROD_WARPS*q) + k: the swizzle term (row & 7) does not depend on q, and tile rows
                    // 4*N_PROD_WARPS apart are ROW_STEP bytes apart
                    asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile + soff[0] + q * ROW_STEP), "r"(il) : "memory");
                    asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile + soff[1] + q * ROW_STEP), "r"(ih) : "memory");
                    asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile + soff[2] + q * ROW_STEP), "r"(ql) : "memory");
                    asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile + soff[3] + q * ROW_STEP), "r"(qh) : "memory");
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(&bar_full[slot]);
#if !SFE_TOEP_EARLY_RELEASE
                if (staged(m)) {                           // the staged rows are split and stored: the slot may be refilled
                    mbar_arrive(&bar_st_empty[q_slot]);
                    if (++q_slot == Q) { q_slot = 0; q_par ^= 1; }
                }
#endif
            }
        }
    } else if (warp == W_MMA) {
        // =================================== MMA issuer ==============================================
        if (lane == 0) {
            const uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(TNCOL >> 3) << 17) | ((128u >> 4) << 24);
            const uint32_t ring_a = smem_u32(ring), b_a = smem_u32(sB);
            // umma_desc() split into its constant high word and the low word of tile 0 (address >> 4 | LBO = 1 at bit 16)
            const uint32_t d_hi = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);
            const uint32_t a_lo0 = ((ring_a & 0x3FFFF) >> 4) | (1u << 16), b_lo0 = ((b_a & 0x3FFFF) >> 4) | (1u << 16);
            // ring bookkeeping in 32-bit increments: this one thread is the pacemaker of the CTA, and 64-bit
            // modulo/divide per K-block (software routines) cost it ~2700 cycles per step
            int ws = 0;                                    // slot / lap parity of the next K-block to wait for
            uint32_t wp = 0;
            int hs = 0;                                    // slot of the oldest K-block of the current step
            uint32_t step = 0;
            PieceIter it(p);
            Piece pc;
            while (it.next(pc)) {
                int ready = 0;                             // K-blocks of this piece known to be in the ring, minus steps retired
                for (long j = pc.j0; j < pc.j1; j++, step++) {
                    for (; ready < KB; ready++) {
                        mbar_wait(&bar_full[ws], wp);
                        if (++ws == R) { ws = 0; wp ^= 1; }
                    }
                    const int a = (int)(step & 1);
                    mbar_wait(&bar_acc_empty[a], ((step >> 1) & 1) ^ 1);
                    asm volatile("tcgen05.fence::after_thread_sync;");
                    const uint32_t d_tmem = tmem + a * ACC_STRIDE;
                    // descriptors differ only in their 14-bit address field: one 32-bit add per operand and MMA
                    uint32_t a_lo = a_lo0 + (uint32_t)hs * (TILE_BYTES >> 4), b_lo = b_lo0;
                    int slot = hs;
                    for (int kb = 0; kb < KB; kb++) {
#pragma unroll
                        for (int ks = 0; ks < 4; ks++) {
                            uint64_t da, db;
                            asm("mov.b64 %0, {%1, %2};" : "=l"(da) : "r"(a_lo + 2 * ks), "r"(d_hi));
                            asm("mov.b64 %0, {%1, %2};" : "=l"(db) : "r"(b_lo + 2 * ks), "r"(d_hi));
                            if (ks == 0)
                                asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
                                             " tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                                             ::"r"(d_tmem), "l"(da), "l"(db), "r"(idesc), "r"(kb) : "memory");
                            else
                                asm volatile("{\n .reg .pred p;\n setp.eq.b32 p, 0, 0;\n"
                                             " tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                                             ::"r"(d_tmem), "l"(da), "l"(db), "r"(idesc) : "memory");
                        }
                        b_lo += (TNCOL * 128) >> 4;
                        a_lo += TILE_BYTES >> 4;
                        if (++slot == R) { slot = 0; a_lo = a_lo0; }
                    }
                    umma_commit(&bar_acc_full[a]);                                   // accumulator ready
                    umma_commit(&bar_empty[hs]);                                     // oldest K-block retired
                    if (++hs == R) hs = 0;
                    ready--;
                }
                // the KB-1 trailing K-blocks of the piece are not reused by the next piece
                for (int t = 0; t < KB - 1; t++) {
                    umma_commit(&bar_empty[hs]);
                    if (++hs == R) hs = 0;
                }
            }
        }
    } else {
        // =================================== epilogue ================================================
        const int plane = lane & 3;                        // 0 I_lo, 1 I_hi, 2 Q_lo, 3 Q_hi
        const bool hi = plane & 1;
        // y = sum_tau w_tau (A_tau - c_tau) * byte_w with the weights pre-scaled
        const float byte_w = hi ? 256.f : 1.f;
        const float w0 = p.w[0] * byte_w, w1 = p.w[1] * byte_w, w2 = p.w[2] * byte_w;
        // the bias of the hi planes (128 * sum of the digit's taps, an exact integer) is removed in int32, before conversion
        const int ci0 = hi ? (int)p.corr[0] : 0, ci1 = hi ? (int)p.corr[1] : 0, ci2 = hi ? (int)p.corr[2] : 0;
        long step = 0;
        // 128-bit stores need 16-byte aligned output rows
        const bool out_vec = (reinterpret_cast<uintptr_t>(p.out) & 15) == 0 && (p.out_pitch & 1) == 0;
        PieceIter it(p);
        Piece pc;
        while (it.next(pc)) {
            const int s = pc.g * 32 + warp * 8 + (lane >> 2);
            for (long j = pc.j0; j < pc.j1; j++, step++) {
                const int a = (int)(step & 1);
                mbar_wait(&bar_acc_full[a], (uint32_t)((step >> 1) & 1));
                asm volatile("tcgen05.fence::after_thread_sync;");
                // one digit (16 columns) at a time: 16 + 16 live registers instead of 48 + 16, so that two CTAs fit an SM
                float f[TNO];
                const uint32_t ta = tmem + ((uint32_t)(warp * 32) << 16) + a * ACC_STRIDE;
#pragma unroll
                for (int c = 0; c < TNT; c++) {
                    uint32_t v[16];
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                                 : "r"(ta + 16 * c));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    const float wc = c == 0 ? w0 : c == 1 ? w1 : w2;
                    const int cc = c == 0 ? ci0 : c == 1 ? ci1 : ci2;
#pragma unroll
                    for (int i = 0; i < TNO; i++) f[i] = c == 0 ? wc * (float)((int)v[i] - cc) : fmaf(wc, (float)((int)v[i] - cc), f[i]);
                }
                asm volatile("tcgen05.fence::before_thread_sync;");
                mbar_arrive(&bar_acc_empty[a]);            // TMEM buffer free for the step after next
#if SFE_TOEP_QUAD_STORE
                // The four lanes of a stream hold the partial sums (I_lo, I_hi, Q_lo, Q_hi) of all 16 outputs.  Two exchange
                // rounds in halves leave every lane with BOTH components of four consecutive outputs: round 1 with the lane
                // of the other byte (lo + 256 hi; the even lane keeps outputs 0-7, the odd one 8-15), round 2 with the lane
                // of the other component (the I lane keeps the first four of its eight, the Q lane the last four).  12
                // shuffles per lane instead of 32, and each lane stores its 32 bytes as two 128-bit words: every sector of
                // the output row is written once, by one instruction (with one lane of four storing 16 float2 one by one,
                // every sector was written four times -- removing the stores alone made the kernel 15 % faster).
                float g8[8];
#pragma unroll
                for (int t = 0; t < 8; t++) {
                    const float send = (plane & 1) ? f[t] : f[t + 8];
                    const float keep = (plane & 1) ? f[t + 8] : f[t];
                    g8[t] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
                }
                float2 o4[4];
#pragma unroll
                for (int t = 0; t < 4; t++) {
                    const float send = (plane & 2) ? g8[t] : g8[t + 4];
                    const float recv = __shfl_xor_sync(0xffffffffu, send, 2);
                    o4[t] = (plane & 2) ? make_float2(recv, g8[t + 4]) : make_float2(g8[t], recv);
                }
                const int k_off = 8 * (plane & 1) + 4 * (plane >> 1);     // first of this lane's four outputs within the step
                const long k_abs = j * TNO + k_off;
                float2 *o = p.out + (long)s * p.out_pitch + k_abs;
                if (s < p.n_streams) {
                    if (k_abs + 4 <= p.n_out && out_vec) {
                        reinterpret_cast<float4 *>(o)[0] = make_float4(o4[0].x, o4[0].y, o4[1].x, o4[1].y);
                        reinterpret_cast<float4 *>(o)[1] = make_float4(o4[2].x, o4[2].y, o4[3].x, o4[3].y);
                    } else {
#pragma unroll
                        for (int t = 0; t < 4; t++)
                            if (k_abs + t < p.n_out) o[t] = o4[t];
                    }
                }
#else
                float2 *o = p.out + (long)s * p.out_pitch + j * TNO;
#pragma unroll
                for (int i = 0; i < TNO; i++) {
                    float fi = f[i];
                    fi += __shfl_xor_sync(0xffffffffu, fi, 1);             // lo + 256*hi
                    const float q = __shfl_down_sync(0xffffffffu, fi, 2); // Q next to I
                    if (plane == 0 && s < p.n_streams && j * TNO + i < p.n_out) o[i] = make_float2(fi, q);
                }
#endif
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    __syncwarp();
    if (warp == W_MMA) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(128));
}

}  // namespace

int toeplitz_max_taps() { return 128 * (MAX_RING - 4 - 1) + 1; }

static int launch_toeplitz_rows(const ToepParams &p_in, cudaStream_t st)
{
    ToepParams p = p_in;
    if (p.n_out <= 0 || p.n_streams <= 0) return SFE_OK;
    p.n_groups = (p.n_streams + 31) / 32;
    p.n_steps = (p.n_out + TNO - 1) / TNO;
    const bool u8 = p.in.fmt == SFE_IN_U8;
    if (p.KB < 1 || p.KB + 2 > MAX_RING || (p.in.fmt != SFE_IN_I16 && !u8)) {
        set_error("toeplitz_i8: unsupported configuration");
        return SFE_ERR_ARG;
    }
    // The raw rows are staged by TMA when the caller's rows can carry a tensor map (16-byte aligned base and pitch).
    // Shared memory per CTA: tap tiles + operand ring (>= KB + 1 tiles) + staging ring; two CTAs per SM when that fits.
    CUtensorMap map;
    memset(&map, 0, sizeof map);
    static const bool no_tma = getenv("SFE_TOEP_NO_TMA") != nullptr;       // tuning hook: producers load from global memory
    // (a box must start on a 16-byte boundary of its row: calls whose first output does not sit on a multiple of four samples
    // -- a stream fed in pieces that are not multiples of 8 samples -- keep the producers' own loads)
    const int amask = u8 ? 7 : 3;            // samples per 16 bytes, minus one
    p.use_tma = !no_tma && (reinterpret_cast<uintptr_t>(p.in.cur) & 15) == 0 && (p.in.pitch & amask) == 0 && p.in.n >= 128 &&
                ((p.n_first - 128L * (p.KB - 1)) & amask) == 0;
    // the map's elements are 32-bit words: one int16 sample or two u8 samples.  An odd u8 row length lets the last word carry
    // one byte pair from behind the row's end (inside its pitch): it sits at sample index n, which only meets zero taps
    // of the outputs a call produces (chain.cu hands such rows over in a staging buffer of its own, so the read never
    // leaves memory the library owns)
    if (p.use_tma && !sfe_plain_map_2d(&map, p.in.cur, u8 ? (p.in.n + 1) / 2 : p.in.n, p.n_streams, u8 ? p.in.pitch / 2 : p.in.pitch,
                                       u8 ? 64 : 128, 32))
        return SFE_ERR_CUDA;
    static const int stage_env = getenv("SFE_TOEP_STAGE") ? atoi(getenv("SFE_TOEP_STAGE")) : 0;   // tuning hook: staging depth
    const size_t b_bytes = ((size_t)(p.KB * TNCOL * 128 + 1023) & ~(size_t)1023);
    auto smem_for = [&](int ring, int stage) { return b_bytes + (size_t)(ring + stage) * TILE_BYTES + 1024; };
    p.n_stage = p.use_tma ? (stage_env > 0 ? stage_env : 3) : 1;
    if (p.n_stage > MAX_STAGE) p.n_stage = MAX_STAGE;
    p.ring = p.KB + 1;
    // spend what two CTAs per SM leave on the operand ring first (up to KB + 4), then fall back to one CTA
    const size_t two_cta = (227 * 1024) / 2 - 2048;
    while (p.ring < p.KB + 4 && p.ring < MAX_RING && smem_for(p.ring + 1, p.n_stage) <= two_cta) p.ring++;
    if (smem_for(p.ring, p.n_stage) > two_cta) {            // one CTA per SM: room for a deep ring
        p.ring = p.KB + 4 <= MAX_RING ? p.KB + 4 : MAX_RING;
        while (smem_for(p.ring, p.n_stage) > 225 * 1024 && p.n_stage > 1) p.n_stage--;
    }
    const size_t smem = smem_for(p.ring, p.n_stage);
    if (smem > 227 * 1024) {
        set_error("toeplitz_i8: tap count too large for shared memory");
        return SFE_ERR_ARG;
    }
    void (*kern)(const CUtensorMap, const ToepParams) = u8 ? toeplitz_i8_kernel<SFE_IN_U8> : toeplitz_i8_kernel<SFE_IN_I16>;
    if (!SFE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))) return SFE_ERR_CUDA;
    int dev = 0, sms = 0;
    if (!SFE_CUDA(cudaGetDevice(&dev)) || !SFE_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev))) return SFE_ERR_CUDA;
    const long F = (long)p.n_groups * p.n_steps;
    long grid = F / 4 > 0 ? F / 4 : 1;                    // at least ~4 steps per CTA
    // two CTAs per SM when their rings fit (320 threads, at most 102 registers each, and 128 TMEM columns each): the second CTA's warps
    // fill the issue slots the first one leaves while its producers, MMA thread and epilogue wait on each other
    static const bool one_cta = getenv("SFE_TOEP_ONE_CTA") != nullptr;   // tuning hook
    const int per_sm = (2 * (smem + 2048) <= 227 * 1024 && !one_cta) ? 2 : 1;
    if (grid > (long)sms * per_sm) grid = (long)sms * per_sm;
    kern<<<(unsigned)grid, N_THREADS, smem, st>>>(map, p);
    count_launch();
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

// ONE long stream (BASELINE.json cfg 2 as it is worded: a single 10 MS/s channel) would fill 1 of the 32 stream rows of
// every tile.  The FIR is time invariant, so the row is cut into pieces of Kp outputs (Q = 8 Kp samples) that run as the
// ROWS of one launch: piece r starts 128 KB samples before its first new sample (its "history" is simply the preceding
// part of the row: overlapping rows, pitch Q), piece 0 keeps the call's real history, and what is left behind the last
// whole piece runs as a third, one-row launch.  Same tiles, same taps, same summation order as the unsplit launch.
int launch_toeplitz(const ToepParams &p, cudaStream_t st)
{
    static const bool no_split = getenv("SFE_TOEP_NO_SPLIT") != nullptr;   // tuning hook
    const long Kp = 1024, Q = TD * Kp, B = 128L * p.KB;
    const long P = p.n_out / Kp;
    if (p.n_streams >= 8 || no_split || P < 4 || p.in.fmt != SFE_IN_I16) return launch_toeplitz_rows(p, st);
    if (p.n_streams > 1) {                   // 2 ... 7 long streams: each one split on its own (rows of different streams are not one pitch apart)
        for (int s = 0; s < p.n_streams; s++) {
            ToepParams one = p;
            one.n_streams = 1;
            one.in.cur = static_cast<const uint32_t *>(p.in.cur) + (long)s * p.in.pitch;
            one.in.hist = static_cast<const uint32_t *>(p.in.hist) + (long)s * p.in.H;
            one.out = p.out + (long)s * p.out_pitch;
            const int rc1 = launch_toeplitz(one, st);
            if (rc1 != SFE_OK) return rc1;
        }
        return SFE_OK;
    }
    ToepParams a = p;                        // piece 0: the real history
    a.n_out = Kp;
    int rc = launch_toeplitz_rows(a, st);
    if (rc != SFE_OK) return rc;
    ToepParams b = p;                        // pieces 1 .. P-2 as rows
    b.in.cur = static_cast<const uint32_t *>(p.in.cur) + (Q - B);
    b.in.hist = b.in.cur;
    b.in.H = 0;
    b.in.pitch = Q;
    b.in.n = B + Q + 16;                     // every sample a piece touches lies inside the original row
    b.n_first = p.n_first + B;
    b.n_streams = (int)(P - 2);
    b.n_out = Kp;
    b.out = p.out + Kp;
    b.out_pitch = Kp;
    rc = launch_toeplitz_rows(b, st);
    if (rc != SFE_OK) return rc;
    ToepParams c = b;                        // the rest: one row from the start of piece P-1
    c.in.cur = static_cast<const uint32_t *>(p.in.cur) + ((P - 1) * Q - B);
    c.in.hist = c.in.cur;
    c.in.n = p.in.n - ((P - 1) * Q - B);
    c.in.pitch = (c.in.n + 3) & ~3L;
    c.n_streams = 1;
    c.n_out = p.n_out - (P - 1) * Kp;
    c.out = p.out + (P - 1) * Kp;
    c.out_pitch = c.n_out;
    return launch_toeplitz_rows(c, st);
}

// Host side: quantise g to 24-bit fixed point, split into three signed base-256 digits and lay the banded
// Toeplitz blocks out exactly as the kernel's shared-memory image (K-major, SWIZZLE_128B).
bool toeplitz_prepare(const std::vector<double> &g, ToepHost &h, int in_fmt)
{
    const int T = (int)g.size();
    h.KB = (T - 1 + 127) / 128 + 1;
    if (h.KB < 2) h.KB = 2;
    double sigma = 0;
    for (double v : g) sigma = fabs(v) > sigma ? fabs(v) : sigma;
    if (!(sigma > 0)) sigma = 1.0;
    const double Zmax = (double)((127 << 16) - (128 << 8) - 128);
    std::vector<int> d0(T), d1(T), d2(T);
    long sum[3] = {0, 0, 0};
    for (int j = 0; j < T; j++) {
        long Z = llround(g[j] / sigma * Zmax);
        auto digit = [](long z) { long m = ((z + 128) % 256 + 256) % 256; return (int)(m - 128); };
        d2[j] = digit(Z);
        Z = (Z - d2[j]) / 256;
        d1[j] = digit(Z);
        Z = (Z - d1[j]) / 256;
        d0[j] = (int)Z;
        if (d0[j] < -128 || d0[j] > 127) return false;
        sum[0] += d0[j];
        sum[1] += d1[j];
        sum[2] += d2[j];
    }
    // x = s/32767 or (b - 128)/127 (source_c_impl.cc:121-132), folded into the weights
    const double Kq = sigma / Zmax / (in_fmt == SFE_IN_U8 ? 127.0 : 32767.0);
    h.w[0] = (float)(Kq * 65536.0);
    h.w[1] = (float)(Kq * 256.0);
    h.w[2] = (float)Kq;
    for (int t = 0; t < 3; t++) h.corr[t] = (float)(128 * sum[t]);   // bias of the hi ^ 0x80 rows, exact in fp32
    h.bmat.assign((size_t)h.KB * TNCOL * 128, 0);
    const std::vector<int> *dig[3] = {&d0, &d1, &d2};
    for (int kb = 0; kb < h.KB; kb++)
        for (int n = 0; n < TNCOL; n++) {
            const int tau = n / TNO, o = n % TNO;
            for (int c = 0; c < 128; c++) {
                const int jt = 128 * (h.KB - 1) + TD * o - (128 * kb + c);
                const int val = (jt >= 0 && jt < T) ? (*dig[tau])[jt] : 0;
                const size_t off = (size_t)kb * TNCOL * 128 + (n >> 3) * 1024 + (n & 7) * 128 + ((((c >> 4) ^ (n & 7)) << 4) | (c & 15));
                h.bmat[off] = (int8_t)val;
            }
        }
    return true;
}

}  // namespace sfe
