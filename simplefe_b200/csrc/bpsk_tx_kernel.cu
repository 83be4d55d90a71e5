// bpsk_tx_kernel.cu -- batched BPSK transmit path: the reference's only real libdsp
// application, examples/bpsk/bpsk.cxx:122-174, for many channels at once.
//
// Reference, per channel: bits -> +-A impulses every `sps` samples (bpsk.cxx:145-159, A = .85f/1.35f, a set bit
// sends -A) -> blkconv with the RRC prototype (bpsk.cxx:125,162: 111 taps / FFT 2048 at 10 samples per symbol)
// -> ring_buffer<float> -> 10-bit DAC words, 4 samples in 5 bytes (bpsk.cxx:76-101).
// The impulse train is 90 % zeros, so the FFT convolver's work collapses to a polyphase interpolator:
//     y[s*sps + r] = sum_m a[s - m] * h[r + m*sps]            (about n_taps/sps MACs per output sample)
// computed here straight from the packed bits; the float waveform never leaves the chip when the caller asks
// for DAC words.  The kernel is bound by its own output stream (12.5 bytes per symbol at sps = 10).
//
// Mapping: a warp owns 256 consecutive symbols of one channel, a thread 8 of them (16 was measured slower: the tile per
// thread limits the resident warps).  The thread derives the symbol signs of
// its group once, from one 64-bit window of the bit stream, as +-1.0f registers, and then walks the phases
// r = 0 .. sps-1: the ~12 pre-scaled taps h[r + m*sps] of the phase come from shared memory with 128-bit loads, every
// tap and output costs one FFMA, and the 8 results go to the warp's own shared-memory tile in time order (group stride
// 8*sps + 2 words: 2-way instead of 32-way bank conflicts).  The warp then packs its tile (DAC words: in place) and
// stores 128-bit words.  Tiles are warp-private, so warps only meet at __syncwarp: while one warp stores, the others
// compute (with CTA-wide tiles 22 % of the stall samples were the barrier between the two passes).
#include "kernels.h"

namespace sfe {

namespace {

constexpr int TX_GROUP = 8;        // symbols per thread
constexpr int TX_WBIT = 64 - TX_GROUP;   // window bit of the thread's first symbol (bit 63 = its last)
constexpr int TX_WSYM = 32 * TX_GROUP;   // symbols per warp
constexpr int TX_MAXM = 32;        // taps per phase, at most (n_taps <= 32 * sps)
constexpr int TX_MAXW = 4;         // warps per CTA, at most

typedef unsigned long long tx_f32x2;
__device__ __forceinline__ tx_f32x2 tx_pack2(float lo, float hi)
{
    tx_f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ tx_f32x2 tx_ffma2(tx_f32x2 a, tx_f32x2 b, tx_f32x2 c)
{
    tx_f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// bit k of the channel's symbol stream; k < 0 reads the history word (bit 63 = symbol -1)
__device__ __forceinline__ uint64_t bits_window(const uint32_t *cur, uint64_t hist, long s_hi, long n_sym)
{
    // returns a 64-bit word whose bit 63 is symbol s_hi, bit 62 symbol s_hi - 1, ...
    uint64_t w = 0;
    const long s_lo = s_hi - 63;
#pragma unroll
    for (int part = 0; part < 3; part++) {                 // at most three 32-bit words overlap the window
        const long word = (s_lo >> 5) + part;              // arithmetic shift: floor for negatives
        uint64_t v;
        if (word >= 0) v = word * 32 < n_sym ? __ldg(cur + word) : 0;   // never read past the row
        else if (word == -1) v = hist >> 32;               // symbols -32 .. -1
        else if (word == -2) v = hist & 0xffffffffull;     // symbols -64 .. -33
        else v = 0;
        const long first = word * 32;                      // symbol index of bit 0 of v
        const long sh = first - s_lo;                      // position of that bit inside the window
        if (sh >= 0 && sh < 64) w |= v << sh;
        else if (sh < 0 && sh > -32) w |= v >> (-sh);
    }
    return w;
}

template <int MT>     // taps per phase, rounded up to a multiple of 4 (the padding taps are zero)
__global__ void __launch_bounds__(32 * TX_MAXW)
bpsk_tx_kernel(const uint32_t *bits, long pitch_words, const uint64_t *hist, int n_hist, long n_sym, int sps, const float *taps,
               int n_taps, float amp, int out_dac10, void *out, long out_pitch_bytes)
{
    extern __shared__ __align__(16) float tx_sm[];
    float *hs = tx_sm;                                     // [sps][MT] pre-scaled taps, hs[r][m] = amp * h[r + m*sps]
    const int gstride = TX_GROUP * sps + 2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *tile = hs + sps * MT + warp * 32 * gstride;     // this warp's [32 groups][16*sps + 2]: samples in time order per group
    const int ch = blockIdx.y;
    const long s0 = ((long)blockIdx.x * (blockDim.x >> 5) + warp) * TX_WSYM;   // first symbol of the warp
    const uint32_t *cur = bits + (long)ch * pitch_words;
    const uint64_t hw = hist[ch];
    const int M = (n_taps + sps - 1) / sps;                // taps per phase

    // Even sps: the taps of phases 2r' and 2r'+1 sit side by side, hs[((r' MT) + m) 2 + (r & 1)], so that one 64-bit
    // register pair (h[2r'][m], h[2r'+1][m]) times a broadcast sign is ONE FFMA2 for two output samples of a symbol.
    const bool paired = (sps & 1) == 0;
    auto tap_at = [&](int r, int m) { return paired ? (((r >> 1) * MT + m) << 1) + (r & 1) : r * MT + m; };
    for (int i = threadIdx.x; i < sps * MT; i += blockDim.x) {
        const int r = i / MT, m = i - r * MT;
        hs[tap_at(r, m)] = (m < M && r + m * sps < n_taps) ? amp * __ldg(taps + r + m * sps) : 0.f;
    }
    __syncthreads();
    if (s0 >= n_sym) return;                               // whole warp

    const long s_first = s0 + (long)lane * TX_GROUP;
    if (s_first < n_sym) {
        // window: bit 63 = the thread's last symbol (s_first + TX_GROUP - 1), bit TX_WBIT = s_first, bit 0 = s_first - TX_WBIT
        const uint64_t w = bits_window(cur, hw, s_first + TX_GROUP - 1, n_sym);
        float *trow = tile + lane * gstride;
        if (s_first - (M - 1) >= -(long)n_hist && s_first + TX_GROUP <= n_sym) {
            // every symbol the group touches exists.  The +-A impulse (bpsk.cxx:150) times a tap is the pre-scaled tap times
            // +-1, and the sign only depends on the symbol i - m: 16 + MT - 1 sign registers shared by all phases, then ONE
            // FFMA per tap and output (the product by +-1 is exact, so this equals adding the sign-flipped tap, as the edge
            // path does).
            float sg[TX_GROUP + MT - 1];                                   // sg[d + MT - 1]: symbol s_first + d
            const uint32_t wl = (uint32_t)w, wh = (uint32_t)(w >> 32);
#pragma unroll
            for (int d = 0; d < TX_GROUP + MT - 1; d++) {
                const int bit = TX_WBIT + 1 - MT + d;                               // compile-time position in the window
                const uint32_t word = bit < 32 ? wl : wh;
                sg[d] = __uint_as_float(0x3F800000u | ((word << (31 - (bit & 31))) & 0x80000000u));
            }
            if (paired) {
                // two phases per round: same FP32-pipe time as scalar FFMAs, half the issue slots (the kernel was issue
                // bound: 85 % of the slots, 62 % of them FFMA).  Each lane of an FFMA2 is the IEEE fma of the scalar loop
                // below, taps in the same order: the samples are bit-identical.
                for (int r = 0; r < sps; r += 2) {
                    tx_f32x2 acc[TX_GROUP];
#pragma unroll
                    for (int i = 0; i < TX_GROUP; i++) acc[i] = 0ull;
                    const ulonglong2 *hr = reinterpret_cast<const ulonglong2 *>(hs + r * MT);   // pairs m, m + 1 per 128-bit load
#pragma unroll
                    for (int m2 = 0; m2 < MT / 2; m2++) {
                        const ulonglong2 h2 = hr[m2];
#pragma unroll
                        for (int u = 0; u < 2; u++) {
                            const int m = 2 * m2 + u;
                            const tx_f32x2 hp = u == 0 ? h2.x : h2.y;
#pragma unroll
                            for (int i = 0; i < TX_GROUP; i++) {
                                const float sgn = sg[i - m + MT - 1];
                                acc[i] = tx_ffma2(hp, tx_pack2(sgn, sgn), acc[i]);
                            }
                        }
                    }
                    tx_f32x2 *tr = reinterpret_cast<tx_f32x2 *>(trow + r);                      // sps and gstride are even: 8-byte aligned
#pragma unroll
                    for (int i = 0; i < TX_GROUP; i++) tr[i * (sps >> 1)] = acc[i];
                }
            } else
            for (int r = 0; r < sps; r++) {
                float acc[TX_GROUP];
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++) acc[i] = 0.f;
                const float4 *hr = reinterpret_cast<const float4 *>(hs + r * MT);
#pragma unroll
                for (int m4 = 0; m4 < MT / 4; m4++) {
                    const float4 h4 = hr[m4];
                    const float hv[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const int m = 4 * m4 + u;
#pragma unroll
                        for (int i = 0; i < TX_GROUP; i++) acc[i] = fmaf(sg[i - m + MT - 1], hv[u], acc[i]);
                    }
                }
                float *tr = trow + r;
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++, tr += sps) *tr = acc[i];
            }
        } else {
            // first symbols of a stream (zero state before the history) and the ragged end
            // (same arithmetic as above, tap by tap, so that a stream cut into calls anywhere gives identical samples)
            for (int r = 0; r < sps; r++) {
                float acc[TX_GROUP];
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++) acc[i] = 0.f;
                for (int m = 0; m < M; m++) {
                    const uint32_t hb = __float_as_uint(hs[tap_at(r, m)]);
#pragma unroll
                    for (int i = 0; i < TX_GROUP; i++) {
                        const long sm = s_first + i - m;                       // symbol s - m
                        const bool valid = sm >= -(long)n_hist && sm < n_sym;  // before the first symbol: zero state
                        const uint32_t sign = (uint32_t)((w >> (TX_WBIT + i - m)) & 1ull) << 31;   // a set bit sends -A (bpsk.cxx:150)
                        acc[i] += valid ? __uint_as_float(hb ^ sign) : 0.f;
                    }
                }
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++) trow[i * sps + r] = acc[i];
            }
        }
    }
    __syncwarp();

    // sample n of the warp's tile (time order) sits at tile[(n / (16 sps)) * gstride + n % (16 sps)]; a quad never straddles a group
    const int qpg = TX_GROUP * sps / 4;                    // quads per group
    const uint32_t qmagic = 0xFFFFFFFFu / (uint32_t)qpg + 1u;   // floor(q / qpg) = umulhi(q, qmagic) for q < 2^16
    auto load_quad = [&](int qd) {
        const int g = (int)__umulhi((uint32_t)qd, qmagic);
        const float2 *src = reinterpret_cast<const float2 *>(tile + g * gstride + (qd - g * qpg) * 4);
        const float2 a = src[0], b = src[1];
        return make_float4(a.x, a.y, b.x, b.y);
    };
    const int n_valid_sym = (int)(n_sym - s0 < TX_WSYM ? n_sym - s0 : TX_WSYM);
    const int n_samp = n_valid_sym * sps;                  // valid samples of this tile
    const long samp0 = s0 * sps;
    char *row0 = static_cast<char *>(out) + (long)ch * out_pitch_bytes;
    if (out_dac10) {
        // bpsk.cxx:76-101: u = ((short)(x*511) + 512) & 0x3FF, 4 words -> 5 bytes, the four 2-bit tops first.
        // Packed in place: quad q is read from bytes >= 16 q of the tile and written to bytes 5 q, so a round of 32 quads
        // only overwrites what this and earlier rounds have read.  Then stored as 128-bit words.
        uint8_t *packed = reinterpret_cast<uint8_t *>(tile);
        const int n_quads = n_samp / 4;
        auto pack_quad = [](const float4 &x, uint32_t (&b)[5]) {
            const float v[4] = {x.x, x.y, x.z, x.w};
            unsigned u[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const short sh = (short)(int)__fmul_rn(v[k], 511.0f);       // C float->short: truncation toward zero
                u[k] = ((unsigned)(sh + 512)) & 0x3FFu;
            }
            b[0] = (u[0] >> 8) | ((u[1] >> 8) << 2) | ((u[2] >> 8) << 4) | ((u[3] >> 8) << 6);
            b[1] = u[0] & 0xFF;
            b[2] = u[1] & 0xFF;
            b[3] = u[2] & 0xFF;
            b[4] = u[3] & 0xFF;
        };
        // four quads (20 bytes = five 32-bit words) per lane and round
        const int n_q4 = n_quads / 4;
        for (int base = 0; base < n_q4; base += 32) {
            const int q4 = base + lane;
            float4 x[4];
            if (q4 < n_q4) {
#pragma unroll
                for (int k = 0; k < 4; k++) x[k] = load_quad(4 * q4 + k);
            }
            __syncwarp();
            if (q4 < n_q4) {
                uint32_t by[20];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    uint32_t b[5];
                    pack_quad(x[k], b);
#pragma unroll
                    for (int j = 0; j < 5; j++) by[5 * k + j] = b[j];
                }
                uint32_t *o = reinterpret_cast<uint32_t *>(packed + q4 * 20);
#pragma unroll
                for (int j = 0; j < 5; j++) o[j] = by[4 * j] | (by[4 * j + 1] << 8) | (by[4 * j + 2] << 16) | (by[4 * j + 3] << 24);
            }
        }
        __syncwarp();
        {   // the last one to three quads of a ragged tile
            const int qd = 4 * n_q4 + lane;
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (qd < n_quads) x = load_quad(qd);
            __syncwarp();
            if (qd < n_quads) {
                uint32_t b[5];
                pack_quad(x, b);
#pragma unroll
                for (int j = 0; j < 5; j++) packed[qd * 5 + j] = (uint8_t)b[j];
            }
        }
        __syncwarp();
        char *dst = row0 + samp0 / 4 * 5;
        const int n_bytes = n_quads * 5;
        if (((reinterpret_cast<uintptr_t>(dst) | (uintptr_t)n_bytes) & 15) == 0) {
            for (int i = lane; i < n_bytes / 16; i += 32) reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(packed)[i];
        } else {
            for (int i = lane; i < n_bytes; i += 32) dst[i] = (char)packed[i];
        }
    } else {
        float *dst = reinterpret_cast<float *>(row0) + samp0;
        if (((reinterpret_cast<uintptr_t>(dst) & 15) | (uintptr_t)(n_samp & 3)) == 0) {
            for (int i = lane; i < n_samp / 4; i += 32) reinterpret_cast<float4 *>(dst)[i] = load_quad(i);
        } else {
            const int spg = TX_GROUP * sps;
            for (int i = lane; i < n_samp; i += 32) dst[i] = tile[(i / spg) * gstride + i % spg];
        }
    }
}

// new history word: the last 64 symbols of (history ++ current)
__global__ void bpsk_tx_hist_kernel(const uint32_t *bits, long pitch_words, const uint64_t *hist_old, uint64_t *hist_new, long n_sym,
                                    int n_channels)
{
    const int ch = blockIdx.x * blockDim.x + threadIdx.x;
    if (ch >= n_channels) return;
    hist_new[ch] = n_sym > 0 ? bits_window(bits + (long)ch * pitch_words, hist_old[ch], n_sym - 1, n_sym) : hist_old[ch];
}

}  // namespace

int bpsk_tx_max_taps(int sps) { return TX_MAXM * sps; }

int launch_bpsk_tx(const uint32_t *bits, long pitch_words, const uint64_t *hist_old, uint64_t *hist_new, int n_hist, long n_sym, int sps,
                   const float *taps, int n_taps, float amp, int n_channels, int out_dac10, void *out, long out_pitch_bytes,
                   cudaStream_t st)
{
    if (n_sym <= 0 || n_channels <= 0) return SFE_OK;
    const int M = (n_taps + sps - 1) / sps;
    const int MT = M <= 4 ? 4 : M <= 8 ? 8 : M <= 12 ? 12 : M <= 16 ? 16 : M <= 24 ? 24 : 32;
    // a warp's tile is 32 groups of 16 sps + 2 floats; two warps per CTA unless the tiles get large (the taps are per CTA)
    const size_t warp_tile = sizeof(float) * 32 * (size_t)(TX_GROUP * sps + 2);
    const long n_warps = (n_sym + TX_WSYM - 1) / TX_WSYM;
    int W = warp_tile <= 20 * 1024 ? 4 : warp_tile <= 40 * 1024 ? 2 : 1;
    if (W > n_warps) W = (int)n_warps;
    const size_t smem = sizeof(float) * (size_t)sps * MT + W * warp_tile;
    auto kern = MT == 4 ? bpsk_tx_kernel<4> : MT == 8 ? bpsk_tx_kernel<8> : MT == 12 ? bpsk_tx_kernel<12> : MT == 16 ? bpsk_tx_kernel<16>
                : MT == 24 ? bpsk_tx_kernel<24> : bpsk_tx_kernel<32>;
    if (!SFE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))) return SFE_ERR_CUDA;
    dim3 grid((unsigned)((n_warps + W - 1) / W), (unsigned)n_channels);
    kern<<<grid, 32 * W, smem, st>>>(bits, pitch_words, hist_old, n_hist, n_sym, sps, taps, n_taps, amp, out_dac10, out, out_pitch_bytes);
    bpsk_tx_hist_kernel<<<(n_channels + 127) / 128, 128, 0, st>>>(bits, pitch_words, hist_old, hist_new, n_sym, n_channels);
    count_launch(2);
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

}  // namespace sfe
