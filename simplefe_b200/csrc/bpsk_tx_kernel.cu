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
// Mapping: a CTA owns up to 2048 consecutive symbols of one channel (as many as keep its tile under 80 KB), a thread
// 16 of them.  The thread derives the symbol signs of its group once, from one 64-bit window of the bit stream, as
// +-1.0f registers, and then walks the phases r = 0 .. sps-1: the ~12 pre-scaled taps h[r + m*sps] of the phase come
// from shared memory with 128-bit loads, every tap and output costs one FFMA, and the 16 results go to a shared-memory
// tile in time order (group stride 16*sps + 2 words: 2-way instead of 32-way bank conflicts).  A second pass packs
// the tile and stores 128-bit words.
#include "kernels.h"

namespace sfe {

namespace {

constexpr int TX_SYM_MAX = 2048;   // symbols per CTA, at most
constexpr int TX_GROUP = 16;       // symbols per thread
constexpr int TX_MAXM = 32;        // taps per phase, at most (n_taps <= 32 * sps)
constexpr int TX_THREADS = TX_SYM_MAX / TX_GROUP;
constexpr int TX_TILE_BYTES = 80 * 1024;

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
__global__ void __launch_bounds__(TX_THREADS)
bpsk_tx_kernel(const uint32_t *bits, long pitch_words, const uint64_t *hist, int n_hist, long n_sym, int sps, int sym_cta, const float *taps,
               int n_taps, float amp, int out_dac10, void *out, long out_pitch_bytes)
{
    extern __shared__ __align__(16) float tx_sm[];
    float *hs = tx_sm;                                     // [sps][MT] pre-scaled taps, hs[r][m] = amp * h[r + m*sps]
    float *tile = hs + sps * MT;                           // [groups][16*sps + 2]: samples in time order per group
    const int gstride = TX_GROUP * sps + 2;
    const int n_groups = sym_cta / TX_GROUP;
    const int ch = blockIdx.y;
    const long s0 = (long)blockIdx.x * sym_cta;
    const uint32_t *cur = bits + (long)ch * pitch_words;
    const uint64_t hw = hist[ch];
    const int M = (n_taps + sps - 1) / sps;                // taps per phase

    for (int i = threadIdx.x; i < sps * MT; i += TX_THREADS) {
        const int r = i / MT, m = i - r * MT;
        hs[i] = (m < M && r + m * sps < n_taps) ? amp * __ldg(taps + r + m * sps) : 0.f;
    }
    __syncthreads();

    for (int grp = threadIdx.x; grp < n_groups; grp += TX_THREADS) {
        const long s_first = s0 + (long)grp * TX_GROUP;
        if (s_first >= n_sym) break;
        // window: bit 63 = symbol s_first + 15, ..., bit 0 = symbol s_first - 48
        const uint64_t w = bits_window(cur, hw, s_first + TX_GROUP - 1, n_sym);
        float *trow = tile + grp * gstride;
        if (s_first - (M - 1) >= -(long)n_hist && s_first + TX_GROUP <= n_sym) {
            // every symbol the group touches exists.  The +-A impulse (bpsk.cxx:150) times a tap is the pre-scaled tap times
            // +-1, and the sign only depends on the symbol i - m: 16 + MT - 1 sign registers shared by all phases, then ONE
            // FFMA per tap and output (the product by +-1 is exact, so this equals adding the sign-flipped tap, as the edge
            // path does).
            float sg[TX_GROUP + MT - 1];                                   // sg[d + MT - 1]: symbol s_first + d
            const uint32_t wl = (uint32_t)w, wh = (uint32_t)(w >> 32);
#pragma unroll
            for (int d = 0; d < TX_GROUP + MT - 1; d++) {
                const int bit = 49 - MT + d;                               // compile-time position in the window
                const uint32_t word = bit < 32 ? wl : wh;
                sg[d] = __uint_as_float(0x3F800000u | ((word << (31 - (bit & 31))) & 0x80000000u));
            }
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
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++) trow[i * sps + r] = acc[i];
            }
        } else {
            // first symbols of a stream (zero state before the history) and the ragged end
            // (same arithmetic as above, tap by tap, so that a stream cut into calls anywhere gives identical samples)
            for (int r = 0; r < sps; r++) {
                float acc[TX_GROUP];
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++) acc[i] = 0.f;
                for (int m = 0; m < M; m++) {
                    const uint32_t hb = __float_as_uint(hs[r * MT + m]);
#pragma unroll
                    for (int i = 0; i < TX_GROUP; i++) {
                        const long sm = s_first + i - m;                       // symbol s - m
                        const bool valid = sm >= -(long)n_hist && sm < n_sym;  // before the first symbol: zero state
                        const uint32_t sign = (uint32_t)((w >> (48 + i - m)) & 1ull) << 31;   // a set bit sends -A (bpsk.cxx:150)
                        acc[i] += valid ? __uint_as_float(hb ^ sign) : 0.f;
                    }
                }
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++) trow[i * sps + r] = acc[i];
            }
        }
    }
    __syncthreads();

    // sample n of the tile (time order) sits at tile[(n / (16 sps)) * gstride + n % (16 sps)]; a quad never straddles a group
    const int qpg = TX_GROUP * sps / 4;                    // quads per group
    const uint32_t qmagic = 0xFFFFFFFFu / (uint32_t)qpg + 1u;   // floor(q / qpg) = umulhi(q, qmagic) for q < 2^16
    auto load_quad = [&](int qd) {
        const int g = (int)__umulhi((uint32_t)qd, qmagic);
        const float2 *src = reinterpret_cast<const float2 *>(tile + g * gstride + (qd - g * qpg) * 4);
        const float2 a = src[0], b = src[1];
        return make_float4(a.x, a.y, b.x, b.y);
    };
    const long n_valid_sym = n_sym - s0 < sym_cta ? n_sym - s0 : sym_cta;
    const long n_samp = n_valid_sym * sps;                 // valid samples of this tile
    const long samp0 = s0 * sps;
    char *row0 = static_cast<char *>(out) + (long)ch * out_pitch_bytes;
    if (out_dac10) {
        // bpsk.cxx:76-101: u = ((short)(x*511) + 512) & 0x3FF, 4 words -> 5 bytes, the four 2-bit tops first.
        // Packed into a second shared-memory tile, then stored as 128-bit words.
        uint8_t *packed = reinterpret_cast<uint8_t *>(tile + n_groups * gstride);
        const long n_quads = n_samp / 4;
        for (long qd = threadIdx.x; qd < n_quads; qd += TX_THREADS) {
            const float4 x = load_quad((int)qd);
            const float v[4] = {x.x, x.y, x.z, x.w};
            unsigned u[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const short sh = (short)(int)__fmul_rn(v[k], 511.0f);       // C float->short: truncation toward zero
                u[k] = ((unsigned)(sh + 512)) & 0x3FFu;
            }
            uint8_t *o = packed + qd * 5;
            o[0] = (uint8_t)((u[0] >> 8) | ((u[1] >> 8) << 2) | ((u[2] >> 8) << 4) | ((u[3] >> 8) << 6));
            o[1] = (uint8_t)(u[0] & 0xFF);
            o[2] = (uint8_t)(u[1] & 0xFF);
            o[3] = (uint8_t)(u[2] & 0xFF);
            o[4] = (uint8_t)(u[3] & 0xFF);
        }
        __syncthreads();
        char *dst = row0 + samp0 / 4 * 5;
        const long n_bytes = n_quads * 5;
        if (((reinterpret_cast<uintptr_t>(dst) | (uintptr_t)n_bytes) & 15) == 0) {
            for (long i = threadIdx.x; i < n_bytes / 16; i += TX_THREADS)
                reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(packed)[i];
        } else {
            for (long i = threadIdx.x; i < n_bytes; i += TX_THREADS) dst[i] = (char)packed[i];
        }
    } else {
        float *dst = reinterpret_cast<float *>(row0) + samp0;
        if (((reinterpret_cast<uintptr_t>(dst) & 15) | (n_samp & 3)) == 0) {
            for (int i = threadIdx.x; i < (int)(n_samp / 4); i += TX_THREADS) reinterpret_cast<float4 *>(dst)[i] = load_quad(i);
        } else {
            const int spg = TX_GROUP * sps;
            for (int i = threadIdx.x; i < (int)n_samp; i += TX_THREADS) dst[i] = tile[(i / spg) * gstride + i % spg];
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
    // symbols per CTA: as many as keep tile + packed copy under TX_TILE_BYTES, whole groups of 4 threads' worth
    const int gstride = TX_GROUP * sps + 2;
    const size_t per_group = sizeof(float) * gstride + (size_t)TX_GROUP * sps / 4 * 5 + 8;
    int n_groups = (int)(TX_TILE_BYTES / per_group);
    if (n_groups > TX_SYM_MAX / TX_GROUP) n_groups = TX_SYM_MAX / TX_GROUP;
    n_groups = n_groups / 4 * 4;
    if (n_groups < 4) n_groups = 4;
    while (n_groups > 4 && (long)(n_groups / 2) * TX_GROUP >= n_sym) n_groups /= 2;   // short calls: no empty threads
    n_groups = (n_groups + 3) / 4 * 4;
    const int sym_cta = n_groups * TX_GROUP;
    const size_t smem = sizeof(float) * ((size_t)sps * MT + (size_t)n_groups * gstride) + (size_t)sym_cta * sps / 4 * 5 + 32;
    auto kern = MT == 4 ? bpsk_tx_kernel<4> : MT == 8 ? bpsk_tx_kernel<8> : MT == 12 ? bpsk_tx_kernel<12> : MT == 16 ? bpsk_tx_kernel<16>
                : MT == 24 ? bpsk_tx_kernel<24> : bpsk_tx_kernel<32>;
    if (!SFE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))) return SFE_ERR_CUDA;
    dim3 grid((unsigned)((n_sym + sym_cta - 1) / sym_cta), (unsigned)n_channels);
    kern<<<grid, TX_THREADS, smem, st>>>(bits, pitch_words, hist_old, n_hist, n_sym, sps, sym_cta, taps, n_taps, amp, out_dac10, out,
                                         out_pitch_bytes);
    bpsk_tx_hist_kernel<<<(n_channels + 127) / 128, 128, 0, st>>>(bits, pitch_words, hist_old, hist_new, n_sym, n_channels);
    count_launch(2);
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

}  // namespace sfe
