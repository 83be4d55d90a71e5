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
// Mapping: a CTA owns 512 symbols of one channel.  A work item is (phase r, group of 16 symbols): its ~12 taps
// h[r + m*sps] sit in registers, the symbol signs come from one 64-bit window of the bit stream, and the 16
// results go to a shared-memory tile in time order.  A second pass packs the tile and stores 128-bit words.
#include "kernels.h"

namespace sfe {

namespace {

constexpr int TX_SYM = 512;        // symbols per CTA
constexpr int TX_GROUP = 16;       // symbols per work item
constexpr int TX_MAXM = 32;        // taps per phase, at most (n_taps <= 32 * sps)
constexpr int TX_THREADS = 256;

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

template <int MT>     // taps per phase, rounded up (the padding taps are zero)
__global__ void __launch_bounds__(TX_THREADS)
bpsk_tx_kernel(const uint32_t *bits, long pitch_words, const uint64_t *hist, int n_hist, long n_sym, int sps, const float *taps,
               int n_taps, float amp, int out_dac10, void *out, long out_pitch_bytes)
{
    extern __shared__ __align__(16) float tile[];          // TX_SYM * sps samples in time order
    const int ch = blockIdx.y;
    const long s0 = (long)blockIdx.x * TX_SYM;
    const uint32_t *cur = bits + (long)ch * pitch_words;
    const uint64_t hw = hist[ch];
    const int n_items = sps * (TX_SYM / TX_GROUP);
    const int M = (n_taps + sps - 1) / sps;                // taps per phase

    for (int item = threadIdx.x; item < n_items; item += TX_THREADS) {
        const int r = item % sps, grp = item / sps;
        const long s_first = s0 + (long)grp * TX_GROUP;
        // window: bit 63 = symbol s_first + 15, ..., bit 0 = symbol s_first - 48
        const uint64_t w = bits_window(cur, hw, s_first + TX_GROUP - 1, n_sym);
        float acc[TX_GROUP];
#pragma unroll
        for (int i = 0; i < TX_GROUP; i++) acc[i] = 0.f;
        if (s_first - (M - 1) >= -(long)n_hist && s_first + TX_GROUP <= n_sym) {
            // every symbol the group touches exists.  The +-A impulse (bpsk.cxx:150) times a tap is the pre-scaled tap times
            // +-1, and the sign only depends on the symbol i - m: 16 + MT - 1 sign registers, then ONE FFMA per tap and output
            // (the product by +-1 is exact, so this equals adding the sign-flipped tap, as the edge path does).
            float sg[TX_GROUP + MT - 1];                                   // sg[d + MT - 1]: symbol s_first + d
#pragma unroll
            for (int d = 0; d < TX_GROUP + MT - 1; d++)
                sg[d] = __uint_as_float(0x3F800000u | ((uint32_t)(w >> (49 - MT + d)) << 31));
#pragma unroll
            for (int m = 0; m < MT; m++) {
                const float hm = (m < M && r + m * sps < n_taps) ? amp * __ldg(taps + r + m * sps) : 0.f;
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++) acc[i] = fmaf(sg[i - m + MT - 1], hm, acc[i]);
            }
        } else {
            // first symbols of a stream (zero state before the history) and the ragged end
            // (same arithmetic as above, tap by tap, so that a stream cut into calls anywhere gives identical samples)
            for (int m = 0; m < M; m++) {
                const uint32_t hb = __float_as_uint(r + m * sps < n_taps ? amp * __ldg(taps + r + m * sps) : 0.f);
#pragma unroll
                for (int i = 0; i < TX_GROUP; i++) {
                    const long sm = s_first + i - m;                       // symbol s - m
                    const bool valid = sm >= -(long)n_hist && sm < n_sym;  // before the first symbol: zero state
                    const uint32_t sign = (uint32_t)((w >> (48 + i - m)) & 1ull) << 31;   // a set bit sends -A (bpsk.cxx:150)
                    acc[i] += valid ? __uint_as_float(hb ^ sign) : 0.f;
                }
            }
        }
#pragma unroll
        for (int i = 0; i < TX_GROUP; i++) tile[(grp * TX_GROUP + i) * sps + r] = acc[i];
    }
    __syncthreads();

    const long n_valid_sym = n_sym - s0 < TX_SYM ? n_sym - s0 : TX_SYM;
    const long n_samp = n_valid_sym * sps;                 // valid samples of this tile
    const long samp0 = s0 * sps;
    char *row0 = static_cast<char *>(out) + (long)ch * out_pitch_bytes;
    if (out_dac10) {
        // bpsk.cxx:76-101: u = ((short)(x*511) + 512) & 0x3FF, 4 words -> 5 bytes, the four 2-bit tops first.
        // Packed into a second shared-memory tile, then stored as 128-bit words.
        uint8_t *packed = reinterpret_cast<uint8_t *>(tile + TX_SYM * sps);
        const long n_quads = n_samp / 4;
        for (long qd = threadIdx.x; qd < n_quads; qd += TX_THREADS) {
            const float4 x = reinterpret_cast<const float4 *>(tile)[qd];
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
            for (long i = threadIdx.x; i < n_samp / 4; i += TX_THREADS)
                reinterpret_cast<float4 *>(dst)[i] = reinterpret_cast<const float4 *>(tile)[i];
        } else {
            for (long i = threadIdx.x; i < n_samp; i += TX_THREADS) dst[i] = tile[i];
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
    const size_t smem = sizeof(float) * TX_SYM * sps + (size_t)TX_SYM * sps / 4 * 5 + 16;
    const int M = (n_taps + sps - 1) / sps;
    auto kern = M <= 4 ? bpsk_tx_kernel<4> : M <= 8 ? bpsk_tx_kernel<8> : M <= 12 ? bpsk_tx_kernel<12> : M <= 16 ? bpsk_tx_kernel<16>
                : M <= 24 ? bpsk_tx_kernel<24> : bpsk_tx_kernel<32>;
    if (!SFE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))) return SFE_ERR_CUDA;
    dim3 grid((unsigned)((n_sym + TX_SYM - 1) / TX_SYM), (unsigned)n_channels);
    kern<<<grid, TX_THREADS, smem, st>>>(bits, pitch_words, hist_old, n_hist, n_sym, sps, taps, n_taps, amp, out_dac10, out, out_pitch_bytes);
    bpsk_tx_hist_kernel<<<(n_channels + 127) / 128, 128, 0, st>>>(bits, pitch_words, hist_old, hist_new, n_sym, n_channels);
    count_launch(2);
    return SFE_CUDA(cudaGetLastError()) ? SFE_OK : SFE_ERR_CUDA;
}

}  // namespace sfe
