// kernels.h -- launch interfaces of the libsfegpu kernels (internal; the public ABI is include/sfe_gpu.h)
#pragma once
#include "schedule.h"
#include "sfe_common.cuh"

namespace sfe {

// ---- K1: FFT block convolver (blkconv_kernel.cu) ------------------------------------------------
enum { BLKCONV_MODE_CONV = 0, BLKCONV_MODE_SPECTRUM = 1 };
struct BlkconvParams {
    StreamView in;        // raw input (+history of ov samples)
    int in_real;          // 1: input/output are float (one real stream per row), else complex
    void *out;            // complex64 (or float) rows
    long out_pitch;
    int n_streams;
    long n_blocks;        // FFT blocks per stream = ceil(n / hop)
    int hop;              // fft_len + 1 - n_taps (blkconv.cxx:47)
    int ov;               // n_taps - 1          (blkconv.cxx:48)
    const float2 *tw;     // per-pass twiddle rows (blkconv_twiddle_table)
    const float2 *hp;     // tap spectrum, kernel (digit-reversed) order, pre-scaled by 1/N
    int mode;
    float scale;          // spectrum mode: 1/N (blkconv.cxx:50)
    // Interpolating epilogue (chain.cu: a one-phase fractional-rate stage behind blkconv): the block's convolved samples
    // stay in shared memory and the kernel writes only out[k] = (1-mu_k) s[pos_k] + mu_k s[pos_k + 1] for the schedule
    // entries k in [lerp_kidx[b], lerp_kidx[b+1]) whose pos lies in block b.  Blocks then start at blk_start + b*hop with
    // hop = fft_len - ov - 1 (one sample of overlap more, for s[pos + 1]) and blk_start = -1 (a held-over output).
    long blk_start;
    const SchedEntry *lerp_sched;   // nullptr: plain convolution output
    const int *lerp_kidx;           // n_blocks + 1 entries
};
int launch_blkconv(int fft_len, const BlkconvParams &p, cudaStream_t st);
bool blkconv_twiddle_table(int fft_len, std::vector<float2> &host);

// ---- K2g: generic polyphase stage, one thread per output (polyphase_kernel.cu) ------------------
struct PolyParams {
    StreamView in;
    int in_real;          // 1: float real rows in / float out (single-object API)
    void *out;
    long out_pitch;
    int n_streams;
    long n_out;
    const float *taps;    // device, n_taps floats (already padded like the reference ctor)
    int n_taps;
    int up;
    int uniform;          // 1: pos = pos0 + k*step, mu = 0
    long pos0, step;
    const SchedEntry *sched;  // device, n_out entries when !uniform
    int exact;            // reference operation order (separate multiply/add)
    float rate;           // the stage's rate (input samples per output): sizes the tile when the schedule is a table
    float lerp_tap;       // != 0: the stage has this single tap (n_taps == 1, up == 1): pure interpolation kernel
    // filled by launch_poly_generic
    int kt, jp, tap_bytes, span_cap;
    int cw_bytes, grp_rows;   // K2s: bytes of combined taps per warp; rows of the grouped flavour's tap table (0 = one output at a time)
};
int launch_poly_generic(const PolyParams &p, cudaStream_t st);

// ---- K2d: decimating FIR on CUDA cores, register tiled (polyphase_kernel.cu) --------------------
// y[k] = sum_j g[j] * x[n_first + D*k - j]
struct DecimParams {
    StreamView in;
    float2 *out;
    long out_pitch;
    int n_streams;
    long n_out;
    const float *gp;      // device: plane-major taps [D][QLP], gp[r][q] = g[D*q + r] (zero padded)
    int D, QLP;
    long n_first;
};
int launch_decim_fir(const DecimParams &p, cudaStream_t st);

// ---- K2t: decimate-by-8 FIR as an int8 Toeplitz contraction on tcgen05 (toeplitz_mma_kernel.cu) ---
struct ToepParams {
    StreamView in;        // int16 IQ rows
    float2 *out;
    long out_pitch;
    int n_streams;
    long n_out;
    long n_first;         // y[k] = sum_j g[j] x[n_first + 8k - j]
    int KB;               // 128-byte K-blocks per window = ceil((T-1)/128) + 1
    const int8_t *bmat;   // device image of the B operand tiles (toeplitz_prepare)
    float w[3];           // digit weights (include max|g| / Zmax and the input scale 1/32767 or 1/127)
    float corr[3];        // 128 * sum_j d_tau[j]
    // filled by launch_toeplitz
    int ring, n_groups;
    int n_stage, use_tma; // staging ring depth; 1: the raw rows are staged by TMA
    long n_steps;
};
struct ToepHost {
    int KB = 0;
    float w[3] = {0, 0, 0}, corr[3] = {0, 0, 0};
    std::vector<int8_t> bmat;
};
bool toeplitz_prepare(const std::vector<double> &g, ToepHost &h, int in_fmt = SFE_IN_I16);
int toeplitz_max_taps();
int launch_toeplitz(const ToepParams &p, cudaStream_t st);

// ---- K3: fused polyphase bank over many int16 IQ streams (fused_bank_kernel.cu) ------------------
// y[k] = sum_u c_ph[u] x[R_k - u];  pos = pos0 + k*step, M = floor(pos/L), ph = pos mod L, R_k = n_first + D*M
struct BankParams {
    StreamView in;        // int16 or u8 IQ rows, 16-byte aligned, pitch a multiple of 16 bytes; history of H samples
    float2 *out;
    long out_pitch;
    int n_streams;
    long n_out;
    long pos0, step;      // output schedule on the L-times upsampled grid of the (virtual) decimated stream
    int L, D;
    long n_first;         // raw index of decimated sample 0 of this call
    int H;                // history depth in samples, multiple of 128 bytes
    const float *table;   // device: n_rows rows of row_len floats (bank_table)
    int row_len, n_rows;  // row_len = 4 * (nc + 2 * pad chunks): bank_row_len()
    int nc;               // chunks (4 samples) of taps per row
    int n_live;           // window slots per lane: the most windows that ever overlap one 32-sample box
    int rows_per_phase;   // 1: one alignment per call (D % spc == 0), spc: row = ph*spc + alignment
    // filled by launch_bank
    int n_groups, n_seg;
    float2 *parts;        // seam exchange: two parts of every output whose window straddles two work items
    int *seam_cnt;        // arrivals per seam (zero between launches)
    int fits32;           // every position of the call fits 31 bits: 32-bit index arithmetic in the kernel
    void *hist_new;       // not null: the kernel also performs the state carry (last H samples of hist ++ cur) into this buffer
};
int bank_nc(int Tc, int spc);         // 16-byte chunks (spc samples each) of taps for a composite filter of Tc taps
int bank_row_len(int Tc, int spc);    // floats per table row, zero pads included
void bank_table(const std::vector<double> &g, int D, const std::vector<float> &r, int L, double scale, int e_only, int spc,
                std::vector<float> &rows, int *row_len, int *n_rows, int *Tc);
size_t bank_smem_bytes(int n_rows, int row_len, int fmt);
int bank_max_live();
int launch_bank(const BankParams &p, cudaStream_t st);

// ---- batched BPSK transmit: bits -> RRC-shaped samples -> float or 10-bit DAC words (bpsk_tx_kernel.cu) ----
int bpsk_tx_max_taps(int sps);
int launch_bpsk_tx(const uint32_t *bits, long pitch_words, const uint64_t *hist_old, uint64_t *hist_new, int n_hist, long n_sym, int sps,
                   const float *taps, int n_taps, float amp, int n_channels, int out_dac10, void *out, long out_pitch_bytes,
                   cudaStream_t st);

// ---- small utilities (util_kernels.cu) -----------------------------------------------------------
// hist_new[s][0..H) <- last H samples of (hist_old ++ cur[0..n)) for every stream row; elem = bytes per sample
// overlap_prev: the kernel launched just before on `st` neither writes `cur`/`hist_old` nor touches `hist_new`, so this one
// may start as soon as that kernel allows it (programmatic dependent launch) instead of after its last CTA
int launch_save_tail(void *hist_new, const void *hist_old, const void *cur, long pitch, long n, int H, int elem, int n_streams,
                     cudaStream_t st, bool overlap_prev = false);
int launch_pack_dac10(const float *in, long n, uint8_t *out, cudaStream_t st);
// d_out[run.k0 + j] <- entry j of every run (schedule.h: SchedRun)
int launch_expand_sched(const SchedRun *d_runs, int n_runs, SchedEntry *d_out, cudaStream_t st);
// kidx[b] = first entry whose pos >= blk_start + b*hop, b = 0 .. n_blocks (entries sorted by pos)
int launch_block_index(const SchedEntry *d_sched, long n_out, long blk_start, int hop, long n_blocks, int *d_kidx, cudaStream_t st);

}  // namespace sfe
