// bpsk_tx.cu -- sfe_bpsk_tx_*: host side of the batched BPSK transmit path (include/sfe_gpu.h section 4).
#include <vector>

#include "kernels.h"

namespace sfe {
bool chain_device_ready(int *dev);
}
using namespace sfe;

struct sfe_bpsk_tx {
    int dev = 0, n_taps = 0, sps = 0, C = 0;
    float amp = 0.f;
    float *d_taps = nullptr;
    uint64_t *d_hist[2] = {nullptr, nullptr};   // last 64 symbols per channel, double-buffered
    int cur = 0;
    long n_seen = 0;                            // symbols sent so far (zero state before them)
};

extern "C" sfe_bpsk_tx *sfe_bpsk_tx_create(const float *taps, int n_taps, int sps, float amp, int n_channels)
{
    if (!taps || n_taps < 1 || sps < 1 || n_channels < 1 || n_taps > bpsk_tx_max_taps(sps) || sps > 64 ||
        n_channels > 65535) {
        set_error("bpsk_tx: need taps, sps in [1, 64], 1 <= n_channels <= 65535 and at most 32 taps per phase");
        return nullptr;
    }
    int dev = 0;
    if (!chain_device_ready(&dev)) return nullptr;
    sfe_bpsk_tx *t = new sfe_bpsk_tx;
    t->dev = dev;
    t->n_taps = n_taps;
    t->sps = sps;
    t->C = n_channels;
    t->amp = amp;
    bool ok = SFE_CUDA(cudaMalloc(&t->d_taps, sizeof(float) * n_taps)) &&
              SFE_CUDA(cudaMemcpy(t->d_taps, taps, sizeof(float) * n_taps, cudaMemcpyHostToDevice));
    for (int i = 0; i < 2 && ok; i++)
        ok = SFE_CUDA(cudaMalloc(&t->d_hist[i], sizeof(uint64_t) * n_channels)) &&
             SFE_CUDA(cudaMemset(t->d_hist[i], 0, sizeof(uint64_t) * n_channels));
    if (!ok) {
        sfe_bpsk_tx_destroy(t);
        return nullptr;
    }
    return t;
}

extern "C" void sfe_bpsk_tx_destroy(sfe_bpsk_tx *t)
{
    if (!t) return;
    cudaSetDevice(t->dev);
    cudaDeviceSynchronize();
    cudaFree(t->d_taps);
    cudaFree(t->d_hist[0]);
    cudaFree(t->d_hist[1]);
    delete t;
}

extern "C" int sfe_bpsk_tx_reset(sfe_bpsk_tx *t)
{
    if (!t) return SFE_ERR_ARG;
    if (!SFE_CUDA(cudaSetDevice(t->dev)) || !SFE_CUDA(cudaDeviceSynchronize())) return SFE_ERR_CUDA;
    t->n_seen = 0;
    for (int i = 0; i < 2; i++)
        if (!SFE_CUDA(cudaMemset(t->d_hist[i], 0, sizeof(uint64_t) * t->C))) return SFE_ERR_CUDA;
    return SFE_OK;
}

extern "C" int sfe_bpsk_tx_process_dev(sfe_bpsk_tx *t, const uint32_t *d_bits, long n_sym, long pitch_words, void *d_out,
                                       long out_pitch, int out_fmt, void *stream)
{
    if (!t || !d_bits || !d_out || n_sym < 0 || pitch_words * 32 < n_sym || (out_fmt != SFE_TX_F32 && out_fmt != SFE_TX_DAC10)) {
        set_error("bpsk_tx_process_dev: bad arguments");
        return SFE_ERR_ARG;
    }
    const long n_samp = n_sym * t->sps;
    if (out_fmt == SFE_TX_DAC10 && (n_samp & 3)) {
        set_error("bpsk_tx_process_dev: DAC words pack 4 samples into 5 bytes, n_sym * sps must be a multiple of 4");
        return SFE_ERR_ARG;
    }
    const long row_bytes = out_fmt == SFE_TX_DAC10 ? n_samp / 4 * 5 : n_samp * 4;
    if (out_pitch < row_bytes || (out_fmt == SFE_TX_F32 && (out_pitch & 3))) {
        set_error("bpsk_tx_process_dev: out_pitch (bytes) smaller than one output row");
        return SFE_ERR_CAP;
    }
    if (n_sym == 0) return SFE_OK;
    if (!SFE_CUDA(cudaSetDevice(t->dev))) return SFE_ERR_CUDA;
    const int n_hist = (int)(t->n_seen < 64 ? t->n_seen : 64);
    const int rc = launch_bpsk_tx(d_bits, pitch_words, t->d_hist[t->cur], t->d_hist[t->cur ^ 1], n_hist, n_sym, t->sps, t->d_taps,
                                  t->n_taps, t->amp, t->C, out_fmt == SFE_TX_DAC10, d_out, out_pitch, (cudaStream_t)stream);
    if (rc != SFE_OK) return rc;
    t->cur ^= 1;
    t->n_seen += n_sym;
    return SFE_OK;
}
