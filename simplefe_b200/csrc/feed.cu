// feed.cu -- sfe_rx_feed_*: the streaming front door.
//
// In the reference a device callback (libsimpleFE/simpleFE.h:42 sfe_callback, registered with sfe_rx_start,
// simpleFE.h:59-62) writes raw I/Q bytes into ring_buffer<unsigned char> under a mutex, prints "O" when the ring is
// full (gr-simplefe/lib/source_c_impl.cc:90-108), and a consumer waits on a condition variable, reads the ring
// through a converting callback and hands floats on (source_c_impl.cc:134-153).  Here the callback is the same,
// the ring is page-locked, and the consumer is the GPU: sfe_ring_read_dev moves whole blocks of bytes to the
// device asynchronously and the chain's first kernel converts them while loading (SFE_IN_U8).
#include <chrono>
#include <cstring>
#include <condition_variable>
#include <mutex>

#include "kernels.h"

using namespace sfe;

struct sfe_rx_feed {
    sfe_chain *chain = nullptr;
    sfe_ring *ring = nullptr;
    int block = 0;                 // complex samples per chain call
    long out_cap = 0;              // outputs per block, upper bound
    void *d_in[2] = {nullptr, nullptr};
    void *d_out = nullptr;
    float2 *h_out = nullptr;       // pinned staging of one block's outputs
    cudaStream_t st = nullptr;
    int flip = 0;
    std::mutex mu;
    std::condition_variable cv;
    long overflows = 0, discarded = 0, bytes_in = 0;
};

extern "C" sfe_rx_feed *sfe_rx_feed_create(const sfe_chain_desc *desc, int ring_bytes, int block_samples)
{
    if (!desc || desc->n_streams != 1 || desc->in_fmt != SFE_IN_U8 || block_samples < 1 || ring_bytes < 2 * block_samples) {
        set_error("rx_feed: needs a one-stream SFE_IN_U8 chain, block_samples >= 1 and a ring of at least one block (2 bytes per sample)");
        return nullptr;
    }
    sfe_rx_feed *f = new sfe_rx_feed;
    f->block = block_samples;
    f->chain = sfe_chain_create(desc);
    f->ring = f->chain ? sfe_ring_create(1, ring_bytes) : nullptr;
    bool ok = f->chain && f->ring;
    if (ok) {
        f->out_cap = sfe_chain_max_out(f->chain, block_samples);
        ok = SFE_CUDA(cudaStreamCreateWithFlags(&f->st, cudaStreamNonBlocking)) &&
             SFE_CUDA(cudaMalloc(&f->d_in[0], 2 * (size_t)block_samples + 16)) &&
             SFE_CUDA(cudaMalloc(&f->d_in[1], 2 * (size_t)block_samples + 16)) &&
             SFE_CUDA(cudaMalloc(&f->d_out, sizeof(float2) * (size_t)f->out_cap)) &&
             SFE_CUDA(cudaHostAlloc((void **)&f->h_out, sizeof(float2) * (size_t)f->out_cap, cudaHostAllocDefault));
    }
    if (!ok) {
        sfe_rx_feed_destroy(f);
        return nullptr;
    }
    return f;
}

extern "C" void sfe_rx_feed_destroy(sfe_rx_feed *f)
{
    if (!f) return;
    if (f->st) cudaStreamSynchronize(f->st);
    sfe_ring_destroy(f->ring);
    sfe_chain_destroy(f->chain);
    cudaFree(f->d_in[0]);
    cudaFree(f->d_in[1]);
    cudaFree(f->d_out);
    if (f->h_out) cudaFreeHost(f->h_out);
    if (f->st) cudaStreamDestroy(f->st);
    delete f;
}

// sfe_callback-compatible producer: source_c_impl::rx_callback / write_data, source_c_impl.cc:84-108
extern "C" int sfe_rx_feed_callback(unsigned char *buffer, int length, void *feed)
{
    sfe_rx_feed *f = static_cast<sfe_rx_feed *>(feed);
    if (!f || length <= 0) return 0;
    if (length & 1) {
        printf("odd number!!!, packet corruption, discard\n");   // source_c_impl.cc:93-96
        std::lock_guard<std::mutex> lk(f->mu);
        f->discarded++;
        return 0;
    }
    std::lock_guard<std::mutex> lk(f->mu);
    if (!sfe_ring_write(f->ring, buffer, length)) {
        fputc('O', stderr);                                      // source_c_impl.cc:98-100
        fflush(stderr);
        f->overflows++;
    } else {
        f->bytes_in += length;
        f->cv.notify_one();
    }
    return 0;
}

// consumer: every whole block that is buffered goes ring -> device -> chain -> out.  Waits up to wait_ms for the
// first block (the reference's work() waits on its condition variable, source_c_impl.cc:142-147).
extern "C" long sfe_rx_feed_pump(sfe_rx_feed *f, void *out, long cap, int wait_ms)
{
    if (!f || !out || cap < 0) return SFE_ERR_ARG;
    const int need = 2 * f->block;
    long written = 0;
    float2 *dst = static_cast<float2 *>(out);
    bool first = true;
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(f->mu);
            if (first && wait_ms > 0)
                f->cv.wait_for(lk, std::chrono::milliseconds(wait_ms), [&] { return sfe_ring_count(f->ring) >= need; });
            first = false;
            if (sfe_ring_count(f->ring) < need || cap - written < f->out_cap) break;
            if (sfe_ring_read_dev(f->ring, f->d_in[f->flip], need, f->st) != need) return SFE_ERR_CUDA;
        }
        long n_out = 0;
        int rc = sfe_chain_process_dev(f->chain, f->d_in[f->flip], f->block, f->block, f->d_out, f->out_cap, &n_out, f->st);
        if (rc != SFE_OK) return rc;
        f->flip ^= 1;
        if (n_out > 0) {
            if (!SFE_CUDA(cudaMemcpyAsync(f->h_out, f->d_out, sizeof(float2) * n_out, cudaMemcpyDeviceToHost, f->st)) ||
                !SFE_CUDA(cudaStreamSynchronize(f->st)))
                return SFE_ERR_CUDA;
            memcpy(dst + written, f->h_out, sizeof(float2) * n_out);
            written += n_out;
        }
    }
    if (!SFE_CUDA(cudaStreamSynchronize(f->st))) return SFE_ERR_CUDA;
    return written;
}

extern "C" long sfe_rx_feed_stat(sfe_rx_feed *f, int which)
{
    if (!f) return SFE_ERR_ARG;
    std::lock_guard<std::mutex> lk(f->mu);
    switch (which) {
        case 0: return f->overflows;
        case 1: return f->discarded;
        case 2: return f->bytes_in;
        case 3: return sfe_ring_count(f->ring);
        case 4: return f->out_cap;
        default: return SFE_ERR_ARG;
    }
}

// =====================================================================================================================
// sfe_tx_feed_*: the transmit side (sink blocks).
//
// In the reference, sink_c/sink_f::work() write float samples into ring_buffer<T> under a mutex, waiting on a condition
// variable for room (gr-simplefe/lib/sink_c_impl.cc:158-174), and the USB thread's tx_callback drains the ring through
// fill_tx_buffer, which packs 4 samples into 5 bytes of 10-bit DAC words, or sends zeros and prints "U" when the ring
// runs dry (sink_c_impl.cc:84-144; the same packing in examples/bpsk/bpsk.cxx:76-101).  Here the packing runs on the
// GPU when the samples are written (sfe_pack_dac10 kernel) and the ring holds the finished wire bytes in pinned memory:
// the USB thread's callback is a memcpy.
namespace sfe {
bool chain_device_ready(int *dev);
}
static int copy_bytes(void *dst, void *src, int n)
{
    memcpy(dst, src, (size_t)n);
    return n;
}
static int same_len(int n) { return n; }

struct sfe_tx_feed {
    sfe_ring *ring = nullptr;      // packed DAC bytes
    int block = 0;                 // most floats per write
    float *h_in = nullptr;         // pinned staging: carry (<= 3 floats) + one write
    uint8_t *h_out = nullptr;      // pinned staging of the packed bytes
    float *d_in = nullptr;
    uint8_t *d_out = nullptr;
    float carry[4];
    int n_carry = 0;               // floats of an unfinished quad kept for the next write
    cudaStream_t st = nullptr;
    std::mutex mu;
    std::condition_variable cv;    // room was made (the reference's m_buf_cond)
    long underflows = 0, bytes_out = 0;
};

extern "C" sfe_tx_feed *sfe_tx_feed_create(int ring_bytes, int block_floats)
{
    if (block_floats < 4 || ring_bytes < (block_floats / 4 + 1) * 5) {
        set_error("tx_feed: block_floats >= 4 and a ring of at least one packed block (5 bytes per 4 samples)");
        return nullptr;
    }
    int dev = 0;
    if (!chain_device_ready(&dev)) return nullptr;
    sfe_tx_feed *f = new sfe_tx_feed;
    f->block = block_floats;
    f->ring = sfe_ring_create(1, ring_bytes);
    const size_t nf = (size_t)block_floats + 4, nb = nf / 4 * 5;
    const bool ok = f->ring && SFE_CUDA(cudaStreamCreateWithFlags(&f->st, cudaStreamNonBlocking)) &&
                    SFE_CUDA(cudaHostAlloc((void **)&f->h_in, sizeof(float) * nf, cudaHostAllocDefault)) &&
                    SFE_CUDA(cudaHostAlloc((void **)&f->h_out, nb, cudaHostAllocDefault)) &&
                    SFE_CUDA(cudaMalloc(&f->d_in, sizeof(float) * nf)) && SFE_CUDA(cudaMalloc(&f->d_out, nb));
    if (!ok) {
        sfe_tx_feed_destroy(f);
        return nullptr;
    }
    return f;
}

extern "C" void sfe_tx_feed_destroy(sfe_tx_feed *f)
{
    if (!f) return;
    if (f->st) cudaStreamSynchronize(f->st);
    sfe_ring_destroy(f->ring);
    if (f->h_in) cudaFreeHost(f->h_in);
    if (f->h_out) cudaFreeHost(f->h_out);
    cudaFree(f->d_in);
    cudaFree(f->d_out);
    if (f->st) cudaStreamDestroy(f->st);
    delete f;
}

// producer: the role of sink_c_impl::work (sink_c_impl.cc:158-174).  n floats (a complex item is two); whole quads
// are packed on the GPU and appended to the ring, up to 3 floats wait for the next write.  Blocks up to wait_ms for
// room (the reference waits on its condition variable without limit); returns n, or 0 when there was no room.
extern "C" long sfe_tx_feed_write(sfe_tx_feed *f, const float *x, long n, int wait_ms)
{
    if (!f || !x || n < 0 || n > f->block) {
        set_error("tx_feed_write: 0 <= n <= block_floats");
        return SFE_ERR_ARG;
    }
    const long tot = f->n_carry + n, quads = tot / 4, nb = quads * 5;
    {
        std::unique_lock<std::mutex> lk(f->mu);
        auto room = [&] { return sfe_ring_space(f->ring) >= nb; };
        if (!room() && wait_ms > 0) f->cv.wait_for(lk, std::chrono::milliseconds(wait_ms), room);
        if (!room()) return 0;
    }
    memcpy(f->h_in, f->carry, sizeof(float) * f->n_carry);
    memcpy(f->h_in + f->n_carry, x, sizeof(float) * n);
    if (quads > 0) {
        if (!SFE_CUDA(cudaMemcpyAsync(f->d_in, f->h_in, sizeof(float) * 4 * quads, cudaMemcpyHostToDevice, f->st))) return SFE_ERR_CUDA;
        const int rc = launch_pack_dac10(f->d_in, 4 * quads, f->d_out, f->st);
        if (rc != SFE_OK) return rc;
        if (!SFE_CUDA(cudaMemcpyAsync(f->h_out, f->d_out, nb, cudaMemcpyDeviceToHost, f->st)) || !SFE_CUDA(cudaStreamSynchronize(f->st)))
            return SFE_ERR_CUDA;
        std::lock_guard<std::mutex> lk(f->mu);
        if (sfe_ring_write(f->ring, f->h_out, (int)nb) != nb) {     // single producer: the room checked above is still there
            set_error("tx_feed_write: ring refused a write it had room for (second producer?)");
            return SFE_ERR_ARG;
        }
    }
    f->n_carry = (int)(tot - 4 * quads);
    memcpy(f->carry, f->h_in + 4 * quads, sizeof(float) * f->n_carry);
    return n;
}

// consumer: sfe_callback signature, sink_c_impl::tx_callback / data_request (sink_c_impl.cc:84-104)
extern "C" int sfe_tx_feed_callback(unsigned char *buffer, int length, void *feed)
{
    sfe_tx_feed *f = static_cast<sfe_tx_feed *>(feed);
    if (!f || !buffer || length <= 0) return 0;
    const int nb = length / 5 * 5;                       // the reference converts length/5 quads (calc_read_len, :107-111)
    std::lock_guard<std::mutex> lk(f->mu);
    if (sfe_ring_count(f->ring) < nb) {
        memset(buffer, 0, (size_t)length);               // :95-97
        fputc('U', stderr);
        fflush(stderr);
        f->underflows++;
    } else {
        sfe_ring_read(f->ring, buffer, (unsigned)nb, copy_bytes, same_len);
        f->bytes_out += nb;
        f->cv.notify_one();
    }
    return 0;
}

extern "C" long sfe_tx_feed_stat(sfe_tx_feed *f, int which)
{
    if (!f) return SFE_ERR_ARG;
    std::lock_guard<std::mutex> lk(f->mu);
    switch (which) {
        case 0: return f->underflows;
        case 1: return f->bytes_out;
        case 2: return sfe_ring_count(f->ring);
        case 3: return (long)sfe_ring_space(f->ring) / 5 * 4 - f->n_carry;   // floats a write can take now
        default: return SFE_ERR_ARG;
    }
}
