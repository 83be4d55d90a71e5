// feed.cu -- sfe_rx_feed_*: the streaming front door.
//
// In the reference a device callback (libsimpleFE/simpleFE.h:42 sfe_callback, registered with sfe_rx_start,
// simpleFE.h:59-62) writes raw I/Q bytes into ring_buffer<unsigned char> under a mutex, prints "O" when the ring is
// full (gr-simplefe/lib/source_c_impl.cc:90-108), and a consumer waits on a condition variable, reads the ring
// through a converting callback and hands floats on (source_c_impl.cc:134-153).  Here the callback is the same,
// the ring is page-locked, and the consumer is the GPU: sfe_ring_read_dev moves whole blocks of bytes to the
// device asynchronously and the chain's first kernel converts them while loading (SFE_IN_U8).
#include <chrono>
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
