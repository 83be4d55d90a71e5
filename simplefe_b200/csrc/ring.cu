// ring.cu -- sfe_ring_*: the GPU-side successor of ring_buffer<T> (libdsp/ringbuf.h:37-142).
//
// Same contract as the reference template (all-or-nothing write, read through a length callback
// and a converting copy callback that is invoked once or twice around the wrap point, no internal
// locking: callers serialise, as bpsk.cxx:108-116,132-170 does with its mutex), but the storage is
// page-locked host memory so that the consumer can be a CUDA copy engine: sfe_ring_read_dev moves
// the next n elements to device memory with at most two cudaMemcpyAsync and returns at once; the
// elements stay reserved (not writable) until the copy's event has completed.
#include <cstring>
#include <deque>

#include "kernels.h"

namespace sfe {
bool chain_device_ready(int *dev);
}
using namespace sfe;

struct sfe_ring {
    int dev = 0;
    int elem = 0;            // bytes per element (sizeof(T))
    int cap = 0;             // capacity in elements
    int count = 0;           // readable elements
    int rd = 0, wr = 0;      // positions in elements
    int inflight = 0;        // elements handed to an async device copy that has not completed yet
    char *buf = nullptr;     // cudaHostAlloc
    std::deque<std::pair<cudaEvent_t, int>> pending;
};

namespace {

// give back the space of device reads whose copies have finished
void reclaim(sfe_ring *r, bool block)
{
    while (!r->pending.empty()) {
        auto &p = r->pending.front();
        cudaError_t e = block ? cudaEventSynchronize(p.first) : cudaEventQuery(p.first);
        if (e == cudaErrorNotReady) {
            (void)cudaGetLastError();
            break;
        }
        r->inflight -= p.second;   // on a real error the copy is dead as well: release the space
        cudaEventDestroy(p.first);
        r->pending.pop_front();
    }
}

}  // namespace

extern "C" sfe_ring *sfe_ring_create(int elem_bytes, int capacity)
{
    if (elem_bytes < 1 || capacity < 0) {
        set_error("ring: need elem_bytes >= 1 and capacity >= 0");
        return nullptr;
    }
    int dev = 0;
    if (!chain_device_ready(&dev)) return nullptr;
    sfe_ring *r = new sfe_ring;
    r->dev = dev;
    r->elem = elem_bytes;
    if (capacity > 0 && sfe_ring_alloc(r, capacity) != SFE_OK) {
        delete r;
        return nullptr;
    }
    return r;
}

// ring_buffer::alloc_buffer  ringbuf.h:60-69: new storage, everything buffered so far is dropped
extern "C" int sfe_ring_alloc(sfe_ring *r, int capacity)
{
    if (!r || capacity < 1) return SFE_ERR_ARG;
    reclaim(r, true);
    if (r->buf) cudaFreeHost(r->buf);
    r->buf = nullptr;
    r->cap = r->count = r->rd = r->wr = r->inflight = 0;
    if (!SFE_CUDA(cudaSetDevice(r->dev)) ||
        !SFE_CUDA(cudaHostAlloc((void **)&r->buf, (size_t)capacity * r->elem, cudaHostAllocDefault)))
        return SFE_ERR_CUDA;
    r->cap = capacity;
    return SFE_OK;
}

extern "C" void sfe_ring_destroy(sfe_ring *r)
{
    if (!r) return;
    reclaim(r, true);
    if (r->buf) cudaFreeHost(r->buf);
    delete r;
}

// ringbuf.h:71-78
extern "C" int sfe_ring_space(sfe_ring *r)
{
    if (!r) return 0;
    reclaim(r, false);
    return r->cap - r->count - r->inflight;
}
extern "C" int sfe_ring_count(sfe_ring *r) { return r ? r->count : 0; }

// ring_buffer::write  ringbuf.h:80-101: all or nothing, returns len or 0
extern "C" int sfe_ring_write(sfe_ring *r, const void *data, int len)
{
    if (!r || !data || len < 0) return 0;
    if (len > sfe_ring_space(r)) {
        reclaim(r, true);                       // space held by copies in flight: wait once, then decide
        if (len > r->cap - r->count - r->inflight) return 0;
    }
    const int first = r->cap - r->wr;
    const char *src = static_cast<const char *>(data);
    if (first < len) {
        memcpy(r->buf + (size_t)r->wr * r->elem, src, (size_t)first * r->elem);
        memcpy(r->buf, src + (size_t)first * r->elem, (size_t)(len - first) * r->elem);
        r->wr = len - first;
    } else {
        memcpy(r->buf + (size_t)r->wr * r->elem, src, (size_t)len * r->elem);
        r->wr += len;   // may equal cap, exactly like the reference (the next write then wraps with first == 0)
    }
    r->count += len;
    return len;
}

// ring_buffer::read  ringbuf.h:106-134: consumes calc_src_len(dst_len) elements through conv
extern "C" int sfe_ring_read(sfe_ring *r, void *dst, unsigned dst_len, sfe_ring_conv_fn conv, sfe_ring_len_fn calc_src_len)
{
    if (!r || !conv || !calc_src_len) return 0;
    const int n = calc_src_len((int)dst_len);
    if (n > r->count) return 0;
    const int first = r->cap - r->rd;
    if (first < n) {
        const int wrote = conv(dst, r->buf + (size_t)r->rd * r->elem, first);
        conv(static_cast<char *>(dst) + wrote, r->buf, n - first);
        r->rd = n - first;
    } else {
        conv(dst, r->buf + (size_t)r->rd * r->elem, n);
        r->rd += n;
    }
    r->count -= n;
    return n;
}

// additive: the next n elements -> device memory, asynchronously on `stream`; returns n or 0
extern "C" int sfe_ring_read_dev(sfe_ring *r, void *d_dst, int n, void *stream)
{
    if (!r || !d_dst || n < 0) return 0;
    if (n > r->count) return 0;
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (r->rd >= r->cap) r->rd = 0;
    const int first = r->cap - r->rd;
    const int n1 = first < n ? first : n;
    bool ok = SFE_CUDA(cudaMemcpyAsync(d_dst, r->buf + (size_t)r->rd * r->elem, (size_t)n1 * r->elem, cudaMemcpyHostToDevice, st));
    if (ok && n1 < n)
        ok = SFE_CUDA(cudaMemcpyAsync(static_cast<char *>(d_dst) + (size_t)n1 * r->elem, r->buf, (size_t)(n - n1) * r->elem,
                                      cudaMemcpyHostToDevice, st));
    cudaEvent_t ev = nullptr;
    ok = ok && SFE_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) && SFE_CUDA(cudaEventRecord(ev, st));
    if (!ok) {
        if (ev) cudaEventDestroy(ev);
        cudaStreamSynchronize(st);   // nothing may still be reading the ring when we report failure
        return 0;
    }
    r->rd = n1 < n ? n - n1 : r->rd + n;
    r->count -= n;
    r->inflight += n;
    r->pending.emplace_back(ev, n);
    return n;
}
