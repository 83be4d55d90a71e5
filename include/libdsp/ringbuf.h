/*
 * ringbuf.h -- drop-in for libdsp/ringbuf.h:37-142 (and gr-simplefe/lib/ringbuf.h:40-145, which is the
 * same template inside namespace gr::simplefe: define SFE_RINGBUF_GR_NAMESPACE before including).
 *
 *   ring_buffer(); ring_buffer(int capacity); void alloc_buffer(int capacity);      :40-69
 *   int get_space(); int get_count();                                               :71-78
 *   int write(const T *data, int len);              all or nothing, len or 0        :80-101
 *   int read(void *dst, unsigned dst_len, int (*conv)(void *dst, void *src, int src_len),
 *            int (*calc_src_len)(int dst_len));     elements consumed or 0          :106-134
 *
 * The storage is page-locked host memory owned by libsfegpu.so (sfe_ring_*), so the consumer may be
 * a GPU copy engine: read_to_device() is the additive call that replaces "read + convert on the CPU"
 * when the conversion is fused into the first GPU kernel (SFE_IN_U8 / SFE_IN_I16).
 * As in the reference there is no internal locking; callers serialise (bpsk.cxx:108-116,132-170).
 * T must be trivially copyable (the reference memcpy's it as well).
 */
#ifndef MY_RINGBUF_H_
#define MY_RINGBUF_H_

#include <stdio.h>
#include <stdlib.h>

#include "../sfe_gpu.h"

#ifdef SFE_RINGBUF_GR_NAMESPACE
namespace gr {
namespace simplefe {
#endif

template <class T>
class ring_buffer
{
public:
    ring_buffer() : m_r(make(0)) {}
    ring_buffer(int capacity) : m_r(make(capacity)) {}
    ~ring_buffer() { sfe_ring_destroy(m_r); }

    void alloc_buffer(int capacity)
    {
        if (sfe_ring_alloc(m_r, capacity) != SFE_OK) die("alloc_buffer");
    }
    int get_space() { return sfe_ring_space(m_r); }
    int get_count() { return sfe_ring_count(m_r); }
    int write(const T *data, int len) { return sfe_ring_write(m_r, data, len); }
    int read(void *dst, unsigned dst_len, int (*conv)(void *dst, void *src, int src_len), int (*calc_src_len)(int dst_len))
    {
        return sfe_ring_read(m_r, dst, dst_len, conv, calc_src_len);
    }
    /* additive: next n elements -> device memory on a CUDA stream, asynchronous; n or 0 */
    int read_to_device(void *d_dst, int n, void *cuda_stream) { return sfe_ring_read_dev(m_r, d_dst, n, cuda_stream); }

private:
    ring_buffer(const ring_buffer &);
    ring_buffer &operator=(const ring_buffer &);
    static sfe_ring *make(int capacity)
    {
        sfe_ring *r = sfe_ring_create((int)sizeof(T), capacity);
        if (!r) die("ring_buffer");
        return r;
    }
    static void die(const char *what)
    {
        fprintf(stderr, "%s: %s\n", what, sfe_last_error());
        abort();
    }
    sfe_ring *m_r;
};

#ifdef SFE_RINGBUF_GR_NAMESPACE
}  // namespace simplefe
}  // namespace gr
#endif

#endif
