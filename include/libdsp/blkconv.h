/*
 * blkconv.h -- drop-in for libdsp/blkconv.h:35-62 (wnmusic/simpleFE): same class name, same public
 * members, global namespace.  The arithmetic runs on the GPU behind the C ABI of libsfegpu.so
 * (include/sfe_gpu.h); there is no CPU implementation behind this header.
 *
 *   blkconv(float *taps, int n_taps, int fft_len)   blkconv.h:38   taps are copied
 *   int    get_blksize()                            blkconv.h:40   fft_len + 1 - n_taps
 *   float *get_process_buf()                        blkconv.h:44   object-owned I/O buffer (pinned host memory)
 *   void   process()                                blkconv.h:48   consumes buf[0..blk), leaves the output there
 *
 * Callers that compile unchanged: libdsp/test/test_blkconv.cxx, examples/bpsk/bpsk.cxx:125-127,162-164.
 * Link with -lsfegpu instead of -lLibdsp -lfftw3f.
 */
#ifndef BLK_CONV_H_
#define BLK_CONV_H_

#include <stdio.h>
#include <stdlib.h>

#include "../sfe_gpu.h"

class blkconv
{
public:
    blkconv(float *taps, int n_taps, int fft_len) : m_h(sfe_blkconv_create(taps, n_taps, fft_len))
    {
        if (!m_h) {
            /* the reference constructor cannot fail; without a GPU this object cannot work at all */
            fprintf(stderr, "blkconv: %s\n", sfe_last_error());
            abort();
        }
    }
    ~blkconv() { sfe_blkconv_destroy(m_h); }

    int get_blksize() { return sfe_blkconv_blksize(m_h); }
    float *get_process_buf() { return sfe_blkconv_buf(m_h); }
    void process()
    {
        if (sfe_blkconv_process(m_h) != SFE_OK) {
            fprintf(stderr, "blkconv::process: %s\n", sfe_last_error());
            abort();
        }
    }

private:
    blkconv(const blkconv &);
    blkconv &operator=(const blkconv &);
    sfe_blkconv *m_h;
};

#endif
