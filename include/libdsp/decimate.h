/*
 * decimate.h -- drop-in for libdsp/decimate.h:33-63: polyphase FIR arbitrary-rate decimator.
 *
 *   decimate(float *taps, int n_taps, int upsample, int blksize)              decimate.h:37
 *   int process(float *in, int n_in, float *out, int out_len, float rate)     decimate.h:45
 *
 * process() returns the number of output samples; on the reference's argument errors
 * (decimate.cxx:75-87: rate < 1, n_in > blksize, out_len < floor(n_in/rate)) it prints the reference's
 * message to stdout and returns 0.  A CUDA failure is fatal (no CPU fallback).
 */
#ifndef DECIMATE_H_
#define DECIMATE_H_

#include <stdio.h>
#include <stdlib.h>

#include "../sfe_gpu.h"

class decimate
{
public:
    decimate(float *taps, int n_taps, int upsample, int blksize)
        : m_h(sfe_decimate_create(taps, n_taps, upsample, blksize))
    {
        if (!m_h) {
            fprintf(stderr, "decimate: %s\n", sfe_last_error());
            abort();
        }
    }
    ~decimate() { sfe_polyphase_destroy(m_h); }

    int process(float *in, int n_in, float *out, int out_len, float rate)
    {
        const int n = sfe_polyphase_process(m_h, in, n_in, out, out_len, rate);
        if (n < 0) {
            fprintf(stderr, "decimate::process: %s\n", sfe_last_error());
            abort();
        }
        return n;
    }
    /* additive: 1 = reference operation order bit for bit (separate multiply/add), 0 = FMA (default) */
    void set_exact(bool on) { sfe_polyphase_set_exact(m_h, on ? 1 : 0); }

private:
    decimate(const decimate &);
    decimate &operator=(const decimate &);
    sfe_polyphase *m_h;
};

#endif
