/*
 * resample.h -- drop-in for libdsp/resample.h:33-61: polyphase FIR rational/arbitrary-rate resampler.
 *
 *   resample(float *taps, int n_taps, int upsample, int blksize)              resample.h:39
 *   int process(float *in, int n_in, float *out, int out_len, float rate)     resample.h:47
 *
 * process() returns the number of output samples; on the reference's argument errors
 * (resample.cxx:91-98: n_in > blksize, rate < 1/upsample, out_len < floor(n_in/rate)) it prints the reference's
 * message to stdout and returns 0.  A CUDA failure is fatal (no CPU fallback).
 */
#ifndef RESAMPLE_H_
#define RESAMPLE_H_

#include <stdio.h>
#include <stdlib.h>

#include "../sfe_gpu.h"

class resample
{
public:
    resample(float *taps, int n_taps, int upsample, int blksize)
        : m_h(sfe_resample_create(taps, n_taps, upsample, blksize))
    {
        if (!m_h) {
            fprintf(stderr, "resample: %s\n", sfe_last_error());
            abort();
        }
    }
    ~resample() { sfe_polyphase_destroy(m_h); }

    int process(float *in, int n_in, float *out, int out_len, float rate)
    {
        const int n = sfe_polyphase_process(m_h, in, n_in, out, out_len, rate);
        if (n < 0) {
            fprintf(stderr, "resample::process: %s\n", sfe_last_error());
            abort();
        }
        return n;
    }
    /* additive: 1 = reference operation order bit for bit (separate multiply/add), 0 = FMA (default) */
    void set_exact(bool on) { sfe_polyphase_set_exact(m_h, on ? 1 : 0); }

private:
    resample(const resample &);
    resample &operator=(const resample &);
    sfe_polyphase *m_h;
};

#endif
