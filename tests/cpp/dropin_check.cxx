// dropin_check.cxx -- exercises the drop-in C++ classes (include/libdsp/*.h) exactly the way the reference's
// callers do (libdsp/test/test_blkconv.cxx:6-34, libdsp/test/test_resample.py:10-25 restated in C++,
// gr-simplefe/lib/qa_simplefe.cc:103-166 ring cases) and prints the results for tests/test_dropin.py.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "blkconv.h"
#include "decimate.h"
#include "resample.h"
#include "ringbuf.h"

static const float TAPS31[] = {4.154052476224024e-06f, 6.836163811385632e-05f, 0.000311697949655354f, 0.0005662626936100423f,
    -0.00019575881015043706f, -0.004200732801109552f, -0.013667905703186989f, -0.027301859110593796f, -0.03584320470690727f,
    -0.02054252102971077f, 0.04132149741053581f, 0.16359668970108032f, 0.33717116713523865f, 0.5246731638908386f,
    0.670900285243988f, 0.7262777090072632f, 0.6709002256393433f, 0.5246731042861938f, 0.33717113733291626f,
    0.16359665989875793f, 0.04132148623466492f, -0.020542524755001068f, -0.03584319353103638f, -0.027301862835884094f,
    -0.013667899183928967f, -0.004200733732432127f, -0.0001957584754563868f, 0.000566263624932617f, 0.0003116970765404403f,
    6.836453394498676e-05f, 4.154052476224024e-06f};

static int conv_copy(void *dst, void *src, int n)
{
    memcpy(dst, src, n * sizeof(float));
    return n * (int)sizeof(float);
}
static int len_floats(int dst_len) { return dst_len / (int)sizeof(float); }
// bytes -> floats with the device rule (b-128)/127, gr-simplefe/lib/source_f_impl.cc:120-129
static int conv_u8(void *dst, void *src, int n)
{
    float *d = (float *)dst;
    unsigned char *s = (unsigned char *)src;
    for (int i = 0; i < n; i++) d[i] = ((int)s[i] - 128) * (1.0f / 127.0f);
    return n * (int)sizeof(float);
}

template <class R> static void run_poly(const char *tag, float rate)
{
    float taps[31];
    memcpy(taps, TAPS31, sizeof taps);
    R r(taps, 31, 4, 128);
    std::vector<float> x(1024), y(512);
    for (int i = 0; i < 1024; i++) x[i] = (float)sin(0.02 * M_PI * i);
    int total = 0;
    double sum = 0, sum2 = 0;
    printf("%s counts", tag);
    for (int b = 0; b < 8; b++) {
        int n = r.process(&x[128 * b], 128, y.data(), 512, rate);
        printf(" %d", n);
        for (int i = 0; i < n; i++) { sum += y[i]; sum2 += (double)y[i] * y[i]; }
        total += n;
    }
    printf("\n%s total %d sum %.6f sum2 %.6f\n", tag, total, sum, sum2);
    int bad = r.process(x.data(), 129, y.data(), 512, rate);   // n_in > blksize: message + 0
    printf("%s bad %d\n", tag, bad);
}

int main()
{
    {   // test_blkconv.cxx scenario
        float taps[] = {1.0f, 1.0f, 1.0f, 1.0f, 1.0f};
        blkconv conv(taps, 5, 32);
        float *in = conv.get_process_buf();
        int len = conv.get_blksize();
        printf("blksize = %d\n", len);
        for (int i = 0; i < len; i++) in[i] = 1.0f;
        conv.process();
        printf("blk1");
        for (int i = 0; i < len; i++) printf(" %.2f", in[i]);
        for (int i = 0; i < len; i++) in[i] = 0.0f;
        conv.process();
        printf("\nblk2");
        for (int i = 0; i < len; i++) printf(" %.2f", in[i]);
        printf("\n");
    }
    run_poly<resample>("resample077", 0.77f);
    run_poly<resample>("resample177", 1.77f);
    run_poly<decimate>("decimate177", 1.77f);
    {   // ring: simple write/read, wrap-around with conversion, refusal when full / short
        ring_buffer<float> rb(8);
        float a[6] = {1, 2, 3, 4, 5, 6}, out[8];
        int w = rb.write(a, 6);   // (separate statements: argument evaluation order is unspecified)
        printf("ring w %d space %d count %d\n", w, rb.get_space(), rb.get_count());
        printf("ring w_full %d\n", rb.write(a, 3));
        int n = rb.read(out, 4 * sizeof(float), conv_copy, len_floats);
        printf("ring r %d : %g %g %g %g\n", n, out[0], out[1], out[2], out[3]);
        w = rb.write(a, 5);
        printf("ring w_wrap %d count %d\n", w, rb.get_count());
        n = rb.read(out, 7 * sizeof(float), conv_copy, len_floats);
        printf("ring r_wrap %d : %g %g %g %g %g %g %g\n", n, out[0], out[1], out[2], out[3], out[4], out[5], out[6]);
        printf("ring r_short %d\n", rb.read(out, 4, conv_copy, len_floats));
        ring_buffer<unsigned char> rc;
        rc.alloc_buffer(5);
        unsigned char b[4] = {128, 255, 1, 0};
        rc.write(b, 4);
        rc.read(out, 3 * sizeof(float), conv_u8, len_floats);
        rc.write(b, 3);
        n = rc.read(out, 4 * sizeof(float), conv_u8, len_floats);
        printf("ring u8 %d : %.6f %.6f %.6f %.6f\n", n, out[0], out[1], out[2], out[3]);
    }
    return 0;
}
