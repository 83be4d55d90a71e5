/*
 * sfe_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement ("port") of the libdsp hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this.  The product (libsfegpu.so) never links, loads or calls anything in oracle/.
 *
 * What is restated, and where it comes from in the reference (/root/reference):
 *   orc_blkconv_*   libdsp/blkconv.cxx:34-75 (ctor), :77-110 (process)   FFT overlap-add, real stream
 *   orc_decimate_*  libdsp/decimate.cxx:37-59 (ctor), :69-129 (process), :132-140 (get_sample)
 *   orc_resample_*  libdsp/resample.cxx:37-69 (ctor), :85-153 (process)
 *   orc_ring_*      libdsp/ringbuf.h:37-142
 *   conversions     gr-simplefe/lib/source_c_impl.cc:121-132 (u8), examples/bpsk/bpsk.cxx:76-101 (dac10)
 *                   int16 rule s*(1.0f/32767.0f): defined by this build
 *
 * Third-party arithmetic: the reference's FFT is FFTW3 single precision (unpinned on Linux,
 * 3.3.5 vendored for Win64 only, blkconv.cxx:60-73,89,103).  FFTW is absent from the image, so
 * the r2c / c2r transforms are restated here from their published definition (forward
 * exp(-2 pi i jk/n), backward unnormalised) and evaluated in DOUBLE with one rounding to
 * float -- a golden that is closer to the exact answer than fftw3f itself.
 *
 * Parity pin: validated (tests/test_oracle.py) against oracle/_ref, i.e. the reference's own
 * decimate.cxx / resample.cxx / blkconv.cxx / ringbuf.h compiled verbatim, and against the
 * known answers of SURVEY.md section 4.1 (tests/golden/).  decimate/resample must match
 * bit for bit; blkconv matches to float rounding of the final c2r result.
 *
 * Build for parity with:  gcc -O2 -ffp-contract=off   (as-written IEEE single arithmetic).
 */
#include "sfe_oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------ */
/* double precision complex FFT, iterative radix-2 decimation in time (power-of-two sizes)      */

static void dfft_pow2(double *re, double *im, int n, int sign)
{
    int i, j, len;
    for (i = 1, j = 0; i < n; i++) {
        int bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) {
            double t = re[i]; re[i] = re[j]; re[j] = t;
            t = im[i]; im[i] = im[j]; im[j] = t;
        }
    }
    for (len = 2; len <= n; len <<= 1) {
        int half = len >> 1, k;
        for (k = 0; k < half; k++) {
            double ang = (double)sign * 2.0 * M_PI * (double)k / (double)len;
            double wr = cos(ang), wi = sin(ang);
            for (i = k; i < n; i += len) {
                int m = i + half;
                double tr = re[m] * wr - im[m] * wi;
                double ti = re[m] * wi + im[m] * wr;
                re[m] = re[i] - tr; im[m] = im[i] - ti;
                re[i] += tr;        im[i] += ti;
            }
        }
    }
}

static void dft_any(double *re, double *im, int n, int sign)
{
    if (n > 1 && (n & (n - 1)) == 0) { dfft_pow2(re, im, n, sign); return; }
    double *or_ = (double *)malloc(sizeof(double) * n), *oi = (double *)malloc(sizeof(double) * n);
    for (int k = 0; k < n; k++) {
        double sr = 0, si = 0;
        for (int j = 0; j < n; j++) {
            double a = (double)sign * 2.0 * M_PI * (double)(((long long)j * k) % n) / (double)n;
            double c = cos(a), s = sin(a);
            sr += re[j] * c - im[j] * s;
            si += re[j] * s + im[j] * c;
        }
        or_[k] = sr; oi[k] = si;
    }
    memcpy(re, or_, sizeof(double) * n); memcpy(im, oi, sizeof(double) * n);
    free(or_); free(oi);
}

/* ------------------------------------------------------------------------------------------ */
/* blkconv: blkconv.cxx:34-110                                                                  */

struct orc_blkconv {
    int fft_len, blk, ov;
    float scaling;     /* 1.0f/fft_len, blkconv.cxx:50 */
    float *buf;        /* fft_len+2 floats: the caller-visible process buffer, blkconv.cxx:44 */
    float *hr, *hi;    /* tap spectrum, fft_len/2+1 bins, rounded to float like fftwf output */
    float *tail;       /* n_taps-1 saved samples, blkconv.cxx:52-55 */
    double *wr, *wi;   /* work */
};

orc_blkconv *orc_blkconv_new(const float *taps, int n_taps, int fft_len)
{
    orc_blkconv *b = (orc_blkconv *)calloc(1, sizeof(*b));
    int nb = fft_len / 2 + 1, i;
    b->fft_len = fft_len;
    b->blk = fft_len + 1 - n_taps;          /* blkconv.cxx:47 */
    b->ov = n_taps - 1;                     /* blkconv.cxx:48 */
    b->scaling = 1.0f / fft_len;            /* blkconv.cxx:50 */
    b->buf = (float *)calloc((size_t)2 * nb, sizeof(float));
    b->hr = (float *)calloc(nb, sizeof(float));
    b->hi = (float *)calloc(nb, sizeof(float));
    b->tail = (float *)calloc(b->ov > 0 ? b->ov : 1, sizeof(float));
    b->wr = (double *)calloc(fft_len, sizeof(double));
    b->wi = (double *)calloc(fft_len, sizeof(double));
    /* spectrum of the zero-padded taps: blkconv.cxx:60-69 */
    for (i = 0; i < fft_len; i++) { b->wr[i] = i < n_taps ? (double)taps[i] : 0.0; b->wi[i] = 0.0; }
    dft_any(b->wr, b->wi, fft_len, -1);
    for (i = 0; i < nb; i++) { b->hr[i] = (float)b->wr[i]; b->hi[i] = (float)b->wi[i]; }
    return b;
}

int orc_blkconv_blksize(orc_blkconv *b) { return b->blk; }
float *orc_blkconv_buf(orc_blkconv *b) { return b->buf; }

void orc_blkconv_process(orc_blkconv *b)
{
    const int n = b->fft_len, nb = n / 2 + 1;
    float *x = b->buf;
    int i;
    for (i = b->blk; i < n; i++) x[i] = 0.0f;                 /* zero pad, blkconv.cxx:85-87 */
    for (i = 0; i < n; i++) { b->wr[i] = (double)x[i]; b->wi[i] = 0.0; }
    dft_any(b->wr, b->wi, n, -1);                             /* r2c, blkconv.cxx:89 */
    /* the r2c result is stored as float by fftwf; the multiply runs in float: blkconv.cxx:92-101 */
    for (i = 0; i < nb; i++) {
        float re = (float)b->wr[i], im = (float)b->wi[i];
        float cr = b->hr[i], ci = b->hi[i];
        float yr = b->scaling * (re * cr - im * ci);
        float yi = b->scaling * (re * ci + im * cr);
        b->wr[i] = (double)yr; b->wi[i] = (double)yi;
    }
    /* c2r reads bins 0..n/2 only and implies Hermitian symmetry; imaginary parts of the DC
       and Nyquist bins are ignored (FFTW c2r definition) */
    b->wi[0] = 0.0;
    if ((n & 1) == 0) b->wi[n / 2] = 0.0;
    for (i = nb; i < n; i++) { b->wr[i] = b->wr[n - i]; b->wi[i] = -b->wi[n - i]; }
    dft_any(b->wr, b->wi, n, +1);                             /* c2r unnormalised, blkconv.cxx:103 */
    for (i = 0; i < n; i++) x[i] = (float)b->wr[i];
    for (i = 0; i < b->ov; i++) {                             /* overlap add, blkconv.cxx:105-109 */
        x[i] = x[i] + b->tail[i];
        b->tail[i] = x[b->blk + i];
    }
}

void orc_blkconv_free(orc_blkconv *b)
{
    if (!b) return;
    free(b->buf); free(b->hr); free(b->hi); free(b->tail); free(b->wr); free(b->wi); free(b);
}

/* ------------------------------------------------------------------------------------------ */
/* the output-time walk shared by decimate and resample: decimate.cxx:73,96-127 and             */
/* resample.cxx:89,119-150.  `fetch(ctx, phase, n)` returns the polyphase output s(phase, n).    */

typedef float (*orc_fetch_fn)(void *ctx, int phase, int n);

typedef struct {
    int pos;          /* m_pos  */
    float mu;         /* m_mu   */
    float held;       /* m_last_remain */
    int have_held;    /* m_is_leftover */
} orc_walk;

static int walk_outputs(orc_walk *w, int up, int n_in, float *out, int out_len, float rate,
                        orc_fetch_fn fetch, void *ctx)
{
    int produced = 0;
    float t = w->pos + w->mu;                                  /* decimate.cxx:73 */
    if (w->have_held) {                                        /* decimate.cxx:96-100 */
        out[produced++] = w->held * (1.0f - w->mu) + w->mu * fetch(ctx, 0, 0);
        w->have_held = 0;
        t += rate * up;
    }
    for (;;) {
        int nxt, ph0, ph1, i0, i1;
        w->pos = (int)floorf(t);                               /* decimate.cxx:106 */
        w->mu = t - w->pos;                                    /* decimate.cxx:107 */
        nxt = w->pos + 1;
        ph0 = w->pos % up; ph1 = nxt % up;
        i0 = w->pos / up;  i1 = nxt / up;
        if (i0 >= n_in || produced >= out_len) break;          /* decimate.cxx:115 */
        if (i1 >= n_in) {                                      /* decimate.cxx:118-121 */
            w->have_held = 1;
            w->held = fetch(ctx, ph0, i0);
            break;
        }
        out[produced++] = fetch(ctx, ph0, i0) * (1.0f - w->mu) + w->mu * fetch(ctx, ph1, i1);
        t += rate * up;
    }
    w->pos -= n_in * up;                                       /* decimate.cxx:127 */
    return produced;
}

/* ------------------------------------------------------------------------------------------ */
/* decimate: decimate.cxx                                                                       */

struct orc_decimate {
    float *taps; int n_taps; int up; int blksize;
    float *hist; int len; float *cur;
    orc_walk w;
};

orc_decimate *orc_decimate_new(const float *taps, int n_taps, int up, int blksize)
{
    orc_decimate *d = (orc_decimate *)calloc(1, sizeof(*d));
    d->n_taps = (n_taps % 2 == 0) ? n_taps + 1 : n_taps;      /* decimate.cxx:42-44 */
    d->taps = (float *)calloc(d->n_taps, sizeof(float));      /* trailing 0 when padded, :45-51 */
    memcpy(d->taps, taps, sizeof(float) * n_taps);
    d->up = up; d->blksize = blksize;
    d->len = d->n_taps + blksize;                             /* decimate.cxx:53 */
    d->hist = (float *)calloc(d->len, sizeof(float));
    return d;
}

static float dec_fetch(void *ctx, int phase, int n)           /* decimate.cxx:132-140 */
{
    orc_decimate *d = (orc_decimate *)ctx;
    float acc = 0.0f;
    int m, j;
    for (m = phase, j = 0; m < d->n_taps; m += d->up, j++) acc += d->taps[m] * d->cur[n - j];
    return acc;
}

int orc_decimate_process(orc_decimate *d, const float *in, int n_in, float *out, int out_len, float rate)
{
    if (rate < 1.0) { printf("rate should be larger than 1.0\n"); return 0; }                 /* :75-78 */
    if (n_in > d->blksize) { printf("number of samples should be less than blksize\n"); return 0; } /* :79-82 */
    if (out_len < floorf(n_in * 1.0f / rate)) { printf("output buffer is not large enough"); return 0; } /* :84-87 */
    memmove(d->hist, d->hist + n_in, sizeof(float) * (d->len - n_in));   /* :89 (memcpy used as memmove) */
    memcpy(d->hist + d->len - n_in, in, sizeof(float) * n_in);           /* :90 */
    d->cur = d->hist + d->len - n_in;                                    /* :91 */
    return walk_outputs(&d->w, d->up, n_in, out, out_len, rate, dec_fetch, d);
}

void orc_decimate_free(orc_decimate *d) { if (d) { free(d->taps); free(d->hist); free(d); } }

/* ------------------------------------------------------------------------------------------ */
/* resample: resample.cxx                                                                       */

struct orc_resample {
    int up, plen, blksize;
    float *ptaps;   /* [up][plen]  ptaps[j][i] = taps[i*up + j], resample.cxx:55-64 */
    float *bank;    /* [up][blksize] all phase outputs of the current call, resample.cxx:49 */
    float *hist;    /* blksize floats, of which plen-1 are used, resample.cxx:52,110-113 */
    orc_walk w;
};

orc_resample *orc_resample_new(const float *taps, int n_taps, int up, int blksize)
{
    orc_resample *r = (orc_resample *)calloc(1, sizeof(*r));
    int i, j;
    r->up = up; r->blksize = blksize;
    r->plen = (n_taps + up - 1) / up;                          /* resample.cxx:43 */
    r->ptaps = (float *)calloc((size_t)up * r->plen, sizeof(float));
    r->bank = (float *)calloc((size_t)up * blksize, sizeof(float));
    r->hist = (float *)calloc(blksize > r->plen ? blksize : r->plen, sizeof(float));
    for (i = 0; i < r->plen; i++)
        for (j = 0; j < up; j++) {
            int n = i * up + j;
            r->ptaps[(size_t)j * r->plen + i] = n < n_taps ? taps[n] : 0.0f;
        }
    return r;
}

static float rs_fetch(void *ctx, int phase, int n)
{
    orc_resample *r = (orc_resample *)ctx;
    return r->bank[(size_t)phase * r->blksize + n];
}

int orc_resample_process(orc_resample *r, const float *in, int n_in, float *out, int out_len, float rate)
{
    int i, j, n;
    if (n_in > r->blksize || rate < 1.0 / r->up) {             /* resample.cxx:91-94 */
        printf("input parameter is wrong, rate <= 1/upsample, n_in <= blksize\n");
        return 0;
    }
    if (out_len < floorf(n_in * 1.0f / rate)) { printf("output buffer is not large enough"); return 0; } /* :95-98 */
    for (i = 0; i < n_in; i++) {                               /* FIR bank, resample.cxx:100-114 */
        for (j = 0; j < r->up; j++) {
            const float *pt = r->ptaps + (size_t)j * r->plen;
            float acc = pt[0] * in[i];
            for (n = 1; n < r->plen; n++) acc += pt[n] * r->hist[n - 1];
            r->bank[(size_t)j * r->blksize + i] = acc;
        }
        for (n = r->plen - 2; n > 0; n--) r->hist[n] = r->hist[n - 1];
        r->hist[0] = in[i];
    }
    return walk_outputs(&r->w, r->up, n_in, out, out_len, rate, rs_fetch, r);
}

void orc_resample_free(orc_resample *r) { if (r) { free(r->ptaps); free(r->bank); free(r->hist); free(r); } }

/* ------------------------------------------------------------------------------------------ */
/* sample format conversions                                                                    */

void orc_i16_to_f32(const int16_t *in, float *out, size_t n)
{
    const float q = 1.0f / 32767.0f;          /* modelled on the u8 rule */
    for (size_t i = 0; i < n; i++) out[i] = (float)in[i] * q;
}

void orc_u8_to_f32(const uint8_t *in, float *out, size_t n)
{
    const float q = 1.0f / 127.0f;            /* source_c_impl.cc:126-129 */
    for (size_t i = 0; i < n; i++) out[i] = (in[i] - 128) * q;
}

int orc_f32_to_dac10(const float *in, int n, uint8_t *out)
{
    /* bpsk.cxx:76-101: 4 samples -> 10-bit offset words -> 5 bytes, high bit-pairs first */
    int j = 0;
    for (int i = 0; i + 3 < n; i += 4) {
        unsigned short u[4];
        for (int k = 0; k < 4; k++) u[k] = ((short)(in[i + k] * 511) + 512) & 0x3FF;
        out[j++] = (uint8_t)((u[0] >> 8) | ((u[1] >> 8) << 2) | ((u[2] >> 8) << 4) | ((u[3] >> 8) << 6));
        for (int k = 0; k < 4; k++) out[j++] = (uint8_t)(u[k] & 0xFF);
    }
    return j;
}

/* ------------------------------------------------------------------------------------------ */
/* ring buffer: libdsp/ringbuf.h:37-142, element size given at run time                         */

struct orc_ring {
    unsigned char *buf; int esz; int cap; int count; int rd; int wr;
};

orc_ring *orc_ring_new(int elem_size, int capacity)
{
    orc_ring *r = (orc_ring *)calloc(1, sizeof(*r));
    r->esz = elem_size; r->cap = capacity;
    r->buf = (unsigned char *)calloc((size_t)(capacity > 0 ? capacity : 1), elem_size);
    return r;
}
void orc_ring_free(orc_ring *r) { if (r) { free(r->buf); free(r); } }
int orc_ring_space(orc_ring *r) { return r->cap - r->count; }      /* ringbuf.h:71-74 */
int orc_ring_count(orc_ring *r) { return r->count; }               /* ringbuf.h:75-78 */

int orc_ring_write(orc_ring *r, const void *data, int len)         /* ringbuf.h:80-101 */
{
    const unsigned char *src = (const unsigned char *)data;
    int first;
    if (len > orc_ring_space(r)) return 0;                         /* all or nothing */
    first = r->cap - r->wr;
    if (first < len) {
        memcpy(r->buf + (size_t)r->wr * r->esz, src, (size_t)first * r->esz);
        memcpy(r->buf, src + (size_t)first * r->esz, (size_t)(len - first) * r->esz);
        r->wr = len - first;
    } else {
        memcpy(r->buf + (size_t)r->wr * r->esz, src, (size_t)len * r->esz);
        r->wr += len;                                              /* may equal cap, as in the reference */
    }
    r->count += len;
    return len;
}

int orc_ring_read(orc_ring *r, void *dst, unsigned dst_len,        /* ringbuf.h:106-134 */
                  int (*conv)(void *, void *, int), int (*calc_src_len)(int))
{
    int need, first;
    if (!conv || !calc_src_len) return 0;
    need = calc_src_len((int)dst_len);
    if (need > r->count) return 0;
    first = r->cap - r->rd;
    if (first < need) {
        int wrote = conv(dst, r->buf + (size_t)r->rd * r->esz, first);
        conv((char *)dst + wrote, r->buf, need - first);
        r->rd = need - first;
    } else {
        conv(dst, r->buf + (size_t)r->rd * r->esz, need);
        r->rd += need;
    }
    r->count -= need;
    return need;
}
