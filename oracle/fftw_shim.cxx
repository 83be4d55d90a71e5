/*
 * fftw_shim.cxx -- TEST INFRASTRUCTURE ONLY (oracle side, never linked into the product).
 *
 * The reference's blkconv (libdsp/blkconv.cxx:41-73,89,103,116-120) calls exactly six
 * FFTW3 single-precision entry points.  No FFTW binary exists for Linux in the reference
 * tree (only Win64 DLLs under contrib/fftw-3.3.5-dll64) nor on this image, so this file
 * implements those six symbols with FFTW's *documented* semantics so that the reference
 * blkconv.cxx can be compiled VERBATIM and linked into oracle/_ref/:
 *
 *   fftwf_malloc / fftwf_free        aligned allocation
 *   fftwf_plan_dft_r2c_1d(n,in,out)  out[k] = sum_j in[j] exp(-2 pi i jk/n), k = 0..n/2,
 *                                    in-place allowed (buffer of 2*(n/2+1) floats)
 *   fftwf_plan_dft_c2r_1d(n,in,out)  out[j] = sum_k X[k] exp(+2 pi i jk/n) with Hermitian
 *                                    extension, UNNORMALISED (n * inverse), may destroy input
 *   fftwf_execute / fftwf_destroy_plan   plans bind the pointers given at plan time
 *
 * Two arithmetic modes, chosen at compile time:
 *   -DSHIM_DOUBLE : butterflies in double, one rounding to float at the end.  This is the
 *                   PARITY build: a golden that is closer to the exact answer than fftw3f.
 *   (default)     : butterflies in float (radix-4/2 Stockham, SoA work arrays so gcc
 *                   vectorises).  This is the TIMING build used for the CPU baseline.
 * Non power-of-two n falls back to an O(n^2) DFT in double (correct, slow).
 */
#include <fftw3.h> /* the reference's vendored header: contrib/fftw-3.3.5-dll64/fftw3.h */
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#ifdef SHIM_DOUBLE
typedef double real_t;
#else
typedef float real_t;
#endif

namespace {

struct cfft {
    /* complex FFT of size m (power of two), Stockham autosort, radix-4 then radix-2 */
    int m;
    std::vector<real_t> wr, wi;         /* w^k = exp(-2 pi i k / m), k < m */
    std::vector<real_t> ar, ai, br, bi; /* ping-pong SoA buffers */
    explicit cfft(int m_) : m(m_), wr(m_), wi(m_), ar(m_), ai(m_), br(m_), bi(m_)
    {
        for (int k = 0; k < m; k++) {
            double a = -2.0 * M_PI * (double)k / (double)m;
            wr[k] = (real_t)cos(a);
            wi[k] = (real_t)sin(a);
        }
    }
    /* forward transform of (ar,ai) in place (result back in ar,ai); sign=-1 fwd, +1 inverse */
    void run(int sign)
    {
        real_t *xr = ar.data(), *xi = ai.data(), *yr = br.data(), *yi = bi.data();
        int n = m, s = 1; /* n: current sub-transform length, s: stride */
        const real_t sg = (real_t)sign;
        while (n >= 4) {
            const int n4 = n / 4;
            for (int p = 0; p < n4; p++) {
                const int tw = p * s; /* twiddle index step m/n = s */
                const real_t w1r = wr[tw], w1i = -sg * wi[tw];
                const real_t w2r = wr[2 * tw], w2i = -sg * wi[2 * tw];
                const real_t w3r = wr[3 * tw], w3i = -sg * wi[3 * tw];
                for (int q = 0; q < s; q++) {
                    const int i0 = q + s * p;
                    const real_t a0r = xr[i0], a0i = xi[i0];
                    const real_t a1r = xr[i0 + s * n4], a1i = xi[i0 + s * n4];
                    const real_t a2r = xr[i0 + 2 * s * n4], a2i = xi[i0 + 2 * s * n4];
                    const real_t a3r = xr[i0 + 3 * s * n4], a3i = xi[i0 + 3 * s * n4];
                    const real_t t0r = a0r + a2r, t0i = a0i + a2i;
                    const real_t t1r = a0r - a2r, t1i = a0i - a2i;
                    const real_t t2r = a1r + a3r, t2i = a1i + a3i;
                    /* t3 = (a1-a3) * (-i) for forward (sg=-1), * (+i) for inverse (sg=+1) */
                    const real_t t3r = -sg * (a1i - a3i), t3i = sg * (a1r - a3r);
                    const int o = q + s * 4 * p;
                    yr[o] = t0r + t2r;
                    yi[o] = t0i + t2i;
                    const real_t b1r = t1r + t3r, b1i = t1i + t3i;
                    const real_t b2r = t0r - t2r, b2i = t0i - t2i;
                    const real_t b3r = t1r - t3r, b3i = t1i - t3i;
                    yr[o + s] = b1r * w1r - b1i * w1i;
                    yi[o + s] = b1r * w1i + b1i * w1r;
                    yr[o + 2 * s] = b2r * w2r - b2i * w2i;
                    yi[o + 2 * s] = b2r * w2i + b2i * w2r;
                    yr[o + 3 * s] = b3r * w3r - b3i * w3i;
                    yi[o + 3 * s] = b3r * w3i + b3i * w3r;
                }
            }
            n = n4;
            s *= 4;
            std::swap(xr, yr);
            std::swap(xi, yi);
        }
        if (n == 2) {
            for (int q = 0; q < s; q++) {
                const real_t a0r = xr[q], a0i = xi[q], a1r = xr[q + s], a1i = xi[q + s];
                yr[q] = a0r + a1r;
                yi[q] = a0i + a1i;
                yr[q + s] = a0r - a1r;
                yi[q + s] = a0i - a1i;
            }
            std::swap(xr, yr);
            std::swap(xi, yi);
        }
        if (xr != ar.data()) {
            memcpy(ar.data(), xr, sizeof(real_t) * m);
            memcpy(ai.data(), xi, sizeof(real_t) * m);
        }
    }
};

} // namespace

struct fftwf_plan_s {
    int n;
    int kind; /* 0 = r2c, 1 = c2r */
    float *rbuf;
    fftwf_complex *cbuf;
    bool pow2;
    cfft *half;                  /* size n/2 complex FFT for the packed real transform */
    std::vector<real_t> er, ei;  /* exp(-2 pi i k / n), k <= n/2 */
};

static bool is_pow2(int n) { return n >= 4 && (n & (n - 1)) == 0; }

static fftwf_plan make_plan(int n, int kind, float *r, fftwf_complex *c)
{
    fftwf_plan p = new fftwf_plan_s;
    p->n = n;
    p->kind = kind;
    p->rbuf = r;
    p->cbuf = c;
    p->pow2 = is_pow2(n);
    p->half = NULL;
    if (p->pow2) {
        p->half = new cfft(n / 2);
        p->er.resize(n / 2 + 1);
        p->ei.resize(n / 2 + 1);
        for (int k = 0; k <= n / 2; k++) {
            double a = -2.0 * M_PI * (double)k / (double)n;
            p->er[k] = (real_t)cos(a);
            p->ei[k] = (real_t)sin(a);
        }
    }
    return p;
}

extern "C" {

void *fftwf_malloc(size_t n)
{
    void *p = NULL;
    if (posix_memalign(&p, 64, n ? n : 64) != 0) return NULL;
    return p;
}

void fftwf_free(void *p) { free(p); }

fftwf_plan fftwf_plan_dft_r2c_1d(int n, float *in, fftwf_complex *out, unsigned)
{
    return make_plan(n, 0, in, out);
}

fftwf_plan fftwf_plan_dft_c2r_1d(int n, fftwf_complex *in, float *out, unsigned)
{
    return make_plan(n, 1, out, in);
}

void fftwf_destroy_plan(fftwf_plan p)
{
    if (!p) return;
    delete p->half;
    delete p;
}

static void slow_r2c(const fftwf_plan p)
{
    const int n = p->n;
    std::vector<double> x(p->rbuf, p->rbuf + n);
    for (int k = 0; k <= n / 2; k++) {
        double sr = 0, si = 0;
        for (int j = 0; j < n; j++) {
            double a = -2.0 * M_PI * (double)(((long long)j * k) % n) / (double)n;
            sr += x[j] * cos(a);
            si += x[j] * sin(a);
        }
        p->cbuf[k][0] = (float)sr;
        p->cbuf[k][1] = (float)si;
    }
}

static void slow_c2r(const fftwf_plan p)
{
    const int n = p->n;
    std::vector<double> xr(n / 2 + 1), xi(n / 2 + 1);
    for (int k = 0; k <= n / 2; k++) {
        xr[k] = p->cbuf[k][0];
        xi[k] = p->cbuf[k][1];
    }
    for (int j = 0; j < n; j++) {
        double s = xr[0];
        for (int k = 1; k <= n / 2; k++) {
            double a = 2.0 * M_PI * (double)(((long long)j * k) % n) / (double)n;
            double term = xr[k] * cos(a) - xi[k] * sin(a);
            s += (2 * k == n) ? term : 2.0 * term;
        }
        p->rbuf[j] = (float)s;
    }
}

void fftwf_execute(const fftwf_plan p)
{
    const int n = p->n, h = n / 2;
    if (!p->pow2) {
        if (p->kind == 0) slow_r2c(p); else slow_c2r(p);
        return;
    }
    cfft &f = *p->half;
    real_t *zr = f.ar.data(), *zi = f.ai.data();
    if (p->kind == 0) {
        /* z[j] = x[2j] + i x[2j+1]; Z = FFT_h(z); untangle into X[0..h] */
        const float *x = p->rbuf;
        for (int j = 0; j < h; j++) {
            zr[j] = (real_t)x[2 * j];
            zi[j] = (real_t)x[2 * j + 1];
        }
        f.run(-1);
        fftwf_complex *X = p->cbuf; /* may alias x: everything needed is in zr/zi now */
        for (int k = 0; k <= h; k++) {
            const int k1 = k % h, k2 = (h - k) % h;
            const real_t ar = zr[k1], ai = zi[k1], br = zr[k2], bi = -zi[k2]; /* conj(Z[h-k]) */
            const real_t er_ = (real_t)0.5 * (ar + br), ei_ = (real_t)0.5 * (ai + bi); /* even part */
            const real_t dr = (real_t)0.5 * (ar - br), di = (real_t)0.5 * (ai - bi);
            /* odd part = -i * d ; X = E + w^k * O */
            const real_t or_ = di, oi_ = -dr;
            const real_t wr = p->er[k], wi = p->ei[k];
            X[k][0] = (float)(er_ + (or_ * wr - oi_ * wi));
            X[k][1] = (float)(ei_ + (or_ * wi + oi_ * wr));
        }
    } else {
        /* inverse of the above, unnormalised: x = n * irfft(X) */
        const fftwf_complex *X = p->cbuf;
        for (int k = 0; k < h; k++) {
            const real_t ar = (real_t)X[k][0], ai = (real_t)X[k][1];
            const real_t br = (real_t)X[h - k][0], bi = -(real_t)X[h - k][1]; /* conj(X[h-k]) */
            const real_t er_ = ar + br, ei_ = ai + bi;          /* 2E */
            const real_t dr = ar - br, di = ai - bi;            /* 2 w^k O */
            const real_t wr = p->er[k], wi = -p->ei[k];         /* w^-k */
            const real_t or_ = dr * wr - di * wi, oi_ = dr * wi + di * wr; /* 2O */
            /* z = E + i O */
            zr[k] = er_ - oi_;
            zi[k] = ei_ + or_;
        }
        f.run(+1);
        float *x = p->rbuf;
        for (int j = 0; j < h; j++) {
            x[2 * j] = (float)zr[j];
            x[2 * j + 1] = (float)zi[j];
        }
    }
}

} /* extern "C" */
