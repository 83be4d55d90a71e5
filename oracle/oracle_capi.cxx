/*
 * oracle_capi.cxx -- TEST INFRASTRUCTURE ONLY.
 *
 * One flat C interface (sfeo_*) over the CPU implementation of the libdsp hot path, built twice:
 *   -DORACLE_IMPL_REF   wraps the reference's OWN classes, compiled verbatim from
 *                       /root/reference/libdsp/{blkconv,decimate,resample}.cxx + ringbuf.h
 *                       (blkconv links oracle/fftw_shim.cxx, see there) -> oracle/_ref/libsfe_ref*.so
 *   -DORACLE_IMPL_PORT  wraps the restatement in oracle/sfe_oracle.c     -> oracle/libsfe_oracle.so
 * Both export identical symbols so oracle/oracle.py can load either.
 *
 * Also holds the chain driver used for the chain oracle and the timed CPU baseline: per stream
 * and per real component (I, Q) one object per stage, exactly the way a libdsp user would
 * compose them (cfg 4/5):
 *     sample conversion -> blkconv (blocks of fft_len+1-n_taps) -> stage A -> stage B
 * where stage A/B are `decimate` or `resample` objects fed in calls of `blk` samples.
 */
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#if defined(ORACLE_IMPL_REF)
#include "blkconv.h"
#include "decimate.h"
#include "resample.h"
#include "ringbuf.h"
typedef blkconv impl_blkconv;
typedef decimate impl_decimate;
typedef resample impl_resample;
static const char *k_kind = "reference";
#elif defined(ORACLE_IMPL_PORT)
#include "sfe_oracle.h"
struct impl_blkconv {
    orc_blkconv *p;
    impl_blkconv(float *taps, int n, int fft) : p(orc_blkconv_new(taps, n, fft)) {}
    ~impl_blkconv() { orc_blkconv_free(p); }
    int get_blksize() { return orc_blkconv_blksize(p); }
    float *get_process_buf() { return orc_blkconv_buf(p); }
    void process() { orc_blkconv_process(p); }
};
struct impl_decimate {
    orc_decimate *p;
    impl_decimate(float *taps, int n, int up, int blk) : p(orc_decimate_new(taps, n, up, blk)) {}
    ~impl_decimate() { orc_decimate_free(p); }
    int process(float *in, int n_in, float *out, int out_len, float rate)
    { return orc_decimate_process(p, in, n_in, out, out_len, rate); }
};
struct impl_resample {
    orc_resample *p;
    impl_resample(float *taps, int n, int up, int blk) : p(orc_resample_new(taps, n, up, blk)) {}
    ~impl_resample() { orc_resample_free(p); }
    int process(float *in, int n_in, float *out, int out_len, float rate)
    { return orc_resample_process(p, in, n_in, out, out_len, rate); }
};
static const char *k_kind = "port";
#else
#error "define ORACLE_IMPL_REF or ORACLE_IMPL_PORT"
#endif

extern "C" {

const char *sfeo_kind(void) { return k_kind; }
const char *sfeo_build_flags(void)
{
#ifdef ORACLE_BUILD_FLAGS
    return ORACLE_BUILD_FLAGS;
#else
    return "";
#endif
}

/* ---- single objects ---------------------------------------------------------------------- */
void *sfeo_blkconv_new(const float *taps, int n_taps, int fft_len)
{ return new impl_blkconv(const_cast<float *>(taps), n_taps, fft_len); }
int sfeo_blkconv_blksize(void *h) { return static_cast<impl_blkconv *>(h)->get_blksize(); }
float *sfeo_blkconv_buf(void *h) { return static_cast<impl_blkconv *>(h)->get_process_buf(); }
void sfeo_blkconv_process(void *h) { static_cast<impl_blkconv *>(h)->process(); }
void sfeo_blkconv_free(void *h) { delete static_cast<impl_blkconv *>(h); }

void *sfeo_decimate_new(const float *taps, int n_taps, int up, int blk)
{ return new impl_decimate(const_cast<float *>(taps), n_taps, up, blk); }
int sfeo_decimate_process(void *h, const float *in, int n_in, float *out, int out_len, float rate)
{ return static_cast<impl_decimate *>(h)->process(const_cast<float *>(in), n_in, out, out_len, rate); }
void sfeo_decimate_free(void *h) { delete static_cast<impl_decimate *>(h); }

void *sfeo_resample_new(const float *taps, int n_taps, int up, int blk)
{ return new impl_resample(const_cast<float *>(taps), n_taps, up, blk); }
int sfeo_resample_process(void *h, const float *in, int n_in, float *out, int out_len, float rate)
{ return static_cast<impl_resample *>(h)->process(const_cast<float *>(in), n_in, out, out_len, rate); }
void sfeo_resample_free(void *h) { delete static_cast<impl_resample *>(h); }

/* ---- sample conversions (same expressions in both builds; see sfe_oracle.c for citations) -- */
/* float -> 10-bit DAC words (examples/bpsk/bpsk.cxx:76-101): the reference's own convert_samples_to_bytes where the
   reference is compiled in (ref_bpsk.cxx), its restatement otherwise.  n a multiple of 4; returns the bytes written. */
#if defined(ORACLE_IMPL_REF)
int sfeo_ref_f32_to_dac10(const float *in, int n, uint8_t *out);
int sfeo_f32_to_dac10(const float *in, int n, uint8_t *out) { return sfeo_ref_f32_to_dac10(in, n, out); }
#else
int sfeo_f32_to_dac10(const float *in, int n, uint8_t *out) { return orc_f32_to_dac10(in, n, out); }
#endif

void sfeo_i16_to_f32(const int16_t *in, float *out, size_t n)
{
    const float q = 1.0f / 32767.0f;
    for (size_t i = 0; i < n; i++) out[i] = (float)in[i] * q;
}
void sfeo_u8_to_f32(const uint8_t *in, float *out, size_t n)
{
    const float q = 1.0f / 127.0f;
    for (size_t i = 0; i < n; i++) out[i] = (in[i] - 128) * q;
}

/* ---- chain driver -------------------------------------------------------------------------- */
struct sfeo_stage {
    int klass;          /* 0 = absent, 1 = decimate class, 2 = resample class */
    const float *taps;
    int n_taps, up, blk;
    float rate;
};
struct sfeo_chain_cfg {
    int in_fmt;         /* 0 = int16 IQ interleaved, 1 = u8 offset-binary IQ, 2 = float32 IQ */
    const float *bc_taps; /* blkconv stage, n_taps == 0 -> absent */
    int bc_ntaps, bc_fft;
    sfeo_stage a, b;
};

static int run_stage(const sfeo_stage &st, std::vector<float> &x, std::vector<float> &y)
{
    /* feed x through one decimate/resample object in calls of st.blk samples */
    const int n = (int)x.size();
    const int out_len = (int)floorf(st.blk / st.rate) + 2;
    y.resize((size_t)((double)n / st.rate) + 2 * (n / st.blk + 2) + 8);
    int produced = 0;
    impl_decimate *d = st.klass == 1 ? new impl_decimate(const_cast<float *>(st.taps), st.n_taps, st.up, st.blk) : NULL;
    impl_resample *r = st.klass == 2 ? new impl_resample(const_cast<float *>(st.taps), st.n_taps, st.up, st.blk) : NULL;
    for (int off = 0; off < n; off += st.blk) {
        const int n_in = n - off < st.blk ? n - off : st.blk;
        if ((size_t)produced + out_len > y.size()) y.resize(y.size() * 2 + out_len);
        produced += d ? d->process(&x[off], n_in, &y[produced], out_len, st.rate)
                      : r->process(&x[off], n_in, &y[produced], out_len, st.rate);
    }
    delete d;
    delete r;
    y.resize(produced);
    return produced;
}

struct scratch { std::vector<float> x, y; };  /* per-thread, reused across streams (no mmap churn) */

static int run_component(const sfeo_chain_cfg &c, const void *iq, long n, int comp, std::vector<float> &res, scratch &sc)
{
    std::vector<float> &x = sc.x, &y = sc.y;
    x.resize((size_t)n);
    if (c.in_fmt == 0) {
        const int16_t *p = static_cast<const int16_t *>(iq);
        const float q = 1.0f / 32767.0f;
        for (long i = 0; i < n; i++) x[i] = (float)p[2 * i + comp] * q;
    } else if (c.in_fmt == 1) {
        const uint8_t *p = static_cast<const uint8_t *>(iq);
        const float q = 1.0f / 127.0f;
        for (long i = 0; i < n; i++) x[i] = (p[2 * i + comp] - 128) * q;
    } else {
        const float *p = static_cast<const float *>(iq);
        for (long i = 0; i < n; i++) x[i] = p[2 * i + comp];
    }
    if (c.bc_ntaps > 0) {
        impl_blkconv bc(const_cast<float *>(c.bc_taps), c.bc_ntaps, c.bc_fft);
        const int blk = bc.get_blksize();
        float *buf = bc.get_process_buf();
        y.resize((size_t)n);
        for (long off = 0; off < n; off += blk) {
            const int m = n - off < blk ? (int)(n - off) : blk;
            memcpy(buf, &x[off], sizeof(float) * m);
            for (int i = m; i < blk; i++) buf[i] = 0.0f;
            bc.process();
            memcpy(&y[off], buf, sizeof(float) * m);
        }
        x.swap(y);
    }
    if (c.a.klass) { run_stage(c.a, x, y); x.swap(y); }
    if (c.b.klass) { run_stage(c.b, x, y); x.swap(y); }
    res.assign(x.begin(), x.end());
    return (int)res.size();
}

/* iq: [n_streams][n][2] in the format c.in_fmt; out: [n_streams][out_cap][2] float; n_out[n_streams].
   returns 0, or -1 when out_cap is too small. */
int sfeo_chain_run(const sfeo_chain_cfg *cfg, const void *iq, int n_streams, long n, float *out,
                   long out_cap, int *n_out, int n_threads)
{
    const sfeo_chain_cfg c = *cfg;
    const size_t in_stride = (size_t)n * 2 * (c.in_fmt == 0 ? 2 : c.in_fmt == 1 ? 1 : 4);
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n_streams) n_threads = n_streams;
    std::vector<int> fail(n_threads, 0);
    auto work = [&](int tid) {
        std::vector<float> ri, rq;
        scratch sc;
        for (int s = tid; s < n_streams; s += n_threads) {
            const char *src = static_cast<const char *>(iq) + in_stride * s;
            const int ni = run_component(c, src, n, 0, ri, sc);
            const int nq = run_component(c, src, n, 1, rq, sc);
            if (ni != nq || ni > out_cap) { fail[tid] = 1; n_out[s] = ni; continue; }
            float *o = out + (size_t)s * out_cap * 2;
            for (int i = 0; i < ni; i++) { o[2 * i] = ri[i]; o[2 * i + 1] = rq[i]; }
            n_out[s] = ni;
        }
    };
    if (n_threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; t++) th.emplace_back(work, t);
        for (auto &t : th) t.join();
    }
    for (int f : fail) if (f) return -1;
    return 0;
}

/* ---- ring_buffer scripted exerciser ---------------------------------------------------------
 * Re-expresses the reference's three CppUnit cases (gr-simplefe/lib/qa_simplefe.cc:103-166) as a
 * data-driven script so the same operations can be replayed on the reference template, the
 * port and the product header (tests/ring_driver.cxx) and compared.
 * ops[i] = {kind, arg}: kind 0 = write `arg` floats (values continue a running counter),
 *                       kind 1 = read with dst_len = arg using the pair-sum converter below.
 * trace receives, per op: return value, count, space, then for reads the converted values.     */
static int conv_copy(void *dst, void *src, int n)
{
    memcpy(dst, src, sizeof(float) * n);
    return (int)(sizeof(float) * n);
}
static int calc_same(int dst_len) { return dst_len; }

int sfeo_ring_script(int capacity, const int *ops, int n_ops, float *trace, int trace_cap)
{
#if defined(ORACLE_IMPL_REF)
    ring_buffer<float> rb(capacity);
#else
    orc_ring *rb = orc_ring_new(sizeof(float), capacity);
#endif
    int t = 0;
    float counter = 0.0f;
    std::vector<float> tmp;
    for (int i = 0; i < n_ops; i++) {
        const int kind = ops[2 * i], arg = ops[2 * i + 1];
        int ret;
        tmp.assign((size_t)(arg > 0 ? arg : 1), 0.0f);
        if (kind == 0) {
            for (int k = 0; k < arg; k++) tmp[k] = counter + (float)k;
#if defined(ORACLE_IMPL_REF)
            ret = rb.write(tmp.data(), arg);
#else
            ret = orc_ring_write(rb, tmp.data(), arg);
#endif
            if (ret) counter += (float)arg;
        } else {
#if defined(ORACLE_IMPL_REF)
            ret = rb.read(tmp.data(), (unsigned)arg, conv_copy, calc_same);
#else
            ret = orc_ring_read(rb, tmp.data(), (unsigned)arg, conv_copy, calc_same);
#endif
        }
        if (t + 3 + (kind == 1 ? ret : 0) > trace_cap) return -1;
        trace[t++] = (float)ret;
#if defined(ORACLE_IMPL_REF)
        trace[t++] = (float)rb.get_count();
        trace[t++] = (float)rb.get_space();
#else
        trace[t++] = (float)orc_ring_count(rb);
        trace[t++] = (float)orc_ring_space(rb);
#endif
        if (kind == 1) for (int k = 0; k < ret; k++) trace[t++] = tmp[k];
    }
#if !defined(ORACLE_IMPL_REF)
    orc_ring_free(rb);
#endif
    return t;
}

} /* extern "C" */
