/*
 * sfe_gpu.h -- C ABI of libsfegpu.so, the B200 (sm_100a) implementation of simpleFE's libdsp
 * sample-stream hot path: blkconv -> decimate -> resample, fed through ring_buffer.
 *
 * Plain pointers and sizes only; no C++/torch types.  Every entry point names the reference
 * interface it stands behind (paths relative to the reference tree, wnmusic/simpleFE).
 * There is NO CPU fallback anywhere in this library: without a usable CUDA device every
 * create/process call fails (NULL / negative status) and sfe_last_error() says why.
 *
 * Two groups:
 *  (1) single-object entry points -- exactly what the reference's C++ classes need, bound by the
 *      drop-in headers include/libdsp/{blkconv,decimate,resample,ringbuf}.h and by the ctypes
 *      `pydsp` module (simplefe_b200/pydsp.py, mirrors libdsp/test/pydsp.i).
 *  (2) batched, additive entry points (sfe_chain_*) -- S independent IQ streams that share one
 *      filter set; this is where the throughput is.  Per stream the result equals a fresh set of
 *      reference objects per real component (I and Q) fed the same samples.
 */
#ifndef SFE_GPU_H_
#define SFE_GPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library itself is built with -fvisibility=hidden */
#endif

#define SFE_OK 0
#define SFE_ERR_ARG (-1)      /* bad argument (message in sfe_last_error) */
#define SFE_ERR_CUDA (-2)     /* CUDA runtime/driver failure: fatal, never papered over */
#define SFE_ERR_NODEV (-3)    /* no CUDA device / not sm_100 */
#define SFE_ERR_CAP (-4)      /* caller-provided capacity too small; the stream state is unchanged */
#define SFE_ERR_HOST (-5)     /* host-side failure (e.g. out of memory); no C++ exception ever crosses this ABI */

/* ---- library ---------------------------------------------------------------------------- */
const char *sfe_version(void);
const char *sfe_last_error(void);          /* thread-local, "" when none */
int sfe_device_count(void);                /* >= 0, or negative status */
int sfe_set_device(int device);            /* device used by objects created afterwards */
int64_t sfe_kernel_launches(void);         /* kernels launched by this library since load */

/* ---- raw memory / stream helpers (for hosts without their own CUDA bindings) -------------- */
void *sfe_pinned_alloc(size_t bytes);      /* cudaHostAlloc, pinned for every device of the box; NULL on failure */
void sfe_pinned_free(void *p);
void *sfe_dev_alloc(size_t bytes);
void sfe_dev_free(void *p);
int sfe_memcpy_h2d(void *dst_dev, const void *src_host, size_t bytes, void *stream);
int sfe_memcpy_d2h(void *dst_host, const void *src_dev, size_t bytes, void *stream);
int sfe_stream_sync(void *stream);         /* NULL = the library's default stream */

/* ---- (1a) blkconv: libdsp/blkconv.h:35-48, blkconv.cxx:34-122 ------------------------------ */
typedef struct sfe_blkconv sfe_blkconv;
/* blkconv::blkconv(float *taps, int n_taps, int fft_len)  blkconv.h:38.  Taps are copied.  Any fft_len >= n_taps
   works, as in the reference (which hands it to FFTW, blkconv.cxx:60,72-73): the block size is fft_len+1-n_taps and
   the stream is the causal convolution whatever fft_len is, so the kernels run overlap-save with an internal
   power-of-two transform (fft_len itself when it is a power of two in [32, 8192]); taps longer than 4097 run as a
   direct FIR.  NULL + error only for bad arguments or without a CUDA device. */
sfe_blkconv *sfe_blkconv_create(const float *taps, int n_taps, int fft_len);
int sfe_blkconv_blksize(sfe_blkconv *h);   /* blkconv::get_blksize  blkconv.h:40 = fft_len+1-n_taps */
float *sfe_blkconv_buf(sfe_blkconv *h);    /* blkconv::get_process_buf blkconv.h:44: pinned host
                                              buffer of fft_len+2 floats owned by the object */
int sfe_blkconv_process(sfe_blkconv *h);   /* blkconv::process blkconv.h:48: consumes buf[0..blk),
                                              leaves the next blk output samples there; returns
                                              after the result is visible in host memory */
void sfe_blkconv_destroy(sfe_blkconv *h);  /* blkconv::~blkconv */

/* ---- (1b) decimate / resample: libdsp/decimate.h:33-46, resample.h:33-47 -------------------- */
typedef struct sfe_polyphase sfe_polyphase;
/* decimate::decimate(float *taps, int n_taps, int upsample, int blksize)  decimate.h:37 */
sfe_polyphase *sfe_decimate_create(const float *taps, int n_taps, int upsample, int blksize);
/* resample::resample(float *taps, int n_taps, int upsample, int blksize)  resample.h:39 */
sfe_polyphase *sfe_resample_create(const float *taps, int n_taps, int upsample, int blksize);
/* int process(float* in, int n_in, float *out, int out_len, float rate)  decimate.h:45 /
   resample.h:47.  Returns the number of output samples.  On the reference's argument errors
   (decimate.cxx:75-87, resample.cxx:91-98) prints the reference's message to stdout and
   returns 0.  A CUDA failure returns a negative status. */
int sfe_polyphase_process(sfe_polyphase *h, const float *in, int n_in, float *out, int out_len,
                          float rate);
void sfe_polyphase_destroy(sfe_polyphase *h);
/* 1 = reproduce the reference's single-precision operation order bit for bit (separate multiply
   and add, ascending taps); 0 (default) = fused multiply-add.  Applies to later process calls. */
int sfe_polyphase_set_exact(sfe_polyphase *h, int exact);

/* ---- (2) batched chain ---------------------------------------------------------------------- */
enum {
    SFE_IN_I16 = 0,   /* int16 I,Q interleaved; f = (float)s * (1.0f/32767.0f)        */
    SFE_IN_U8 = 1,    /* device wire format, offset binary; f = (b-128)*(1.0f/127.0f)
                         gr-simplefe/lib/source_c_impl.cc:121-132                                  */
    SFE_IN_F32 = 2    /* float32 I,Q interleaved (complex64)                                       */
};
enum { SFE_KLASS_NONE = 0, SFE_KLASS_DECIMATE = 1, SFE_KLASS_RESAMPLE = 2 };
enum {
    SFE_FLAG_EXACT = 1,        /* polyphase stages in reference operation order (slow path)        */
    SFE_FLAG_NO_FUSE = 2,      /* run blkconv and stage A as separate kernels even when fusable    */
    SFE_FLAG_NO_TENSOR = 4,    /* never use the tensor-core Toeplitz kernel                        */
    SFE_FLAG_NO_BANK = 8       /* never fold the whole chain into one polyphase bank (K3)          */
};

typedef struct {
    int klass;            /* SFE_KLASS_*; NONE = stage absent                                      */
    const float *taps;    /* copied                                                                */
    int n_taps;
    int up;               /* `upsample` ctor argument                                              */
    int blk;              /* `blksize` ctor argument = samples per reference process() call        */
    float rate;           /* `rate` process() argument                                             */
} sfe_stage_desc;

typedef struct {
    int in_fmt;           /* SFE_IN_*                                                              */
    int n_streams;        /* S complex (IQ) streams, all sharing the filters below                 */
    const float *bc_taps; /* blkconv stage (blkconv.h:38); bc_ntaps == 0 -> absent                 */
    int bc_ntaps;
    int bc_fft;
    sfe_stage_desc a;     /* first polyphase stage (e.g. decimate)                                 */
    sfe_stage_desc b;     /* second polyphase stage (e.g. resample)                                */
    int flags;            /* SFE_FLAG_*                                                            */
} sfe_chain_desc;

typedef struct sfe_chain sfe_chain;

sfe_chain *sfe_chain_create(const sfe_chain_desc *desc);
void sfe_chain_destroy(sfe_chain *c);
/* forget all stream state (overlap, histories, resampler phase): next call starts a new stream */
int sfe_chain_reset(sfe_chain *c);
/* upper bound of output samples per stream for a call with n_in input samples per stream */
long sfe_chain_max_out(sfe_chain *c, long n_in);
/* Streaming call, device buffers.  d_in: S rows of n_in complex samples in desc.in_fmt, row pitch
   in_pitch (in complex samples).  d_out: S rows of complex64, row pitch out_pitch >= n_out.
   *n_out receives the per-stream output count of this call (identical for all streams: the
   output schedule is data independent).  Work is enqueued on `stream` (a cudaStream_t, NULL =
   library default stream) and the call returns without synchronising.
   Per stream the call is equivalent to feeding the reference objects: blkconv in blocks of
   fft_len+1-n_taps, stage a in ceil(n_in/a.blk) process() calls of a.blk samples
   (out_len = floor(a.blk/a.rate)+2), stage b likewise over stage a's output. */
int sfe_chain_process_dev(sfe_chain *c, const void *d_in, long n_in, long in_pitch, void *d_out,
                          long out_pitch, long *n_out, void *stream);
/* Same, host buffers (pinned for full speed): the S streams are cut into groups whose H2D copy,
   kernels and D2H copy are double-buffered on separate CUDA streams (the GPU-side successor of
   ring_buffer's producer/consumer decoupling).  Returns after the result is in h_out. */
int sfe_chain_process_host(sfe_chain *c, const void *h_in, long n_in, long in_pitch, void *h_out,
                           long out_pitch, long *n_out);
/* describe the kernels the last process call used, e.g. "toeplitz_i8+polyphase_generic" */
const char *sfe_chain_plan(sfe_chain *c);
/* Per-kernel timing for roofline reports: sfe_chain_process_dev brackets the main kernel of every op with
   CUDA events on the caller's stream in every `on`-th call (1 = every call, 0 = off).  A bracketed call costs
   two event records per op and keeps the state-carry kernel from overlapping the main kernel's tail. */
int sfe_chain_n_ops(sfe_chain *c);
const char *sfe_chain_op_name(sfe_chain *c, int op);
int sfe_chain_profile_enable(sfe_chain *c, int on);
/* mean duration in ms of op's main kernel over the calls since enable / the previous read
   (synchronises on the recorded events); *n_samples = launches averaged */
int sfe_chain_profile_read(sfe_chain *c, int op, float *mean_ms, int *n_samples);

/* ---- (2b) the whole box: the batched chain over every GPU of one node (additive) ---------------------- */
/* The path shards by stream and by nothing else -- all state is per object (libdsp/blkconv.h:51-61, decimate.h:48-59,
   resample.h:49-59) -- so device g of G owns the contiguous stream range [g*S/G, (g+1)*S/G) with its own chain, its own
   host worker thread, and its own pinned-host/device double buffering on two CUDA streams.  No collective, no inter-GPU
   traffic.  n_devices: how many of the visible devices to use (0 = all; devices 0..n-1).  desc.n_streams is the TOTAL. */
typedef struct sfe_box sfe_box;
sfe_box *sfe_box_create(const sfe_chain_desc *desc, int n_devices);
void sfe_box_destroy(sfe_box *b);
int sfe_box_n_devices(sfe_box *b);
/* streams [*first, *first + *count) run on the returned device ordinal; index < sfe_box_n_devices */
int sfe_box_stream_range(sfe_box *b, int index, int *first, int *count);
int sfe_box_reset(sfe_box *b);
long sfe_box_max_out(sfe_box *b, long n_in);
const char *sfe_box_plan(sfe_box *b);
/* Host buffers holding ALL S streams (rows as in sfe_chain_process_host; allocate them with sfe_pinned_alloc, which
   pins for every device): each device's worker runs its stream range through sfe_chain_process_host at the same time.
   Returns when every device has finished and the whole result is in h_out. */
int sfe_box_process_host(sfe_box *b, const void *h_in, long n_in, long in_pitch, void *h_out, long out_pitch, long *n_out);
/* Device-resident: d_in[g] / d_out[g] are the row blocks of device ordinal g's stream range, resident on that device;
   streams[g] the cudaStream_t to enqueue on (streams == NULL: each device's default stream).  Enqueues on every
   device and returns without synchronising (sfe_box_sync waits for all devices). */
int sfe_box_process_dev(sfe_box *b, const void *const *d_in, long n_in, long in_pitch, void *const *d_out, long out_pitch,
                        long *n_out, void *const *streams);
int sfe_box_sync(sfe_box *b);

/* ---- (3) ring_buffer<T>: libdsp/ringbuf.h:37-142 (same body in gr-simplefe/lib/ringbuf.h:40-145) ------ */
/* Byte-oriented twin of the reference template; include/libdsp/ringbuf.h forwards ring_buffer<T> to
   it with elem_bytes = sizeof(T).  Storage is page-locked host memory.  No internal locking: callers
   serialise, as the reference's users do (bpsk.cxx:108-116,132-170). */
typedef struct sfe_ring sfe_ring;
typedef int (*sfe_ring_conv_fn)(void *dst, void *src, int src_len);   /* returns BYTES written to dst */
typedef int (*sfe_ring_len_fn)(int dst_len);                           /* elements needed for dst_len */
sfe_ring *sfe_ring_create(int elem_bytes, int capacity);   /* ring_buffer() / ring_buffer(int)  :40-53;
                                                              capacity 0 = no storage yet             */
int sfe_ring_alloc(sfe_ring *r, int capacity);             /* alloc_buffer  :60-69                    */
void sfe_ring_destroy(sfe_ring *r);                        /* ~ring_buffer  :54-57                    */
int sfe_ring_space(sfe_ring *r);                           /* get_space     :71-74                    */
int sfe_ring_count(sfe_ring *r);                           /* get_count     :75-78                    */
int sfe_ring_write(sfe_ring *r, const void *data, int len);/* write         :80-101: len or 0         */
int sfe_ring_read(sfe_ring *r, void *dst, unsigned dst_len, sfe_ring_conv_fn conv,
                  sfe_ring_len_fn calc_src_len);           /* read          :106-134: elements or 0   */
/* additive: move the next n elements to device memory with <= 2 cudaMemcpyAsync on `stream` and
   return at once (n, or 0 when fewer are buffered); their space becomes writable again when the
   copy has completed.  This is the producer/consumer decoupling of the reference with a copy
   engine as the consumer. */
int sfe_ring_read_dev(sfe_ring *r, void *d_dst, int n, void *stream);

/* ---- (4) batched BPSK transmit path: examples/bpsk/bpsk.cxx:122-174 for many channels (additive) -------- */
/* Per channel: symbol bits -> impulses +-amp every sps samples (a set bit sends -amp, bpsk.cxx:150) ->
   pulse shaping with `taps` (the reference runs blkconv(RRC_TAPS_111, 111, 2048), bpsk.cxx:56-59,125,162) ->
   float samples (what the reference writes into ring_buffer<float>, bpsk.cxx:164) or, fused, the 10-bit DAC
   words of convert_samples_to_bytes (bpsk.cxx:76-101).  The waveform is the causal convolution from zero
   state and continues across calls. */
enum { SFE_TX_F32 = 0, SFE_TX_DAC10 = 1 };
typedef struct sfe_bpsk_tx sfe_bpsk_tx;
sfe_bpsk_tx *sfe_bpsk_tx_create(const float *taps, int n_taps, int sps, float amp, int n_channels);
void sfe_bpsk_tx_destroy(sfe_bpsk_tx *t);
int sfe_bpsk_tx_reset(sfe_bpsk_tx *t);
/* d_bits: n_channels rows of packed symbols, bit j of word w = symbol 32 w + j (the reference consumes
   rand() words LSB first, bpsk.cxx:148-150); row pitch in 32-bit words.  n_sym symbols per channel; for
   SFE_TX_DAC10, n_sym * sps must be a multiple of 4.  Output rows: n_sym * sps floats, or n_sym * sps / 4 * 5
   bytes; out_pitch in BYTES.  Enqueues on `stream`, returns without synchronising. */
int sfe_bpsk_tx_process_dev(sfe_bpsk_tx *t, const uint32_t *d_bits, long n_sym, long pitch_words, void *d_out,
                            long out_pitch, int out_fmt, void *stream);

/* ---- (5) streaming front door: device callback -> pinned ring -> GPU chain (additive) --------------------- */
/* sfe_rx_feed_callback has the signature of libsimpleFE's sfe_callback (libsimpleFE/simpleFE.h:42) and the
   behaviour of gr-simplefe's rx path (gr-simplefe/lib/source_c_impl.cc:84-108): pass it, with the feed as
   userdata, to sfe_rx_start (simpleFE.h:59-62).  Odd-length packets are discarded with the reference's message,
   a full ring prints "O" on stderr and drops the packet.  The chain must be one stream of SFE_IN_U8. */
typedef struct sfe_rx_feed sfe_rx_feed;
sfe_rx_feed *sfe_rx_feed_create(const sfe_chain_desc *desc, int ring_bytes, int block_samples);
void sfe_rx_feed_destroy(sfe_rx_feed *f);
int sfe_rx_feed_callback(unsigned char *buffer, int length, void *feed);
/* consumer side (the role of source_c_impl::work, source_c_impl.cc:134-153): runs every whole block that is
   buffered through the chain and appends the complex64 outputs to `out` (host memory, cap in samples); waits
   up to wait_ms for the first block.  Returns the number of output samples, or a negative status. */
long sfe_rx_feed_pump(sfe_rx_feed *f, void *out, long cap, int wait_ms);
/* which: 0 overflows ("O"), 1 discarded odd packets, 2 bytes accepted, 3 bytes waiting in the ring,
   4 the most output samples one block can produce (pump needs at least that much room in `out`) */
long sfe_rx_feed_stat(sfe_rx_feed *f, int which);

/* ---- (6) transmit side: float samples -> GPU DAC packing -> pinned byte ring -> device callback (additive) --------------- */
/* The counterpart of the reference's sink blocks (gr-simplefe/lib/sink_c_impl.cc:84-174, sink_f_impl.cc): work() writes
   float samples (a complex item is two floats), the USB thread's sfe_callback drains packed 10-bit DAC words.  The
   packing (fill_tx_buffer, sink_c_impl.cc:118-144 = bpsk.cxx:76-101) runs on the GPU at write time; the ring holds wire
   bytes in pinned memory.  One producer thread, one consumer thread. */
typedef struct sfe_tx_feed sfe_tx_feed;
sfe_tx_feed *sfe_tx_feed_create(int ring_bytes, int block_floats);   /* block_floats: most floats per write */
void sfe_tx_feed_destroy(sfe_tx_feed *f);
/* n <= block_floats floats; waits up to wait_ms for room in the ring (the reference's work() waits on a condition variable,
   sink_c_impl.cc:165-169).  Returns n, 0 when there still was no room, or a negative status.  Up to 3 floats of an
   unfinished quad are kept for the next write. */
long sfe_tx_feed_write(sfe_tx_feed *f, const float *x, long n, int wait_ms);
/* sfe_callback signature (libsimpleFE/simpleFE.h:42) for sfe_tx_start, `feed` as userdata: fills length/5*5 bytes, or
   zeros the buffer and prints "U" on stderr when the ring holds fewer (sink_c_impl.cc:91-104) */
int sfe_tx_feed_callback(unsigned char *buffer, int length, void *feed);
/* which: 0 underflows ("U"), 1 bytes delivered, 2 bytes waiting, 3 floats a write can take now */
long sfe_tx_feed_stat(sfe_tx_feed *f, int which);

/* ---- (7) a process() call that allocates nothing ---------------------------------------------------------------------- */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
   (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  One decimate / resample process() call
   (the arithmetic of (1b)) on the calling thread's CURRENT CUDA device, with every buffer the caller's: device pointers for
   the taps (n_taps floats as given to the reference constructor), the input (n_in floats), the output (room for
   min(out_len, floor(n_in/rate) + 2) floats) and the history -- the last n_hist input samples BEFORE this call, normally the
   tail of the previous call's input buffer; samples older than that read as zero, so a stream needs
   n_hist >= (n_taps - 1) / upsample + 1 from its second call on, and NULL / 0 is a fresh object.  The three scalars a
   reference object carries between calls (decimate.h:55-59) live in the caller's sfe_walk_state: zero it for a fresh
   object; it is advanced on success and untouched on failure.  exact != 0 is sfe_polyphase_set_exact.
   The call's schedule table is the only scratch memory, and it sits in a static device array inside the library
   (8 bytes x SFE_NOALLOC_MAX_OUT = 2 MB on every GPU that loads libsfegpu, until exit): no cudaMalloc, cudaFree or
   allocating library routine runs.  A call that would produce more than SFE_NOALLOC_MAX_OUT outputs is refused with
   SFE_ERR_CAP (message in sfe_last_error; feed the stream in smaller calls).  All calls on a device share the array:
   calls that overlap on the GPU (two streams, two host threads) race on it -- issue them on one stream, or serialise them.
   Asynchronous on `stream` (0 = the legacy default stream).  Returns the number of outputs written, 0 after printing the
   reference's message for its own argument checks, or a negative status. */
#define SFE_NOALLOC_MAX_OUT (1 << 18)
typedef struct sfe_walk_state { int pos; float mu; int held; } sfe_walk_state;
int sfe_polyphase_process_noalloc(int klass, const float *d_taps, int n_taps, int upsample, sfe_walk_state *state,
                                  const float *d_hist, int n_hist, const float *d_in, int n_in, float *d_out, int out_len,
                                  float rate, int exact, void *stream);

/* ---- sample format helpers (device wire formats) ---------------------- */
/* float -> 10-bit DAC words, 4 samples -> 5 bytes: examples/bpsk/bpsk.cxx:76-101,
   gr-simplefe/lib/sink_c_impl.cc:118-144.  n must be a multiple of 4; device pointers. */
int sfe_pack_dac10_dev(const float *d_in, long n, uint8_t *d_out, void *stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* SFE_GPU_H_ */
