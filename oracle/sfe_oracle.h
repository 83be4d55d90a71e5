/* sfe_oracle.h -- TEST INFRASTRUCTURE ONLY: C interface of the CPU restatement (see sfe_oracle.c). */
#ifndef SFE_ORACLE_H_
#define SFE_ORACLE_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_blkconv orc_blkconv;
typedef struct orc_decimate orc_decimate;
typedef struct orc_resample orc_resample;
typedef struct orc_ring orc_ring;

orc_blkconv *orc_blkconv_new(const float *taps, int n_taps, int fft_len);
int orc_blkconv_blksize(orc_blkconv *b);
float *orc_blkconv_buf(orc_blkconv *b);
void orc_blkconv_process(orc_blkconv *b);
void orc_blkconv_free(orc_blkconv *b);

orc_decimate *orc_decimate_new(const float *taps, int n_taps, int up, int blksize);
int orc_decimate_process(orc_decimate *d, const float *in, int n_in, float *out, int out_len, float rate);
void orc_decimate_free(orc_decimate *d);

orc_resample *orc_resample_new(const float *taps, int n_taps, int up, int blksize);
int orc_resample_process(orc_resample *r, const float *in, int n_in, float *out, int out_len, float rate);
void orc_resample_free(orc_resample *r);

orc_ring *orc_ring_new(int elem_size, int capacity);
void orc_ring_free(orc_ring *r);
int orc_ring_space(orc_ring *r);
int orc_ring_count(orc_ring *r);
int orc_ring_write(orc_ring *r, const void *data, int len);
int orc_ring_read(orc_ring *r, void *dst, unsigned dst_len,
                  int (*conv)(void *, void *, int), int (*calc_src_len)(int));

void orc_i16_to_f32(const int16_t *in, float *out, size_t n);
void orc_u8_to_f32(const uint8_t *in, float *out, size_t n);
int orc_f32_to_dac10(const float *in, int n, uint8_t *out);

#ifdef __cplusplus
}
#endif
#endif
