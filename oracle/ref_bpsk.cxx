/*
 * ref_bpsk.cxx -- TEST INFRASTRUCTURE ONLY (part of oracle/_ref, built by `make ref` where /root/reference is mounted).
 *
 * Compiles the reference's OWN transmitter, examples/bpsk/bpsk.cxx, verbatim and in place (nothing is copied into this
 * repository) so that the checkers can call its file-local functions:
 *   convert_samples_to_bytes  bpsk.cxx:76-101   float -> 10-bit DAC words, 4 samples -> 5 bytes
 *   calc_num_samples          bpsk.cxx:70-73    samples a TX packet of `bytes` consumes
 *   tx_callback               bpsk.cxx:104-119  the USB thread's side of ring_buffer<float> ("U" on underflow)
 * bpsk.cxx's main() is renamed (it opens the USB device); the libsimpleFE functions it names are out of scope for this
 * repository and are satisfied by stubs that are never called.
 */
#include <cstdint>
#include <cstdlib>

#define main sfeo_reference_bpsk_main
#include "examples/bpsk/bpsk.cxx"
#undef main

extern "C" {
static void unreachable() { abort(); }
sfe *sfe_init() { unreachable(); return 0; }
void sfe_close(sfe *) { unreachable(); }
int sfe_set_sample_rate(sfe *, unsigned) { unreachable(); return -1; }
void sfe_tx_enable(sfe *, int, int) { unreachable(); }
int sfe_tx_start(sfe *, sfe_callback *, void *) { unreachable(); return -1; }
void sfe_stop_tx(sfe *) { unreachable(); }
void sfe_reset_board(sfe *) { unreachable(); }
unsigned get_real_sample_rate(sfe *) { unreachable(); return 0; }

int sfeo_ref_f32_to_dac10(const float *in, int n, uint8_t *out) { return convert_samples_to_bytes(out, const_cast<float *>(in), n); }
int sfeo_ref_dac10_samples(int bytes) { return calc_num_samples(bytes); }
/* ring_buffer<float>(capacity) filled with x[0..n), then tx_callback for packets of `pkt` bytes until the ring runs dry:
   what the USB thread would have put on the wire.  Returns the bytes written to out. */
long sfeo_ref_tx_packets(const float *x, int n, int capacity, int pkt, uint8_t *out, long out_cap)
{
    ring_buffer<float> rb(capacity);
    if (rb.write(x, n) != n) return -1;
    long w = 0;
    while (rb.get_count() >= calc_num_samples(pkt) && w + pkt <= out_cap) {
        tx_callback(out + w, pkt, &rb);
        w += pkt;
    }
    return w;
}
}
