/* gpu_source_c.h -- GNU Radio source block that delivers the GPU chain's output.
 *
 * Shape of the reference's blocks (gr-simplefe/include/simplefe/source_c.h:36-50, lib/source_c_impl.h:34-54):
 * a gr::sync_block with no inputs and one gr_complex output, created through a static make() that returns a
 * boost::shared_ptr, with the device's receive callback as a static member.  The reference's source_c converts
 * the ADC bytes to complex floats in work() (source_c_impl.cc:121-153); this block hands the same bytes to the
 * B200 chain (u8 wire format -> blkconv -> decimate -> resample, include/sfe_gpu.h section 5) and work() returns
 * the chain's output samples.
 *
 * Opening the USB device is the caller's business here (libsimpleFE is out of scope for this repository):
 *     gr::simplefe::gpu_source_c::sptr src = gr::simplefe::gpu_source_c::make(&desc, ring_bytes, block_samples);
 *     sfe_rx_start(dev, gr::simplefe::gpu_source_c::rx_callback, src->callback_data());   // simpleFE.h:59-62
 * GNU Radio, Boost and SWIG are absent from this image: the block is compile- and run-checked against the minimal
 * stand-in headers under tests/cpp/stubs (tests/test_gr_block.py).
 */
#ifndef INCLUDED_SIMPLEFE_GPU_SOURCE_C_H
#define INCLUDED_SIMPLEFE_GPU_SOURCE_C_H

#include <gnuradio/sync_block.h>
#include <boost/shared_ptr.hpp>
#include <vector>

#include "../sfe_gpu.h"

namespace gr {
namespace simplefe {

class gpu_source_c : virtual public gr::sync_block
{
public:
    typedef boost::shared_ptr<gpu_source_c> sptr;

    /* desc: one stream of SFE_IN_U8 (the device's offset-binary I/Q bytes); ring_bytes: pinned ring between the
       USB thread and the GPU (the reference holds 4 transfers, source_c_impl.cc:73-75); block_samples: samples
       per chain call.  Throws std::runtime_error when there is no CUDA device (no CPU fallback). */
    static sptr make(const sfe_chain_desc *desc, int ring_bytes, int block_samples);

    /* the same from plain vectors, the form SWIG and a GRC <make> string can call (grc/simplefe_gpu_source_c.xml):
       an empty tap vector leaves that component out; decimate then resample, blksize 4096 each as in
       examples/bpsk/bpsk.cxx:193-199; ring of 16 blocks */
    static sptr make(const std::vector<float> &blkconv_taps, int fft_len, const std::vector<float> &decim_taps, float decim_rate,
                     const std::vector<float> &resamp_taps, int resamp_up, float resamp_rate, int block_samples);

    /* signature of libsimpleFE's sfe_callback (simpleFE.h:42); `data` is callback_data() */
    static int rx_callback(unsigned char *buffer, int length, void *data);
    virtual void *callback_data() = 0;
    /* overflow ("O") and discarded odd-packet counters of the feed (source_c_impl.cc:84-108) */
    virtual long overflows() = 0;
    virtual long discarded() = 0;
};

}  // namespace simplefe
}  // namespace gr

#endif /* INCLUDED_SIMPLEFE_GPU_SOURCE_C_H */
