/* gpu_source_f.h -- GNU Radio source block for ONE real channel through the GPU chain.
 *
 * Shape of the reference's source_f (gr-simplefe/include/simplefe/source_f.h:36-52, lib/source_f_impl.cc:40-150): a
 * gr::sync_block with no input and one float output, make(sample_rate, channel) enabling only the I or the Q converter
 * (sfe_rx_enable(rx_i, rx_q), source_f_impl.cc:80), whose bytes arrive one per sample and become floats by
 * (b-128)*(1/127) (source_f_impl.cc:120-129).  Here the bytes run through the B200 chain as the real part of a stream
 * (the imaginary part is the format's zero, 128) and work() returns the real part of the chain's output.
 */
#ifndef INCLUDED_SIMPLEFE_GPU_SOURCE_F_H
#define INCLUDED_SIMPLEFE_GPU_SOURCE_F_H

#include <gnuradio/sync_block.h>
#include <boost/shared_ptr.hpp>
#include <vector>

#include "../sfe_gpu.h"

namespace gr {
namespace simplefe {

class gpu_source_f : virtual public gr::sync_block
{
public:
    typedef boost::shared_ptr<gpu_source_f> sptr;
    /* desc: one stream of SFE_IN_U8; ring_bytes / block_samples as gpu_source_c; channel: 0 = I, 1 = Q (for the caller's
       sfe_rx_enable).  Throws std::runtime_error when there is no CUDA device. */
    static sptr make(const sfe_chain_desc *desc, int ring_bytes, int block_samples, int channel);
    /* sfe_callback signature (simpleFE.h:42): `length` bytes = `length` samples of the enabled channel */
    static int rx_callback(unsigned char *buffer, int length, void *data);
    virtual void *callback_data() = 0;
    virtual long overflows() = 0;
    virtual int channel() = 0;
};

}  // namespace simplefe
}  // namespace gr

#endif /* INCLUDED_SIMPLEFE_GPU_SOURCE_F_H */
