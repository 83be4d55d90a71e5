/* gpu_sink.h -- GNU Radio sink blocks gr::simplefe::gpu_sink_c / gpu_sink_f.
 *
 * Shapes of the reference's sinks (gr-simplefe/include/simplefe/sink_c.h:36-52, sink_f.h:36-52, lib/sink_c_impl.h,
 * lib/sink_c_impl.cc:47-174, lib/sink_f_impl.cc): a gr::sync_block with one input (gr_complex / float) and no output,
 * created through a static make() returning a boost::shared_ptr; work() hands the items to a ring (waiting for room)
 * and the device's transmit callback drains it as packed 10-bit DAC words (fill_tx_buffer, sink_c_impl.cc:118-144).
 * Here work() gives the samples to the GPU, which packs them (include/sfe_gpu.h section 6), and the callback copies
 * finished wire bytes out of a pinned ring.
 *
 * Opening the USB device stays with the caller (libsimpleFE is out of scope for this repository):
 *     gr::simplefe::gpu_sink_c::sptr snk = gr::simplefe::gpu_sink_c::make(ring_bytes, max_items);
 *     sfe_tx_enable(dev, 1, 1);                                                        // sink_c_impl.cc:86
 *     sfe_tx_start(dev, gr::simplefe::gpu_sink_c::tx_callback, snk->callback_data());  // sink_c_impl.cc:87
 * Compile- and run-checked against the stand-in headers under tests/cpp/stubs (GNU Radio, Boost, SWIG are absent).
 */
#ifndef INCLUDED_SIMPLEFE_GPU_SINK_H
#define INCLUDED_SIMPLEFE_GPU_SINK_H

#include <gnuradio/sync_block.h>
#include <boost/shared_ptr.hpp>

#include "../sfe_gpu.h"

namespace gr {
namespace simplefe {

class gpu_sink_c : virtual public gr::sync_block
{
public:
    typedef boost::shared_ptr<gpu_sink_c> sptr;
    /* ring_bytes: pinned ring of packed DAC bytes between work() and the USB thread (the reference holds 4 transfers of
       floats, sink_c_impl.cc:80-82); max_items: most items one work() call brings.  Throws std::runtime_error when there
       is no CUDA device (no CPU fallback). */
    static sptr make(int ring_bytes, int max_items);
    /* signature of libsimpleFE's sfe_callback (simpleFE.h:42); `data` is callback_data() */
    static int tx_callback(unsigned char *buffer, int length, void *data);
    virtual void *callback_data() = 0;
    virtual long underflows() = 0;   /* "U" count (sink_c_impl.cc:95-98) */
};

class gpu_sink_f : virtual public gr::sync_block
{
public:
    typedef boost::shared_ptr<gpu_sink_f> sptr;
    /* channel: 0 = I, 1 = Q, what the caller passes to sfe_tx_enable (sink_f_impl.cc: tx_i / tx_q from `channel`) */
    static sptr make(int ring_bytes, int max_items, int channel);
    static int tx_callback(unsigned char *buffer, int length, void *data);
    virtual void *callback_data() = 0;
    virtual long underflows() = 0;
    virtual int channel() = 0;
};

}  // namespace simplefe
}  // namespace gr

#endif /* INCLUDED_SIMPLEFE_GPU_SINK_H */
