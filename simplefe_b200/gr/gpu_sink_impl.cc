// gpu_sink_impl.cc -- gr::simplefe::gpu_sink_c / gpu_sink_f over the C ABI (include/sfe_gpu.h section 6).
// Mirrors gr-simplefe/lib/sink_c_impl.cc and sink_f_impl.cc: the constructor sets the block's signature (one input, no
// output, sink_c_impl.cc:48-50), tx_callback forwards to the feed (:84-104), work() waits for room, hands the items over
// and consumes them (:158-174).
#include <simplefe/gpu_sink.h>

#include <gnuradio/io_signature.h>

#include <stdexcept>
#include <string>

namespace gr {
namespace simplefe {

namespace {

// what both sinks share: `floats_per_item` floats of every input item go to the feed
class tx_side
{
protected:
    sfe_tx_feed *m_feed;
    int m_max_floats;

    tx_side(int ring_bytes, int max_floats) : m_feed(sfe_tx_feed_create(ring_bytes, max_floats)), m_max_floats(max_floats)
    {
        if (!m_feed) throw std::runtime_error(std::string("gpu_sink: ") + sfe_last_error());
    }
    ~tx_side() { sfe_tx_feed_destroy(m_feed); }

    void push(const float *x, long n)
    {
        while (n > 0) {
            const long m = n < m_max_floats ? n : m_max_floats;
            // the reference waits on its condition variable until the ring has room (sink_c_impl.cc:165-169)
            long w;
            while ((w = sfe_tx_feed_write(m_feed, x, m, 100)) == 0) {
            }
            if (w < 0) throw std::runtime_error(std::string("gpu_sink: ") + sfe_last_error());
            x += m;
            n -= m;
        }
    }
};

class gpu_sink_c_impl : public gpu_sink_c, private tx_side
{
public:
    gpu_sink_c_impl(int ring_bytes, int max_items)
        : gr::sync_block("gpu_sink_c", gr::io_signature::make(1, 1, sizeof(gr_complex)), gr::io_signature::make(0, 0, 0)),
          tx_side(ring_bytes, 2 * max_items)
    {
    }
    void *callback_data() { return m_feed; }
    long underflows() { return sfe_tx_feed_stat(m_feed, 0); }

    int work(int noutput_items, gr_vector_const_void_star &input_items, gr_vector_void_star &output_items)
    {
        (void)output_items;
        push(reinterpret_cast<const float *>(input_items[0]), 2L * noutput_items);   // a complex item is (I, Q)
        consume_each(noutput_items);
        return 0;
    }
};

class gpu_sink_f_impl : public gpu_sink_f, private tx_side
{
    int m_channel;

public:
    gpu_sink_f_impl(int ring_bytes, int max_items, int channel)
        : gr::sync_block("gpu_sink_f", gr::io_signature::make(1, 1, sizeof(float)), gr::io_signature::make(0, 0, 0)),
          tx_side(ring_bytes, max_items), m_channel(channel)
    {
    }
    void *callback_data() { return m_feed; }
    long underflows() { return sfe_tx_feed_stat(m_feed, 0); }
    int channel() { return m_channel; }

    int work(int noutput_items, gr_vector_const_void_star &input_items, gr_vector_void_star &output_items)
    {
        (void)output_items;
        push(static_cast<const float *>(input_items[0]), noutput_items);
        consume_each(noutput_items);
        return 0;
    }
};

}  // namespace

gpu_sink_c::sptr gpu_sink_c::make(int ring_bytes, int max_items) { return gpu_sink_c::sptr(new gpu_sink_c_impl(ring_bytes, max_items)); }
int gpu_sink_c::tx_callback(unsigned char *buffer, int length, void *data) { return sfe_tx_feed_callback(buffer, length, data); }

gpu_sink_f::sptr gpu_sink_f::make(int ring_bytes, int max_items, int channel)
{
    if (channel != 0 && channel != 1) throw std::out_of_range("channel is 0 (I) or 1 (Q)\n");
    return gpu_sink_f::sptr(new gpu_sink_f_impl(ring_bytes, max_items, channel));
}
int gpu_sink_f::tx_callback(unsigned char *buffer, int length, void *data) { return sfe_tx_feed_callback(buffer, length, data); }

}  // namespace simplefe
}  // namespace gr
