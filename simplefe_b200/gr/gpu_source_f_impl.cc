// gpu_source_f_impl.cc -- gr::simplefe::gpu_source_f over the C ABI (include/sfe_gpu.h section 5).
// Mirrors gr-simplefe/lib/source_f_impl.cc: no input, one float output (:45-47), rx_callback writes the packet into
// the ring ("O" when it is full, :92-106), work() blocks until it can fill the request (:131-147).  One real channel
// rides through the chain as the real part of an IQ stream whose imaginary bytes are 128, the format's zero.
#include <simplefe/gpu_source_f.h>

#include <gnuradio/io_signature.h>

#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace gr {
namespace simplefe {

class gpu_source_f_impl : public gpu_source_f
{
    sfe_rx_feed *m_feed;
    int m_channel;
    std::vector<float> m_backlog;
    std::vector<gr_complex> m_tmp;
    std::vector<unsigned char> m_pair;   // callback scratch: (b, 128) pairs; one USB thread, as in the reference

public:
    gpu_source_f_impl(const sfe_chain_desc *desc, int ring_bytes, int block_samples, int channel)
        : gr::sync_block("gpu_source_f", gr::io_signature::make(0, 0, 0), gr::io_signature::make(1, 1, sizeof(float))),
          m_feed(sfe_rx_feed_create(desc, ring_bytes, block_samples)), m_channel(channel)
    {
        if (!m_feed) throw std::runtime_error(std::string("gpu_source_f: ") + sfe_last_error());
        m_tmp.resize((size_t)sfe_rx_feed_stat(m_feed, 4) * 2);
    }
    ~gpu_source_f_impl() { sfe_rx_feed_destroy(m_feed); }

    void *callback_data() { return this; }
    long overflows() { return sfe_rx_feed_stat(m_feed, 0); }
    int channel() { return m_channel; }

    int write_data(unsigned char *buffer, int length)   // source_f_impl::write_data, :92-106
    {
        if (length <= 0) return 0;
        m_pair.resize(2 * (size_t)length);
        for (int i = 0; i < length; i++) {
            m_pair[2 * i] = buffer[i];
            m_pair[2 * i + 1] = 128;
        }
        return sfe_rx_feed_callback(m_pair.data(), 2 * length, m_feed);
    }

    int work(int noutput_items, gr_vector_const_void_star &input_items, gr_vector_void_star &output_items)
    {
        (void)input_items;
        float *out = (float *)output_items[0];
        while ((int)m_backlog.size() < noutput_items) {
            const long n = sfe_rx_feed_pump(m_feed, m_tmp.data(), (long)m_tmp.size(), 100);
            if (n < 0) throw std::runtime_error(std::string("gpu_source_f: ") + sfe_last_error());
            for (long i = 0; i < n; i++) m_backlog.push_back(m_tmp[i].real());
        }
        memcpy(out, m_backlog.data(), sizeof(float) * (size_t)noutput_items);
        m_backlog.erase(m_backlog.begin(), m_backlog.begin() + noutput_items);
        return noutput_items;
    }
};

gpu_source_f::sptr gpu_source_f::make(const sfe_chain_desc *desc, int ring_bytes, int block_samples, int channel)
{
    if (channel != 0 && channel != 1) throw std::out_of_range("channel is 0 (I) or 1 (Q)\n");
    return gpu_source_f::sptr(new gpu_source_f_impl(desc, ring_bytes, block_samples, channel));
}

int gpu_source_f::rx_callback(unsigned char *buffer, int length, void *data)
{
    return static_cast<gpu_source_f_impl *>(data)->write_data(buffer, length);
}

}  // namespace simplefe
}  // namespace gr
