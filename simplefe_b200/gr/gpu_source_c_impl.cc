// gpu_source_c_impl.cc -- implementation of gr::simplefe::gpu_source_c over the C ABI (include/sfe_gpu.h section 5).
// Mirrors gr-simplefe/lib/source_c_impl.cc: the constructor sets the block's signature (no input, one gr_complex
// output, :45-48), rx_callback forwards to the feed (:84-108), work() blocks until it can fill the request (:134-153).
#include <simplefe/gpu_source_c.h>

#include <gnuradio/io_signature.h>

#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace gr {
namespace simplefe {

class gpu_source_c_impl : public gpu_source_c
{
    sfe_rx_feed *m_feed;
    std::vector<gr_complex> m_backlog;   // chain outputs not yet handed to the scheduler
    std::vector<gr_complex> m_tmp;

public:
    gpu_source_c_impl(const sfe_chain_desc *desc, int ring_bytes, int block_samples)
        : gr::sync_block("gpu_source_c", gr::io_signature::make(0, 0, 0), gr::io_signature::make(1, 1, sizeof(gr_complex))),
          m_feed(sfe_rx_feed_create(desc, ring_bytes, block_samples))
    {
        if (!m_feed) throw std::runtime_error(std::string("gpu_source_c: ") + sfe_last_error());
        m_tmp.resize((size_t)sfe_rx_feed_stat(m_feed, 4) * 2);   // room for two blocks' worth of outputs per pump
    }
    ~gpu_source_c_impl() { sfe_rx_feed_destroy(m_feed); }

    void *callback_data() { return m_feed; }
    long overflows() { return sfe_rx_feed_stat(m_feed, 0); }
    long discarded() { return sfe_rx_feed_stat(m_feed, 1); }

    // Where all the action really happens
    int work(int noutput_items, gr_vector_const_void_star &input_items, gr_vector_void_star &output_items)
    {
        (void)input_items;
        gr_complex *out = (gr_complex *)output_items[0];
        while ((int)m_backlog.size() < noutput_items) {   // the reference waits on its condition variable (:141-147)
            const long n = sfe_rx_feed_pump(m_feed, m_tmp.data(), (long)m_tmp.size(), 100);
            if (n < 0) throw std::runtime_error(std::string("gpu_source_c: ") + sfe_last_error());
            m_backlog.insert(m_backlog.end(), m_tmp.begin(), m_tmp.begin() + n);
        }
        memcpy(out, m_backlog.data(), sizeof(gr_complex) * (size_t)noutput_items);
        m_backlog.erase(m_backlog.begin(), m_backlog.begin() + noutput_items);
        // Tell runtime system how many output items we produced.
        return noutput_items;
    }
};

gpu_source_c::sptr gpu_source_c::make(const sfe_chain_desc *desc, int ring_bytes, int block_samples)
{
    return gpu_source_c::sptr(new gpu_source_c_impl(desc, ring_bytes, block_samples));
}

gpu_source_c::sptr gpu_source_c::make(const std::vector<float> &blkconv_taps, int fft_len, const std::vector<float> &decim_taps,
                                      float decim_rate, const std::vector<float> &resamp_taps, int resamp_up, float resamp_rate,
                                      int block_samples)
{
    sfe_chain_desc d;
    memset(&d, 0, sizeof d);
    d.in_fmt = SFE_IN_U8;
    d.n_streams = 1;
    if (!blkconv_taps.empty()) {
        d.bc_taps = blkconv_taps.data();
        d.bc_ntaps = (int)blkconv_taps.size();
        d.bc_fft = fft_len;
    }
    sfe_stage_desc *next = &d.a;   // the chain's stages are filled front to back
    if (!decim_taps.empty()) {
        next->klass = SFE_KLASS_DECIMATE;
        next->taps = decim_taps.data();
        next->n_taps = (int)decim_taps.size();
        next->up = 1;
        next->blk = 4096;
        next->rate = decim_rate;
        next = &d.b;
    }
    if (!resamp_taps.empty()) {
        next->klass = SFE_KLASS_RESAMPLE;
        next->taps = resamp_taps.data();
        next->n_taps = (int)resamp_taps.size();
        next->up = resamp_up;
        next->blk = 4096;
        next->rate = resamp_rate;
    }
    return make(&d, 16 * 2 * block_samples, block_samples);
}

int gpu_source_c::rx_callback(unsigned char *buffer, int length, void *data) { return sfe_rx_feed_callback(buffer, length, data); }

}  // namespace simplefe
}  // namespace gr
