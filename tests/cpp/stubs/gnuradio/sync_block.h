// Minimal stand-in for <gnuradio/sync_block.h> (GNU Radio 3.7 is absent from this image): just enough of
// gr::sync_block for an out-of-tree block to compile and for a test to call work().  Test infrastructure only.
#pragma once
#include <complex>
#include <string>
#include <vector>
#include <gnuradio/io_signature.h>

typedef std::complex<float> gr_complex;
typedef std::vector<const void *> gr_vector_const_void_star;
typedef std::vector<void *> gr_vector_void_star;

namespace gr {
class sync_block
{
    std::string d_name;
    io_signature::sptr d_in, d_out;
    long d_consumed = 0;

public:
    sync_block() {}
    sync_block(const std::string &name, io_signature::sptr in, io_signature::sptr out) : d_name(name), d_in(in), d_out(out) {}
    virtual ~sync_block() {}
    virtual int work(int noutput_items, gr_vector_const_void_star &input_items, gr_vector_void_star &output_items) = 0;
    void consume_each(int n) { d_consumed += n; }   // (a sync_block's scheduler bookkeeping; the check reads it back)
    long consumed() const { return d_consumed; }
    const std::string &name() const { return d_name; }
    io_signature::sptr input_signature() const { return d_in; }
    io_signature::sptr output_signature() const { return d_out; }
};
}  // namespace gr
