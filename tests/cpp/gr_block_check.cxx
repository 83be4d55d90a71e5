// gr_block_check.cxx -- drives gr::simplefe::gpu_source_c the way the GNU Radio scheduler and libsimpleFE would:
// a producer thread delivers ISO-packet sized byte buffers to the block's rx_callback (simpleFE.h:42), the main
// thread calls work() for fixed-size requests, and the produced samples are written to a file that
// tests/test_gr_block.py compares with the reference classes fed the same bytes.
// usage: gr_block_check chain.bin in.u8 out.c64 n_items
#include <simplefe/gpu_source_c.h>

#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

// The chain description comes from a small binary file the Python test writes, so both sides share the taps exactly:
//   int32 bc_ntaps, int32 bc_fft, bc_ntaps floats, then for stage a and stage b:
//   int32 klass, int32 n_taps, int32 up, int32 blk, float rate, n_taps floats          (klass 0 = stage absent)
struct ChainFile {
    std::vector<float> bc, ta, tb;
    sfe_chain_desc d;
};

static bool read_exact(FILE *f, void *p, size_t n) { return fread(p, 1, n, f) == n; }

static bool read_stage(FILE *f, sfe_stage_desc *s, std::vector<float> *t)
{
    int32_t h[4];
    float rate;
    if (!read_exact(f, h, sizeof h) || !read_exact(f, &rate, 4)) return false;
    t->resize((size_t)h[1]);
    if (h[1] > 0 && !read_exact(f, t->data(), 4 * (size_t)h[1])) return false;
    s->klass = h[0];
    s->taps = t->data();
    s->n_taps = h[1];
    s->up = h[2];
    s->blk = h[3];
    s->rate = rate;
    return true;
}

static bool read_chain(const char *path, ChainFile *c)
{
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    memset(&c->d, 0, sizeof c->d);
    int32_t h[2];
    bool ok = read_exact(f, h, sizeof h);
    if (ok) {
        c->bc.resize((size_t)h[0]);
        ok = h[0] == 0 || read_exact(f, c->bc.data(), 4 * (size_t)h[0]);
        c->d.bc_taps = h[0] ? c->bc.data() : NULL;
        c->d.bc_ntaps = h[0];
        c->d.bc_fft = h[1];
    }
    ok = ok && read_stage(f, &c->d.a, &c->ta) && read_stage(f, &c->d.b, &c->tb);
    fclose(f);
    c->d.in_fmt = SFE_IN_U8;
    c->d.n_streams = 1;
    return ok;
}

int main(int argc, char **argv)
{
    if (argc < 5) {
        fprintf(stderr, "usage: gr_block_check chain.bin in.u8 out.c64 n_items\n");
        return 2;
    }
    ChainFile cf;
    if (!read_chain(argv[1], &cf)) return 2;
    std::vector<unsigned char> raw;
    {
        FILE *f = fopen(argv[2], "rb");
        if (!f) return 2;
        unsigned char buf[65536];
        size_t n;
        while ((n = fread(buf, 1, sizeof buf, f)) > 0) raw.insert(raw.end(), buf, buf + n);
        fclose(f);
    }
    const int n_items = atoi(argv[4]);
    const int block = 4096, ring_bytes = 16 * 2 * block;
    // a fifth argument "vec" goes through the vector factory (the one a GRC flow graph calls) when the chain has that shape
    const bool vec = argc > 5 && !strcmp(argv[5], "vec");
    gr::simplefe::gpu_source_c::sptr src;
    if (vec) {
        const bool a_dec = cf.d.a.klass == SFE_KLASS_DECIMATE;
        const sfe_stage_desc &rs = a_dec ? cf.d.b : cf.d.a;
        src = gr::simplefe::gpu_source_c::make(cf.bc, cf.d.bc_fft, a_dec ? cf.ta : std::vector<float>(), cf.d.a.rate,
                                               a_dec ? cf.tb : cf.ta, rs.up, rs.rate, block);
    } else {
        src = gr::simplefe::gpu_source_c::make(&cf.d, ring_bytes, block);
    }
    if (src->name() != "gpu_source_c" || src->output_signature()->sizeof_stream_item(0) != (int)sizeof(gr_complex) ||
        src->input_signature()->max_streams() != 0)
        return 3;

    std::thread usb([&] {   // the libusb event thread of libsimpleFE (simpleFE.c:416-418)
        const size_t pkt = 1024;
        for (size_t pos = 0; pos < raw.size(); pos += pkt) {
            const int len = (int)(raw.size() - pos < pkt ? raw.size() - pos : pkt);
            // a real device cannot wait and would overflow ("O", source_c_impl.cc:98-100); the test wants every byte in
            while (sfe_rx_feed_stat((sfe_rx_feed *)src->callback_data(), 3) + len > ring_bytes) std::this_thread::yield();
            gr::simplefe::gpu_source_c::rx_callback(raw.data() + pos, len, src->callback_data());
        }
    });
    FILE *fo = fopen(argv[3], "wb");
    std::vector<gr_complex> out(257);
    gr_vector_const_void_star in_items;
    gr_vector_void_star out_items(1);
    int produced = 0;
    while (produced < n_items) {
        const int want = n_items - produced < (int)out.size() ? n_items - produced : (int)out.size();
        out_items[0] = out.data();
        const int got = src->work(want, in_items, out_items);
        if (got != want) return 4;
        fwrite(out.data(), sizeof(gr_complex), (size_t)got, fo);
        produced += got;
    }
    fclose(fo);
    usb.join();
    printf("produced %d items, overflows %ld, discarded %ld\n", produced, src->overflows(), src->discarded());
    return 0;
}
