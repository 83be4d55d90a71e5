// gr_sink_check.cxx -- drives gr::simplefe::gpu_sink_c / gpu_sink_f / gpu_source_f the way the GNU Radio scheduler and
// libsimpleFE would (stand-in headers under tests/cpp/stubs; test infrastructure only).
//   gr_sink_check sink_c|sink_f in.f32 out.bytes pkt_bytes        work() in ragged item counts on the main thread, a
//                                                                  "USB" thread calling tx_callback for pkt_bytes packets
//   gr_sink_check source_f chain.bin in.u8 out.f32 n_items         one channel's bytes through rx_callback, work() -> floats
// tests/test_gr_block.py compares the files with the reference's own packing / classes.
#include <simplefe/gpu_sink.h>
#include <simplefe/gpu_source_f.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static std::vector<unsigned char> slurp(const char *path)
{
    std::vector<unsigned char> v;
    FILE *f = fopen(path, "rb");
    if (!f) return v;
    unsigned char buf[65536];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) v.insert(v.end(), buf, buf + n);
    fclose(f);
    return v;
}

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

template <class Sink> static int run_sink(Sink snk, const std::vector<unsigned char> &raw, int floats_per_item, const char *out_path, int pkt)
{
    const float *x = reinterpret_cast<const float *>(raw.data());
    const long n_items = (long)(raw.size() / 4) / floats_per_item;
    const long total_bytes = (n_items * floats_per_item) / 4 * 5;          // whole quads reach the wire
    if (snk->input_signature()->sizeof_stream_item(0) != (int)(4 * floats_per_item) || snk->output_signature()->max_streams() != 0) return 3;
    std::vector<unsigned char> wire;
    std::atomic<bool> done(false);
    std::thread usb([&] {   // libsimpleFE's event thread asking for TX packets (simpleFE.c tx path)
        std::vector<unsigned char> buf((size_t)pkt);
        while ((long)wire.size() < total_bytes / pkt * pkt) {
            // a real device cannot wait and would underflow ("U", sink_c_impl.cc:95-98); the test wants every byte
            while (sfe_tx_feed_stat((sfe_tx_feed *)snk->callback_data(), 2) < pkt && !done.load()) std::this_thread::yield();
            if (sfe_tx_feed_stat((sfe_tx_feed *)snk->callback_data(), 2) < pkt) break;
            Sink::element_type::tx_callback(buf.data(), pkt, snk->callback_data());
            wire.insert(wire.end(), buf.begin(), buf.begin() + pkt / 5 * 5);
        }
    });
    gr_vector_const_void_star in_items(1);
    gr_vector_void_star out_items;
    long off = 0;
    const int sizes[] = {1, 257, 3, 1000, 64, 2, 4095};
    for (int k = 0; off < n_items; k++) {
        long n = sizes[k % 7];
        if (n > n_items - off) n = n_items - off;
        in_items[0] = x + off * floats_per_item;
        if (snk->work((int)n, in_items, out_items) != 0) return 4;
        off += n;
    }
    if (snk->consumed() != n_items) return 5;
    done.store(true);
    usb.join();
    // an underflow on purpose: the ring now holds less than a packet
    std::vector<unsigned char> buf((size_t)pkt, 0xAA);
    const long before = snk->underflows();
    Sink::element_type::tx_callback(buf.data(), pkt, snk->callback_data());
    bool zeroed = true;
    for (unsigned char b : buf) zeroed = zeroed && b == 0;
    if (snk->underflows() != before + 1 || !zeroed) return 6;
    FILE *fo = fopen(out_path, "wb");
    fwrite(wire.data(), 1, wire.size(), fo);
    fclose(fo);
    printf("items %ld, wire bytes %zu, underflows %ld\n", n_items, wire.size(), snk->underflows());
    return 0;
}

int main(int argc, char **argv)
{
    if (argc < 6) return 2;
    const std::string mode = argv[1];
    if (mode == "sink_c" || mode == "sink_f") {
        const std::vector<unsigned char> raw = slurp(argv[2]);
        const int pkt = atoi(argv[4]);
        if (mode == "sink_c") {
            gr::simplefe::gpu_sink_c::sptr s = gr::simplefe::gpu_sink_c::make(16384, 1024);
            if (s->name() != "gpu_sink_c") return 3;
            return run_sink(s, raw, 2, argv[3], pkt);
        }
        gr::simplefe::gpu_sink_f::sptr s = gr::simplefe::gpu_sink_f::make(16384, 2048, 1);
        if (s->name() != "gpu_sink_f" || s->channel() != 1) return 3;
        return run_sink(s, raw, 1, argv[3], pkt);
    }
    if (mode != "source_f") return 2;
    // chain file: as gr_block_check.cxx
    sfe_chain_desc d;
    memset(&d, 0, sizeof d);
    std::vector<float> bc, ta, tb;
    {
        FILE *f = fopen(argv[2], "rb");
        if (!f) return 2;
        int32_t h[2];
        bool ok = read_exact(f, h, sizeof h);
        if (ok) {
            bc.resize((size_t)h[0]);
            ok = h[0] == 0 || read_exact(f, bc.data(), 4 * (size_t)h[0]);
            d.bc_taps = h[0] ? bc.data() : NULL;
            d.bc_ntaps = h[0];
            d.bc_fft = h[1];
        }
        ok = ok && read_stage(f, &d.a, &ta) && read_stage(f, &d.b, &tb);
        fclose(f);
        if (!ok) return 2;
        d.in_fmt = SFE_IN_U8;
        d.n_streams = 1;
    }
    const std::vector<unsigned char> raw = slurp(argv[3]);
    const int n_items = atoi(argv[5]);
    const int block = 4096, ring_bytes = 16 * 2 * block;
    gr::simplefe::gpu_source_f::sptr src = gr::simplefe::gpu_source_f::make(&d, ring_bytes, block, 0);
    if (src->name() != "gpu_source_f" || src->output_signature()->sizeof_stream_item(0) != (int)sizeof(float)) return 3;
    std::thread usb([&] {
        std::vector<unsigned char> pkt(512);
        for (size_t pos = 0; pos < raw.size(); pos += 511) {            // odd packet lengths are fine on one channel
            const int len = (int)(raw.size() - pos < 511 ? raw.size() - pos : 511);
            memcpy(pkt.data(), raw.data() + pos, (size_t)len);
            for (;;) {                                                   // the test wants every byte in: retry instead of "O"
                const long before = src->overflows();
                gr::simplefe::gpu_source_f::rx_callback(pkt.data(), len, src->callback_data());
                if (src->overflows() == before) break;
                std::this_thread::yield();
            }
        }
    });
    FILE *fo = fopen(argv[4], "wb");
    std::vector<float> out(257);
    gr_vector_const_void_star in_items;
    gr_vector_void_star out_items(1);
    int produced = 0;
    while (produced < n_items) {
        const int want = n_items - produced < (int)out.size() ? n_items - produced : (int)out.size();
        out_items[0] = out.data();
        if (src->work(want, in_items, out_items) != want) return 4;
        fwrite(out.data(), sizeof(float), (size_t)want, fo);
        produced += want;
    }
    fclose(fo);
    usb.join();
    printf("produced %d items\n", produced);
    return 0;
}
