// box.cu -- sfe_box_*: the batched chain over ALL GPUs of one box, inside the library (include/sfe_gpu.h section 2b).
//
// The path shards by stream and nothing else: every piece of state the reference keeps is per object
// (libdsp/blkconv.h:51-61, decimate.h:48-59, resample.h:49-59), so device g of G owns the contiguous stream range
// [g*S/G, (g+1)*S/G) with its own sfe_chain, its own pinned-host/device double buffering
// (sfe_chain_process_host: two CUDA streams per device) and its own host worker thread.  There is no data-path
// collective and no inter-GPU traffic; NCCL is not linked.  A C++ caller of include/libdsp reaches the whole box
// with one call instead of one process per GPU.
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "kernels.h"

using namespace sfe;

namespace {

struct Worker {
    int dev = 0, s0 = 0, ns = 0;          // device ordinal, first stream, stream count
    sfe_chain *ch = nullptr;
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<int(Worker &)> job;
    bool has_job = false, done = false, quit = false;
    int rc = SFE_OK;
    std::string err;
    long n_out = 0;

    void loop()
    {
        sfe_set_device(dev);               // thread-local: objects created on this thread live on `dev`
        std::unique_lock<std::mutex> lk(m);
        for (;;) {
            cv.wait(lk, [&] { return has_job || quit; });
            if (quit) break;
            std::function<int(Worker &)> j = std::move(job);
            has_job = false;
            lk.unlock();
            int r;
            try {
                r = j(*this);
            } catch (const std::exception &e) {
                set_error(std::string("host-side failure: ") + e.what());
                r = SFE_ERR_HOST;
            }
            lk.lock();
            rc = r;
            err = r < 0 ? sfe_last_error() : "";
            done = true;
            cv.notify_all();
        }
    }
    void post(std::function<int(Worker &)> j)
    {
        std::lock_guard<std::mutex> lk(m);
        job = std::move(j);
        has_job = true;
        done = false;
        cv.notify_all();
    }
    int wait()
    {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return done; });
        return rc;
    }
};

}  // namespace

struct sfe_box {
    int S = 0, in_fmt = 0;
    std::vector<Worker *> w;
};

namespace {

// run `f` on every worker at once, wait for all; first failure wins
int run_all(sfe_box *b, const std::function<int(Worker &)> &f)
{
    for (Worker *w : b->w) w->post(f);
    int rc = SFE_OK;
    for (Worker *w : b->w) {
        const int r = w->wait();
        if (r < 0 && rc == SFE_OK) {
            rc = r;
            set_error("device " + std::to_string(w->dev) + ": " + w->err);
        }
    }
    return rc;
}

}  // namespace

extern "C" sfe_box *sfe_box_create(const sfe_chain_desc *desc, int n_devices)
{
    SFE_GUARDED(nullptr, {
        if (!desc || desc->n_streams < 1) {
            set_error("box: bad descriptor");
            return nullptr;
        }
        const int have = sfe_device_count();
        if (have <= 0) {
            set_error("box: no CUDA device available (libsfegpu has no CPU fallback)");
            return nullptr;
        }
        int G = n_devices > 0 ? n_devices : have;
        if (G > have) {
            set_error("box: " + std::to_string(n_devices) + " devices requested, " + std::to_string(have) + " visible");
            return nullptr;
        }
        if (G > desc->n_streams) G = desc->n_streams;    // fewer streams than GPUs: the rest stay idle ("replicas only" does not apply)
        sfe_box *b = new sfe_box;
        b->S = desc->n_streams;
        b->in_fmt = desc->in_fmt;
        const int base = b->S / G, extra = b->S % G;     // contiguous, balanced: the first S % G devices carry one more
        for (int g = 0; g < G; g++) {
            Worker *w = new Worker;
            w->dev = g;
            w->s0 = g * base + (g < extra ? g : extra);
            w->ns = base + (g < extra ? 1 : 0);
            w->th = std::thread([w] { w->loop(); });
            b->w.push_back(w);
        }
        const sfe_chain_desc d0 = *desc;
        const int rc = run_all(b, [d0](Worker &w) {
            sfe_chain_desc d = d0;
            d.n_streams = w.ns;
            w.ch = sfe_chain_create(&d);
            return w.ch ? SFE_OK : SFE_ERR_CUDA;
        });
        if (rc != SFE_OK) {
            const std::string keep = sfe_last_error();
            sfe_box_destroy(b);
            set_error(keep);
            return nullptr;
        }
        return b;
    })
}

extern "C" void sfe_box_destroy(sfe_box *b)
{
    if (!b) return;
    for (Worker *w : b->w) {
        w->post([](Worker &x) {
            if (x.ch) sfe_chain_destroy(x.ch);
            x.ch = nullptr;
            return SFE_OK;
        });
        w->wait();
        {
            std::lock_guard<std::mutex> lk(w->m);
            w->quit = true;
            w->cv.notify_all();
        }
        if (w->th.joinable()) w->th.join();
        delete w;
    }
    delete b;
}

extern "C" int sfe_box_n_devices(sfe_box *b) { return b ? (int)b->w.size() : SFE_ERR_ARG; }

extern "C" int sfe_box_stream_range(sfe_box *b, int index, int *first, int *count)
{
    if (!b || index < 0 || index >= (int)b->w.size()) return SFE_ERR_ARG;
    if (first) *first = b->w[index]->s0;
    if (count) *count = b->w[index]->ns;
    return b->w[index]->dev;
}

extern "C" const char *sfe_box_plan(sfe_box *b) { return b && !b->w.empty() ? sfe_chain_plan(b->w[0]->ch) : ""; }

extern "C" int sfe_box_reset(sfe_box *b)
{
    if (!b) return SFE_ERR_ARG;
    SFE_GUARDED(SFE_ERR_HOST, return run_all(b, [](Worker &w) { return sfe_chain_reset(w.ch); });)
}

extern "C" long sfe_box_max_out(sfe_box *b, long n_in)
{
    return b && !b->w.empty() ? sfe_chain_max_out(b->w[0]->ch, n_in) : SFE_ERR_ARG;
}

extern "C" int sfe_box_process_host(sfe_box *b, const void *h_in, long n_in, long in_pitch, void *h_out, long out_pitch, long *n_out)
{
    if (!b || !h_in || !h_out || n_in < 0 || in_pitch < n_in) {
        set_error("box_process_host: bad arguments");
        return SFE_ERR_ARG;
    }
    SFE_GUARDED(SFE_ERR_HOST, {
        const int eb = fmt_bytes(b->in_fmt);
        const int rc = run_all(b, [=](Worker &w) {
            const char *src = static_cast<const char *>(h_in) + (size_t)w.s0 * in_pitch * eb;
            char *dst = static_cast<char *>(h_out) + sizeof(float2) * (size_t)w.s0 * out_pitch;
            return sfe_chain_process_host(w.ch, src, n_in, in_pitch, dst, out_pitch, &w.n_out);
        });
        if (n_out) *n_out = b->w[0]->n_out;   // identical on every device: the output schedule is data independent
        return rc;
    })
}

extern "C" int sfe_box_process_dev(sfe_box *b, const void *const *d_in, long n_in, long in_pitch, void *const *d_out, long out_pitch,
                                   long *n_out, void *const *streams)
{
    if (!b || !d_in || !d_out || n_in < 0 || in_pitch < n_in) {
        set_error("box_process_dev: bad arguments");
        return SFE_ERR_ARG;
    }
    SFE_GUARDED(SFE_ERR_HOST, {
        const int rc = run_all(b, [=](Worker &w) {
            const int g = w.dev;
            return sfe_chain_process_dev(w.ch, d_in[g], n_in, in_pitch, d_out[g], out_pitch, &w.n_out, streams ? streams[g] : nullptr);
        });
        if (n_out) *n_out = b->w[0]->n_out;
        return rc;
    })
}

extern "C" int sfe_box_sync(sfe_box *b)
{
    if (!b) return SFE_ERR_ARG;
    SFE_GUARDED(SFE_ERR_HOST, return run_all(b, [](Worker &) { return SFE_CUDA(cudaDeviceSynchronize()) ? SFE_OK : SFE_ERR_CUDA; });)
}
