// Process-wide batcher of single-region calls (pairhmm_cuda_config::batch_window_us > 0; SURVEY.md 8b "Threading").
//
// Mutect2Cpp calls the PairHMM once per active region from each of its T worker threads (M2/main.cpp:199,570-574), and one
// region (~200 reads x ~10 haplotypes, 1e8 DP cells) cannot fill 148 SMs: a launch group per caller leaves the device
// running many small groups side by side, each with its own launch, copy and tail costs. The reference's answer to "one
// thread's region is too small a unit" is its cross-thread work sharing -- computeLikelihoodsNative_concurrent
// (IntelPairHmm.cc:296-330) publishes a region's testcases, idle threads claim them by an atomic index
// (PairHMMConcurrentControl.h:15-53, M2/main.cpp:396-417). The GPU-side counterpart merges the other way round: the
// regions that different host threads submit while the device is busy with the previous group (or within
// batch_window_us of the first one when it is idle) are packed into ONE chunk -- one set of launches -- by a dispatcher
// thread that owns a single-device backend handle; every caller's results are scattered straight into that caller's
// array (finish_chunk does that per region anyway) and the caller is woken. The group size adapts by itself: a lone
// caller gets a group per region, sixteen busy callers get groups of ~16 regions.
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <thread>

#include "engine.h"

namespace phmm {

int submit_chunk(pairhmm_cuda_handle *h, Slot &s, const RegionView *regs, size_t n, int threads);
int finish_chunk(pairhmm_cuda_handle *h, Slot &s, int threads);

struct BatchRequest {
    RegionView rv;
    int status = PAIRHMM_OK;
    bool done = false;
    std::string error;
    std::condition_variable cv;
};

struct Batcher {
    pairhmm_cuda_config key{};                // configuration the callers share (device included)
    pairhmm_cuda_handle *backend = nullptr;   // single-device handle driven by the dispatcher thread only
    int window_us = 0;
    int refs = 0;
    std::mutex m;
    std::condition_variable cv_work;
    std::deque<BatchRequest *> queue;
    std::thread th;
    bool stop = false;
    pairhmm_cuda_stats stats{};               // snapshot of the backend's counters, taken under the lock
    uint64_t chunks = 0, regions = 0;

    void finish(std::vector<BatchRequest *> &reqs, int status) {
        std::lock_guard<std::mutex> lk(m);
        stats = backend->stats;
        for (BatchRequest *r : reqs) {
            r->status = status;
            if (status != PAIRHMM_OK) r->error = backend->last_error;
            r->done = true;
            r->cv.notify_one();
        }
        reqs.clear();
    }

    void loop() {
        cudaSetDevice(backend->device);
        std::vector<BatchRequest *> inflight[2];
        std::vector<RegionView> views;
        int next_slot = 0, oldest = 0;
        const int64_t chunk_cells = backend->cfg.chunk_cells > 0 ? backend->cfg.chunk_cells : (int64_t)100000000000ll;
        for (;;) {
            std::unique_lock<std::mutex> lk(m);
            const bool any_inflight = !inflight[0].empty() || !inflight[1].empty();
            if (!any_inflight) cv_work.wait(lk, [&] { return stop || !queue.empty(); });
            if (stop && queue.empty() && !any_inflight) return;
            if (!queue.empty() && inflight[next_slot].empty()) {
                // idle device and a lone request: give the other threads a moment to join it (they tend to arrive together)
                if (!any_inflight && queue.size() == 1 && window_us > 1)
                    cv_work.wait_for(lk, std::chrono::microseconds(window_us), [&] { return stop || queue.size() > 1; });
                std::vector<BatchRequest *> &reqs = inflight[next_slot];
                views.clear();
                int64_t cells = 0;
                while (!queue.empty()) {
                    BatchRequest *r = queue.front();
                    int64_t rl = 0, hl = 0;
                    for (int q = 0; q < r->rv.n_reads; ++q) rl += std::max(0, r->rv.reads[q].len);
                    for (int q = 0; q < r->rv.n_haps; ++q) hl += std::max(0, r->rv.haps[q].len);
                    if (!reqs.empty() && cells + rl * hl > chunk_cells) break;
                    cells += rl * hl;
                    reqs.push_back(r); views.push_back(r->rv);
                    queue.pop_front();
                }
                lk.unlock();
                Slot &s = backend->slots[next_slot];
                const int st = submit_chunk(backend, s, views.data(), views.size(), pack_threads(backend));
                chunks += 1; regions += views.size();
                if (st != PAIRHMM_OK) { set_busy(backend, s, false); finish(reqs, st); continue; }
                if (inflight[next_slot ^ 1].empty()) oldest = next_slot;
                next_slot ^= 1;
                continue;
            }
            lk.unlock();
            if (any_inflight) {      // both slots taken, or nothing new: retire the older chunk (blocks until its results are home)
                if (inflight[oldest].empty()) oldest ^= 1;
                const int st = finish_chunk(backend, backend->slots[oldest], pack_threads(backend));
                finish(inflight[oldest], st);
                oldest ^= 1;
            }
        }
    }
};

namespace {
std::mutex g_batchers_mutex;
std::vector<Batcher *> g_batchers;

pairhmm_cuda_config batcher_key(const pairhmm_cuda_handle *h) {
    pairhmm_cuda_config k = h->cfg;
    k.device = h->device; k.batch_window_us = 0; k.device_mask = 0;
    return k;
}
}  // namespace

Batcher *batcher_acquire(pairhmm_cuda_handle *h) {
    std::lock_guard<std::mutex> lk(g_batchers_mutex);
    const pairhmm_cuda_config key = batcher_key(h);
    for (Batcher *b : g_batchers)
        if (!std::memcmp(&b->key, &key, sizeof key)) { ++b->refs; return b; }
    pairhmm_cuda_handle *backend = nullptr;
    if (pairhmm_cuda_create(&key, &backend) != PAIRHMM_OK) return nullptr;   // the handle then launches for itself
    Batcher *b = new Batcher();
    b->key = key; b->backend = backend; b->window_us = h->cfg.batch_window_us; b->refs = 1;
    b->th = std::thread([b] { b->loop(); });
    g_batchers.push_back(b);
    return b;
}

void batcher_release(Batcher *b) {
    if (!b) return;
    {
        std::lock_guard<std::mutex> lk(g_batchers_mutex);
        if (--b->refs > 0) return;
        for (size_t i = 0; i < g_batchers.size(); ++i) if (g_batchers[i] == b) { g_batchers.erase(g_batchers.begin() + (long)i); break; }
    }
    { std::lock_guard<std::mutex> lk(b->m); b->stop = true; }
    b->cv_work.notify_all();
    if (b->th.joinable()) b->th.join();
    pairhmm_cuda_destroy(b->backend);
    delete b;
}

int batcher_compute(pairhmm_cuda_handle *h, const RegionView &rv) {
    // a caller's bad region must not fail the chunk it would share with other callers' regions
    const int st = check_region_args(h, rv);
    if (st != PAIRHMM_OK) return st;
    Batcher *b = h->batcher;
    BatchRequest req;
    req.rv = rv;
    {
        std::unique_lock<std::mutex> lk(b->m);
        b->queue.push_back(&req);
        b->cv_work.notify_one();
        req.cv.wait(lk, [&] { return req.done; });
        h->stats = b->stats;       // the shared backend's counters (header: statistics of a batched handle)
    }
    if (req.status != PAIRHMM_OK) return fail(h, req.status, req.error);
    return PAIRHMM_OK;
}

}  // namespace phmm
