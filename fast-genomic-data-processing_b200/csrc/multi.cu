// Multi-device handles (pairhmm_cuda_config::device_mask): one process, one handle, N GPUs.
//
// The reference is ONE process with T worker threads (M2/main.cpp:199,570-574) whose PairHMM work is a list of independent
// testcases (IntelPairHmm.cc:332-351 takes one testcase; the only sharing it does is the atomic-index hand-out of
// computeLikelihoodsNative_concurrent, IntelPairHmm.cc:296-330). The multi-GPU form of that here:
//   * a batch call's regions are split over the devices by DP cells with the LPT rule -- whole regions first (a region's
//     reads and haplotypes cross PCIe once), a region's reads are cut into parts only when the region alone exceeds
//     1 / (4 x devices) of the batch (BASELINE configs[2]: one 2000 x 32 region);
//   * every device has its own single-device child handle (two slots, pinned arena, streams: engine.cu) and ONE host worker
//     thread that drives it; the packing loops of all children draw on the same process-wide pool (hostpool.h);
//   * every part's results are scattered straight into the caller's one result array (read-major rows of a region are
//     contiguous, so a part is a row range): that is the gather. No collective, no peer copy, no NCCL -- by design.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <numeric>
#include <thread>

#include "engine.h"

namespace phmm {

// ---- the plan ---------------------------------------------------------------------------------------------------
std::vector<DevicePart> plan_devices(const uint64_t *read_len_sum, const uint64_t *hap_len_sum, const int32_t *n_reads,
                                     int64_t n_regions, int n_devices) {
    std::vector<DevicePart> parts;
    if (n_regions <= 0 || n_devices <= 0) return parts;
    parts.reserve((size_t)n_regions + 8u * (size_t)n_devices);
    long double total = 0;
    for (int64_t r = 0; r < n_regions; ++r) total += (long double)read_len_sum[r] * (long double)hap_len_sum[r];
    // a region heavier than this is cut into parts of whole reads (equal read counts: reads of a region are of similar length)
    const long double limit = n_devices > 1 ? total / (4.0L * n_devices) : total + 1;
    std::vector<long double> cost;
    cost.reserve(parts.capacity());
    for (int64_t r = 0; r < n_regions; ++r) {
        const long double c = (long double)read_len_sum[r] * (long double)hap_len_sum[r];
        const int32_t nr = std::max(0, n_reads[r]);
        int32_t k = 1;
        if (n_devices > 1 && c > limit && nr > 1) k = (int32_t)std::min<long double>((long double)nr, std::ceil(c / limit));
        for (int32_t p = 0; p < k; ++p) {
            const int32_t a = (int32_t)((int64_t)nr * p / k), b = (int32_t)((int64_t)nr * (p + 1) / k);
            parts.push_back(DevicePart{r, a, b - a, 0});
            cost.push_back(nr > 0 ? c * (long double)(b - a) / (long double)nr : 0);
        }
    }
    if (n_devices == 1) return parts;
    // longest processing time first onto the least loaded device; ties keep the input order (deterministic)
    std::vector<size_t> order(parts.size());
    std::iota(order.begin(), order.end(), (size_t)0);
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return cost[a] > cost[b]; });
    std::vector<long double> load((size_t)n_devices, 0);
    for (size_t i : order) {
        int best = 0;
        for (int d = 1; d < n_devices; ++d) if (load[(size_t)d] < load[(size_t)best]) best = d;
        parts[i].device = best;
        load[(size_t)best] += cost[i];
    }
    return parts;
}

// ---- one host worker thread per device ----------------------------------------------------------------------------
struct DeviceWorker {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<int()> job;
    bool has_job = false, done = true, stop = false;
    int status = PAIRHMM_OK;

    void start() { th = std::thread([this] { loop(); }); }
    void loop() {
        std::unique_lock<std::mutex> lk(m);
        for (;;) {
            cv.wait(lk, [&] { return has_job || stop; });
            if (stop) return;
            std::function<int()> f = std::move(job);
            has_job = false;
            lk.unlock();
            const int st = f();
            lk.lock();
            status = st; done = true;
            cv.notify_all();
        }
    }
    void post(std::function<int()> f) {
        std::lock_guard<std::mutex> lk(m);
        job = std::move(f); has_job = true; done = false;
        cv.notify_all();
    }
    int wait() {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return done; });
        return status;
    }
    void shutdown() {
        { std::lock_guard<std::mutex> lk(m); stop = true; }
        cv.notify_all();
        if (th.joinable()) th.join();
    }
};

int multi_create(pairhmm_cuda_handle *h) {
    std::vector<int> devs;
    for (int d = 0; d < 32; ++d) if (h->cfg.device_mask & (1u << d)) devs.push_back(d);
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return PAIRHMM_ERR_NO_DEVICE; }
    for (int d : devs) if (d >= n) return fail(h, PAIRHMM_ERR_ARG, "device_mask names a CUDA ordinal this process does not see");
    // the children share the box's host threads: an equal share of the pool each unless the caller fixed pack_threads
    unsigned hc = std::thread::hardware_concurrency();
    const int share = (int)std::max(2u, (hc ? hc : 16u) / (unsigned)devs.size());
    for (int d : devs) {
        pairhmm_cuda_config c = h->cfg;
        c.device_mask = 0; c.device = d; c.batch_window_us = 0;
        if (c.pack_threads <= 0) c.pack_threads = std::min(share, 16);
        pairhmm_cuda_handle *child = nullptr;
        const int st = pairhmm_cuda_create(&c, &child);
        if (st != PAIRHMM_OK) return st;
        h->children.push_back(child);
        h->workers.push_back(new DeviceWorker());
        h->workers.back()->start();
    }
    h->device = devs.front();
    h->sm_count = h->children.front()->sm_count;
    return PAIRHMM_OK;
}

void multi_destroy(pairhmm_cuda_handle *h) {
    for (auto *w : h->workers) { w->shutdown(); delete w; }
    h->workers.clear();
    for (auto *c : h->children) pairhmm_cuda_destroy(c);
    h->children.clear();
}

namespace {

void propagate_call_state(pairhmm_cuda_handle *h) {
    for (auto *c : h->children) { c->ep_on = h->ep_on; c->ep = h->ep; c->model = h->model; c->no_output_ok = h->no_output_ok; }
}

// per device the RegionViews of its parts (in input order) and, optionally, the offset of every view's results in the
// caller's flat result arrays (region-major, read-major: resident batches)
struct Split { std::vector<std::vector<RegionView>> views; std::vector<std::vector<size_t>> out_off; };

Split split_regions(pairhmm_cuda_handle *h, const RegionView *regions, int64_t n_regions, bool want_offsets) {
    const int nd = (int)h->children.size();
    std::vector<uint64_t> rl((size_t)n_regions, 0), hl((size_t)n_regions, 0);
    std::vector<int32_t> nr((size_t)n_regions, 0);
    std::vector<size_t> region_off;
    if (want_offsets) region_off.resize((size_t)n_regions + 1, 0);
    for (int64_t i = 0; i < n_regions; ++i) {
        const RegionView &r = regions[i];
        nr[(size_t)i] = r.n_reads;
        if (r.reads || r.raw) for (int q = 0; q < r.n_reads; ++q) rl[(size_t)i] += (uint64_t)std::max(0, r.raw ? r.raw[q].len : r.reads[q].len);
        if (r.haps) for (int q = 0; q < r.n_haps; ++q) hl[(size_t)i] += (uint64_t)std::max(0, r.haps[q].len);
        if (want_offsets) region_off[(size_t)i + 1] = region_off[(size_t)i] + (size_t)std::max(0, r.n_reads) * (size_t)std::max(0, r.n_haps);
    }
    const std::vector<DevicePart> parts = plan_devices(rl.data(), hl.data(), nr.data(), n_regions, nd);
    Split s;
    s.views.resize((size_t)nd);
    if (want_offsets) s.out_off.resize((size_t)nd);
    for (const DevicePart &p : parts) {
        const RegionView &r = regions[p.region];
        RegionView v = r;
        const size_t row0 = (size_t)p.first_read * (size_t)std::max(0, r.n_haps);
        if (r.reads) v.reads = r.reads + p.first_read;
        if (r.raw) v.raw = r.raw + p.first_read;
        v.n_reads = p.n_reads;
        if (r.out) v.out = r.out + row0;
        if (r.resc) v.resc = r.resc + row0;
        if (r.summ) v.summ = r.summ + p.first_read;
        s.views[(size_t)p.device].push_back(v);
        if (want_offsets) s.out_off[(size_t)p.device].push_back(region_off[(size_t)p.region] + row0);
    }
    return s;
}

int collect(pairhmm_cuda_handle *h) {
    int status = PAIRHMM_OK;
    for (size_t d = 0; d < h->children.size(); ++d) {
        const int st = h->workers[d]->wait();
        if (st != PAIRHMM_OK && status == PAIRHMM_OK) {
            status = st;
            fail(h, st, "device " + std::to_string(h->children[d]->device) + ": " + h->children[d]->last_error);
        }
    }
    return status;
}

}  // namespace

int multi_compute_batch(pairhmm_cuda_handle *h, const RegionView *regions, int64_t n_regions) {
    const auto t0 = std::chrono::steady_clock::now();
    // earlier single-region submits live on the children: retire them first (as the single-device batch call does)
    for (auto &kv : h->ticket_of) { const int st = pairhmm_cuda_wait(h->children[(size_t)kv.second.first], kv.second.second); if (st != PAIRHMM_OK) return st; }
    h->ticket_of.clear();
    h->resident_loaded = false;
    propagate_call_state(h);
    auto split = std::make_shared<Split>(split_regions(h, regions, n_regions, false));
    for (size_t d = 0; d < h->children.size(); ++d) {
        pairhmm_cuda_handle *c = h->children[d];
        h->workers[d]->post([c, split, d] { return compute_batch_impl(c, split->views[d].data(), (int64_t)split->views[d].size()); });
    }
    const int status = collect(h);
    h->stats.calls += 1;
    h->stats.host_ms_total += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return status;
}

int multi_submit(pairhmm_cuda_handle *h, const RegionView &rv, pairhmm_cuda_ticket *ticket) {
    const int d = h->next_child;
    h->next_child = (h->next_child + 1) % (int)h->children.size();
    pairhmm_cuda_ticket ct = 0;
    const int st = pairhmm_cuda_submit(h->children[(size_t)d], rv.reads, rv.n_reads, rv.haps, rv.n_haps, rv.out, rv.resc, &ct);
    if (st != PAIRHMM_OK) return fail(h, st, h->children[(size_t)d]->last_error);
    *ticket = h->next_ticket++;
    h->ticket_of[*ticket] = {d, ct};
    h->stats.calls += 1;
    return PAIRHMM_OK;
}

int multi_wait(pairhmm_cuda_handle *h, pairhmm_cuda_ticket ticket) {
    if (ticket <= 0 || ticket >= h->next_ticket) return fail(h, PAIRHMM_ERR_STATE, "wait: unknown ticket");
    auto it = h->ticket_of.find(ticket);
    if (it == h->ticket_of.end()) return PAIRHMM_OK;      // already retired
    const int d = it->second.first;
    const int st = pairhmm_cuda_wait(h->children[(size_t)d], it->second.second);
    h->ticket_of.erase(it);
    return st == PAIRHMM_OK ? st : fail(h, st, h->children[(size_t)d]->last_error);
}

int multi_resident_load(pairhmm_cuda_handle *h, const RegionView *views, size_t n) {
    h->resident_loaded = false;
    h->no_output_ok = true;
    propagate_call_state(h);
    h->no_output_ok = false;
    auto split = std::make_shared<Split>(split_regions(h, views, (int64_t)n, true));
    for (size_t d = 0; d < h->children.size(); ++d) {
        pairhmm_cuda_handle *c = h->children[d];
        h->workers[d]->post([c, split, d] {
            return resident_load_views(c, split->views[d].data(), split->views[d].size(), split->out_off[d].data());
        });
    }
    const int status = collect(h);
    for (auto *c : h->children) c->no_output_ok = false;
    h->resident_loaded = status == PAIRHMM_OK;
    return status;
}

int multi_resident_run(pairhmm_cuda_handle *h, float *kernel_ms, float *float_kernel_ms) {
    if (!h->resident_loaded) return fail(h, PAIRHMM_ERR_STATE, "resident_run: nothing loaded (or the batch was replaced by a later compute call)");
    auto ms = std::make_shared<std::vector<std::pair<float, float>>>(h->children.size());
    for (size_t d = 0; d < h->children.size(); ++d) {
        pairhmm_cuda_handle *c = h->children[d];
        h->workers[d]->post([c, ms, d] { return pairhmm_cuda_resident_run(c, &(*ms)[d].first, &(*ms)[d].second); });
    }
    const int status = collect(h);
    float a = 0.f, b = 0.f;
    for (auto &p : *ms) { a = std::max(a, p.first); b = std::max(b, p.second); }   // the devices run side by side
    if (kernel_ms) *kernel_ms = a;
    if (float_kernel_ms) *float_kernel_ms = b;
    return status;
}

int multi_resident_fetch(pairhmm_cuda_handle *h, double *log10_out, uint8_t *rescued_out) {
    if (!h->resident_loaded) return fail(h, PAIRHMM_ERR_STATE, "resident_fetch: nothing loaded (or the batch was replaced by a later compute call)");
    for (size_t d = 0; d < h->children.size(); ++d) {
        pairhmm_cuda_handle *c = h->children[d];
        h->workers[d]->post([c, log10_out, rescued_out] { return pairhmm_cuda_resident_fetch(c, log10_out, rescued_out); });
    }
    return collect(h);
}

void multi_stats(pairhmm_cuda_handle *h, pairhmm_cuda_stats *out) {
    pairhmm_cuda_stats s = h->stats;      // calls, host_ms_total of the multi-device handle itself
    for (auto *c : h->children) {
        const pairhmm_cuda_stats &k = c->stats;
        s.regions += k.regions; s.reads += k.reads; s.unique_reads += k.unique_reads; s.pairs += k.pairs;
        s.unique_pairs += k.unique_pairs; s.cells += k.cells; s.rescued_pairs += k.rescued_pairs; s.h2d_bytes += k.h2d_bytes;
        s.d2h_bytes += k.d2h_bytes; s.kernel_launches += k.kernel_launches; s.cells_swept += k.cells_swept;
        s.gpu_ms_float = std::max(s.gpu_ms_float, k.gpu_ms_float); s.gpu_ms_double = std::max(s.gpu_ms_double, k.gpu_ms_double);
        s.gpu_ms_total = std::max(s.gpu_ms_total, k.gpu_ms_total); s.host_ms_pack = std::max(s.host_ms_pack, k.host_ms_pack);
        s.host_ms_scatter = std::max(s.host_ms_scatter, k.host_ms_scatter);
    }
    *out = s;
}

}  // namespace phmm

extern "C" int pairhmm_cuda_device_plan(const uint64_t *read_len_sum, const uint64_t *hap_len_sum, const int32_t *n_reads,
                                        int64_t n_regions, int n_devices, int64_t max_parts, int64_t *part_region,
                                        int32_t *part_first_read, int32_t *part_n_reads, int32_t *part_device, int64_t *n_parts) {
    if (n_regions < 0 || n_devices < 1 || n_devices > 32 || !n_parts ||
        (n_regions > 0 && (!read_len_sum || !hap_len_sum || !n_reads || !part_region || !part_first_read || !part_n_reads || !part_device)))
        return PAIRHMM_ERR_ARG;
    const std::vector<phmm::DevicePart> parts = phmm::plan_devices(read_len_sum, hap_len_sum, n_reads, n_regions, n_devices);
    *n_parts = (int64_t)parts.size();
    if ((int64_t)parts.size() > max_parts) return PAIRHMM_ERR_ARG;
    for (size_t i = 0; i < parts.size(); ++i) {
        part_region[i] = parts[i].region; part_first_read[i] = parts[i].first_read; part_n_reads[i] = parts[i].n_reads;
        part_device[i] = parts[i].device;
    }
    return PAIRHMM_OK;
}
