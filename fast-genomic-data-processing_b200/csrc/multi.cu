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
#include "hostpool.h"

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
    // (twice the equal share: packing comes in bursts, one chunk at a time per device, and a burst that ends sooner starts
    // the device's next chunk sooner; the pool hands its threads to whichever jobs are open)
    const int share = (int)std::max(4u, 2u * (hc ? hc : 16u) / (unsigned)devs.size());
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
    // the length sums walk every read descriptor of the batch (5 million for BASELINE configs[4]: 240 MB): on the pool,
    // not on the calling thread alone (20 ms of a 130 ms eight-device call)
    HostPool::instance().parallel_for((size_t)n_regions, 32, 64, [&](size_t i) {
        const RegionView &r = regions[i];
        nr[i] = r.n_reads;
        uint64_t a = 0, b = 0;
        if (r.reads || r.raw) for (int q = 0; q < r.n_reads; ++q) a += (uint64_t)std::max(0, r.raw ? r.raw[q].len : r.reads[q].len);
        if (r.haps) for (int q = 0; q < r.n_haps; ++q) b += (uint64_t)std::max(0, r.haps[q].len);
        rl[i] = a; hl[i] = b;
    });
    if (want_offsets)
        for (int64_t i = 0; i < n_regions; ++i)
            region_off[(size_t)i + 1] = region_off[(size_t)i] + (size_t)std::max(0, regions[i].n_reads) * (size_t)std::max(0, regions[i].n_haps);
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
        // whole regions carry their length sums down, so that the device's chunk cutter does not walk the reads again
        if (p.n_reads == r.n_reads) { v.read_len_sum = (int64_t)rl[(size_t)p.region]; v.hap_len_sum = (int64_t)hl[(size_t)p.region]; }
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

// pairhmm_cuda_compute_flat on a multi-device handle. The generic path (views of the whole batch on the calling thread,
// then the split) walks 48 bytes per read twice before any device starts: 14 ms of a 130 ms eight-device call on BASELINE
// configs[4]. Here the split works on the flat length arrays (4 bytes per read, on the pool) and every device's worker
// builds the views of ITS regions only, side by side with the other workers.
int multi_compute_flat(pairhmm_cuda_handle *h, const FlatBatch &fb) {
    const auto t0 = std::chrono::steady_clock::now();
    if (fb.n_reads < 0 || fb.n_haps < 0 || fb.n_regions < 0) return fail(h, PAIRHMM_ERR_ARG, "negative count");
    if (fb.n_reads && (!fb.read_bases || !fb.base_q || !fb.ins_q || !fb.del_q || !fb.gcp || !fb.read_off || !fb.read_len))
        return fail(h, PAIRHMM_ERR_ARG, "null read array");
    if (fb.n_haps && (!fb.hap_bases || !fb.hap_off || !fb.hap_len)) return fail(h, PAIRHMM_ERR_ARG, "null haplotype array");
    if (fb.n_regions && (!fb.reg_read_start || !fb.reg_hap_start)) return fail(h, PAIRHMM_ERR_ARG, "null region array");
    const size_t nreg = (size_t)fb.n_regions;
    std::vector<uint64_t> rl(nreg, 0), hl(nreg, 0);
    std::vector<int32_t> nr(nreg, 0);
    std::vector<size_t> out_off(nreg + 1, 0);
    for (size_t r = 0; r < nreg; ++r) {
        const int64_t r0 = fb.reg_read_start[r], r1 = fb.reg_read_start[r + 1], h0 = fb.reg_hap_start[r], h1 = fb.reg_hap_start[r + 1];
        if (r0 < 0 || r1 < r0 || r1 > fb.n_reads || h0 < 0 || h1 < h0 || h1 > fb.n_haps || r1 - r0 > 0x7fffffff || h1 - h0 > 0x7fffffff)
            return fail(h, PAIRHMM_ERR_ARG, "region bounds out of range");
        nr[r] = (int32_t)(r1 - r0);
        out_off[r + 1] = out_off[r] + (size_t)(r1 - r0) * (size_t)(h1 - h0);
        if ((r1 - r0) * (h1 - h0) > 0 && !fb.log10_out) return fail(h, PAIRHMM_ERR_ARG, "null output");
    }
    HostPool::instance().parallel_for(nreg, 32, 256, [&](size_t r) {
        uint64_t a = 0, b = 0;
        for (int64_t q = fb.reg_read_start[r]; q < fb.reg_read_start[r + 1]; ++q) a += (uint64_t)std::max(0, fb.read_len[q]);
        for (int64_t q = fb.reg_hap_start[r]; q < fb.reg_hap_start[r + 1]; ++q) b += (uint64_t)std::max(0, fb.hap_len[q]);
        rl[r] = a; hl[r] = b;
    });
    for (auto &kv : h->ticket_of) { const int st = pairhmm_cuda_wait(h->children[(size_t)kv.second.first], kv.second.second); if (st != PAIRHMM_OK) return st; }
    h->ticket_of.clear();
    h->resident_loaded = false;
    propagate_call_state(h);
    const int nd = (int)h->children.size();
    auto parts = std::make_shared<std::vector<std::vector<DevicePart>>>((size_t)nd);
    for (const DevicePart &p : plan_devices(rl.data(), hl.data(), nr.data(), fb.n_regions, nd)) (*parts)[(size_t)p.device].push_back(p);
    auto sums = std::make_shared<std::pair<std::vector<uint64_t>, std::vector<uint64_t>>>(std::move(rl), std::move(hl));
    auto offs = std::make_shared<std::vector<size_t>>(std::move(out_off));
    for (size_t d = 0; d < (size_t)nd; ++d) {
        pairhmm_cuda_handle *c = h->children[d];
        h->workers[d]->post([c, parts, sums, offs, fb, d] {
            const std::vector<DevicePart> &mine = (*parts)[d];
            // per worker thread, kept between calls (page-faulting 48 B per read in on every call costs more than filling it)
            static thread_local std::vector<phmm_read> reads;
            static thread_local std::vector<phmm_hap> haps;
            static thread_local std::vector<RegionView> views;
            static thread_local std::vector<size_t> r_at, h_at;
            r_at.assign(mine.size() + 1, 0); h_at.assign(mine.size() + 1, 0);
            for (size_t i = 0; i < mine.size(); ++i) {
                r_at[i + 1] = r_at[i] + (size_t)mine[i].n_reads;
                h_at[i + 1] = h_at[i] + (size_t)(fb.reg_hap_start[mine[i].region + 1] - fb.reg_hap_start[mine[i].region]);
            }
            reads.resize(r_at.back()); haps.resize(h_at.back()); views.resize(mine.size());
            phmm_read *rp = reads.data(); phmm_hap *hp = haps.data(); RegionView *vp = views.data();   // pool threads have their own thread_locals
            const size_t *ra = r_at.data(), *ha = h_at.data();
            auto build = [&, rp, hp, vp, ra, ha](size_t i) {
                const DevicePart &p = mine[i];
                const int64_t r0 = fb.reg_read_start[p.region] + p.first_read, h0 = fb.reg_hap_start[p.region];
                const int32_t nh = (int32_t)(fb.reg_hap_start[p.region + 1] - h0);
                for (int32_t q = 0; q < p.n_reads; ++q) {
                    const int64_t o = fb.read_off[r0 + q];
                    rp[ra[i] + (size_t)q] = phmm_read{fb.read_bases + o, fb.base_q + o, fb.ins_q + o, fb.del_q + o, fb.gcp + o, fb.read_len[r0 + q]};
                }
                for (int32_t q = 0; q < nh; ++q) hp[ha[i] + (size_t)q] = phmm_hap{fb.hap_bases + fb.hap_off[h0 + q], fb.hap_len[h0 + q]};
                const size_t o = (*offs)[(size_t)p.region] + (size_t)p.first_read * (size_t)nh;
                RegionView v{rp + ra[i], nullptr, p.n_reads, hp + ha[i], nh, fb.log10_out ? fb.log10_out + o : nullptr,
                             fb.rescued_out ? fb.rescued_out + o : nullptr, nullptr};
                if ((int64_t)p.n_reads == fb.reg_read_start[p.region + 1] - fb.reg_read_start[p.region]) {
                    v.read_len_sum = (int64_t)sums->first[(size_t)p.region]; v.hap_len_sum = (int64_t)sums->second[(size_t)p.region];
                }
                vp[i] = v;
            };
            // the views are built block by block as the chunk cutter reaches them: the first chunk needs ~50 regions, the rest
            // are filled in under the kernels of the chunks already submitted (building all 3000 first held the GPU back 4 ms)
            size_t built = 0, block = 64;
            const int threads = pack_threads(c);
            const std::function<void(int64_t)> ensure = [&](int64_t i) {
                while ((size_t)i >= built) {
                    const size_t b0 = built, b1 = std::min(mine.size(), built + block);
                    HostPool::instance().parallel_for(b1 - b0, threads, 16, [&, b0](size_t k) { build(b0 + k); });
                    built = b1; block = std::min<size_t>(block * 2, 2048);
                }
            };
            return compute_batch_impl(c, vp, (int64_t)mine.size(), &ensure);
        });
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
