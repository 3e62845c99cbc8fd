// The engine driven from C++ exactly as Mutect2Cpp would drive it, with nothing but the C ABI of include/pairhmm_cuda.h:
// T host threads (M2/main.cpp:199 keeps one Mutect2Engine, hence one PairHMM, per worker thread), one handle each, one
// blocking pairhmm_cuda_compute per active region (VectorLoglessPairHMM::computeLog10Likelihoods, PHLCE.cpp:88-91).
// Synthetic regions shaped like BASELINE configs[1] (~200 reads x 4-16 haplotypes, 150-base reads, 250-500-base
// haplotypes); the per-region results are compared with one pairhmm_cuda_compute_batch call over the same regions.
//
//   bash examples/build.sh && examples/region_calls [regions=1200] [--submit] [--batch-window US] [--devices MASK]
// --batch-window US: the handles are created with config.batch_window_us = US, i.e. the blocking calls of all threads go
//   through the library's process-wide batcher (regions of different threads share launch groups);
// --devices MASK: multi-device handles (config.device_mask) -- the batch call is split over the devices, per-region calls
//   go to them in turn.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

#include "pairhmm_cuda.h"

struct Region {
    std::vector<uint8_t> bytes;                  // all byte arrays of the region, back to back
    std::vector<phmm_read> reads;
    std::vector<phmm_hap> haps;
    std::vector<double> out, out_batch;
    uint64_t cells = 0;
};

static void make_region(Region &g, uint64_t seed) {
    std::mt19937_64 rng(seed);
    auto uni = [&](int lo, int hi) { return lo + (int)(rng() % (uint64_t)(hi - lo + 1)); };
    const char acgt[4] = {'A', 'C', 'G', 'T'};
    const int n_reads = std::clamp(200 + uni(-40, 40), 50, 400), n_haps = uni(4, 16), H0 = uni(250, 500), R = 150;
    std::vector<std::vector<uint8_t>> haps(n_haps);
    haps[0].resize(H0);
    for (auto &b : haps[0]) b = (uint8_t)acgt[rng() & 3];
    for (int h = 1; h < n_haps; ++h) {                       // variants: a few SNVs and one short indel
        haps[h] = haps[0];
        for (int k = uni(1, 3); k > 0; --k) haps[h][rng() % haps[h].size()] = (uint8_t)acgt[rng() & 3];
        const size_t p = 1 + rng() % (haps[h].size() - 8);
        if (rng() & 1) haps[h].insert(haps[h].begin() + p, (size_t)uni(1, 5), (uint8_t)acgt[rng() & 3]);
        else haps[h].erase(haps[h].begin() + p, haps[h].begin() + p + uni(1, 5));
    }
    size_t total = 0;
    for (auto &h : haps) total += h.size();
    g.bytes.resize(total + (size_t)n_reads * R * 5);
    uint8_t *w = g.bytes.data();
    uint64_t hap_cols = 0;
    for (auto &h : haps) { std::memcpy(w, h.data(), h.size()); g.haps.push_back(phmm_hap{w, (int32_t)h.size()}); w += h.size(); hap_cols += h.size(); }
    std::normal_distribution<double> qd(32.0, 6.0);
    for (int r = 0; r < n_reads; ++r) {
        const auto &src = haps[rng() % haps.size()];
        const size_t st = rng() % (src.size() - R);
        uint8_t *b = w, *bq = w + R, *iq = w + 2 * R, *dq = w + 3 * R, *gc = w + 4 * R;
        const bool foreign = rng() % 50 == 0;                // a mis-mapped read: underflows in float, takes the FP64 pass
        for (int i = 0; i < R; ++i) {
            int q = (int)std::lround(qd(rng)); q = std::clamp(q, 2, 41); if (q < 18) q = 6;      // PHLCE.cpp:262-263
            bq[i] = (uint8_t)q; iq[i] = dq[i] = 45; gc[i] = 10;
            const bool err = foreign || (rng() % 100000) < (uint64_t)(100000.0 * std::pow(10.0, -q / 10.0));
            b[i] = err ? (uint8_t)acgt[rng() & 3] : src[st + i];
        }
        g.reads.push_back(phmm_read{b, bq, iq, dq, gc, R});
        w += 5 * R;
    }
    g.out.assign((size_t)n_reads * n_haps, 0.0); g.out_batch = g.out;
    g.cells = (uint64_t)n_reads * R * hap_cols;
}

int main(int argc, char **argv) {
    int n_regions = 1200; bool use_submit = false; int window_us = 0; unsigned mask = 0;
    for (int i = 1; i < argc; ++i) {
        if (!std::strcmp(argv[i], "--submit")) use_submit = true;
        else if (!std::strcmp(argv[i], "--batch-window") && i + 1 < argc) window_us = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--devices") && i + 1 < argc) mask = (unsigned)std::strtoul(argv[++i], nullptr, 0);
        else n_regions = std::atoi(argv[i]);
    }
    if (pairhmm_cuda_device_count() < 1) { std::fprintf(stderr, "no sm_100 device: the engine has no CPU fallback\n"); return 2; }
    std::vector<Region> regs((size_t)n_regions);
    {
        std::vector<std::thread> th;
        for (int t = 0; t < 16; ++t) th.emplace_back([&, t] { for (int g = t; g < n_regions; g += 16) make_region(regs[g], 0xB200 + g); });
        for (auto &x : th) x.join();
    }
    uint64_t cells = 0;
    for (auto &g : regs) cells += g.cells;
    pairhmm_cuda_config cfg; pairhmm_cuda_default_config(&cfg);
    cfg.device_mask = mask;
    {   // reference result: the same regions in one batch call
        pairhmm_cuda_handle *h = nullptr;
        if (pairhmm_cuda_create(&cfg, &h)) return 3;
        std::vector<phmm_region> pr;
        for (auto &g : regs) pr.push_back(phmm_region{g.reads.data(), (int32_t)g.reads.size(), g.haps.data(), (int32_t)g.haps.size(), g.out_batch.data(), nullptr});
        for (int rep = 0; rep < 2; ++rep) {
            auto t0 = std::chrono::steady_clock::now();
            if (pairhmm_cuda_compute_batch(h, pr.data(), (int64_t)pr.size())) { std::fprintf(stderr, "%s\n", pairhmm_cuda_last_error(h)); return 4; }
            const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (rep) std::printf("{\"mode\": \"one batch call\", \"regions\": %d, \"cells\": %.4g, \"seconds\": %.4f, \"gcups\": %.1f}\n", n_regions, (double)cells, dt, cells / dt / 1e9);
        }
        pairhmm_cuda_destroy(h);
    }
    cfg.pack_threads = window_us > 0 ? 0 : 1;
    cfg.batch_window_us = window_us;
    for (int T : {1, 2, 4, 8, 16, 32}) {
        std::vector<pairhmm_cuda_handle *> hs((size_t)T, nullptr);
        for (auto &h : hs) { if (pairhmm_cuda_create(&cfg, &h)) return 3; pairhmm_cuda_initialize(h, 150, 512); }
        std::atomic<int> errors{0};
        double dt = 0;
        for (int rep = 0; rep < 2; ++rep) {                  // the first pass warms the handles' buffers
            for (auto &g : regs) std::fill(g.out.begin(), g.out.end(), 0.0);
            std::vector<std::thread> th;
            auto t0 = std::chrono::steady_clock::now();
            for (int t = 0; t < T; ++t) th.emplace_back([&, t] {
                pairhmm_cuda_handle *h = hs[t];
                pairhmm_cuda_ticket prev = 0;
                for (int g = t; g < n_regions; g += T) {
                    Region &r = regs[g];
                    if (!use_submit) {
                        if (pairhmm_cuda_compute(h, r.reads.data(), (int)r.reads.size(), r.haps.data(), (int)r.haps.size(), r.out.data(), nullptr)) ++errors;
                    } else {                                   // overlap this thread's next region with the current one
                        pairhmm_cuda_ticket tk = 0;
                        if (pairhmm_cuda_submit(h, r.reads.data(), (int)r.reads.size(), r.haps.data(), (int)r.haps.size(), r.out.data(), nullptr, &tk)) ++errors;
                        if (prev && pairhmm_cuda_wait(h, prev)) ++errors;
                        prev = tk;
                    }
                }
                if (prev && pairhmm_cuda_wait(h, prev)) ++errors;
            });
            for (auto &x : th) x.join();
            dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }
        bool same = true;
        for (auto &g : regs) same = same && !std::memcmp(g.out.data(), g.out_batch.data(), g.out.size() * sizeof(double));
        std::printf("{\"mode\": \"%s per region\", \"threads\": %d, \"regions\": %d, \"seconds\": %.4f, \"gcups\": %.1f, \"mean_call_ms\": %.3f, "
                    "\"errors\": %d, \"identical_to_batch_call\": %s}\n", use_submit ? "submit/wait" : window_us > 0 ? "compute (shared batcher)" : "compute", T, n_regions, dt,
                    cells / dt / 1e9, dt * T / n_regions * 1e3, errors.load(), same ? "true" : "false");
        std::fflush(stdout);
        for (auto &h : hs) pairhmm_cuda_destroy(h);
    }
    return 0;
}
