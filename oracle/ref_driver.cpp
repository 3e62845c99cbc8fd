// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// C entry points around the *reference's own* PairHMM sources, which oracle/Makefile compiles
// unmodified from /root/reference into oracle/_ref/libref_pairhmm.so. Nothing from the reference
// is copied into this repository; this file only calls its public functions:
//   initNative                              M2/intel/pairhmm/IntelPairHmm.cc:202
//   computeLikelihoodsNative_concurrent_i   M2/intel/pairhmm/IntelPairHmm.cc:332
//   g_compute_full_prob_float (fn pointer)  M2/intel/pairhmm/IntelPairHmm.cc:50
//   ReadForPairHMM / initializeFloatVector  M2/haplotypecaller/ReadForPairHMM.cpp:18,103
//   testcase                                M2/intel/pairhmm/pairhmm_common.h:45
//   Context<float|double> tables            M2/intel/pairhmm/Context.h:36-210
// (M2 = deepmutect/Mutect2Cpp-master/src). Used by tests/ (to pin the C oracle and to make the
// golden fixtures) and by bench.py's cpu_baseline / --impl reference arm.
#include <cstdint>
#include <cstring>
#include <vector>
#include <memory>
#include <thread>
#include <chrono>
#include <algorithm>
#include <numeric>

#include "intel/pairhmm/IntelPairHmm.h"
#include "intel/pairhmm/Context.h"
#include "intel/common/avx.h"

extern float (*g_compute_full_prob_float)(testcase *tc);

namespace {
struct Flat {
    const uint8_t *bases, *base_q, *ins_q, *del_q, *gcp;
    const int64_t *off;
    const int32_t *len;
};

std::shared_ptr<ReadForPairHMM> make_read(const Flat &f, int r) {
    const int64_t o = f.off[r];
    return std::make_shared<ReadForPairHMM>(f.len[r], f.base_q + o, f.ins_q + o, f.del_q + o,
                                            (const char *)(f.gcp + o), f.bases + o);
}
}  // namespace

extern "C" {

// 1 when the reference's own CPUID test selects its AVX-512 kernels, else 0 (AVX-256).
int ref_pairhmm_isa(void) { return is_avx512_supported() ? 512 : 256; }

// Copies the reference's lookup tables out (built by Context<>'s constructors).
void ref_pairhmm_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d, int mm_count) {
    Context<float> cf;
    Context<double> cd;
    std::memcpy(ph2pr_f, Context<float>::ph2pr, 128 * sizeof(float));
    std::memcpy(ph2pr_d, Context<double>::ph2pr, 128 * sizeof(double));
    std::memcpy(mm_f, Context<float>::matchToMatchProb, mm_count * sizeof(float));
    std::memcpy(mm_d, Context<double>::matchToMatchProb, mm_count * sizeof(double));
}

// Runs the reference dispatch (float kernel -> <1e-28f -> double kernel -> log10) over a flat pair
// list with n_threads std::threads, one static cost-balanced slice each (this mirrors Mutect2Cpp's
// one-engine-per-thread model, M2/main.cpp:199,570-574). Returns wall seconds of the compute loop.
// tables_in_timed_region: 0 = per-read float tables prebuilt before the clock starts, 1 = inside.
// rescued (nullable): 1 where the float kernel's result was < 1e-28f (costs one more float pass,
// outside the timed region).
double ref_pairhmm_compute(const uint8_t *read_bases, const uint8_t *base_q, const uint8_t *ins_q,
                           const uint8_t *del_q, const uint8_t *gcp, const int64_t *read_off,
                           const int32_t *read_len, int n_reads, const uint8_t *hap_bases,
                           const int64_t *hap_off, const int32_t *hap_len, int n_haps,
                           const uint32_t *pair_read, const uint32_t *pair_hap, int64_t n_pairs,
                           int use_double, int n_threads, int tables_in_timed_region, double *out,
                           uint8_t *rescued) {
    (void)n_haps;
    Flat f{read_bases, base_q, ins_q, del_q, gcp, read_off, read_len};
    if (n_threads < 1) n_threads = 1;
    initNative(use_double != 0, 1);

    std::vector<std::shared_ptr<ReadForPairHMM>> reads(n_reads);
    for (int r = 0; r < n_reads; ++r) reads[r] = make_read(f, r);
    if (!tables_in_timed_region)
        for (auto &r : reads) r->initializeFloatVector();

    std::vector<testcase> tcs;
    tcs.reserve(n_pairs);
    for (int64_t p = 0; p < n_pairs; ++p)
        tcs.emplace_back(hap_len[pair_hap[p]], hap_bases + hap_off[pair_hap[p]], reads[pair_read[p]]);
    std::vector<double> result(n_pairs);

    // LPT-ish static split: pairs sorted by cost descending, dealt to the least loaded thread.
    std::vector<std::vector<int64_t>> slice(n_threads);
    {
        std::vector<int64_t> order(n_pairs);
        std::iota(order.begin(), order.end(), 0);
        auto cost = [&](int64_t p) { return (int64_t)read_len[pair_read[p]] * hap_len[pair_hap[p]]; };
        std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return cost(a) > cost(b); });
        std::vector<int64_t> load(n_threads, 0);
        for (int64_t p : order) {
            int t = (int)(std::min_element(load.begin(), load.end()) - load.begin());
            slice[t].push_back(p);
            load[t] += cost(p);
        }
        for (auto &s : slice) std::sort(s.begin(), s.end());
    }

    auto t0 = std::chrono::steady_clock::now();
    if (tables_in_timed_region) {
        // each read's table is built by whichever thread owns reads r = t (mod n_threads)
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t)
            th.emplace_back([&, t] { for (int r = t; r < n_reads; r += n_threads) reads[r]->initializeFloatVector(); });
        for (auto &x : th) x.join();
    }
    {
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t)
            th.emplace_back([&, t] {
                initNative(use_double != 0, 1);  // FTZ is per thread (IntelPairHmm.cc:230)
                for (int64_t p : slice[t]) computeLikelihoodsNative_concurrent_i(tcs, result, (unsigned long)p);
            });
        for (auto &x : th) x.join();
    }
    auto t1 = std::chrono::steady_clock::now();

    std::memcpy(out, result.data(), n_pairs * sizeof(double));
    if (rescued) {
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t)
            th.emplace_back([&, t] {
                initNative(use_double != 0, 1);
                for (int64_t p : slice[t])
                    rescued[p] = use_double ? 1 : (g_compute_full_prob_float(&tcs[p]) < MIN_ACCEPTED ? 1 : 0);
            });
        for (auto &x : th) x.join();
    }
    return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
