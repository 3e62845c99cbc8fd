// TEST INFRASTRUCTURE ONLY. C entry point over the reference's OWN Smith-Waterman (Intel GKL sources under
// M2/intel/smithwaterman, compiled unmodified from where they lie by oracle/Makefile): smithwaterman_initial() +
// SmithWaterman_align() per pair, on `threads` host threads; returns the wall seconds of the loop.
#include <chrono>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#include "intel/smithwaterman/IntelSmithWaterman.h"

extern "C" double ref_sw_align_batch(const uint8_t *ref_bases, const int64_t *ref_off, const int32_t *ref_len,
                                     const uint8_t *alt_bases, const int64_t *alt_off, const int32_t *alt_len, int64_t n,
                                     int match, int mismatch, int open, int extend, int strategy, char *cigars,
                                     int cigar_stride, int32_t *offsets, int threads) {
    static bool init = (smithwaterman_initial(), true);
    (void)init;
    if (threads < 1) threads = 1;
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int t = 0; t < threads; ++t)
        th.emplace_back([=] {
            for (int64_t p = t; p < n; p += threads) {
                std::memset(cigars + p * cigar_stride, 0, (size_t)cigar_stride);
                offsets[p] = SmithWaterman_align(const_cast<uint8_t *>(ref_bases + ref_off[p]), ref_len[p],
                                                 const_cast<uint8_t *>(alt_bases + alt_off[p]), alt_len[p],
                                                 reinterpret_cast<uint8_t *>(cigars + p * cigar_stride), cigar_stride, match,
                                                 mismatch, open, extend, (uint8_t)strategy);
            }
        });
    for (auto &x : th) x.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
