#include "engine.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <thread>

#include "hostcopy.h"
#include "hostpool.h"
#include "tables.h"

using namespace phmm;

namespace phmm {

template <class T> cudaError_t PinnedBuf<T>::reserve(size_t n) {
    if (n <= cap) return cudaSuccess;
    size_t want = std::max<size_t>(n + n / 4, 1024);
    T *q = nullptr;
    cudaError_t e = cudaHostAlloc((void **)&q, want * sizeof(T), cudaHostAllocDefault);
    if (e != cudaSuccess) return e;
    if (p) cudaFreeHost(p);
    p = q; cap = want;
    return cudaSuccess;
}
template <class T> void PinnedBuf<T>::release() { if (p && !view) cudaFreeHost(p); p = nullptr; cap = 0; view = false; }
template <class T> cudaError_t DevBuf<T>::reserve(size_t n) {
    if (n <= cap) return cudaSuccess;
    size_t want = std::max<size_t>(n + n / 4, 1024);
    T *q = nullptr;
    cudaError_t e = cudaMalloc((void **)&q, want * sizeof(T));
    if (e != cudaSuccess) return e;
    if (p) cudaFree(p);
    p = q; cap = want;
    return cudaSuccess;
}
template <class T> void DevBuf<T>::release() { if (p && !view) cudaFree(p); p = nullptr; cap = 0; view = false; }

void Slot::release() {
    h_meta.release(); d_meta.release();
    h_rd.release(); h_hapbytes.release(); h_flag.release(); h_reads.release(); h_haps.release();
    h_hapoff.release(); h_haplist.release(); h_happlan.release(); d_happlan.release(); d_long_scratch.release(); h_units.release(); h_out.release();
    h_counters.release(); h_res.release(); d_res.release();
    h_rows.release(); h_summ.release(); d_rows.release(); d_summ.release(); h_mapq.release(); d_mapq.release();
    d_rd.release(); d_hapbytes.release(); d_flag.release(); d_reads.release(); d_haps.release();
    d_hapoff.release(); d_haplist.release(); d_hapw.release(); d_counters.release(); d_units.release();
    d_rescue.release(); d_pair_table.release(); d_fres.release(); d_out.release();
    if (ev_begin) cudaEventDestroy(ev_begin);
    if (ev_float_done) cudaEventDestroy(ev_float_done);
    if (ev_fwd_done) cudaEventDestroy(ev_fwd_done);
    if (ev_end) cudaEventDestroy(ev_end);
    if (ev_fork) cudaEventDestroy(ev_fork);
    if (ev_join) cudaEventDestroy(ev_join);
    if (stream2) cudaStreamDestroy(stream2);
    if (stream) cudaStreamDestroy(stream);
    ev_begin = ev_float_done = ev_fwd_done = ev_end = ev_fork = ev_join = nullptr; stream = stream2 = nullptr;
}

}  // namespace phmm

namespace {

using clk = std::chrono::steady_clock;
inline double ms_since(clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); }

std::mutex g_err_mutex;   // plan_region runs on several threads and may report through fail()
// PAIRHMM_CUDA_TRACE=1: per-chunk timeline of a batch on stderr (host ms since the batch began; GPU ms since the
// base event recorded at the same moment)
bool trace_on() { static const bool on = std::getenv("PAIRHMM_CUDA_TRACE") != nullptr; return on; }
clk::time_point g_trace_t0_[64];          // per device: several devices of one multi-device handle trace side by side
cudaEvent_t g_trace_base_[64] = {nullptr};
#define g_trace_t0 g_trace_t0_[h->device & 63]
#define g_trace_base g_trace_base_[h->device & 63]

}  // namespace
// plan_region runs on several pool threads and may report through fail(): the text is written under a lock (the status
// itself travels through an atomic in pack_chunk), and only the first error of a call is kept by its caller
int phmm::fail(pairhmm_cuda_handle *h, int status, const std::string &msg) {
    if (h) { std::lock_guard<std::mutex> lk(g_err_mutex); h->last_error = msg; }
    return status;
}
namespace {
int fail_cuda(pairhmm_cuda_handle *h, cudaError_t e, const char *what) {
    cudaGetLastError();  // clear sticky-less error state
    return fail(h, e == cudaErrorMemoryAllocation ? PAIRHMM_ERR_OOM : PAIRHMM_ERR_CUDA,
                std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(h, expr)                                                    \
    do {                                                               \
        cudaError_t _e = (expr);                                       \
        if (_e != cudaSuccess) return fail_cuda((h), _e, #expr);       \
    } while (0)

// counters layout in d_counters, one word per (K,KL) bin each: work queues of the FP32 launches,
// rescue counts, work queues of the FP64 launches
// (zeroed per run), then the device copy of Slot::rescue_base
constexpr int kQueueF = 0, kRescueCount = kNumBins, kQueueD = kNumBins + kNumBinsW32, kQueueLong = kNumBins + 2 * kNumBinsW32,
              kRescuedPairs = kNumBins + 2 * kNumBinsW32 + 1;
// ... the same for the W = 16 lists of the FP64 pass (units of two reads) and the number of pairing candidates
constexpr int kRescueCount16 = kRescuedPairs + 1, kQueueD16 = kRescueCount16 + kNumBinsW16, kPendingCount = kQueueD16 + kNumBinsW16;
constexpr int kZeroedWords = kPendingCount + 1;
constexpr int kRescueBaseDev = kZeroedWords, kRescueBase16Dev = kRescueBaseDev + kNumBinsW32 + 1;
constexpr int kCounterWords = kRescueBase16Dev + kNumBinsW16 + 1 + 3;

// Runs fn(i) for i in [0,n) on the calling thread plus up to `threads` - 1 threads of the process-wide pool
// (dynamic chunks of `grain`; hostpool.h).
template <class F> void parallel_for(size_t n, int threads, size_t grain, F fn) {
    HostPool::instance().parallel_for(n, threads, grain, fn);
}

inline uint64_t mix64(uint64_t x) {
    x ^= x >> 32; x *= 0xd6e8feb86659fd93ull; x ^= x >> 32; x *= 0xd6e8feb86659fd93ull; x ^= x >> 32;
    return x;
}
// De-duplication hash of a byte string: four independent multiply-xor lanes (one multiply per 8 bytes, no serial
// avalanche per word -- hashing the five arrays of every read was the largest part of the host packing time), one
// avalanche at the end. Equal hashes are always confirmed by memcmp (same_read), so quality only affects speed.
inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
uint64_t hash_bytes(const uint8_t *p, size_t n, uint64_t h) {
    constexpr uint64_t M = 0x9e3779b97f4a7c15ull;
    uint64_t a = h, b = h ^ 0x243f6a8885a308d3ull, c = h ^ 0x13198a2e03707344ull, d = h ^ 0xa4093822299f31d0ull;
    size_t i = 0;
    for (; i + 32 <= n; i += 32) {
        uint64_t w[4]; std::memcpy(w, p + i, 32);
        a = (a ^ w[0]) * M; b = (b ^ w[1]) * M; c = (c ^ w[2]) * M; d = (d ^ w[3]) * M;
        a ^= a >> 29; b ^= b >> 29; c ^= c >> 29; d ^= d >> 29;
    }
    for (; i + 8 <= n; i += 8) { uint64_t w; std::memcpy(&w, p + i, 8); a = (a ^ w) * M; a ^= a >> 29; }
    uint64_t w = 0;
    if (i < n) std::memcpy(&w, p + i, n - i);
    return mix64(a ^ rotl64(b, 17) ^ rotl64(c, 31) ^ rotl64(d, 47) ^ w ^ ((uint64_t)n << 56));
}
// a read of either kind: processed (five arrays) or raw (bases, base_q, optional ins_q / del_q, mapping quality)
struct ReadRef { const uint8_t *bases, *base_q, *ins_q, *del_q, *gcp; int32_t len, mapq; };
inline ReadRef read_ref(const RegionView &rv, size_t i) {
    if (rv.raw) { const phmm_raw_read &r = rv.raw[i]; return ReadRef{r.bases, r.base_q, r.ins_q, r.del_q, nullptr, r.len, r.mapq}; }
    const phmm_read &r = rv.reads[i];
    return ReadRef{r.bases, r.base_q, r.ins_q, r.del_q, r.gcp, r.len, 0};
}
uint64_t hash_read(const ReadRef &r) {
    uint64_t h = 0x243f6a8885a308d3ull ^ (uint64_t)r.len ^ ((uint64_t)(uint32_t)r.mapq << 32);
    // bases and base qualities only: two reads that agree in both and differ in their indel / gap-continuation qualities
    // are rare (Mutect2Cpp derives those from the bases, PHLCE.cpp:123-157,284-292), and same_read() compares all five
    // arrays anyway -- hashing 2 arrays instead of 5 halves the plan phase of the packer
    h = hash_bytes(r.bases, r.len, h); h = hash_bytes(r.base_q, r.len, h);
    return h;
}
inline bool same_plane(const uint8_t *a, const uint8_t *b, int32_t n) { return (!a || !b) ? a == b || (!a && !b) : !std::memcmp(a, b, n); }
bool same_read(const ReadRef &a, const ReadRef &b) {
    return a.len == b.len && a.mapq == b.mapq && !std::memcmp(a.bases, b.bases, a.len) && !std::memcmp(a.base_q, b.base_q, a.len) &&
           same_plane(a.ins_q, b.ins_q, a.len) && same_plane(a.del_q, b.del_q, a.len) && same_plane(a.gcp, b.gcp, a.len);
}

constexpr uint32_t kNoRead = 0xffffffffu;

// getErrorModelAdjustedQual (PairHMMLikelihoodCalculationEngine.cpp:45-48; MathUtils::fastRound MathUtils.cpp:65-68;
// INITIAL_QSCORE 40, MIN_ADJUSTED_QSCORE 10), host libm as in the reference
uint8_t pcr_adjusted_qual(int repeat_length, int rate_factor) {
    const double d = 40.0 - std::exp((double)repeat_length / ((double)rate_factor * M_PI));
    const int r = d > 0.0 ? (int)(d + 0.5) : (int)(d - 0.5);
    return (uint8_t)(char)std::max(10, r + 1);
}

// haplotypes per work unit of a region when the batch is large enough to fill the GPU
// runs of (nearly) equal length: the same number of units as runs of at most g haplotypes, without a short last one
// (a region of 32 haplotypes cut at 5 would give 6 units of 5 and one of 2: the warps that draw the short ones idle)
uint32_t balanced_unit_haps(int n_haps, uint32_t g) {
    if (n_haps <= 0 || g == 0) return std::max(1u, g);
    const uint32_t n = (uint32_t)n_haps, units = (n + g - 1) / g;
    return (n + units - 1) / units;
}
uint32_t default_unit_haps(int n_haps, int32_t hmax) {
    uint32_t g = (uint32_t)std::max(1, n_haps);
    const uint32_t cap_cols = 16384u / (uint32_t)(hmax + 1);
    g = std::min(g, std::max(1u, cap_cols));
    return balanced_unit_haps(n_haps, std::min(g, (uint32_t)kMaxUnitHaps));
}

}  // namespace
// what plan_region refuses, without planning (the shared batcher validates a caller's region before it joins a chunk
// of other callers' regions, so that one bad region cannot fail theirs)
int phmm::check_region_args(pairhmm_cuda_handle *h, const RegionView &rv) {
    if (rv.n_reads < 0 || rv.n_haps < 0 || (rv.n_reads > 0 && !rv.reads && !rv.raw) || (rv.n_haps > 0 && !rv.haps))
        return fail(h, PAIRHMM_ERR_ARG, "null pointer or negative count in region");
    if (rv.n_reads > 0 && rv.n_haps > 0 && !rv.out && !h->no_output_ok)
        return fail(h, PAIRHMM_ERR_ARG, "null output array");
    if ((uint32_t)rv.n_haps > 0xffffu) return fail(h, PAIRHMM_ERR_ARG, "more than 65535 haplotypes in one region");
    for (int i = 0; i < rv.n_haps; ++i)
        if (rv.haps[i].len < 0 || (rv.haps[i].len > 0 && !rv.haps[i].bases)) return fail(h, PAIRHMM_ERR_ARG, "bad haplotype");
    for (int i = 0; i < rv.n_reads; ++i) {
        const ReadRef r = read_ref(rv, (size_t)i);
        if (r.len < 0 || (r.len > 0 && (!r.bases || !r.base_q || (!rv.raw && (!r.ins_q || !r.del_q || !r.gcp)))))
            return fail(h, PAIRHMM_ERR_ARG, "bad read");
    }
    return PAIRHMM_OK;
}
namespace {
// Pass A for one region: read de-duplication (VectorLoglessPairHMM.cpp:83-103) and sizes.
int plan_region(pairhmm_cuda_handle *h, const RegionView &rv, RegionPlan &pl, uint32_t group_hint) {
    pl = RegionPlan();
    if (rv.n_reads < 0 || rv.n_haps < 0 || (rv.n_reads > 0 && !rv.reads && !rv.raw) || (rv.n_haps > 0 && !rv.haps))
        return fail(h, PAIRHMM_ERR_ARG, "null pointer or negative count in region");
    if (rv.n_reads > 0 && rv.n_haps > 0 && !rv.out && !h->no_output_ok)
        return fail(h, PAIRHMM_ERR_ARG, "null output array");
    pl.uniq_of_read.assign(rv.n_reads, kNoRead);
    const bool dedup = h->cfg.dedup_reads != 0;
    std::vector<uint32_t> table;   // open addressing: local unique index + 1
    std::vector<uint64_t> hashes;
    size_t mask = 0;
    if (dedup && rv.n_reads > 1) {
        size_t cap = 16; while (cap < (size_t)rv.n_reads * 2) cap <<= 1;
        table.assign(cap, 0); mask = cap - 1; hashes.reserve(rv.n_reads);
    }
    int32_t hmax = 0;
    for (int i = 0; i < rv.n_haps; ++i) {
        const phmm_hap &hp = rv.haps[i];
        if (hp.len < 0 || (hp.len > 0 && !hp.bases)) return fail(h, PAIRHMM_ERR_ARG, "bad haplotype");
        pl.hap_bytes += (uint64_t)hp.len;
        pl.hap_words += hap_words(hp.len);
        hmax = std::max(hmax, hp.len);
    }
    for (int i = 0; i < rv.n_reads; ++i) {
        const ReadRef r = read_ref(rv, (size_t)i);
        if (r.len < 0 || (r.len > 0 && (!r.bases || !r.base_q || (!rv.raw && (!r.ins_q || !r.del_q || !r.gcp)))))
            return fail(h, PAIRHMM_ERR_ARG, "bad read");
        pl.read_len_sum += (uint64_t)r.len;
        if (r.len == 0) { pl.identity = false; continue; }   // undefined in the reference; reported as -inf, see scatter
        uint32_t u = kNoRead;
        if (!table.empty()) {
            const uint64_t hv = hash_read(r);
            size_t s = (size_t)hv & mask;
            while (table[s]) {
                const uint32_t cand = table[s] - 1;
                if (hashes[cand] == hv && same_read(read_ref(rv, pl.uniq_src[cand]), r)) { u = cand; break; }
                s = (s + 1) & mask;
            }
            if (u == kNoRead) { table[s] = (uint32_t)pl.uniq_src.size() + 1; hashes.push_back(hv); }
        }
        if (u == kNoRead) {
            u = (uint32_t)pl.uniq_src.size();
            pl.uniq_src.push_back((uint32_t)i);
            pl.read_bytes += (uint64_t)r.len;
            if (rv.raw) {                           // constant by construction (buildGapContinuationPenalties)
                pl.gcp_min = pl.gcp_max = (uint8_t)(h->model.constant_gcp & 127);
                pl.any_indel_q |= r.ins_q != nullptr || r.del_q != nullptr;
            } else {                                // gap-continuation range, as the kernel sees it (&127)
                // the usual case is one constant (buildGapContinuationPenalties): compare 8 bytes at a time against the first
                const uint8_t c0 = r.gcp[0] & 127;
                const uint64_t bc = 0x0101010101010101ull * c0, m7 = 0x7f7f7f7f7f7f7f7full;
                uint64_t diff = 0;
                int q = 0;
                for (; q + 8 <= r.len; q += 8) { uint64_t w; std::memcpy(&w, r.gcp + q, 8); diff |= (w & m7) ^ bc; }
                for (; q < r.len; ++q) diff |= (uint64_t)((r.gcp[q] & 127) ^ c0);
                if (!diff) { pl.gcp_min = std::min(pl.gcp_min, c0); pl.gcp_max = std::max(pl.gcp_max, c0); }
                else for (q = 0; q < r.len; ++q) {
                    const uint8_t c = r.gcp[q] & 127;
                    pl.gcp_min = std::min(pl.gcp_min, c); pl.gcp_max = std::max(pl.gcp_max, c);
                }
            }
        }
        pl.uniq_of_read[i] = u;
        pl.identity &= u == (uint32_t)i;
    }
    pl.n_uniq = (uint32_t)pl.uniq_src.size();
    pl.hmax = hmax;
    // haplotypes per work unit: all of them unless the chunk is too small to fill the GPU
    // (group_hint) or the unit would get very long (tail balance)
    uint32_t g = default_unit_haps(rv.n_haps, hmax);
    if (group_hint) g = std::min(g, group_hint);
    if (h->cfg.reserved[5] > 0) g = (uint32_t)std::min<int>(std::min(h->cfg.reserved[5], (int)kMaxUnitHaps), std::max(1, rv.n_haps));
    g = balanced_unit_haps(rv.n_haps, g);
    pl.group = g;
    pl.n_groups = rv.n_haps ? ((uint32_t)rv.n_haps + g - 1) / g : 0;
    return PAIRHMM_OK;
}

// Lane-group width for a read of length R under the given FP32 variant. Narrow groups (more rows per
// lane) exist for the uniform-gcp kernels only; forced_w (config.reserved[1]) overrides the choice.
int choose_width(int R, Variant v, int forced_w, bool f64_w16) {
    // use_double chunks with a uniform gap-continuation penalty: the wide FP64 kernel at two reads per warp (W = 16) where
    // the rows per lane fit its registers -- twice the rows per lane halve the per-cell share of shuffles / selects / loop
    if (f64_w16 && R <= 16 * kMaxKF64W16 && bin_of_read(R, 16) >= 0) return 16;
    if (!is_packed_f32(v)) return 32;
    if (forced_w && bin_of_read(R, forced_w) >= 0) return forced_w;
    // the narrowest group that fits wins: more rows per lane amortise the per-column overhead (shuffles,
    // table address, loop) over more cells
    if (bin_of_read(R, 8) >= 0) return 8;
    if (bin_of_read(R, 16) >= 0) return 16;
    return 32;
}

// Pass A2 for one region, once the chunk's kernel variant is known: order the unique reads by kernel bin
// (stable, so equal-length reads keep their input order) and form the bin runs.
void bin_region(const RegionView &rv, RegionPlan &pl, Variant v, int forced_w, bool f64_w16) {
    const uint32_t n = pl.n_uniq;
    std::vector<uint16_t> bin(n);
    for (uint32_t j = 0; j < n; ++j) {
        const int R = read_ref(rv, pl.uniq_src[j]).len;
        if (R > kMaxReadLen) { bin[j] = (uint16_t)kLongBin; continue; }      // swept in row stripes by long.cu, one unit per pair
        bin[j] = (uint16_t)bin_of_read(R, choose_width(R, v, forced_w, f64_w16));
        pl.pairs32[bin_of_read(R, 32)] += (uint32_t)rv.n_haps;
        if (R <= 16 * kMaxKF64W16) pl.pairs16[bin_of_read(R, 16) - kBinBase[1]] += (uint32_t)rv.n_haps;
    }
    std::vector<uint32_t> order(n), inv(n);
    for (uint32_t j = 0; j < n; ++j) order[j] = j;
    bool sorted = true;
    for (uint32_t j = 1; j < n && sorted; ++j) sorted = bin[j - 1] <= bin[j];
    if (!sorted) {
        pl.identity = false;
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return bin[a] < bin[b]; });
        std::vector<uint32_t> src(n);
        for (uint32_t j = 0; j < n; ++j) { src[j] = pl.uniq_src[order[j]]; inv[order[j]] = j; }
        pl.uniq_src.swap(src);
        for (auto &u : pl.uniq_of_read) if (u != kNoRead) u = inv[u];
    }
    pl.runs.clear();
    for (uint32_t j = 0; j < n;) {
        const uint16_t b = bin[order[j]];
        uint32_t e = j;
        while (e < n && bin[order[e]] == b) ++e;
        pl.runs.push_back(RegionPlan::BinRun{b, j, e - j, 0});
        j = e;
    }
}

constexpr double kPfxMinSharedFrac = 0.12;   // config.reserved[4] == 1: share only if the plan skips more than this part of the cells

// Pass A3 (Variant::F32_PFX only): order the region's haplotypes by (length, bytes) and plan, per work unit, which
// sweeps start from a checkpoint. The GPU analogue of the reference's trie (M2/trie/buildTreeUtils.cpp:155-258:
// buildTreeWithHaplotype_same_height shares 32-column prefix blocks between haplotypes of one length; here the unit is
// a tile of kTile columns). In sorted order the prefix two neighbours share is the longest either shares with anything
// before it, so a stack of checkpoints (ascending tile counts, all on the current haplotype's path) is enough:
// haplotype j drops the checkpoints beyond what it shares with j-1, resumes from the deepest one left, and leaves one at
// the boundary it shares with j+1 if that lies inside its own sweep and a slot is free.
void plan_prefix(const RegionView &rv, RegionPlan &pl) {
    const uint32_t n = (uint32_t)rv.n_haps;
    pl.hap_order.resize(n);
    for (uint32_t j = 0; j < n; ++j) pl.hap_order[j] = j;
    std::stable_sort(pl.hap_order.begin(), pl.hap_order.end(), [&](uint32_t a, uint32_t b) {
        const phmm_hap &x = rv.haps[a], &y = rv.haps[b];
        if (x.len != y.len) return x.len < y.len;
        return x.len > 0 && std::memcmp(x.bases, y.bases, (size_t)x.len) < 0;
    });
    pl.hap_plan.assign(n, pfx_plan(0, 0, 0, 0));
    pl.shared_cols = 0;
    const uint32_t group = std::max(1u, pl.group);
    std::vector<uint32_t> shared(n + 1, 0);        // tiles position j shares with position j-1 of the same unit
    for (uint32_t j = 1; j < n; ++j) {
        if (j % group == 0) continue;
        const phmm_hap &x = rv.haps[pl.hap_order[j - 1]], &y = rv.haps[pl.hap_order[j]];
        if (x.len != y.len || x.len <= 0 || hap_tiles(x.len) > kPfxMaxTiles) continue;
        int32_t lcp = 0;
        while (lcp < x.len && x.bases[lcp] == y.bases[lcp]) ++lcp;
        shared[j] = lcp == x.len ? ((uint32_t)lcp + kTile - 1u) / kTile : (uint32_t)lcp / kTile;
    }
    struct Ck { uint32_t tiles, slot; } stack[kPfxSlots];
    int depth = 0;
    for (uint32_t j = 0; j < n; ++j) {
        if (j % group == 0) depth = 0;
        while (depth && stack[depth - 1].tiles > shared[j]) --depth;
        uint32_t ts = 0, rs = 0, tv = 0, ss = 0;
        if (depth) { ts = stack[depth - 1].tiles; rs = stack[depth - 1].slot + 1; }
        const uint32_t next = (j + 1 < n && (j + 1) % group != 0) ? shared[j + 1] : 0;
        if (next > ts && depth < kPfxSlots) {
            uint32_t used = 0;
            for (int d = 0; d < depth; ++d) used |= 1u << stack[d].slot;
            while (used & (1u << ss)) ++ss;
            tv = next;
            stack[depth++] = Ck{next, ss};
        }
        pl.hap_plan[j] = pfx_plan(ts, rs, tv, ss);
        pl.shared_cols += std::min<uint64_t>((uint64_t)rv.haps[pl.hap_order[j]].len, (uint64_t)ts * kTile);
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// chunk = set of regions packed into one slot
// ------------------------------------------------------------------------------------------------
namespace phmm {

int pack_chunk(pairhmm_cuda_handle *h, Slot &s, const RegionView *regs, size_t n_regs, int threads) {
    auto t0 = clk::now();
    s.regions.assign(regs, regs + n_regs);
    s.plans.resize(n_regs);
    // group hint: if there are few reads, split haplotype runs so that every warp slot gets work
    uint64_t total_reads = 0;
    for (size_t i = 0; i < n_regs; ++i) total_reads += (uint64_t)std::max(0, regs[i].n_reads);
    // the packed kernels keep 8 warps per SM resident (2 CTAs of 4 warps): aim at three waves of warp units. Counted in
    // warps, not reads -- a warp takes 2-8 reads (tools/unit_sweep.py: C3 2699 -> 3162, C4 2954 -> 3093 GCUPS in the FP32
    // pass against the former reads-based rule, which cut the runs of mid-sized chunks too short or raggedly)
    // Other chunks in flight on this device (other host threads' handles, this handle's other slot): the more there are,
    // the fewer units this chunk needs to keep the device full, and the longer -- cheaper per cell -- they may be.
    const int others = device_chunks_in_flight(h->device);
    const uint64_t want_units = std::max<uint64_t>((uint64_t)h->sm_count * 8u * 3u / (uint64_t)(1 + std::max(0, others)), 64u);
    uint32_t group_hint = 0;
    int chunk_w = h->cfg.reserved[1];       // lane-group width forced for this chunk (0 = the narrowest that fits, per read)
    if (total_reads && total_reads < (uint64_t)h->sm_count * 8u * 3u * 8u) {
        double warps = 0.0;
        const int forced = h->cfg.reserved[1];
        uint64_t warps_at[kNumWidths] = {0, 0, 0};
        bool fits_at[kNumWidths] = {true, true, true};
        int rmax = 1;
        uint32_t maxh = 1;
        double tiles = 0.0, nhaps = 0.0;
        for (size_t i = 0; i < n_regs; ++i) {
            maxh = std::max<uint32_t>(maxh, (uint32_t)std::max(1, regs[i].n_haps));
            if (regs[i].haps) for (int q = 0; q < regs[i].n_haps; ++q) { tiles += hap_tiles(std::max(0, regs[i].haps[q].len)); nhaps += 1.0; }
            if (regs[i].n_reads <= 0 || (!regs[i].reads && !regs[i].raw)) continue;
            uint32_t n_ok = 0;
            for (int r = 0; r < regs[i].n_reads; ++r) {
                const int len = read_ref(regs[i], (size_t)r).len;
                if (len < 1 || len > kMaxReadLen) continue;
                ++n_ok; rmax = std::max(rmax, len);
                warps += 1.0 / (double)reads_per_warp(Variant::F32_PK, choose_width(len, Variant::F32_PK, forced, false));
            }
            for (int wi = 0; wi < kNumWidths; ++wi) {
                const uint32_t per_warp = (uint32_t)reads_per_warp(Variant::F32_PK, kWidths[wi]);
                warps_at[wi] += (n_ok + per_warp - 1) / per_warp;
            }
        }
        for (int wi = 0; wi < kNumWidths; ++wi) fits_at[wi] = bin_of_read(rmax, kWidths[wi]) >= 0;
        // A chunk so small that ONE wave of warps takes all of it even with a unit per (warp of reads, haplotype) -- a
        // region per call -- is as slow as its slowest unit: (tiles of one haplotype + W fill steps) x rows per lane. Wide
        // lane groups make that short (W = 32, K = 5 sweeps a 150-base read in 0.57 of the time of W = 16, K = 10), so
        // the width that is best for throughput (the narrowest) is not taken here.
        const uint64_t slots = (uint64_t)h->sm_count * 8u;
        if (!forced && warps > 0.0 && others == 0) {
            double best = 0.0;
            const double t_avg = nhaps > 0.0 ? tiles / nhaps : 1.0;
            for (int wi = 0; wi < kNumWidths; ++wi) {
                if (!fits_at[wi] || warps_at[wi] * (uint64_t)maxh > slots) continue;
                const double t = (t_avg + (double)kWidths[wi]) * (double)((rmax + kWidths[wi] - 1) / kWidths[wi]);
                if (best == 0.0 || t < best) { best = t; chunk_w = kWidths[wi]; group_hint = 1; }
            }
        }
        if (!group_hint && warps > 0.0 && warps < (double)want_units) {
            const uint64_t splits = (uint64_t)std::ceil((double)want_units / warps);   // units wanted per warp of reads
            group_hint = (uint32_t)std::max<uint64_t>(1, maxh / std::min<uint64_t>(splits, maxh));
        }
    }
    std::atomic<int> status{PAIRHMM_OK};
    parallel_for(n_regs, threads, 8, [&](size_t i) {
        int st = plan_region(h, s.regions[i], s.plans[i], group_hint);
        if (st != PAIRHMM_OK) status.store(st);
    });
    if (status.load() != PAIRHMM_OK) return status.load();

    // kernel variant of the chunk: uniform gap-continuation penalty -> constant-bank kernels
    uint8_t gmin = 255, gmax = 0;
    for (size_t i = 0; i < n_regs; ++i) { gmin = std::min(gmin, s.plans[i].gcp_min); gmax = std::max(gmax, s.plans[i].gcp_max); }
    s.uniform_gcp = gmin == gmax;
    s.gcp_value = gmin;
    // use_double bins at W = 32 for the FP64 kernels; reserved[2] selects the scalar uniform-gcp kernel (cross-check)
    s.float_variant = (h->cfg.use_double || h->cfg.exact_fp32) ? (h->cfg.exact_fp32 ? Variant::F32_EXACT : Variant::F32_GEN)
                      : (s.uniform_gcp && !h->cfg.reserved[0]) ? (h->cfg.reserved[2] ? Variant::F32_UG : Variant::F32_PK)
                      : Variant::F32_GEN;
    // prefix sharing (config.reserved[4]) reorders a region's haplotypes inside the chunk; the fused epilogue reports
    // haplotype indices and breaks ties by position, so it keeps the caller's order
    const bool pfx_candidate = s.float_variant == Variant::F32_PK && h->cfg.reserved[4] && !h->ep_on;
    const int forced_w = chunk_w;
    static const bool f64_w16_on = std::getenv("PAIRHMM_CUDA_F64_W32") == nullptr;
    const bool f64_w16 = f64_w16_on && h->cfg.use_double && s.uniform_gcp && !h->cfg.reserved[0] && !h->cfg.reserved[2] && !h->cfg.exact_fp32;
    const double ms_plan = ms_since(t0);
    parallel_for(n_regs, threads, 8, [&](size_t i) {
        bin_region(s.regions[i], s.plans[i], s.float_variant, forced_w, f64_w16);
        if (pfx_candidate) plan_prefix(s.regions[i], s.plans[i]);
    });
    if (pfx_candidate) {
        // the sharing kernel's dirty and checkpoint steps cost ~10 % of its rate (tools/prefix_perf.py): reserved[4] == 1
        // takes it only when the plan saves more than that, reserved[4] >= 2 always (testing)
        uint64_t all = 0, shared = 0;
        for (size_t i = 0; i < n_regs; ++i) { all += s.plans[i].read_bytes * s.plans[i].hap_bytes; shared += s.plans[i].read_bytes * s.plans[i].shared_cols; }
        if (h->cfg.reserved[4] >= 2 || (double)shared > kPfxMinSharedFrac * (double)all) s.float_variant = Variant::F32_PFX;
        else for (size_t i = 0; i < n_regs; ++i) { s.plans[i].hap_order.clear(); s.plans[i].hap_plan.clear(); s.plans[i].shared_cols = 0; }
    }
    const double ms_bin = ms_since(t0);

    // prefix sums
    uint64_t rbytes = 0, hbytes = 0, cells = 0, pairs = 0, reads_in = 0, swept = 0;
    uint32_t nu = 0, nh = 0, hw = kLeadWords, hl = 0, up = 0;
    std::vector<uint32_t> bin_count(kNumBins + 1, 0);
    uint32_t bin_pairs[kNumBinsW32] = {0};
    uint32_t bin_pairs16[kNumBinsW16] = {0};
    int32_t hap_max = 0;
    for (size_t i = 0; i < n_regs; ++i) {
        RegionPlan &p = s.plans[i];
        p.first_uniq = nu; p.read_byte_off = rbytes; p.first_hap = nh; p.hap_byte_off = hbytes;
        p.hap_word_off = hw; p.hap_list_off = hl; p.upair_off = up;
        hap_max = std::max(hap_max, p.hmax);
        for (auto &run : p.runs) {
            run.unit_off = bin_count[run.bin];
            if (run.bin == kLongBin) { bin_count[kLongBin] += run.n_reads * (uint32_t)s.regions[i].n_haps; continue; }   // one unit per pair
            int W, K, KL; bin_shape(run.bin, W, K, KL);
            const uint32_t per_warp = (uint32_t)reads_per_warp(s.float_variant, W);
            bin_count[run.bin] += (run.n_reads + per_warp - 1) / per_warp * p.n_groups;
        }
        for (int k = 0; k < kNumBinsW32; ++k) bin_pairs[k] += p.pairs32[k];
        for (int k = 0; k < kNumBinsW16; ++k) bin_pairs16[k] += p.pairs16[k];
        if ((uint32_t)s.regions[i].n_haps > 0xffffu)
            return fail(h, PAIRHMM_ERR_ARG, "more than 65535 haplotypes in one region");
        const uint64_t up_next = (uint64_t)up + (uint64_t)p.n_uniq * (uint64_t)s.regions[i].n_haps;
        if (up_next > 0xfffffff0ull || rbytes + p.read_bytes > 0xfffffff0ull || (uint64_t)hw + p.hap_words > 0xfffffff0ull)
            return fail(h, PAIRHMM_ERR_ARG, "chunk too large for 32-bit device indices (lower chunk_cells)");
        nu += p.n_uniq; rbytes += p.read_bytes; nh += (uint32_t)s.regions[i].n_haps; hbytes += p.hap_bytes;
        hw += p.hap_words; hl += (uint32_t)s.regions[i].n_haps; up = (uint32_t)up_next;
        cells += p.read_len_sum * p.hap_bytes;
        swept += p.read_bytes * (p.hap_bytes - p.shared_cols);
        pairs += (uint64_t)s.regions[i].n_reads * (uint64_t)s.regions[i].n_haps;
        reads_in += (uint64_t)s.regions[i].n_reads;
    }
    s.bin_start[0] = 0;
    s.rescue_base[0] = 0;
    for (int k = 0; k <= kNumBins; ++k) s.bin_start[k + 1] = s.bin_start[k] + bin_count[k];
    for (int k = 0; k < kNumBinsW32; ++k) s.rescue_base[k + 1] = s.rescue_base[k] + bin_pairs[k];
    // FP64 pass with two reads per warp (rescue_scan_kernel): the lists of the W = 16 bins (a unit takes two reads' pairs)
    // and the pairing candidates lie behind the W = 32 lists. Only where the chunk could fill the GPU that way (the scan
    // decides by the number of pairs that did underflow); config.reserved[7]: 1 = never, 2 = always (testing)
    {
        const bool wide64 = s.uniform_gcp && !h->cfg.use_double && !h->cfg.exact_fp32 && !h->cfg.reserved[0] && !h->cfg.reserved[2];
        const uint32_t slots16 = (uint32_t)h->sm_count * 8u;
        s.pair_mode = (!wide64 || h->cfg.reserved[7] == 1) ? 0u : h->cfg.reserved[7] == 2 ? 2u : (up >= 4u * slots16 ? 1u : 0u);
        s.rescue_base16[0] = up + 1;
        for (int k = 0; k < kNumBinsW16; ++k) s.rescue_base16[k + 1] = s.rescue_base16[k] + (s.pair_mode ? (bin_pairs16[k] + 1) / 2 : 0u);
        s.pending_off = s.rescue_base16[kNumBinsW16];
    }
    const uint32_t n_units = s.bin_start[kNumBins + 1];
    s.hap_max = hap_max;
    // the packed FP32 kernels and the wide FP64 kernel take the (uniform) gap-continuation penalty as a launch constant and
    // never read the gcp plane: it is neither staged nor sent (a fifth of the read bytes). The long-read kernel reads it.
    // the read kind of the chunk = that of its first region that has reads (an empty leading region, e.g. an empty
    // std::vector's data(), says nothing); mixing kinds is refused below
    bool chunk_raw = false;
    for (size_t i = 0; i < n_regs; ++i)
        if (s.regions[i].n_reads > 0) { chunk_raw = s.regions[i].raw != nullptr; break; }
    s.skip_gcp = is_packed_f32(s.float_variant) && s.bin_start[kNumBins + 1] == s.bin_start[kNumBins] && !chunk_raw;
    s.plane = (rbytes + 15) & ~15ull;
    s.n_uniq = nu; s.n_haps = nh; s.n_hap_words = hw; s.n_haplist = hl; s.n_upairs = up; s.hap_bytes = hbytes;
    s.cells = cells; s.pairs = pairs; s.reads_in = reads_in; s.cells_swept = swept;

    s.ep_on = h->ep_on; s.ep = h->ep;
    s.raw_on = chunk_raw; s.model = h->model; s.raw_indel_q = false;
    for (size_t i = 0; i < n_regs; ++i) {
        if ((s.regions[i].raw != nullptr) != s.raw_on && s.regions[i].n_reads > 0)
            return fail(h, PAIRHMM_ERR_ARG, "raw and processed reads mixed in one call");
        s.raw_indel_q |= s.plans[i].any_indel_q;
    }
    // read planes that cross PCIe: raw reads send bases + base qualities (+ their own indel qualities if any read has them),
    // processed reads all but the gcp plane when the kernels take the penalty as a constant
    s.sent_planes = s.raw_on ? (s.raw_indel_q ? 4 : 2) : (s.skip_gcp ? 4 : 5);

    cudaError_t e;
#define RES(buf, n) if ((e = s.buf.reserve(n)) != cudaSuccess) return fail_cuda(h, e, "reserve " #buf)
    {   // ONE arena for everything that goes to the device, same layout on both sides (256-byte aligned parts): descriptors,
        // haplotype bytes, work units, the counter block (zeroed work queues + the rescue list bases), per-read side tables
        // and, last, the read planes -- a chunk crosses PCIe in ONE copy (a per-region call once made nine driver calls
        // here; with T host threads the driver's lock, not the GPU, was what a region call waited for). The device arena
        // has room for all five planes, the host arena for the planes that are sent.
        const bool pfx = s.float_variant == Variant::F32_PFX;
        uint64_t off = 0;
        auto part = [&](uint64_t bytes) { const uint64_t o = off; off = (off + bytes + 255) & ~255ull; return o; };
        const uint64_t o_reads = part(sizeof(DevRead) * ((uint64_t)nu + 1)), o_hapoff = part(sizeof(uint64_t) * ((uint64_t)nh + 1)),
                       o_haps = part(sizeof(DevHap) * ((uint64_t)nh + 1)), o_haplist = part(sizeof(uint32_t) * ((uint64_t)hl + 1)),
                       o_happlan = part(pfx ? sizeof(uint2) * ((uint64_t)hl + 1) : 0), o_units = part(sizeof(DevUnit) * ((uint64_t)n_units + 1)),
                       o_hapbytes = part(hbytes + 16), o_ctr = part(sizeof(uint32_t) * kCounterWords),
                       o_mapq = part(s.raw_on ? sizeof(int32_t) * ((uint64_t)nu + 1) : 0),
                       o_rows = part(s.ep_on ? sizeof(RowInfo) * ((uint64_t)nu + 1) : 0), o_rd = part(0);
        s.meta_bytes = o_rd + s.plane * (uint64_t)s.sent_planes;
        s.ctr_off = o_ctr;
        RES(h_meta, s.meta_bytes + 16); RES(d_meta, o_rd + s.plane * 5 + 16);
        s.h_reads.set_view(s.h_meta.p + o_reads); s.d_reads.set_view(s.d_meta.p + o_reads);
        s.h_hapoff.set_view(s.h_meta.p + o_hapoff); s.d_hapoff.set_view(s.d_meta.p + o_hapoff);
        s.h_haps.set_view(s.h_meta.p + o_haps); s.d_haps.set_view(s.d_meta.p + o_haps);
        s.h_haplist.set_view(s.h_meta.p + o_haplist); s.d_haplist.set_view(s.d_meta.p + o_haplist);
        s.h_happlan.set_view(s.h_meta.p + o_happlan); s.d_happlan.set_view(s.d_meta.p + o_happlan);
        s.h_units.set_view(s.h_meta.p + o_units); s.d_units.set_view(s.d_meta.p + o_units);
        s.h_hapbytes.set_view(s.h_meta.p + o_hapbytes); s.d_hapbytes.set_view(s.d_meta.p + o_hapbytes);
        s.h_counters.set_view(s.h_meta.p + o_ctr); s.d_counters.set_view(s.d_meta.p + o_ctr);
        s.h_mapq.set_view(s.h_meta.p + o_mapq); s.d_mapq.set_view(s.d_meta.p + o_mapq);
        s.h_rows.set_view(s.h_meta.p + o_rows); s.d_rows.set_view(s.d_meta.p + o_rows);
        s.h_rd.set_view(s.h_meta.p + o_rd); s.d_rd.set_view(s.d_meta.p + o_rd);
        std::memset(s.h_counters.p, 0, sizeof(uint32_t) * kCounterWords);
        std::memcpy(s.h_counters.p + kRescueBaseDev, s.rescue_base, sizeof(uint32_t) * (kNumBinsW32 + 1));
        std::memcpy(s.h_counters.p + kRescueBase16Dev, s.rescue_base16, sizeof(uint32_t) * (kNumBinsW16 + 1));
    }
    {   // results: the log10 values and, behind them, the rescued flags -- one buffer on either side, ONE copy back
        const uint64_t flag_off = sizeof(double) * ((uint64_t)up + 1);
        s.res_bytes = flag_off + (uint64_t)up + 1;
        RES(h_res, s.res_bytes); RES(d_res, s.res_bytes);
        s.h_out.set_view(s.h_res.p); s.d_out.set_view(s.d_res.p);
        s.h_flag.set_view(s.h_res.p + flag_off); s.d_flag.set_view(s.d_res.p + flag_off);
    }
    RES(d_hapw, (size_t)hw + 4 * kTile); RES(d_fres, up + 1); RES(d_rescue, s.pair_mode ? (size_t)s.pending_off + up + 1 : (size_t)up + 1);
    if (s.pair_mode && !s.d_pair_table.p) {
        RES(d_pair_table, kPairSlots);
        if ((e = cudaMemsetAsync(s.d_pair_table.p, 0, sizeof(uint32_t) * kPairSlots, s.stream)) != cudaSuccess)
            return fail_cuda(h, e, "zero d_pair_table");
    }
    if (s.ep_on) { RES(h_summ, nu + 1); RES(d_summ, nu + 1); }
#undef RES

    const double ms_reserve = ms_since(t0);
    // pass B: copy bytes, fill descriptors and units
    parallel_for(n_regs, threads, 4, [&](size_t i) {
        const RegionView &rv = s.regions[i];
        const RegionPlan &p = s.plans[i];
        uint64_t ro = p.read_byte_off;
        // Reads that follow one another in the caller's arrays (the flat entry points; a BAM slice) are copied as ONE run per
        // plane: a region of 200 reads is 4 memcpys of 30 KB instead of 800 of 150 bytes (the per-call cost of the small
        // copies, not memory bandwidth, was what the copy phase spent its time on)
        ReadRef run0{}; uint64_t run_ro = 0, run_len = 0;
        auto flush = [&] {
            if (!run_len) return;
            // the next reader of the arena is the copy engine: streaming stores (hostcopy.h; runs below 512 bytes are memcpy'd)
            uint8_t *d = s.h_rd.p + run_ro;
            copy_stream(d, run0.bases, run_len); copy_stream(d + s.plane, run0.base_q, run_len);
            if (!s.raw_on) {
                copy_stream(d + 2 * s.plane, run0.ins_q, run_len); copy_stream(d + 3 * s.plane, run0.del_q, run_len);
                if (!s.skip_gcp) copy_stream(d + 4 * s.plane, run0.gcp, run_len);
            } else if (s.raw_indel_q) {
                // planes 2-4 are filled on the device unless some read of the chunk has its own indel qualities
                if (run0.ins_q) copy_stream(d + 2 * s.plane, run0.ins_q, run_len); else std::memset(d + 2 * s.plane, 45, run_len);
                if (run0.del_q) copy_stream(d + 3 * s.plane, run0.del_q, run_len); else std::memset(d + 3 * s.plane, 45, run_len);
            }
            run_len = 0;
        };
        for (uint32_t j = 0; j < p.n_uniq; ++j) {
            const ReadRef r = read_ref(rv, p.uniq_src[j]);
            const bool follows = run_len && r.bases == run0.bases + run_len && r.base_q == run0.base_q + run_len &&
                                 (s.raw_on ? ((r.ins_q != nullptr) == (run0.ins_q != nullptr) && (!r.ins_q || r.ins_q == run0.ins_q + run_len) &&
                                              (r.del_q != nullptr) == (run0.del_q != nullptr) && (!r.del_q || r.del_q == run0.del_q + run_len))
                                           : (r.ins_q == run0.ins_q + run_len && r.del_q == run0.del_q + run_len &&
                                              (s.skip_gcp || r.gcp == run0.gcp + run_len)));
            if (!follows) { flush(); run0 = r; run_ro = ro; }
            run_len += (uint64_t)r.len;
            if (s.raw_on) s.h_mapq.p[p.first_uniq + j] = r.mapq;
            s.h_reads.p[p.first_uniq + j] = DevRead{(uint32_t)ro, r.len};
            if (s.ep_on) s.h_rows.p[p.first_uniq + j] = RowInfo{p.upair_off + j * (uint32_t)rv.n_haps, (uint32_t)rv.n_haps};
            ro += (uint64_t)r.len;
        }
        flush();
        // work units: per bin run, 32/W consecutive reads per warp, one unit per haplotype group
        const uint32_t nhap = (uint32_t)rv.n_haps;
        for (const auto &run : p.runs) {
            DevUnit *ud = s.h_units.p + s.bin_start[run.bin] + run.unit_off;
            if (run.bin == kLongBin) {           // long reads: one unit per (read, haplotype) pair
                for (uint32_t j = 0; j < run.n_reads; ++j)
                    for (uint32_t q = 0; q < nhap; ++q)
                        *ud++ = DevUnit{p.first_uniq + run.first + j, p.hap_list_off + q, unit_shape(1u, 1u, nhap),
                                        p.upair_off + (run.first + j) * nhap + q};
                continue;
            }
            int W, K, KL; bin_shape(run.bin, W, K, KL);
            const uint32_t per_warp = (uint32_t)reads_per_warp(s.float_variant, W);
            for (uint32_t j = 0; j < run.n_reads; j += per_warp) {
                const uint32_t nr = std::min(per_warp, run.n_reads - j), first = run.first + j;
                for (uint32_t g = 0; g < p.n_groups; ++g) {
                    const uint32_t h0 = g * p.group;
                    *ud++ = DevUnit{p.first_uniq + first, p.hap_list_off + h0,
                                    unit_shape(std::min(p.group, nhap - h0), nr, nhap),
                                    p.upair_off + first * nhap + h0};
                }
            }
        }
        uint64_t ho = p.hap_byte_off;
        uint32_t wo = p.hap_word_off;
        const bool pfx = s.float_variant == Variant::F32_PFX;
        for (int q = 0; q < rv.n_haps; ++q) {
            const phmm_hap &hp = rv.haps[pfx ? p.hap_order[q] : (uint32_t)q];
            if (pfx) s.h_happlan.p[p.hap_list_off + q] = p.hap_plan[q];
            if (hp.len) std::memcpy(s.h_hapbytes.p + ho, hp.bases, hp.len);
            s.h_hapoff.p[p.first_hap + q] = ho;
            s.h_haps.p[p.first_hap + q] = DevHap{wo, hp.len};
            s.h_haplist.p[p.hap_list_off + q] = p.first_hap + (uint32_t)q;
            ho += (uint64_t)hp.len; wo += hap_words(hp.len);
        }
        copy_stream_fence();
    });
    // A chunk of few units (a handful of waves of warps) ends when its slowest warp does: hand the long units out first
    // (LPT). Large chunks keep the packing order -- their tail is a fraction of a per cent and sorting a million units is not free.
    // (up to ~14 waves of resident warps: beyond that the tail is below one per cent, and the sort -- serial, 3 ms for the
    // 42 000 units of a full chunk -- was most of what the copy phase of a large chunk cost)
    if (n_units > (uint32_t)h->sm_count * 8u && n_units <= 16384u) {      // more than one wave of resident warps
        std::vector<std::pair<uint32_t, DevUnit>> tmp;
        for (int b = 0; b < kNumBins; ++b) {
            const uint32_t n = s.bin_start[b + 1] - s.bin_start[b];
            if (n < 2) continue;
            DevUnit *u = s.h_units.p + s.bin_start[b];
            tmp.resize(n);
            bool varied = false;
            for (uint32_t k = 0; k < n; ++k) {
                uint32_t cost = 0;
                for (uint32_t q = 0, nh = u[k].shape & 0xffu; q < nh; ++q) cost += hap_tiles(s.h_haps.p[s.h_haplist.p[u[k].hap_list + q]].len);
                tmp[k] = {cost, u[k]};
                varied |= cost != tmp[0].first;
            }
            if (!varied) continue;
            std::stable_sort(tmp.begin(), tmp.end(), [](const auto &x, const auto &y) { return x.first > y.first; });
            for (uint32_t k = 0; k < n; ++k) u[k] = tmp[k].second;
        }
    }
    h->stats.host_ms_pack += ms_since(t0);
    if (trace_on())
        std::fprintf(stderr, "[trace d%d] pack %u reads: plan %.1f | bin %.1f | prefix+reserve %.1f | copy %.1f ms (%d threads)\n", h->device, s.n_uniq,
                     ms_plan, ms_bin - ms_plan, ms_reserve - ms_bin, ms_since(t0) - ms_reserve, threads);
    return PAIRHMM_OK;
}

// H2D of a packed chunk + haplotype word expansion. Asynchronous on the slot's stream.
int upload_chunk(pairhmm_cuda_handle *h, Slot &s) {
    cudaStream_t st = s.stream;
    // descriptors, haplotype bytes, work units, counters, side tables and the read planes: one copy of the arena
    if (s.meta_bytes) CU(h, cudaMemcpyAsync(s.d_meta.p, s.h_meta.p, s.meta_bytes, cudaMemcpyHostToDevice, st));
    if (s.n_uniq && s.raw_on) {
        // raw reads: bases + base qualities (+ insertion / deletion qualities if any read has them) cross PCIe; the default
        // indel qualities (ReadUtils.h:40) and the constant gap-continuation plane are written here
        if (!s.raw_indel_q) CU(h, cudaMemsetAsync(s.d_rd.p + 2 * s.plane, 45, 2 * s.plane, st));
        CU(h, cudaMemsetAsync(s.d_rd.p + 4 * s.plane, s.model.constant_gcp & 0xff, s.plane, st));
        PrepReadsArgs pa{};
        pa.rd = s.d_rd.p; pa.plane = s.plane; pa.reads = s.d_reads.p; pa.mapq = s.d_mapq.p; pa.n = s.n_uniq;
        pa.pcr_on = s.model.pcr_rate_factor > 0; pa.base_q_threshold = s.model.base_quality_threshold;
        for (int i = 0; i <= 20 && pa.pcr_on; ++i) pa.pcr_table[i] = pcr_adjusted_qual(i, s.model.pcr_rate_factor);
        CU(h, launch_prep_reads(pa, h->sm_count, st));
        h->stats.kernel_launches += 1;
    }
    CU(h, launch_prep_haps(s.d_hapbytes.p, s.d_hapoff.p, s.d_haps.p, s.n_haps, s.d_hapw.p, st));
    h->stats.kernel_launches += 1;
    h->stats.h2d_bytes += s.meta_bytes;
    return PAIRHMM_OK;
}

// FP32 pass per (W,K,KL) bin -> finalize -> FP64 rescue per W=32 bin. Asynchronous on the slot's stream.
static int blocks_for(pairhmm_cuda_handle *h, Variant v, int bin) {
    int &n = h->blocks_per_sm[(int)v][bin];
    if (n == 0) { int W, K, KL; bin_shape(bin, W, K, KL); n = std::max(1, forward_blocks_per_sm(v, W, K, KL)); }
    if (h->cfg.reserved[3] > 0) return std::min(n, h->cfg.reserved[3]);   // tuning: cap the resident CTAs per SM
    return n;
}

int run_kernels(pairhmm_cuda_handle *h, Slot &s, bool replay) {
    cudaStream_t st = s.stream;
    // the counter block (zeroed work queues, rescue list bases) came with the arena; a resident replay sends it again
    if (replay) CU(h, cudaMemcpyAsync(s.d_counters.p, s.h_counters.p, sizeof(uint32_t) * kCounterWords, cudaMemcpyHostToDevice, st));
    static const bool force_timing = std::getenv("PAIRHMM_CUDA_TIMING") != nullptr;
    s.timed = !h->light_call || force_timing || trace_on();
    // The two slots' kernels must not share the GPU: run side by side they finish together, both slots come free at the
    // same moment and the GPU idles while the host packs. Chained on the previous chunk's end event, chunk i is done
    // (and its slot reusable) while chunk i+1 still computes; the copies in front of this point are not held back.
    if (h->prev_end && h->prev_end != s.ev_end) CU(h, cudaStreamWaitEvent(st, h->prev_end, 0));
    if (s.timed) CU(h, cudaEventRecord(s.ev_begin, st));
    const HostTables &t = host_tables();
    KernelArgs a{};
    a.rd = s.d_rd.p; a.plane = s.plane; a.reads = s.d_reads.p; a.hapw = s.d_hapw.p; a.haps = s.d_haps.p;
    a.hap_list = s.d_haplist.p; a.hap_plan = s.d_happlan.p; a.fres = s.d_fres.p; a.out = s.d_out.p; a.flag = s.d_flag.p;
    a.log10_init_d = h->log10_init_d;
    // uniform gap-continuation constants, computed exactly as the per-row path does (ph2pr[c], 1 - ph2pr[c])
    const int gc = s.gcp_value & 127;
    a.u_xy_f = t.ph2pr_f[gc]; a.u_gapm_f = 1.0f - t.ph2pr_f[gc];
    std::memcpy(&a.u_xy_bits, &a.u_xy_f, sizeof a.u_xy_bits);
    a.u_xy_d = t.ph2pr_d[gc]; a.u_gapm_d = 1.0 - t.ph2pr_d[gc];
    const bool all_double = h->cfg.use_double != 0;
    // FP64 pass: the wide kernel when the chunk's gap-continuation penalty is uniform (what lets the FP32 pass use its
    // packed or scalar uniform kernel), else the general scalar one
    const Variant vf = s.float_variant;
    const Variant vd = (s.uniform_gcp && !h->cfg.reserved[2] && !h->cfg.reserved[0] && !h->cfg.exact_fp32) ? Variant::F64_PK : Variant::F64_GEN;
    auto grid_for = [&](Variant v, int bin, uint64_t work_items) {
        const uint64_t full = (uint64_t)h->sm_count * (uint64_t)blocks_for(h, v, bin);
        return (int)std::min<uint64_t>(full, (work_items + 3) / 4);
    };
    if (all_double && s.timed) CU(h, cudaEventRecord(s.ev_fwd_done, st));
    if (!all_double) {
        a.ph2pr = h->d_ph2pr_f; a.mm = h->d_mm_f;
        for (int b = 0; b < kNumBins; ++b) {
            const uint32_t n = s.bin_start[b + 1] - s.bin_start[b];
            if (!n) continue;
            int W, K, KL; bin_shape(b, W, K, KL);
            a.units = s.d_units.p + s.bin_start[b]; a.n_units = n; a.n_units_dev = nullptr;
            a.queue = s.d_counters.p + kQueueF + b;
            CU(h, launch_forward(vf, W, K, KL, a, grid_for(vf, b, n), st));
            h->stats.kernel_launches += 1;
        }
        if (s.timed) CU(h, cudaEventRecord(s.ev_fwd_done, st));
        FinalizeArgs f{s.d_fres.p, s.d_out.p, s.d_flag.p, s.n_upairs, h->log10_init_f, s.d_counters.p + kRescuedPairs};
        // a chunk of fewer pairs than would ever make the scan group them (max_pairs = rescued / (4 x resident FP64 warps) < 2):
        // finalize and scan in one launch (a region per call: one launch of five less)
        static const bool no_fuse = std::getenv("PAIRHMM_CUDA_NO_FUSED_SCAN") != nullptr;
        const bool fused_scan = !no_fuse && !s.pair_mode && h->cfg.reserved[6] <= 0 && s.n_upairs < 8u * (uint32_t)h->sm_count * 12u;
        if (!fused_scan) {
            CU(h, launch_finalize(f, st));
            if (s.n_upairs) h->stats.kernel_launches += 1;
        }
        RescueScanArgs r{s.d_units.p, s.bin_start[kNumBins], s.d_reads.p, s.d_fres.p, s.d_rescue.p,
                         s.d_counters.p + kRescueBaseDev, s.d_counters.p + kRescueCount,
                         // 0 = chosen on the device from the number of underflowed pairs finalize_kernel counted
                         h->cfg.reserved[6] > 0 ? (uint32_t)h->cfg.reserved[6] : 0u, s.d_counters.p + kRescuedPairs,
                         (uint32_t)h->sm_count * 12u};
        r.pair_mode = s.pair_mode;
        r.pending = s.d_rescue.p + s.pending_off; r.pending_count = s.d_counters.p + kPendingCount;
        r.pair_table = s.d_pair_table.p;
        r.rescue_base16 = s.d_counters.p + kRescueBase16Dev; r.rescue_count16 = s.d_counters.p + kRescueCount16;
        r.fp64_warp_slots16 = (uint32_t)h->sm_count * 8u;
        r.n_pairs = s.n_upairs;
        if (fused_scan) CU(h, launch_finalize_scan_small(r, f, st));
        else CU(h, launch_rescue_scan(r, st));
        if (r.n_units) h->stats.kernel_launches += 1;
        if (r.n_units && r.pair_mode) {
            CU(h, launch_rescue_flush(r, st));
            h->stats.kernel_launches += 1;
        }
    }
    if (s.timed) CU(h, cudaEventRecord(s.ev_float_done, st));
    a.ph2pr = h->d_ph2pr_d; a.mm = h->d_mm_d;
    a.rescue_filter = all_double ? 0u : 1u;
    // Units of two reads first (the W = 16 instance of the wide FP64 kernel), the units of one read beside them on the slot's
    // second stream: both are persistent grids that leave when their list is empty, so the second launch moves into the SMs
    // as the first one's warps run dry -- in series each launch would end in its own tail.
    cudaStream_t st32 = st;
    if (!all_double && s.pair_mode) {
        CU(h, cudaEventRecord(s.ev_fork, st));
        for (int b = 0; b < kNumBinsW16; ++b) {
            const uint64_t cap = (uint64_t)s.rescue_base16[b + 1] - s.rescue_base16[b];
            if (!cap) continue;
            int W, K, KL; bin_shape(kBinBase[1] + b, W, K, KL);
            a.queue = s.d_counters.p + kQueueD16 + b;
            a.units = s.d_rescue.p + s.rescue_base16[b]; a.n_units = 0;
            a.n_units_dev = s.d_counters.p + kRescueCount16 + b;
            CU(h, launch_forward(Variant::F64_PK, W, K, KL, a, grid_for(Variant::F64_PK, kBinBase[1] + b, cap), st));
            h->stats.kernel_launches += 1;
        }
        st32 = s.stream2;
        CU(h, cudaStreamWaitEvent(st32, s.ev_fork, 0));
    }
    for (int b = 0; b < (all_double ? kNumBins : kNumBinsW32); ++b) {
        int W, K, KL; bin_shape(b, W, K, KL);
        // use_double: the FP32 pass' queues are free, one per bin of any width
        a.queue = s.d_counters.p + (all_double ? kQueueF : kQueueD) + b;
        int grid;
        if (all_double) {      // every unit straight to the FP64 kernels (the chunk was binned for them: choose_width)
            const uint32_t n = s.bin_start[b + 1] - s.bin_start[b];
            if (!n) continue;
            a.units = s.d_units.p + s.bin_start[b]; a.n_units = n; a.n_units_dev = nullptr;
            grid = grid_for(vd, b, n);
        } else {
            // the rescue count is only known on the device: a persistent grid bounded by the bin's pairs
            const uint64_t cap = (uint64_t)s.rescue_base[b + 1] - s.rescue_base[b];
            if (!cap) continue;
            a.units = s.d_rescue.p + s.rescue_base[b]; a.n_units = 0;
            a.n_units_dev = s.d_counters.p + kRescueCount + b;
            grid = grid_for(vd, b, cap);
        }
        CU(h, launch_forward(vd, W, K, KL, a, grid, st32));
        h->stats.kernel_launches += 1;
    }
    if (st32 != st) {
        CU(h, cudaEventRecord(s.ev_join, st32));
        CU(h, cudaStreamWaitEvent(st, s.ev_join, 0));
    }
    if (const uint32_t n_long = s.bin_start[kNumBins + 1] - s.bin_start[kNumBins]) {
        // reads longer than 384 bases: row stripes, one warp per pair, float + threshold + double + log10 in one kernel;
        // it runs after finalize, whose output for these entries (computed from unwritten FP32 sums) it replaces
        LongArgs l{};
        l.rd = s.d_rd.p; l.plane = s.plane; l.reads = s.d_reads.p; l.hap_bytes = s.d_hapbytes.p; l.hap_byte_off = s.d_hapoff.p;
        l.haps = s.d_haps.p; l.hap_list = s.d_haplist.p; l.units = s.d_units.p + s.bin_start[kNumBins]; l.n_units = n_long;
        l.queue = s.d_counters.p + kQueueLong;
        l.ph2pr_f = h->d_ph2pr_f; l.mm_f = h->d_mm_f; l.ph2pr_d = h->d_ph2pr_d; l.mm_d = h->d_mm_d;
        l.log10_init_f = h->log10_init_f; l.log10_init_d = h->log10_init_d;
        l.out = s.d_out.p; l.flag = s.d_flag.p; l.use_double = all_double ? 1 : 0;
        const int grid = (int)std::min<uint64_t>((uint64_t)h->sm_count * 4u, ((uint64_t)n_long + 3) / 4);
        l.scratch_per_warp = ((uint64_t)(s.hap_max + 1) * 3 * sizeof(double) + 255) & ~255ull;
        cudaError_t e = s.d_long_scratch.reserve((size_t)l.scratch_per_warp * 4u * (size_t)grid);
        if (e != cudaSuccess) return fail_cuda(h, e, "reserve d_long_scratch");
        l.scratch = s.d_long_scratch.p;
        CU(h, launch_forward_long(l, grid, st));
        h->stats.kernel_launches += 1;
    }
    if (s.ep_on && s.n_uniq) {       // fused per-read epilogue on the finished matrix
        EpilogueArgs e{s.d_out.p, s.d_rows.p, s.d_reads.p, s.d_summ.p, s.n_uniq, s.ep.likelihood_cap, s.ep.max_error_per_base,
                       std::log(10.0), std::isinf(s.ep.likelihood_cap) ? 0 : 1, s.ep.natural_log ? 1 : 0};
        CU(h, launch_epilogue(e, st));
        h->stats.kernel_launches += 1;
    }
    CU(h, cudaEventRecord(s.ev_end, st));
    h->prev_end = s.ev_end;
    return PAIRHMM_OK;
}

}  // namespace phmm

namespace phmm {

// D2H of results; asynchronous.
int download_chunk(pairhmm_cuda_handle *h, Slot &s) {
    if (s.n_upairs) {
        // log10 values and rescued flags lie in one buffer: one copy (the few bytes between them included)
        CU(h, cudaMemcpyAsync(s.h_res.p, s.d_res.p, s.res_bytes - 1, cudaMemcpyDeviceToHost, s.stream));
        h->stats.d2h_bytes += (sizeof(double) + 1) * (uint64_t)s.n_upairs;
    }
    if (s.ep_on && s.n_uniq) {
        CU(h, cudaMemcpyAsync(s.h_summ.p, s.d_summ.p, sizeof(phmm_read_summary) * s.n_uniq, cudaMemcpyDeviceToHost, s.stream));
        h->stats.d2h_bytes += sizeof(phmm_read_summary) * (uint64_t)s.n_uniq;
    }
    return PAIRHMM_OK;
}

// PAIRHMM_CUDA_TRACE: what the scan made of the chunk's underflowed pairs (the stream is idle)
static void trace_rescue_counts(pairhmm_cuda_handle *h, Slot &s) {
    std::vector<uint32_t> c(kCounterWords);
    if (cudaMemcpy(c.data(), s.d_counters.p, sizeof(uint32_t) * kCounterWords, cudaMemcpyDeviceToHost) != cudaSuccess) return;
    uint64_t u32 = 0, u16 = 0;
    for (int b = 0; b < kNumBinsW32; ++b) u32 += c[kRescueCount + b];
    for (int b = 0; b < kNumBinsW16; ++b) u16 += c[kRescueCount16 + b];
    std::fprintf(stderr, "[trace d%d] rescue: %u pairs underflowed of %u; units of one read %llu, of two reads %llu; candidates %u (mode %u)\n",
                 h->device, c[kRescuedPairs], s.n_upairs, (unsigned long long)u32, (unsigned long long)u16, c[kPendingCount], s.pair_mode);
}

// Waits for the slot's stream, then maps unique (read, hap) results back to every input read
// (VectorLoglessPairHMM.cpp:123-125: read-major, duplicates share their unique read's row).
int finish_chunk(pairhmm_cuda_handle *h, Slot &s, int threads) {
    CU(h, cudaStreamSynchronize(s.stream));
    if (h->prev_end == s.ev_end) h->prev_end = nullptr;      // retired: the next chunk has nothing to wait for
    float ms_f = 0.f, ms_all = 0.f;
    if (s.timed) {
        cudaEventElapsedTime(&ms_f, s.ev_begin, s.ev_float_done);
        cudaEventElapsedTime(&ms_all, s.ev_begin, s.ev_end);
    }
    if (trace_on()) trace_rescue_counts(h, s);
    if (trace_on() && g_trace_base) {
        float b0 = 0.f, b1 = 0.f;
        cudaEventElapsedTime(&b0, g_trace_base, s.ev_begin); cudaEventElapsedTime(&b1, g_trace_base, s.ev_end);
        std::fprintf(stderr, "[trace d%d] chunk done: host %.1f ms | gpu kernels %.1f .. %.1f ms (%.1f) | cells %.3g\n",
                     h->device, ms_since(g_trace_t0), b0, b1, b1 - b0, (double)s.cells);
    }
    h->stats.gpu_ms_float += ms_f; h->stats.gpu_ms_double += ms_all - ms_f; h->stats.gpu_ms_total += ms_all;
    auto t0 = clk::now();
    std::atomic<uint64_t> rescued{0}, uniq_pairs{0};
    const bool stream_out = s.n_upairs >= (1u << 18);
    parallel_for(s.regions.size(), threads, 16, [&](size_t i) {
        const RegionView &rv = s.regions[i];
        const RegionPlan &p = s.plans[i];
        const size_t nh = (size_t)rv.n_haps;
        uint64_t resc = 0;
        // every read unique, none empty, none moved by the binning: the region's block of the chunk's matrix IS the caller's
        // matrix -- one copy instead of one per read (streaming stores for a large chunk: the caller reads it much later)
        const bool whole = p.identity && p.hap_order.empty() && rv.n_reads > 0 && nh > 0;
        if (whole) {
            const size_t n = (size_t)rv.n_reads * nh;
            if (stream_out) {
                copy_stream(rv.out, s.h_out.p + p.upair_off, n * sizeof(double));
                if (rv.resc) copy_stream(rv.resc, s.h_flag.p + p.upair_off, n);
                copy_stream_fence();
            } else {
                std::memcpy(rv.out, s.h_out.p + p.upair_off, n * sizeof(double));
                if (rv.resc) std::memcpy(rv.resc, s.h_flag.p + p.upair_off, n);
            }
            if (s.ep_on && rv.summ) std::memcpy(rv.summ, s.h_summ.p + p.first_uniq, sizeof(phmm_read_summary) * (size_t)rv.n_reads);
        }
        for (int r = 0; r < rv.n_reads && !whole; ++r) {
            const uint32_t u = p.uniq_of_read[r];
            double *o = rv.out + (size_t)r * nh;
            uint8_t *f = rv.resc ? rv.resc + (size_t)r * nh : nullptr;
            if (u == kNoRead) {   // zero-length read: no alignment exists
                for (size_t q = 0; q < nh; ++q) o[q] = -std::numeric_limits<double>::infinity();
                if (f) std::memset(f, 1, nh);
                if (s.ep_on && rv.summ) {
                    const double ninf = -std::numeric_limits<double>::infinity();
                    rv.summ[r] = phmm_read_summary{ninf, ninf, nh ? 0 : -1, 1, {0, 0, 0}};
                }
                continue;
            }
            if (s.ep_on && rv.summ) rv.summ[r] = s.h_summ.p[p.first_uniq + u];
            const size_t src = (size_t)p.upair_off + (size_t)u * nh;
            if (p.hap_order.empty()) {
                std::memcpy(o, s.h_out.p + src, nh * sizeof(double));
                if (f) std::memcpy(f, s.h_flag.p + src, nh);
            } else {                   // prefix sharing: the chunk holds the haplotypes in sorted order
                for (size_t q = 0; q < nh; ++q) o[p.hap_order[q]] = s.h_out.p[src + q];
                if (f) for (size_t q = 0; q < nh; ++q) f[p.hap_order[q]] = s.h_flag.p[src + q];
            }
        }
        for (size_t q = 0; q < (size_t)p.n_uniq * nh; ++q) resc += s.h_flag.p[p.upair_off + q];
        rescued += resc; uniq_pairs += (uint64_t)p.n_uniq * nh;
    });
    h->stats.host_ms_scatter += ms_since(t0);
    h->stats.regions += s.regions.size(); h->stats.reads += s.reads_in; h->stats.unique_reads += s.n_uniq;
    h->stats.pairs += s.pairs; h->stats.unique_pairs += uniq_pairs.load(); h->stats.cells += s.cells;
    h->stats.cells_swept += s.cells_swept;
    h->stats.rescued_pairs += rescued.load();
    set_busy(h, s, false);
    return PAIRHMM_OK;
}

static std::atomic<int> g_chunks_in_flight[64];
void set_busy(pairhmm_cuda_handle *h, Slot &s, bool busy) {
    s.busy = busy;
    if (busy != s.counted) { s.counted = busy; g_chunks_in_flight[h->device & 63].fetch_add(busy ? 1 : -1, std::memory_order_relaxed); }
}
int device_chunks_in_flight(int device) { return g_chunks_in_flight[device & 63].load(std::memory_order_relaxed); }

int pack_threads(const pairhmm_cuda_handle *h) {
    if (h->cfg.pack_threads > 0) return h->cfg.pack_threads;
    unsigned hc = std::thread::hardware_concurrency();
    return (int)std::max(1u, std::min(hc ? hc : 1u, 16u));
}

int submit_chunk(pairhmm_cuda_handle *h, Slot &s, const RegionView *regs, size_t n, int threads) {
    h->resident_loaded = false;      // the slots are re-packed: a batch staged by _resident_load is gone
    int st = pack_chunk(h, s, regs, n, threads);
    if (st != PAIRHMM_OK) return st;
    set_busy(h, s, true);
    if ((st = upload_chunk(h, s)) != PAIRHMM_OK) return st;
    if ((st = run_kernels(h, s, false)) != PAIRHMM_OK) return st;
    return download_chunk(h, s);
}

}  // namespace phmm

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

int pairhmm_cuda_abi_version(void) { return PAIRHMM_CUDA_ABI_VERSION; }

int pairhmm_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int d = 0; d < n; ++d) {
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, d) == cudaSuccess && p.major == 10) ++ok;
    }
    return ok;
}

void pairhmm_cuda_default_config(pairhmm_cuda_config *cfg) {
    if (!cfg) return;
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->device = -1; cfg->use_double = 0; cfg->exact_fp32 = 0; cfg->dedup_reads = 1; cfg->pack_threads = 0;
    cfg->chunk_cells = 0; cfg->device_mask = 0; cfg->batch_window_us = 0;
}

int pairhmm_cuda_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d) {
    const HostTables &t = host_tables();
    if (ph2pr_f) std::memcpy(ph2pr_f, t.ph2pr_f, sizeof t.ph2pr_f);
    if (mm_f) std::memcpy(mm_f, t.mm_f, sizeof t.mm_f);
    if (ph2pr_d) std::memcpy(ph2pr_d, t.ph2pr_d, sizeof t.ph2pr_d);
    if (mm_d) std::memcpy(mm_d, t.mm_d, sizeof t.mm_d);
    return PAIRHMM_OK;
}

int pairhmm_cuda_prefix_plan(const phmm_hap *haps, int n_haps, int unit_haps, uint32_t *order, uint32_t *t_start,
                             int32_t *restore_slot, uint32_t *t_save, uint32_t *save_slot, uint64_t *shared_cols) {
    if (n_haps < 0 || (n_haps > 0 && (!haps || !order || !t_start || !restore_slot || !t_save || !save_slot))) return PAIRHMM_ERR_ARG;
    int32_t hmax = 0;
    for (int i = 0; i < n_haps; ++i) {
        if (haps[i].len < 0 || (haps[i].len > 0 && !haps[i].bases)) return PAIRHMM_ERR_ARG;
        hmax = std::max(hmax, haps[i].len);
    }
    RegionView rv{};
    rv.haps = haps; rv.n_haps = n_haps;
    RegionPlan pl;
    pl.group = unit_haps > 0 ? balanced_unit_haps(n_haps, (uint32_t)std::min(unit_haps, (int)kMaxUnitHaps)) : default_unit_haps(n_haps, hmax);
    plan_prefix(rv, pl);
    for (int q = 0; q < n_haps; ++q) {
        const uint2 w = pl.hap_plan[q];
        order[q] = pl.hap_order[q];
        t_start[q] = w.x & 0xffffu; t_save[q] = w.x >> 16;
        restore_slot[q] = (int32_t)(w.y & 0xffu) - 1; save_slot[q] = (w.y >> 8) & 0xffu;
    }
    if (shared_cols) *shared_cols = pl.shared_cols;
    return PAIRHMM_OK;
}

int pairhmm_cuda_create(const pairhmm_cuda_config *cfg, pairhmm_cuda_handle **out) {
    if (!out) return PAIRHMM_ERR_ARG;
    *out = nullptr;
    pairhmm_cuda_config c;
    if (cfg) c = *cfg; else pairhmm_cuda_default_config(&c);
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return PAIRHMM_ERR_NO_DEVICE; }
    if (c.device_mask) {          // multi-device handle: one child handle + one host worker per device (multi.cu)
        auto *m = new (std::nothrow) pairhmm_cuda_handle();
        if (!m) return PAIRHMM_ERR_OOM;
        m->cfg = c;
        std::memset(&m->stats, 0, sizeof m->stats);
        const int st = multi_create(m);
        if (st != PAIRHMM_OK) { pairhmm_cuda_destroy(m); return st; }
        *out = m;
        return PAIRHMM_OK;
    }
    int dev = c.device;
    if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) return PAIRHMM_ERR_NO_DEVICE; }
    if (dev >= n) return PAIRHMM_ERR_ARG;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return PAIRHMM_ERR_NO_DEVICE;
    if (prop.major != 10) return PAIRHMM_ERR_NO_DEVICE;   // kernels are built for sm_100a only
    auto *h = new (std::nothrow) pairhmm_cuda_handle();
    if (!h) return PAIRHMM_ERR_OOM;
    h->cfg = c; h->device = dev; h->sm_count = prop.multiProcessorCount;
    std::memset(&h->stats, 0, sizeof h->stats);
    auto bail = [&](int st) { pairhmm_cuda_destroy(h); return st; };
    if (cudaSetDevice(dev) != cudaSuccess) return bail(PAIRHMM_ERR_CUDA);
    const HostTables &t = host_tables();
    h->log10_init_f = log10f(ldexpf(1.f, 120));     // Context.h:184
    h->log10_init_d = log10(ldexp(1.0, 1020));      // Context.h:143
#define UP(dst, src) \
    if (cudaMalloc((void **)&h->dst, sizeof t.src) != cudaSuccess || \
        cudaMemcpy(h->dst, t.src, sizeof t.src, cudaMemcpyHostToDevice) != cudaSuccess) return bail(PAIRHMM_ERR_CUDA)
    UP(d_ph2pr_f, ph2pr_f); UP(d_mm_f, mm_f); UP(d_ph2pr_d, ph2pr_d); UP(d_mm_d, mm_d);
#undef UP
    for (auto &s : h->slots) {
        if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&s.stream2, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&s.ev_fork, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&s.ev_join, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreate(&s.ev_begin) != cudaSuccess || cudaEventCreate(&s.ev_float_done) != cudaSuccess ||
            cudaEventCreate(&s.ev_fwd_done) != cudaSuccess ||
            cudaEventCreate(&s.ev_end) != cudaSuccess)
            return bail(PAIRHMM_ERR_CUDA);
    }
    if (forward_blocks_per_sm(Variant::F32_GEN, 32, 1, 0) < 1) return bail(PAIRHMM_ERR_CUDA);   // kernels loadable?
    if (c.batch_window_us > 0) h->batcher = batcher_acquire(h);      // process-wide, per device (batcher.cu)
    *out = h;
    return PAIRHMM_OK;
}

int pairhmm_cuda_destroy(pairhmm_cuda_handle *h) {
    if (!h) return PAIRHMM_OK;
    if (h->is_multi() || h->cfg.device_mask) { multi_destroy(h); delete h; return PAIRHMM_OK; }
    if (h->batcher) { batcher_release(h->batcher); h->batcher = nullptr; }
    cudaSetDevice(h->device);
    for (auto &s : h->slots) { if (s.stream) cudaStreamSynchronize(s.stream); set_busy(h, s, false); s.release(); }
    if (h->d_ph2pr_f) cudaFree(h->d_ph2pr_f);
    if (h->d_mm_f) cudaFree(h->d_mm_f);
    if (h->d_ph2pr_d) cudaFree(h->d_ph2pr_d);
    if (h->d_mm_d) cudaFree(h->d_mm_d);
    delete h;
    return PAIRHMM_OK;
}

int pairhmm_cuda_initialize(pairhmm_cuda_handle *h, int read_max_len, int hap_max_len) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (read_max_len <= 0 || hap_max_len <= 0)   // PairHMM.cpp:16 asserts haplotypeMaxLength > 0
        return fail(h, PAIRHMM_ERR_ARG, "initialize: maximum lengths must be positive");
    h->read_max_len = read_max_len; h->hap_max_len = hap_max_len;
    for (auto *c : h->children) { c->read_max_len = read_max_len; c->hap_max_len = hap_max_len; }
    return PAIRHMM_OK;
}

const char *pairhmm_cuda_last_error(pairhmm_cuda_handle *h) { return h ? h->last_error.c_str() : "null handle"; }

int pairhmm_cuda_stats_get(pairhmm_cuda_handle *h, pairhmm_cuda_stats *out) {
    if (!h || !out) return PAIRHMM_ERR_ARG;
    if (h->is_multi()) { multi_stats(h, out); return PAIRHMM_OK; }
    *out = h->stats;
    return PAIRHMM_OK;
}
int pairhmm_cuda_stats_reset(pairhmm_cuda_handle *h) {
    if (!h) return PAIRHMM_ERR_ARG;
    std::memset(&h->stats, 0, sizeof h->stats);
    for (auto *c : h->children) std::memset(&c->stats, 0, sizeof c->stats);
    return PAIRHMM_OK;
}

int pairhmm_cuda_submit(pairhmm_cuda_handle *h, const phmm_read *reads, int n_reads, const phmm_hap *haps,
                        int n_haps, double *log10_out, uint8_t *rescued_out, pairhmm_cuda_ticket *ticket) {
    if (!h || !ticket) return PAIRHMM_ERR_ARG;
    if (h->is_multi())
        return multi_submit(h, RegionView{reads, nullptr, n_reads, haps, n_haps, log10_out, rescued_out, nullptr}, ticket);
    auto t0 = clk::now();
    CU(h, cudaSetDevice(h->device));
    int si = !h->slots[0].busy ? 0 : (!h->slots[1].busy ? 1 : -1);
    if (si < 0) {   // both in flight: retire the older one first
        si = h->slot_ticket[0] < h->slot_ticket[1] ? 0 : 1;
        int st = finish_chunk(h, h->slots[si], 1);
        if (st != PAIRHMM_OK) return st;
        h->slot_ticket[si] = 0;
    }
    RegionView rv{reads, nullptr, n_reads, haps, n_haps, log10_out, rescued_out, nullptr};
    h->light_call = true;
    int st = submit_chunk(h, h->slots[si], &rv, 1, 1);
    h->light_call = false;
    if (st != PAIRHMM_OK) { set_busy(h, h->slots[si], false); return st; }
    h->slot_ticket[si] = *ticket = h->next_ticket++;
    h->stats.calls += 1;
    h->stats.host_ms_total += ms_since(t0);
    return PAIRHMM_OK;
}

int pairhmm_cuda_wait(pairhmm_cuda_handle *h, pairhmm_cuda_ticket ticket) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (h->is_multi()) return multi_wait(h, ticket);
    auto t0 = clk::now();
    if (ticket <= 0 || ticket >= h->next_ticket) return fail(h, PAIRHMM_ERR_STATE, "wait: unknown ticket");
    for (int si = 0; si < 2; ++si)
        if (h->slot_ticket[si] == ticket) {
            CU(h, cudaSetDevice(h->device));
            int st = finish_chunk(h, h->slots[si], 1);
            h->slot_ticket[si] = 0;
            h->stats.host_ms_total += ms_since(t0);
            return st;
        }
    return PAIRHMM_OK;   // already retired by a later submit
}

int pairhmm_cuda_compute(pairhmm_cuda_handle *h, const phmm_read *reads, int n_reads, const phmm_hap *haps,
                         int n_haps, double *log10_out, uint8_t *rescued_out) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (n_reads == 0) return PAIRHMM_OK;   // VectorLoglessPairHMM.cpp:46-47
    if (h->batcher)     // merged with the regions other host threads submit meanwhile (config.batch_window_us)
        return batcher_compute(h, RegionView{reads, nullptr, n_reads, haps, n_haps, log10_out, rescued_out, nullptr});
    if (h->is_multi() && (int64_t)n_reads * n_haps >= 32768) {      // a region worth several devices: cut by reads
        const RegionView rv{reads, nullptr, n_reads, haps, n_haps, log10_out, rescued_out, nullptr};
        return multi_compute_batch(h, &rv, 1);
    }
    pairhmm_cuda_ticket t;
    int st = pairhmm_cuda_submit(h, reads, n_reads, haps, n_haps, log10_out, rescued_out, &t);
    if (st != PAIRHMM_OK) return st;
    return pairhmm_cuda_wait(h, t);
}

}  // extern "C"

// the batch pipeline over regions of either read kind (RegionView::reads or ::raw)
int phmm::compute_batch_impl(pairhmm_cuda_handle *h, const RegionView *regions, int64_t n_regions,
                             const std::function<void(int64_t)> *ensure) {
    if (!h || n_regions < 0 || (n_regions > 0 && !regions)) return PAIRHMM_ERR_ARG;
    if (h->is_multi()) return multi_compute_batch(h, regions, n_regions);
    auto t0 = clk::now();
    CU(h, cudaSetDevice(h->device));
    for (int si = 0; si < 2; ++si)
        if (h->slots[si].busy) { int st = finish_chunk(h, h->slots[si], 1); if (st) return st; h->slot_ticket[si] = 0; }
    const int threads = pack_threads(h);
    if (trace_on()) {
        if (!g_trace_base) cudaEventCreate(&g_trace_base);
        g_trace_t0 = clk::now();
        cudaEventRecord(g_trace_base, h->slots[0].stream);
    }
    const int64_t chunk_cells = h->cfg.chunk_cells > 0 ? h->cfg.chunk_cells : (int64_t)100000000000ll;
    std::vector<RegionView> views;
    int64_t i = 0;
    int si = 0;
    int status = PAIRHMM_OK;
    int chunk_no = 0;
    int ladder = 0;
    std::vector<uint64_t> sums;
    size_t summed = 0, sum_block = 128;
    while (i < n_regions && status == PAIRHMM_OK) {
        views.clear();
        // The GPU idles while the first chunk is packed, so that one is small (1/16 of a full chunk); the following ones
        // double until the full size is reached: packing chunk i+1 then fits under the kernels of chunk i both when the
        // host packs ~3x faster than the kernels compute (one device, 16 uncontended threads) and when it packs ~1.7x
        // faster (eight devices sharing the host's threads). A FIXED ladder on purpose: sizes that followed the measured
        // packing rate differed from call to call, and the slots' grow-only pinned buffers were re-allocated (tens of ms
        // each) in a handle's second and third large call.
        if (chunk_no > 0 && ladder < 4) ++ladder;
        const int64_t chunk_limit = chunk_cells >> (4 - ladder);
        ++chunk_no;
        int64_t cells = 0;
        uint64_t pairs = 0, rbytes = 0;
        while (i < n_regions) {
            if (ensure) (*ensure)(i);
            const RegionView &r = regions[i];
            if (r.n_reads > 0 && !r.reads && !r.raw) { status = fail(h, PAIRHMM_ERR_ARG, "null read array"); break; }
            if (r.n_haps > 0 && !r.haps) { status = fail(h, PAIRHMM_ERR_ARG, "null haplotype array"); break; }
            uint64_t rl = 0, hl = 0;
            if (r.read_len_sum >= 0 && r.hap_len_sum >= 0) { rl = (uint64_t)r.read_len_sum; hl = (uint64_t)r.hap_len_sum; }
            else {
                // the length sums walk every read descriptor (48 bytes each): block by block on the pool, ahead of the cutter
                if ((size_t)i >= summed) {
                    const size_t b0 = summed, b1 = std::min((size_t)n_regions, summed + sum_block);
                    if (sums.size() < 2 * (size_t)n_regions) sums.resize(2 * (size_t)n_regions);
                    parallel_for(b1 - b0, threads, 32, [&, b0](size_t k) {
                        const RegionView &q = regions[b0 + k];
                        uint64_t a = 0, b = 0;
                        if (q.reads || q.raw) for (int t = 0; t < q.n_reads; ++t) a += (uint64_t)std::max(0, q.raw ? q.raw[t].len : q.reads[t].len);
                        if (q.haps) for (int t = 0; t < q.n_haps; ++t) b += (uint64_t)std::max(0, q.haps[t].len);
                        sums[2 * (b0 + k)] = a; sums[2 * (b0 + k) + 1] = b;
                    });
                    summed = b1; sum_block = std::min<size_t>(sum_block * 2, 4096);
                }
                rl = sums[2 * (size_t)i]; hl = sums[2 * (size_t)i + 1];
            }
            const uint64_t p = (uint64_t)std::max(0, r.n_reads) * (uint64_t)std::max(0, r.n_haps);
            if (!views.empty() && (cells + (int64_t)(rl * hl) > chunk_limit || pairs + p > 0x30000000ull || rbytes + rl > 0xc0000000ull)) break;
            views.push_back(r);
            cells += (int64_t)(rl * hl); pairs += p; rbytes += rl;
            ++i;
        }
        Slot &s = h->slots[si];
        if (trace_on()) std::fprintf(stderr, "[trace d%d] host %.1f ms: chunk of %zu regions cut\n", h->device, ms_since(g_trace_t0), views.size());
        if (s.busy) status = finish_chunk(h, s, threads);
        if (trace_on()) std::fprintf(stderr, "[trace d%d] host %.1f ms: slot free, packing\n", h->device, ms_since(g_trace_t0));
        if (status == PAIRHMM_OK) status = submit_chunk(h, s, views.data(), views.size(), threads);
        if (trace_on()) std::fprintf(stderr, "[trace d%d] host %.1f ms: submitted\n", h->device, ms_since(g_trace_t0));
        si ^= 1;
    }
    for (int k = 0; k < 2; ++k) {
        Slot &s = h->slots[(si + k) & 1];
        if (s.busy) { int st = finish_chunk(h, s, threads); if (status == PAIRHMM_OK) status = st; }
    }
    h->stats.calls += 1;
    h->stats.host_ms_total += ms_since(t0);
    return status;
}

extern "C" {

static std::vector<RegionView> views_of(const phmm_region *regions, int64_t n_regions, phmm_read_summary *const *summaries) {
    std::vector<RegionView> v((size_t)std::max<int64_t>(0, n_regions));
    for (int64_t i = 0; i < n_regions; ++i) {
        const phmm_region &r = regions[i];
        v[i] = RegionView{r.reads, nullptr, r.n_reads, r.haps, r.n_haps, r.log10_out, r.rescued_out, summaries ? summaries[i] : nullptr};
    }
    return v;
}

int pairhmm_cuda_compute_batch(pairhmm_cuda_handle *h, const phmm_region *regions, int64_t n_regions) {
    if (!h || n_regions < 0 || (n_regions > 0 && !regions)) return PAIRHMM_ERR_ARG;
    const std::vector<RegionView> v = views_of(regions, n_regions, nullptr);
    return compute_batch_impl(h, v.data(), n_regions);
}

static int check_epilogue(pairhmm_cuda_handle *h, const pairhmm_cuda_epilogue *ep) {
    // AlleleLikelihoods.h:373: the cap is a log10 difference below the best allele, i.e. negative; -inf switches it off
    if (!ep || std::isnan(ep->likelihood_cap) || !(ep->likelihood_cap < 0.0))
        return fail(h, PAIRHMM_ERR_ARG, "epilogue: likelihood_cap must be negative (or -inf)");
    if (std::isnan(ep->max_error_per_base) || ep->max_error_per_base < 0.0)
        return fail(h, PAIRHMM_ERR_ARG, "epilogue: max_error_per_base must be >= 0");
    return PAIRHMM_OK;
}

int pairhmm_cuda_compute_batch_normalized(pairhmm_cuda_handle *h, const phmm_region *regions, int64_t n_regions,
                                          const pairhmm_cuda_epilogue *ep, phmm_read_summary *const *summary_out) {
    if (!h) return PAIRHMM_ERR_ARG;
    int st = check_epilogue(h, ep);
    if (st != PAIRHMM_OK) return st;
    // chunks of earlier plain submits must not pick the epilogue up: retire them first
    for (int si = 0; si < 2; ++si)
        if (h->slots[si].busy) { CU(h, cudaSetDevice(h->device)); st = finish_chunk(h, h->slots[si], 1); if (st) return st; h->slot_ticket[si] = 0; }
    if (n_regions < 0 || (n_regions > 0 && !regions)) return PAIRHMM_ERR_ARG;
    const std::vector<RegionView> v = views_of(regions, n_regions, summary_out);
    h->ep_on = true; h->ep = *ep;
    st = compute_batch_impl(h, v.data(), n_regions);
    h->ep_on = false;
    return st;
}

int pairhmm_cuda_compute_normalized(pairhmm_cuda_handle *h, const phmm_read *reads, int n_reads, const phmm_hap *haps,
                                    int n_haps, const pairhmm_cuda_epilogue *ep, double *log10_out, uint8_t *rescued_out,
                                    phmm_read_summary *summary_out) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (n_reads == 0) return check_epilogue(h, ep);   // VectorLoglessPairHMM.cpp:46-47
    const phmm_region r{reads, n_reads, haps, n_haps, log10_out, rescued_out};
    phmm_read_summary *const sp = summary_out;
    return pairhmm_cuda_compute_batch_normalized(h, &r, 1, ep, summary_out ? &sp : nullptr);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// flat / pair-list / resident entry points
// ------------------------------------------------------------------------------------------------
namespace {

struct FlatViews {
    std::vector<phmm_read> reads;
    std::vector<phmm_hap> haps;
    std::vector<phmm_region> regions;
};

int build_flat_views(pairhmm_cuda_handle *h, FlatViews &v, const uint8_t *read_bases, const uint8_t *base_q,
                     const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *gcp, const int64_t *read_off,
                     const int32_t *read_len, int64_t n_reads, const uint8_t *hap_bases, const int64_t *hap_off,
                     const int32_t *hap_len, int64_t n_haps, const int64_t *reg_read_start,
                     const int64_t *reg_hap_start, int64_t n_regions, double *log10_out, uint8_t *rescued_out) {
    if (n_reads < 0 || n_haps < 0 || n_regions < 0) return fail(h, PAIRHMM_ERR_ARG, "negative count");
    if (n_reads && (!read_bases || !base_q || !ins_q || !del_q || !gcp || !read_off || !read_len))
        return fail(h, PAIRHMM_ERR_ARG, "null read array");
    if (n_haps && (!hap_bases || !hap_off || !hap_len)) return fail(h, PAIRHMM_ERR_ARG, "null haplotype array");
    if (n_regions && (!reg_read_start || !reg_hap_start)) return fail(h, PAIRHMM_ERR_ARG, "null region array");
    v.reads.resize(n_reads); v.haps.resize(n_haps); v.regions.resize(n_regions);
    parallel_for((size_t)n_reads, pack_threads(h), 65536, [&](size_t i) {
        const int64_t o = read_off[i];
        v.reads[i] = phmm_read{read_bases + o, base_q + o, ins_q + o, del_q + o, gcp + o, read_len[i]};
    });
    for (int64_t i = 0; i < n_haps; ++i) v.haps[i] = phmm_hap{hap_bases + hap_off[i], hap_len[i]};
    int64_t out_off = 0;
    for (int64_t r = 0; r < n_regions; ++r) {
        const int64_t r0 = reg_read_start[r], r1 = reg_read_start[r + 1], h0 = reg_hap_start[r], h1 = reg_hap_start[r + 1];
        if (r0 < 0 || r1 < r0 || r1 > n_reads || h0 < 0 || h1 < h0 || h1 > n_haps || r1 - r0 > 0x7fffffff || h1 - h0 > 0x7fffffff)
            return fail(h, PAIRHMM_ERR_ARG, "region bounds out of range");
        v.regions[r] = phmm_region{v.reads.data() + r0, (int32_t)(r1 - r0), v.haps.data() + h0, (int32_t)(h1 - h0),
                                   log10_out ? log10_out + out_off : nullptr, rescued_out ? rescued_out + out_off : nullptr};
        out_off += (r1 - r0) * (h1 - h0);
    }
    return PAIRHMM_OK;
}

}  // namespace

extern "C" {

int pairhmm_cuda_compute_flat(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                              const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *gcp,
                              const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                              const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len,
                              int64_t n_haps, const int64_t *reg_read_start, const int64_t *reg_hap_start,
                              int64_t n_regions, double *log10_out, uint8_t *rescued_out) {
    if (!h) return PAIRHMM_ERR_ARG;
    auto t0 = clk::now();
    if (h->is_multi()) {     // every device's worker builds the views of its own share (multi.cu)
        const FlatBatch fb{read_bases, base_q, ins_q, del_q, gcp, read_off, read_len, n_reads, hap_bases, hap_off, hap_len, n_haps,
                           reg_read_start, reg_hap_start, n_regions, log10_out, rescued_out};
        return multi_compute_flat(h, fb);
    }
    static thread_local FlatViews v;     // kept between calls: 48 B per read, page-faulting it in every call costs more than filling it
    int st = build_flat_views(h, v, read_bases, base_q, ins_q, del_q, gcp, read_off, read_len, n_reads, hap_bases,
                              hap_off, hap_len, n_haps, reg_read_start, reg_hap_start, n_regions, log10_out, rescued_out);
    if (st != PAIRHMM_OK) return st;
    for (auto &r : v.regions)
        if ((int64_t)r.n_reads * r.n_haps > 0 && !r.log10_out) return fail(h, PAIRHMM_ERR_ARG, "null output");
    h->stats.host_ms_total += ms_since(t0);   // view building counts as host time of the call
    return pairhmm_cuda_compute_batch(h, v.regions.data(), n_regions);
}

int pairhmm_cuda_compute_flat_normalized(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                                         const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *gcp,
                                         const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                                         const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len,
                                         int64_t n_haps, const int64_t *reg_read_start, const int64_t *reg_hap_start,
                                         int64_t n_regions, const pairhmm_cuda_epilogue *ep, double *log10_out,
                                         uint8_t *rescued_out, phmm_read_summary *summary_out) {
    if (!h) return PAIRHMM_ERR_ARG;
    static thread_local FlatViews v;
    int st = build_flat_views(h, v, read_bases, base_q, ins_q, del_q, gcp, read_off, read_len, n_reads, hap_bases,
                              hap_off, hap_len, n_haps, reg_read_start, reg_hap_start, n_regions, log10_out, rescued_out);
    if (st != PAIRHMM_OK) return st;
    for (auto &r : v.regions)
        if ((int64_t)r.n_reads * r.n_haps > 0 && !r.log10_out) return fail(h, PAIRHMM_ERR_ARG, "null output");
    std::vector<phmm_read_summary *> sp;
    if (summary_out) {
        sp.resize((size_t)n_regions);
        for (int64_t r = 0; r < n_regions; ++r) sp[r] = summary_out + reg_read_start[r];
    }
    return pairhmm_cuda_compute_batch_normalized(h, v.regions.data(), n_regions, ep, summary_out ? sp.data() : nullptr);
}

// ---- raw reads: quality pre-pass on the device ----
void pairhmm_cuda_default_read_model(pairhmm_cuda_read_model *m) {
    if (!m) return;
    m->pcr_rate_factor = 3;           // CONSERVATIVE, LikelihoodEngineArgumentCollection.h:30
    m->base_quality_threshold = 18;   // BASE_QUALITY_SCORE_THRESHOLD
    m->constant_gcp = 10;             // gcpHMM
    m->reserved = 0;
}

static int check_read_model(pairhmm_cuda_handle *h, const pairhmm_cuda_read_model *m) {
    if (!m || m->pcr_rate_factor < 0 || m->pcr_rate_factor > 3)
        return fail(h, PAIRHMM_ERR_ARG, "read model: pcr_rate_factor must be 0 (none), 1, 2 or 3");
    if (m->base_quality_threshold < 6 || m->base_quality_threshold > 127)       // PHLCE.cpp:39
        return fail(h, PAIRHMM_ERR_ARG, "read model: base_quality_threshold must be in [6, 127]");
    if (m->constant_gcp < 0 || m->constant_gcp > 127) return fail(h, PAIRHMM_ERR_ARG, "read model: constant_gcp must be in [0, 127]");
    return PAIRHMM_OK;
}

int pairhmm_cuda_compute_batch_raw(pairhmm_cuda_handle *h, const phmm_raw_region *regions, int64_t n_regions,
                                   const pairhmm_cuda_read_model *model) {
    if (!h || n_regions < 0 || (n_regions > 0 && !regions)) return PAIRHMM_ERR_ARG;
    int st = check_read_model(h, model);
    if (st != PAIRHMM_OK) return st;
    std::vector<RegionView> v((size_t)n_regions);
    for (int64_t i = 0; i < n_regions; ++i) {
        const phmm_raw_region &r = regions[i];
        v[i] = RegionView{nullptr, r.reads, r.n_reads, r.haps, r.n_haps, r.log10_out, r.rescued_out, nullptr};
    }
    h->model = *model;
    return compute_batch_impl(h, v.data(), n_regions);
}

int pairhmm_cuda_compute_raw(pairhmm_cuda_handle *h, const phmm_raw_read *reads, int n_reads, const phmm_hap *haps,
                             int n_haps, const pairhmm_cuda_read_model *model, double *log10_out, uint8_t *rescued_out) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (n_reads == 0) return check_read_model(h, model);   // VectorLoglessPairHMM.cpp:46-47
    const phmm_raw_region r{reads, n_reads, haps, n_haps, log10_out, rescued_out};
    return pairhmm_cuda_compute_batch_raw(h, &r, 1, model);
}

int pairhmm_cuda_compute_flat_raw(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                                  const uint8_t *ins_q, const uint8_t *del_q, const int32_t *mapq,
                                  const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                                  const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len,
                                  int64_t n_haps, const int64_t *reg_read_start, const int64_t *reg_hap_start,
                                  int64_t n_regions, const pairhmm_cuda_read_model *model, double *log10_out,
                                  uint8_t *rescued_out) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (n_reads < 0 || n_haps < 0 || n_regions < 0) return fail(h, PAIRHMM_ERR_ARG, "negative count");
    if (n_reads && (!read_bases || !base_q || !mapq || !read_off || !read_len)) return fail(h, PAIRHMM_ERR_ARG, "null read array");
    if (n_haps && (!hap_bases || !hap_off || !hap_len)) return fail(h, PAIRHMM_ERR_ARG, "null haplotype array");
    if (n_regions && (!reg_read_start || !reg_hap_start)) return fail(h, PAIRHMM_ERR_ARG, "null region array");
    static thread_local std::vector<phmm_raw_read> reads;
    static thread_local std::vector<phmm_hap> haps;
    reads.resize((size_t)n_reads); haps.resize((size_t)n_haps);
    phmm_raw_read *rp = reads.data();   // a worker thread naming `reads` would get ITS thread_local instance
    parallel_for((size_t)n_reads, pack_threads(h), 65536, [&, rp](size_t i) {
        const int64_t o = read_off[i];
        rp[i] = phmm_raw_read{read_bases + o, base_q + o, ins_q ? ins_q + o : nullptr, del_q ? del_q + o : nullptr, read_len[i], mapq[i]};
    });
    for (int64_t i = 0; i < n_haps; ++i) haps[i] = phmm_hap{hap_bases + hap_off[i], hap_len[i]};
    std::vector<phmm_raw_region> regs((size_t)n_regions);
    int64_t out_off = 0;
    for (int64_t r = 0; r < n_regions; ++r) {
        const int64_t r0 = reg_read_start[r], r1 = reg_read_start[r + 1], h0 = reg_hap_start[r], h1 = reg_hap_start[r + 1];
        if (r0 < 0 || r1 < r0 || r1 > n_reads || h0 < 0 || h1 < h0 || h1 > n_haps || r1 - r0 > 0x7fffffff || h1 - h0 > 0x7fffffff)
            return fail(h, PAIRHMM_ERR_ARG, "region bounds out of range");
        if ((r1 - r0) * (h1 - h0) > 0 && !log10_out) return fail(h, PAIRHMM_ERR_ARG, "null output");
        regs[r] = phmm_raw_region{reads.data() + r0, (int32_t)(r1 - r0), haps.data() + h0, (int32_t)(h1 - h0),
                                  log10_out ? log10_out + out_off : nullptr, rescued_out ? rescued_out + out_off : nullptr};
        out_off += (r1 - r0) * (h1 - h0);
    }
    return pairhmm_cuda_compute_batch_raw(h, regs.data(), n_regions, model);
}

// The pre-pass alone: pack the reads as one haplotype-less region, run prep_reads_kernel, copy the planes back.
int pairhmm_cuda_prepare_reads_flat(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                                    const uint8_t *ins_q, const uint8_t *del_q, const int32_t *mapq,
                                    const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                                    const pairhmm_cuda_read_model *model, uint8_t *out_base_q, uint8_t *out_ins_q,
                                    uint8_t *out_del_q, uint8_t *out_gcp) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (h->is_multi())
        return pairhmm_cuda_prepare_reads_flat(h->children[0], read_bases, base_q, ins_q, del_q, mapq, read_off, read_len, n_reads,
                                               model, out_base_q, out_ins_q, out_del_q, out_gcp);
    int st = check_read_model(h, model);
    if (st != PAIRHMM_OK) return st;
    if (n_reads < 0 || n_reads > 0x7fffffff) return fail(h, PAIRHMM_ERR_ARG, "bad read count");
    if (n_reads == 0) return PAIRHMM_OK;
    if (!read_bases || !base_q || !mapq || !read_off || !read_len || !out_base_q || !out_ins_q || !out_del_q || !out_gcp)
        return fail(h, PAIRHMM_ERR_ARG, "null array");
    CU(h, cudaSetDevice(h->device));
    for (int si = 0; si < 2; ++si)
        if (h->slots[si].busy) { st = finish_chunk(h, h->slots[si], 1); if (st) return st; h->slot_ticket[si] = 0; }
    std::vector<phmm_raw_read> reads((size_t)n_reads);
    for (int64_t i = 0; i < n_reads; ++i) {
        const int64_t o = read_off[i];
        reads[i] = phmm_raw_read{read_bases + o, base_q + o, ins_q ? ins_q + o : nullptr, del_q ? del_q + o : nullptr, read_len[i], mapq[i]};
    }
    const bool dedup = h->cfg.dedup_reads != 0;
    h->cfg.dedup_reads = 0;               // every read keeps its own row: unique index == input index
    h->model = *model;
    h->resident_loaded = false;
    RegionView rv{nullptr, reads.data(), (int32_t)n_reads, nullptr, 0, nullptr, nullptr, nullptr};
    Slot &s = h->slots[0];
    h->no_output_ok = true;
    st = pack_chunk(h, s, &rv, 1, pack_threads(h));
    h->no_output_ok = false;
    h->cfg.dedup_reads = dedup ? 1 : 0;
    if (st != PAIRHMM_OK) return st;
    if ((st = upload_chunk(h, s)) != PAIRHMM_OK) return st;
    std::vector<uint8_t> tmp((size_t)s.plane * 4);
    if (s.n_uniq) CU(h, cudaMemcpyAsync(tmp.data(), s.d_rd.p + s.plane, s.plane * 4, cudaMemcpyDeviceToHost, s.stream));
    CU(h, cudaStreamSynchronize(s.stream));
    const RegionPlan &p = s.plans[0];
    for (int64_t i = 0; i < n_reads; ++i) {
        const uint32_t u = p.uniq_of_read[(size_t)i];
        if (u == 0xffffffffu) continue;   // zero-length read
        const DevRead dr = s.h_reads.p[p.first_uniq + u];
        const int64_t o = read_off[i];
        std::memcpy(out_base_q + o, tmp.data() + dr.off, dr.len);
        std::memcpy(out_ins_q + o, tmp.data() + s.plane + dr.off, dr.len);
        std::memcpy(out_del_q + o, tmp.data() + 2 * s.plane + dr.off, dr.len);
        std::memcpy(out_gcp + o, tmp.data() + 3 * s.plane + dr.off, dr.len);
    }
    return PAIRHMM_OK;
}

int pairhmm_cuda_compute_pairs(pairhmm_cuda_handle *h, const phmm_read *reads, int n_reads, const phmm_hap *haps,
                               int n_haps, const uint32_t *read_idx, const uint32_t *hap_idx, size_t n_pairs,
                               double *log10_out, uint8_t *rescued_out) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (n_pairs == 0) return PAIRHMM_OK;
    if (!reads || !haps || !read_idx || !hap_idx || !log10_out || n_reads <= 0 || n_haps <= 0)
        return fail(h, PAIRHMM_ERR_ARG, "compute_pairs: null pointer or empty tables");
    for (size_t p = 0; p < n_pairs; ++p)
        if (read_idx[p] >= (uint32_t)n_reads || hap_idx[p] >= (uint32_t)n_haps)
            return fail(h, PAIRHMM_ERR_ARG, "compute_pairs: index out of range");
    // full cross product in read-major order = one region
    bool cross = n_pairs == (size_t)n_reads * (size_t)n_haps;
    for (size_t p = 0; cross && p < n_pairs; ++p)
        cross = read_idx[p] == p / (size_t)n_haps && hap_idx[p] == p % (size_t)n_haps;
    if (cross) return pairhmm_cuda_compute(h, reads, n_reads, haps, n_haps, log10_out, rescued_out);
    // general list: one single-read region per distinct read, carrying that read's haplotypes
    std::vector<size_t> start((size_t)n_reads + 1, 0), pos(n_pairs);
    for (size_t p = 0; p < n_pairs; ++p) start[read_idx[p] + 1]++;
    for (int r = 0; r < n_reads; ++r) start[r + 1] += start[r];
    std::vector<size_t> fill(start.begin(), start.end() - 1);
    std::vector<phmm_hap> hl(n_pairs);
    for (size_t p = 0; p < n_pairs; ++p) { pos[p] = fill[read_idx[p]]++; hl[pos[p]] = haps[hap_idx[p]]; }
    std::vector<double> tmp(n_pairs);
    std::vector<uint8_t> tmpf(n_pairs);
    std::vector<phmm_region> regs;
    for (int r = 0; r < n_reads; ++r)
        if (start[r + 1] > start[r])
            regs.push_back(phmm_region{reads + r, 1, hl.data() + start[r], (int32_t)(start[r + 1] - start[r]),
                                       tmp.data() + start[r], tmpf.data() + start[r]});
    int st = pairhmm_cuda_compute_batch(h, regs.data(), (int64_t)regs.size());
    if (st != PAIRHMM_OK) return st;
    for (size_t p = 0; p < n_pairs; ++p) { log10_out[p] = tmp[pos[p]]; if (rescued_out) rescued_out[p] = tmpf[pos[p]]; }
    return PAIRHMM_OK;
}

}  // extern "C"

// stage RegionViews on slot 0 of a single-device handle (H2D + layout); out_off[i] = offset of view i's results in the
// caller's flat arrays at _resident_fetch
int phmm::resident_load_views(pairhmm_cuda_handle *h, const RegionView *views, size_t n, const size_t *out_off) {
    CU(h, cudaSetDevice(h->device));
    for (int si = 0; si < 2; ++si)
        if (h->slots[si].busy) { int st = finish_chunk(h, h->slots[si], 1); if (st) return st; h->slot_ticket[si] = 0; }
    h->resident_loaded = false;
    Slot &s = h->slots[0];
    h->no_output_ok = true;
    int st = pack_chunk(h, s, views, n, pack_threads(h));
    h->no_output_ok = false;
    if (st != PAIRHMM_OK) return st;
    if ((st = upload_chunk(h, s)) != PAIRHMM_OK) return st;
    CU(h, cudaStreamSynchronize(s.stream));
    // the views point into the caller's arrays, which may die now: keep only the shapes (plans hold everything scatter needs)
    for (auto &rv : s.regions) { rv.reads = nullptr; rv.raw = nullptr; rv.haps = nullptr; rv.out = nullptr; rv.resc = nullptr; }
    h->resident_out_off.assign(out_off, out_off + n);
    h->resident_loaded = true;
    return PAIRHMM_OK;
}

extern "C" {

int pairhmm_cuda_resident_load(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                               const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *gcp,
                               const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                               const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len,
                               int64_t n_haps, const int64_t *reg_read_start, const int64_t *reg_hap_start,
                               int64_t n_regions) {
    if (!h) return PAIRHMM_ERR_ARG;
    FlatViews v;
    int st = build_flat_views(h, v, read_bases, base_q, ins_q, del_q, gcp, read_off, read_len, n_reads, hap_bases,
                              hap_off, hap_len, n_haps, reg_read_start, reg_hap_start, n_regions, nullptr, nullptr);
    if (st != PAIRHMM_OK) return st;
    std::vector<RegionView> views(n_regions);
    std::vector<size_t> out_off((size_t)n_regions + 1, 0);
    for (int64_t i = 0; i < n_regions; ++i) {
        views[i] = RegionView{v.regions[i].reads, nullptr, v.regions[i].n_reads, v.regions[i].haps, v.regions[i].n_haps, nullptr, nullptr, nullptr};
        out_off[(size_t)i + 1] = out_off[(size_t)i] + (size_t)v.regions[i].n_reads * (size_t)v.regions[i].n_haps;
    }
    if (h->is_multi()) return multi_resident_load(h, views.data(), views.size());
    return resident_load_views(h, views.data(), views.size(), out_off.data());
}

int pairhmm_cuda_resident_run(pairhmm_cuda_handle *h, float *kernel_ms, float *float_kernel_ms) {
    if (!h) return PAIRHMM_ERR_ARG;
    if (h->is_multi()) return multi_resident_run(h, kernel_ms, float_kernel_ms);
    if (!h->resident_loaded) return fail(h, PAIRHMM_ERR_STATE, "resident_run: nothing loaded (or the batch was replaced by a later compute call)");
    CU(h, cudaSetDevice(h->device));
    Slot &s = h->slots[0];
    if (s.busy) return fail(h, PAIRHMM_ERR_STATE, "resident_run: a submitted chunk is in flight on this handle");
    int st = run_kernels(h, s, true);
    if (st != PAIRHMM_OK) return st;
    CU(h, cudaStreamSynchronize(s.stream));
    if (trace_on()) trace_rescue_counts(h, s);
    float a = 0.f, b = 0.f;
    CU(h, cudaEventElapsedTime(&a, s.ev_begin, s.ev_end));
    CU(h, cudaEventElapsedTime(&b, s.ev_begin, s.ev_fwd_done));   // FP32 forward kernels only (no finalize)
    if (kernel_ms) *kernel_ms = a;
    if (float_kernel_ms) *float_kernel_ms = b;
    h->stats.gpu_ms_total += a; h->stats.gpu_ms_float += b; h->stats.gpu_ms_double += a - b;
    h->stats.cells += s.cells; h->stats.pairs += s.pairs; h->stats.cells_swept += s.cells_swept;
    return PAIRHMM_OK;
}

int pairhmm_cuda_resident_fetch(pairhmm_cuda_handle *h, double *log10_out, uint8_t *rescued_out) {
    if (!h || !log10_out) return PAIRHMM_ERR_ARG;
    if (h->is_multi()) return multi_resident_fetch(h, log10_out, rescued_out);
    if (!h->resident_loaded) return fail(h, PAIRHMM_ERR_STATE, "resident_fetch: nothing loaded (or the batch was replaced by a later compute call)");
    CU(h, cudaSetDevice(h->device));
    Slot &s = h->slots[0];
    if (s.busy) return fail(h, PAIRHMM_ERR_STATE, "resident_fetch: a submitted chunk is in flight on this handle");
    for (size_t i = 0; i < s.regions.size(); ++i) {
        const size_t off = h->resident_out_off[i];
        s.regions[i].out = log10_out + off; s.regions[i].resc = rescued_out ? rescued_out + off : nullptr;
    }
    int st = download_chunk(h, s);
    if (st != PAIRHMM_OK) return st;
    const uint64_t c = h->stats.cells, p = h->stats.pairs, cs = h->stats.cells_swept;   // finish_chunk would count the batch again
    const double g0 = h->stats.gpu_ms_total, g1 = h->stats.gpu_ms_float, g2 = h->stats.gpu_ms_double;
    st = finish_chunk(h, s, pack_threads(h));
    h->stats.cells = c; h->stats.pairs = p; h->stats.cells_swept = cs;
    h->stats.gpu_ms_total = g0; h->stats.gpu_ms_float = g1; h->stats.gpu_ms_double = g2;
    return st;
}

// Raw FP32-pass sums (2^120-scaled, before log10) of the last _resident_run for every input pair, in the
// same order as the results (region-major, read-major). For bit-exactness tests of the FP32 kernel.
int pairhmm_cuda_resident_fetch_raw(pairhmm_cuda_handle *h, float *fres_out, int64_t n) {
    if (!h || !fres_out) return PAIRHMM_ERR_ARG;
    if (h->is_multi()) return fail(h, PAIRHMM_ERR_STATE, "resident_fetch_raw: a single-device test hook");
    if (!h->resident_loaded) return fail(h, PAIRHMM_ERR_STATE, "resident_fetch_raw: nothing loaded");
    Slot &s = h->slots[0];
    if (n != (int64_t)s.pairs) return fail(h, PAIRHMM_ERR_ARG, "resident_fetch_raw: n != number of pairs");
    CU(h, cudaSetDevice(h->device));
    std::vector<float> tmp(s.n_upairs);
    CU(h, cudaMemcpyAsync(tmp.data(), s.d_fres.p, sizeof(float) * s.n_upairs, cudaMemcpyDeviceToHost, s.stream));
    CU(h, cudaStreamSynchronize(s.stream));
    size_t off = 0;
    for (size_t i = 0; i < s.regions.size(); ++i) {
        const RegionView &rv = s.regions[i];
        const RegionPlan &p = s.plans[i];
        const size_t nh = (size_t)rv.n_haps;
        for (int r = 0; r < rv.n_reads; ++r, off += nh) {
            const uint32_t u = p.uniq_of_read[r];
            for (size_t q = 0; q < nh; ++q)
                fres_out[off + (p.hap_order.empty() ? q : (size_t)p.hap_order[q])] =
                    (u == kNoRead) ? 0.f : tmp[(size_t)p.upair_off + (size_t)u * nh + q];
        }
    }
    return PAIRHMM_OK;
}

}  // extern "C"
