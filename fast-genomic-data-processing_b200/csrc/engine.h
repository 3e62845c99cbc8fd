// Host side of libpairhmm_cuda: turns regions (reads x haplotypes) into warp work units, stages them
// through pinned memory, launches the kernels of kernels.cu and scatters results back.
// Replaces the compute half of VectorLoglessPairHMM (M2/utils/pairhmm/VectorLoglessPairHMM.cpp:43-148)
// and all of gkl_pairhmm's dispatch (M2/intel/pairhmm/IntelPairHmm.cc:259-351).
#pragma once
#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../../include/pairhmm_cuda.h"
#include "kernels.cuh"

namespace phmm {

// Growable pinned-host / device arrays (never shrink; 1.25x growth).
// A buffer may also be a VIEW into another one (Slot::h_meta / d_meta): it then owns nothing and is never reserved.
template <class T> struct PinnedBuf {
    T *p = nullptr; size_t cap = 0; bool view = false;
    cudaError_t reserve(size_t n);
    void release();
    void set_view(void *q) { if (!view) release(); p = static_cast<T *>(q); cap = 0; view = true; }
};
template <class T> struct DevBuf {
    T *p = nullptr; size_t cap = 0; bool view = false;
    cudaError_t reserve(size_t n);
    void release();
    void set_view(void *q) { if (!view) release(); p = static_cast<T *>(q); cap = 0; view = true; }
};

// One region as the packer sees it (pointers borrowed from the caller).
struct RegionView {
    const phmm_read *reads;           // processed reads (the kernels' input), or
    const phmm_raw_read *raw;         // reads as the BAM record holds them: the device runs the quality pre-pass (exactly one is set)
    int32_t n_reads;
    const phmm_hap *haps; int32_t n_haps;
    double *out; uint8_t *resc;
    phmm_read_summary *summ;          // per-read epilogue results (nullable; only with an active epilogue)
    int64_t read_len_sum = -1, hap_len_sum = -1;   // sums of the lengths if a caller up the stack has them already (else -1)
};

// Sizes/offsets of one region inside a packed chunk (filled by pass A + prefix sums).
struct RegionPlan {
    uint32_t n_uniq = 0;            // unique reads with len > 0
    uint32_t first_uniq = 0;        // global unique-read id of this region's first unique read
    uint64_t read_bytes = 0, read_byte_off = 0;
    uint64_t read_len_sum = 0;      // bases of every input read (duplicates included): the region's DP cells = this x hap_bytes
    uint32_t first_hap = 0;         // global hap id
    uint64_t hap_bytes = 0, hap_byte_off = 0;
    uint32_t hap_words = 0, hap_word_off = 0;
    uint32_t hap_list_off = 0;
    uint32_t group = 0, n_groups = 0;   // haplotypes per work unit / units per read
    uint32_t upair_off = 0;         // first unique-pair index
    // unique reads are ordered by kernel bin; a run = the reads of one (W,K,KL) bin
    struct BinRun { uint16_t bin; uint32_t first, n_reads, unit_off; };
    std::vector<BinRun> runs;
    uint32_t pairs32[kNumBinsW32] = {0};   // unique pairs per W=32 bin (bounds that bin's FP64 rescue list)
    uint32_t pairs16[kNumBinsW16] = {0};   // the same per W=16 bin, reads the W=16 FP64 kernel takes (two of them per unit)
    uint8_t gcp_min = 255, gcp_max = 0;    // range of the (masked) gap-continuation bytes of the unique reads
    std::vector<uint32_t> uniq_of_read;  // per input read: local unique index, or 0xffffffff for len==0
    std::vector<uint32_t> uniq_src;      // per local unique read: index of its first occurrence
    bool any_indel_q = false;            // raw reads: some read has its own insertion / deletion qualities
    bool identity = true;                // unique read j is input read j for every j (no duplicates, no empty reads, no re-ordering)
    // prefix sharing (Variant::F32_PFX): the region's haplotypes are laid out ordered by (length, bytes)
    std::vector<uint32_t> hap_order;     // position in the chunk -> index in the caller's list (empty = identity)
    std::vector<uint2> hap_plan;         // sweep plan per position (kernels.cuh: pfx_plan)
    uint64_t shared_cols = 0;            // haplotype columns not swept again thanks to a checkpoint
    int32_t hmax = 0;                    // longest haplotype of the region
};

// Everything one in-flight chunk needs: pinned staging, device mirrors, stream, events.
struct Slot {
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;   // FP64 pass: the units of one read run beside the units of two reads (run_kernels)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_fwd_done = nullptr, ev_float_done = nullptr, ev_end = nullptr;
    // host staging
    PinnedBuf<uint8_t> h_rd, h_hapbytes, h_flag;
    PinnedBuf<DevRead> h_reads;
    PinnedBuf<DevHap> h_haps;
    PinnedBuf<uint64_t> h_hapoff;
    PinnedBuf<uint32_t> h_haplist;
    PinnedBuf<uint2> h_happlan;       // F32_PFX: sweep plan per hap_list entry
    // The descriptors, the haplotype bytes and the work units of a chunk (h_reads, h_hapoff, h_haps, h_haplist, h_happlan,
    // h_units, h_hapbytes and their device mirrors) are views into one pinned / one device arena with the same layout:
    // they cross PCIe in ONE copy (a per-region call made seven of them)
    PinnedBuf<uint8_t> h_meta;
    uint64_t meta_bytes = 0;          // bytes of the arena that cross PCIe (everything up to the last sent read plane)
    uint64_t ctr_off = 0;             // offset of the counter block inside the arena (resident replays re-send it)
    int sent_planes = 5;              // read planes that are staged and sent
    PinnedBuf<uint32_t> h_counters;   // view: zeroed work queues / rescue counts + the rescue list bases
    PinnedBuf<uint8_t> h_res;         // results: doubles, then the rescued flags (h_out / h_flag are views)
    DevBuf<uint8_t> d_res;
    uint64_t res_bytes = 0;
    bool timed = false;               // this chunk's kernels are bracketed by timing events
    PinnedBuf<DevUnit> h_units;
    PinnedBuf<double> h_out;
    PinnedBuf<int32_t> h_mapq;        // raw reads: mapping quality per unique read
    PinnedBuf<RowInfo> h_rows;        // epilogue: row of every unique read in the result matrix
    PinnedBuf<phmm_read_summary> h_summ;
    // device
    DevBuf<uint8_t> d_rd, d_hapbytes, d_flag;
    DevBuf<DevRead> d_reads;
    DevBuf<DevHap> d_haps;
    DevBuf<uint64_t> d_hapoff;
    DevBuf<uint32_t> d_haplist, d_hapw, d_counters;
    DevBuf<uint2> d_happlan;
    DevBuf<uint8_t> d_meta;
    DevBuf<uint8_t> d_long_scratch;   // stripe hand-over lines of the long-read kernel
    int32_t hap_max = 0;              // longest haplotype of the chunk
    DevBuf<DevUnit> d_units, d_rescue;
    DevBuf<float> d_fres;
    DevBuf<double> d_out;
    DevBuf<int32_t> d_mapq;
    DevBuf<RowInfo> d_rows;
    DevBuf<phmm_read_summary> d_summ;
    // shape of the packed chunk
    std::vector<RegionView> regions;
    std::vector<RegionPlan> plans;
    uint64_t plane = 0;               // bytes per read plane
    uint32_t n_uniq = 0, n_haps = 0, n_hap_words = 0, n_haplist = 0, n_upairs = 0;
    uint64_t hap_bytes = 0;
    uint32_t bin_start[kNumBins + 2] = {0};   // units of bin b are [bin_start[b], bin_start[b+1]); bin kLongBin = long reads (long.cu)
    uint32_t rescue_base[kNumBinsW32 + 1] = {0}; // W=32 bin b appends its rescue units at d_rescue + rescue_base[b]
    uint32_t rescue_base16[kNumBinsW16 + 1] = {0}; // ... and W=16 bin b its units of two reads (behind the W=32 lists)
    uint32_t pair_mode = 0;                   // RescueScanArgs::pair_mode of this chunk (0: no W=16 lists, no pending area)
    uint32_t pending_off = 0;                 // first entry of the pairing candidates' area in d_rescue
    DevBuf<uint32_t> d_pair_table;            // kPairSlots words, zero between chunks
    Variant float_variant = Variant::F32_GEN; // kernels this chunk was binned for
    bool uniform_gcp = false;                 // every gcp byte of the chunk equals gcp_value -> UG kernels
    uint8_t gcp_value = 0;
    bool skip_gcp = false;                    // the chunk's kernels never read the gcp plane: not staged, not sent
    uint64_t cells = 0, pairs = 0, reads_in = 0;
    uint64_t cells_swept = 0;         // DP cells the FP32 pass sweeps: unique reads x haplotype columns not shared
    bool busy = false;
    bool counted = false;             // this chunk is in the device's in-flight count (set_busy)
    bool ep_on = false;               // this chunk runs the fused epilogue (pairhmm_cuda_compute_*_normalized)
    pairhmm_cuda_epilogue ep{};
    bool raw_on = false;              // this chunk holds raw reads: quality pre-pass on the device (pairhmm_cuda_compute_*_raw)
    bool raw_indel_q = false;         //   ... and at least one of them brings its own insertion / deletion qualities
    pairhmm_cuda_read_model model{};
    void release();
};

struct DeviceWorker;     // multi.cu: the host worker thread of one device of a multi-device handle
struct Batcher;          // batcher.cu: process-wide merger of single-region calls (config.batch_window_us)

// one part of a multi-device plan: reads [first_read, first_read + n_reads) of a region, on device `device`
struct DevicePart { int64_t region; int32_t first_read, n_reads, device; };
// LPT split of a batch over n_devices (pairhmm_cuda_device_plan); cost of region r = read_len_sum[r] * hap_len_sum[r]
std::vector<DevicePart> plan_devices(const uint64_t *read_len_sum, const uint64_t *hap_len_sum, const int32_t *n_reads,
                                     int64_t n_regions, int n_devices);

// ---- internal entry points shared by engine.cu / multi.cu / batcher.cu (single-device handles) ----
// ensure (nullable): called with i before regions[i] is first read -- a caller that fills the array lazily (multi.cu builds a
// device's views block by block, under the kernels of the chunks already submitted) completes regions[0 .. i] there
int compute_batch_impl(pairhmm_cuda_handle *h, const RegionView *regions, int64_t n_regions,
                       const std::function<void(int64_t)> *ensure = nullptr);
int resident_load_views(pairhmm_cuda_handle *h, const RegionView *views, size_t n, const size_t *out_off);
int check_region_args(pairhmm_cuda_handle *h, const RegionView &rv);   // the validation plan_region does, without the plan
int fail(pairhmm_cuda_handle *h, int status, const std::string &msg);
int pack_threads(const pairhmm_cuda_handle *h);
// marks a slot's chunk as in flight / retired and keeps the per-device count of chunks in flight (all handles of the
// process): the packer shapes a small chunk for latency when the device is otherwise idle and for throughput when other
// host threads keep it busy
void set_busy(pairhmm_cuda_handle *h, Slot &s, bool busy);
int device_chunks_in_flight(int device);
// multi-device handles (multi.cu)
int multi_create(pairhmm_cuda_handle *h);
void multi_destroy(pairhmm_cuda_handle *h);
int multi_compute_batch(pairhmm_cuda_handle *h, const RegionView *regions, int64_t n_regions);
// the arguments of pairhmm_cuda_compute_flat
struct FlatBatch {
    const uint8_t *read_bases, *base_q, *ins_q, *del_q, *gcp; const int64_t *read_off; const int32_t *read_len; int64_t n_reads;
    const uint8_t *hap_bases; const int64_t *hap_off; const int32_t *hap_len; int64_t n_haps;
    const int64_t *reg_read_start, *reg_hap_start; int64_t n_regions; double *log10_out; uint8_t *rescued_out;
};
int multi_compute_flat(pairhmm_cuda_handle *h, const FlatBatch &fb);
int multi_submit(pairhmm_cuda_handle *h, const RegionView &rv, pairhmm_cuda_ticket *ticket);
int multi_wait(pairhmm_cuda_handle *h, pairhmm_cuda_ticket ticket);
int multi_resident_load(pairhmm_cuda_handle *h, const RegionView *views, size_t n);
int multi_resident_run(pairhmm_cuda_handle *h, float *kernel_ms, float *float_kernel_ms);
int multi_resident_fetch(pairhmm_cuda_handle *h, double *log10_out, uint8_t *rescued_out);
void multi_stats(pairhmm_cuda_handle *h, pairhmm_cuda_stats *out);
// shared batcher (batcher.cu): one blocking region of handle h through the process-wide batcher of h's device
int batcher_compute(pairhmm_cuda_handle *h, const RegionView &rv);
Batcher *batcher_acquire(pairhmm_cuda_handle *h);
void batcher_release(Batcher *b);

}  // namespace phmm

struct pairhmm_cuda_handle {
    pairhmm_cuda_config cfg;
    int device = 0;
    int sm_count = 0;
    std::string last_error;
    pairhmm_cuda_stats stats;
    // device tables (per handle; 70 KB)
    float *d_ph2pr_f = nullptr, *d_mm_f = nullptr;
    double *d_ph2pr_d = nullptr, *d_mm_d = nullptr;
    float log10_init_f = 0.f;
    double log10_init_d = 0.0;
    int blocks_per_sm[phmm::kNumVariants][phmm::kNumBins] = {{0}};   // [Variant][bin], 0 = not queried yet
    int read_max_len = 0, hap_max_len = 0;
    phmm::Slot slots[2];
    int64_t next_ticket = 1;
    int64_t slot_ticket[2] = {0, 0};
    cudaEvent_t prev_end = nullptr;   // end event of the most recently submitted chunk still in flight (kernels of the next one wait on it)
    bool light_call = false;          // set around a single-region submit: no timing events (stats gpu_ms_* stay 0 for it
                                      // unless PAIRHMM_CUDA_TIMING is set), fewer driver calls
    bool resident_loaded = false;
    std::vector<size_t> resident_out_off;  // per region of the resident batch: offset of its results in the caller's arrays
    bool no_output_ok = false;        // resident_load / prepare_reads pack regions without result arrays
    // ---- multi-device handle (cfg.device_mask != 0): one single-device child handle + one host worker thread per device
    std::vector<pairhmm_cuda_handle *> children;
    std::vector<phmm::DeviceWorker *> workers;            // owned (multi_destroy)
    std::unordered_map<int64_t, std::pair<int, int64_t>> ticket_of;   // ticket of this handle -> (child, child's ticket)
    int next_child = 0;               // single-region calls go to the devices in turn
    bool is_multi() const { return !children.empty(); }
    // ---- shared batcher (cfg.batch_window_us > 0)
    phmm::Batcher *batcher = nullptr; // process-wide, per device; not owned
    bool ep_on = false;               // set for the duration of a *_normalized call
    pairhmm_cuda_epilogue ep{};
    pairhmm_cuda_read_model model{};  // of the current *_raw call
};
