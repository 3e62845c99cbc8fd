// Device-side data layout and launch entry points of the PairHMM engine (sm_100a).
//
// Layout in HBM (all built by the host packer / the prep kernel):
//   read planes  uint8  rd[5][plane]    bases | base_q | ins_q | del_q | gcp, raw bytes as the caller
//                                       gave them (masking &127 and base->code happen on device)
//   DevRead      {off,len}              byte offset of a read inside every plane
//   hap words    uint32 hapw[]          one word per DP column, in tiles of kTile columns. Words
//                                       [0,kLeadWords) = pipeline fill (PAD.. + one SEP) and drain (PAD..);
//                                       then per haplotype: H code words, PADs up to the next tile boundary, and
//                                       one tile [SEP, PAD..] (a SEP is always followed by a PAD: two flagged
//                                       columns in a row reset a lane's state, forward2.cuh). A code word holds the
//                                       base index b (A0 C1 T2 G3 N4) pre-multiplied for every rows-per-lane
//                                       bucket (see code_word()), so a lane turns it into K match bits with
//                                       one funnel shift of its per-base row masks.
//   DevHap       {off,len}              word offset (multiple of kTile) / H
//   hap_list     uint32[]               haplotype ids, one run per work unit
//   DevUnit      {read,hap_list,nh,out} one warp-task: read x nh haplotypes streamed back to back
//   results      float fres[] (FP32 pass), double out[] (final log10), uint8 flag[] (rescued)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/pairhmm_cuda.h"

namespace phmm {

struct DevRead { uint32_t off; int32_t len; };
struct DevHap  { uint32_t off; int32_t len; };
// One warp task: `nreads` consecutive unique reads (one per lane group of W lanes) against the same run of
// nh haplotypes. shape = nh | nreads << 8 | out_stride << 16; result of (read j, hap i) goes to
// out + j*out_stride + i.
struct alignas(16) DevUnit { uint32_t read; uint32_t hap_list; uint32_t shape; uint32_t out; };
inline __host__ __device__ uint32_t unit_shape(uint32_t nh, uint32_t nreads, uint32_t stride) { return nh | (nreads << 8) | (stride << 16); }

// flagged words carry base index 5 in bits 25..27: the packed kernel runs them through its column code with the
// all-zero row of its match-factor table (forward2.cuh); the scalar kernels ignore those bits
constexpr uint32_t kWordPad = 0xCA000000u;   // flagged, no haplotype change (pipeline fill/drain, alignment)
constexpr uint32_t kWordSep = 0x8A000000u;   // flagged, column 0 of the next haplotype
#ifndef PHMM_TILE
#define PHMM_TILE 4
#endif
constexpr int kTile = PHMM_TILE;             // alignment of the word layout = columns per step of the packed kernel
constexpr int kScalarTile = 2;               // columns per step of the scalar kernels (forward.cuh)
static_assert(PHMM_TILE % 2 == 0, "the scalar kernels walk the layout two columns at a time");
constexpr int kMaxUnitHaps = 32;             // haplotypes streamed by one work unit (one per lane in the per-warp table)
constexpr int kLeadWords = 64 * kTile;       // [0, 32*kTile-1) PAD, [32*kTile-1] SEP, rest PAD
constexpr int kMaxRowsPerLane = 12;          // K at W = 32: one warp covers reads up to 384 rows (kMaxReadLen)
constexpr int kMaxReadLen = 32 * kMaxRowsPerLane;
constexpr int kThreads = 128;                // 4 warps per CTA, one per SM sub-partition

// words a haplotype of length H occupies in hapw (code words + alignment PADs + the [SEP, PAD..] tile)
static_assert(PHMM_TILE >= 2, "a SEP needs a PAD behind it in its own tile");
inline __host__ __device__ uint32_t hap_tiles(int32_t H) { return ((uint32_t)H + kTile - 1u) / kTile + 1u; }
inline uint32_t hap_words(int32_t H) { return hap_tiles(H) * kTile; }

// match-mask geometry of a rows-per-lane bucket: rows are split into groups of kGroupBits(K) rows, each
// group keeps 5 (one per haplotype base) bit fields of that width in one 32-bit register
__host__ __device__ constexpr int group_bits(int K) { return K <= 6 ? K : 6; }
// bit position inside a code word of the field that holds b * group_bits(K)
__host__ __device__ constexpr int code_field(int K) {
    return group_bits(K) == 5 ? 0 : group_bits(K) == 4 ? 5 : group_bits(K) == 6 ? 10
         : group_bits(K) == 3 ? 15 : group_bits(K) == 2 ? 20 : 25;
}

struct KernelArgs {
    const uint8_t *rd;          // read planes
    uint64_t plane;             // plane stride in bytes
    const DevRead *reads;
    const uint32_t *hapw;
    const DevHap *haps;
    const uint32_t *hap_list;
    const DevUnit *units;
    const uint32_t *n_units_dev;  // if non-null the unit count is read from here (rescue lists)
    uint32_t n_units;
    uint32_t *queue;              // work counter for this launch (zeroed by the host)
    // tables (precision of the launch)
    const void *ph2pr;            // T[128]
    const void *mm;               // T[8256]
    // uniform gap-continuation constants (UG kernels): ph2pr[gcp] and 1 - ph2pr[gcp]
    float u_xy_f, u_gapm_f;
    uint32_t u_xy_bits;           // bit pattern of u_xy_f again (packed kernel: source of the per-column Y coefficient)
    double u_xy_d, u_gapm_d;
    // FP32 pass output
    float *fres;                  // scaled forward sum per unique pair
    // FP64 pass outputs
    double *out;                  // log10 likelihood
    uint8_t *flag;                // 1 = went through the FP64 kernel
    double log10_init_d;          // log10(2^1020), host libm (Context.h:143)
    uint32_t rescue_filter;       // FP64 launches over rescue units: a unit names a read and its whole haplotype run, only
                                  // the pairs with fres < 1e-28f are swept
    const uint2 *hap_plan;        // F32_PFX: sweep plan per hap_list entry (pfx_plan() below)
};

// Prefix sharing (the reference's trie, M2/trie/buildTreeUtils.cpp:155-258 + avx-pairhmm-template.h:475-721,
// shares the DP columns of a common prefix between SAME-LENGTH haplotypes -- the initial Y row is 2^s/H, template.h:115, so
// only equal H gives equal columns). Here: the host orders a region's haplotypes by (length, bytes), and for every
// haplotype of a work unit names the tile its sweep starts at (t_start tiles are shared with an earlier haplotype of the
// unit), the checkpoint slot that holds the lane state after those tiles, and at most one tile boundary t_save at which this
// sweep leaves a checkpoint for a later one. A checkpoint = a lane's M/X/Y rows, its last-row sums and the diagonal input
// of the next column, kept per thread in local memory; lanes save / restore at their own (skewed) step.
constexpr int kPfxSlots = 4;                 // checkpoints alive at once per lane (depth of the shared-prefix stack)
// x = t_start | t_save << 16 (tiles from the haplotype's first tile; t_save == 0: no checkpoint is written)
// y = restore slot + 1 (0 = fresh start, t_start == 0) | save slot << 8
inline __host__ __device__ uint2 pfx_plan(uint32_t t_start, uint32_t restore_slot_plus1, uint32_t t_save, uint32_t save_slot) {
    return make_uint2(t_start | (t_save << 16), restore_slot_plus1 | (save_slot << 8));
}
constexpr uint32_t kPfxMaxTiles = 0xffffu;   // haplotypes with more tiles are swept without sharing

// Between the passes: every read with at least one pair below the reference's threshold (fres < 1e-28f,
// IntelPairHmm.cc:342) becomes ONE unit of the FP64 pass -- the read against its unit's whole haplotype run, the FP64
// kernel sweeps only the underflowed haplotypes (KernelArgs::rescue_filter). Units go to the list of the read's W = 32
// bin: rescue_units[rescue_base[bin] + atomicAdd(&rescue_count[bin], 1)].
struct RescueScanArgs {
    const DevUnit *units;         // every FP32 work unit of the chunk
    uint32_t n_units;
    const DevRead *reads;
    const float *fres;
    DevUnit *rescue_units;
    const uint32_t *rescue_base;
    uint32_t *rescue_count;
    uint32_t max_pairs;           // underflowed pairs per FP64 unit (>= 1; kMaxUnitHaps = the read's whole row); 0 = chosen
                                  // on the device from the chunk's number of underflowed pairs:
    const uint32_t *rescued_pairs;  //   counted by finalize_kernel
    uint32_t fp64_warp_slots;     //   warps of the FP64 kernel resident on the device
    // Read pairing for the wide FP64 kernel at W = 16 (two reads per warp, rows of up to kMaxKF64W16 per lane): see
    // rescue_scan_kernel. The lists of the W = 16 bins lie in rescue_units too, behind those of the W = 32 bins.
    uint32_t pair_mode;           // 0 = off; 1 = when the chunk has enough underflowed pairs to fill the GPU with pairs of reads;
                                  // 2 = always (testing)
    DevUnit *pending;             // candidates (a read x a run of its haplotypes) that wait for a partner
    uint32_t *pending_count;
    uint32_t *pair_table;         // kPairSlots words, all zero between chunks (rescue_flush_kernel leaves them so)
    const uint32_t *rescue_base16;
    uint32_t *rescue_count16;
    uint32_t fp64_warp_slots16;   // warps of the W = 16 FP64 kernel resident on the device
    uint32_t n_pairs;             // unique pairs of the chunk
};
constexpr int kMaxUnitReads = 8;  // reads of one work unit: 2 * 32 / 8 (packed kernel, 8-lane groups)

// Fused per-read epilogue (include/pairhmm_cuda.h, pairhmm_cuda_epilogue): one thread per unique read normalises its
// row of the result matrix in place and writes the read's summary.
struct RowInfo { uint32_t off, nh; };      // first unique-pair index of the read's row, haplotypes in the row
struct EpilogueArgs {
    double *out;                  // result matrix, rows of nh doubles
    const RowInfo *rows;
    const DevRead *reads;
    phmm_read_summary *summ;
    uint32_t n;                   // unique reads
    double cap;                   // likelihood_cap
    double err_rate;              // max_error_per_base
    double ln10;                  // std::log(10) of the host libm
    int32_t do_cap, natural_log;
};

// Read quality pre-pass on the device (include/pairhmm_cuda.h, phmm_raw_read): rewrites the base / insertion / deletion
// quality planes in place from the raw bytes; one CTA per read at a time, one thread per base.
struct PrepReadsArgs {
    uint8_t *rd;                  // read planes (bases | base_q | ins_q | del_q | gcp)
    uint64_t plane;
    const DevRead *reads;
    const int32_t *mapq;          // per read
    uint32_t n;
    int32_t pcr_on, base_q_threshold;
    uint8_t pcr_table[24];        // getErrorModelAdjustedQual(repeat length 0..20), host libm
};

// Reads longer than kMaxReadLen (long.cu): one work unit = one (read, haplotype) pair, swept in row stripes by one warp;
// float pass, threshold, double pass and log10 in one kernel. unit.hap_list = index of the pair's entry in hap_list,
// unit.out = index of its result.
struct LongArgs {
    const uint8_t *rd; uint64_t plane; const DevRead *reads;
    const uint8_t *hap_bytes; const uint64_t *hap_byte_off; const DevHap *haps; const uint32_t *hap_list;
    const DevUnit *units; uint32_t n_units; uint32_t *queue;
    const float *ph2pr_f, *mm_f; const double *ph2pr_d, *mm_d;
    float log10_init_f; double log10_init_d;
    uint8_t *scratch; uint64_t scratch_per_warp;   // 3 lines of (longest haplotype) doubles per resident warp
    double *out; uint8_t *flag;
    int32_t use_double;
};

struct FinalizeArgs {
    const float *fres;
    double *out;
    uint8_t *flag;
    uint32_t n;
    float log10_init_f;           // log10f(2^120), host libm (Context.h:184)
    uint32_t *rescued_pairs;      // += pairs left to the FP64 pass (one atomic per warp); read by the rescue scan
};

// Work bins. A read of length R is swept by a group of W lanes (W = 32, 16 or 8; 32/W reads per warp),
// K = ceil(R/W) rows per lane, its last row in slot KL = (R-1) % K of lane (R-1)/K of the group. One kernel
// instantiation per (W, K, KL). Narrow groups mean more rows per lane, i.e. the per-column overhead
// (shuffles, boundary selects, loop) is amortised over more cells, at the price of registers.
constexpr int kNumWidths = 3;
constexpr int kWidths[kNumWidths] = {32, 16, 8};
constexpr int kMaxK[kNumWidths] = {kMaxRowsPerLane, 12, 12};   // rows per lane limit per width
constexpr int kMaxKF64W16 = 10;     // the wide FP64 kernel at W = 16 (two reads per warp): rows per lane its registers hold
constexpr int tri(int k) { return k * (k + 1) / 2; }
constexpr int kBinBase[kNumWidths + 1] = {0, tri(12), tri(12) + tri(12), tri(12) + tri(12) + tri(12)};
constexpr int kNumBins = kBinBase[kNumWidths];           // 78 + 78 + 78 = 234
constexpr int kNumBinsW32 = kBinBase[1];
constexpr int kNumBinsW16 = kBinBase[2] - kBinBase[1];
constexpr uint32_t kPairSlots = 1u << 15;   // RescueScanArgs::pair_table
constexpr int kLongBin = kNumBins;                       // pseudo-bin of the reads longer than kMaxReadLen (long.cu)

inline int width_index(int W) { return W == 32 ? 0 : W == 16 ? 1 : W == 8 ? 2 : -1; }
inline int bin_of(int W, int K, int KL) { return kBinBase[width_index(W)] + K * (K - 1) / 2 + KL; }
// bin of a read at a given width, -1 if it does not fit
inline int bin_of_read(int R, int W) {
    const int wi = width_index(W);
    const int K = (R + W - 1) / W;
    return (wi < 0 || R < 1 || K > kMaxK[wi]) ? -1 : bin_of(W, K, (R - 1) % K);
}
inline void bin_shape(int bin, int &W, int &K, int &KL) {
    int wi = 0;
    while (wi + 1 < kNumWidths && bin >= kBinBase[wi + 1]) ++wi;
    W = kWidths[wi];
    const int r = bin - kBinBase[wi];
    K = 1;
    while (tri(K) <= r) ++K;
    KL = r - tri(K - 1);
}

// F32_PK: packed (two reads per lane, FFMA2) uniform-gcp kernel of forward2.cuh -- the production FP32 path;
// F32_UG: scalar uniform-gcp kernel (W = 32 only; kept as the cross-check of the packed one);
// F32_GEN / F32_EXACT / F64_GEN: per-row gap-continuation registers, W = 32 only;
// F64_PK: the wide kernel of forward2.cuh on doubles (uniform gcp, W = 32) -- the production FP64 rescue path;
// F32_PFX: F32_PK with prefix sharing between the same-length haplotypes of a work unit (config.reserved[4]).
enum class Variant { F32_UG = 0, F32_GEN = 1, F32_EXACT = 2, F64_GEN = 3, F32_PK = 4, F64_PK = 5, F32_PFX = 6 };
constexpr int kNumVariants = 7;
inline bool is_packed_f32(Variant v) { return v == Variant::F32_PK || v == Variant::F32_PFX; }
// reads swept by one warp
inline int reads_per_warp(Variant v, int W) { return (is_packed_f32(v) ? 2 : 1) * (32 / W); }

// Launchers. All asynchronous on `st`; return the cudaError of the launch.
cudaError_t launch_prep_haps(const uint8_t *hap_bytes, const uint64_t *byte_off, const DevHap *haps,
                             uint32_t n_haps, uint32_t *hapw, cudaStream_t st);
cudaError_t launch_forward(Variant v, int W, int K, int KL, const KernelArgs &a, int grid_blocks, cudaStream_t st);
cudaError_t launch_finalize(const FinalizeArgs &a, cudaStream_t st);
cudaError_t launch_forward_long(const LongArgs &a, int grid_blocks, cudaStream_t st);
cudaError_t launch_rescue_scan(const RescueScanArgs &a, cudaStream_t st);
cudaError_t launch_rescue_flush(const RescueScanArgs &a, cudaStream_t st);
// finalize + scan in one launch; valid when the scan would make single-pair units (kernels.cu)
cudaError_t launch_finalize_scan_small(const RescueScanArgs &a, const FinalizeArgs &f, cudaStream_t st);   // after the scan, when a.pair_mode != 0
cudaError_t launch_epilogue(const EpilogueArgs &a, cudaStream_t st);
cudaError_t launch_prep_reads(const PrepReadsArgs &a, int sm_count, cudaStream_t st);
// resident CTAs per SM of the given instantiation (occupancy API), for sizing persistent grids
int forward_blocks_per_sm(Variant v, int W, int K, int KL);
// whether (variant, W, K) was instantiated (the general / exact / FP64 kernels exist at W = 32 only)
bool forward_available(Variant v, int W, int K);

}  // namespace phmm
