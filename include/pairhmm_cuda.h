/* libpairhmm_cuda -- C ABI of the B200-native PairHMM read-likelihood engine.
 *
 * This is the drop-in boundary for deepmutect / Mutect2Cpp's PairHMM plugin point. Citations are
 * relative to /root/reference/deepmutect/Mutect2Cpp-master/src ("M2/").
 *
 *   reference interface                                          replaced by
 *   ------------------------------------------------------------ -------------------------------
 *   initNative(use_double, max_threads)                          pairhmm_cuda_create
 *       M2/intel/pairhmm/IntelPairHmm.h:37, IntelPairHmm.cc:202
 *   PairHMM::initialize(readMaxLength, haplotypeMaxLength)       pairhmm_cuda_initialize
 *       M2/utils/pairhmm/PairHMM.h:22-37, PairHMM.cpp:14-27
 *   VectorLoglessPairHMM::computeLog10Likelihoods (+_trie,       pairhmm_cuda_compute
 *   _trie_unique): reads x haplotypes of one region, read-major  (async: _submit / _wait)
 *   out[r*nHaps+h]  M2/utils/pairhmm/VectorLoglessPairHMM.cpp:43-148,150-286
 *   computeLikelihoodsNative(testcases, likelihoodArray)         pairhmm_cuda_compute_pairs
 *       M2/intel/pairhmm/IntelPairHmm.h:39, IntelPairHmm.cc:259-293,332-351
 *   (many regions from many host threads; the reference has no   pairhmm_cuda_compute_batch
 *   equivalent -- it loops; here one call = few large launches)
 *   doneNative (commented JNI original, IntelPairHmm.cc:192-200) pairhmm_cuda_destroy
 *
 * Numerics contract: FP32 forward pass scaled by 2^120; a pair whose float
 * sum is < 1e-28f (MIN_ACCEPTED, M2/intel/pairhmm/pairhmm_common.h:41) is recomputed in FP64 scaled
 * by 2^1020; value returned = log10 likelihood as double. Qualities are masked with &127
 * (M2/haplotypecaller/ReadForPairHMM.cpp:35); bases other than A,C,G,T,N behave as 'A'
 * (pairhmm_common.cc:26); 'N' on either side matches everything (avx-pairhmm-template.h:38,48-51).
 *
 * Ownership: all inputs are borrowed for the duration of the call only (for _submit: until _submit
 * returns -- the engine has copied them to its pinned staging by then); outputs must stay valid
 * until the call (or the matching _wait) returns. No C++ types and no exceptions cross this ABI.
 * Every function returns 0 on success or a negative pairhmm_cuda_status; the text of the last
 * error on a handle is available from pairhmm_cuda_last_error.
 *
 * Threading: one handle per host thread (as Mutect2Cpp keeps one PairHMM per worker thread,
 * M2/main.cpp:199); calls on different handles are fully concurrent. A handle owns its CUDA
 * streams, pinned staging and device buffers. Host-side packing loops of all handles share one
 * process-wide pool of worker threads (config.pack_threads bounds a call's share of it).
 */
#ifndef PAIRHMM_CUDA_H
#define PAIRHMM_CUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PAIRHMM_CUDA_ABI_VERSION 2   /* 2: pairhmm_cuda_config gained device_mask / batch_window_us (multi-GPU handles, shared batcher) */

typedef enum pairhmm_cuda_status {
    PAIRHMM_OK = 0,
    PAIRHMM_ERR_ARG = -1,      /* null pointer, negative size or count (reads and haplotypes may have any length) */
    PAIRHMM_ERR_CUDA = -2,     /* a CUDA runtime call failed (text in last_error)                    */
    PAIRHMM_ERR_NO_DEVICE = -3,/* no usable sm_100 device: the engine has NO CPU fallback            */
    PAIRHMM_ERR_OOM = -4,      /* host or device allocation failed                                   */
    PAIRHMM_ERR_STATE = -5     /* e.g. wait on an unknown ticket                                     */
} pairhmm_cuda_status;

/* One read as VectorLoglessPairHMM.cpp:71-77 gathers it: five byte arrays of length len. */
typedef struct phmm_read {
    const uint8_t *bases;  /* getBasesNoCopy()                                   */
    const uint8_t *base_q; /* getBaseQualitiesNoCopy()                           */
    const uint8_t *ins_q;  /* ReadUtils::getBaseInsertionQualities (default 45)  */
    const uint8_t *del_q;  /* ReadUtils::getBaseDeletionQualities  (default 45)  */
    const uint8_t *gcp;    /* gap continuation penalties, (*gcp)[read]           */
    int32_t len;
} phmm_read;

/* One haplotype: HaplotypeDataHolder{haplotypeBases,length}, M2/utils/pairhmm/HaplotypeDataHolder.h:13 */
typedef struct phmm_hap {
    const uint8_t *bases;
    int32_t len;
} phmm_hap;

/* One active region = one computeLog10Likelihoods call. log10_out / rescued_out are
 * [n_reads * n_haps], read-major; rescued_out may be NULL. */
typedef struct phmm_region {
    const phmm_read *reads;
    int32_t n_reads;
    const phmm_hap *haps;
    int32_t n_haps;
    double *log10_out;
    uint8_t *rescued_out;
} phmm_region;

typedef struct pairhmm_cuda_config {
    int32_t device;          /* CUDA ordinal; -1 = current device                                  */
    int32_t use_double;      /* initNative's use_double: 1 = every pair through the FP64 kernel     */
    int32_t exact_fp32;      /* 1 = FP32 pass without FMA contraction (bit-identical to the         */
                             /*     reference's AVX-256 build); 0 = contracted (as its AVX-512 build) */
    int32_t dedup_reads;     /* 1 = de-duplicate identical reads per region (VLPH.cpp:83-103)       */
    int32_t pack_threads;    /* host threads used to pack a batch (0 = auto)                        */
    int64_t chunk_cells;     /* target DP cells per internal launch group (0 = default)             */
    int32_t reserved[8];     /* [0] != 0: never use the uniform-gcp kernels (testing);               */
                             /* [1] = 8|16|32: force that lane-group width where it fits (tuning)    */
                             /* [2] != 0: scalar instead of packed (FFMA2) uniform-gcp kernels (testing) */
                             /* [3] > 0: at most that many resident CTAs per SM (tuning)              */
                             /* [5] > 0: at most that many haplotypes per work unit (tuning; default: all of a region's, fewer */
                             /*     when the batch is too small to fill the GPU otherwise)                  */
                             /* [6] > 0: underflowed pairs per unit of the FP64 pass (tuning; default by chunk size) */
                             /* [7]: FP64 pass with two reads per warp where the rows fit: 0 = when the launch group has enough   */
                             /*     underflowed pairs, 1 = never, 2 = always (testing); results do not depend on it                 */
                             /* [4] != 0: prefix sharing -- same-length haplotypes of a region that share  */
                             /*     a prefix are swept from a checkpoint (the reference's trie,            */
                             /*     VectorLoglessPairHMM.cpp:35,150-286; M2/trie/buildTreeUtils.cpp:155);  */
                             /*     results are bit-identical to the per-pair sweep. Applies to the packed */
                             /*     FP32 pass; calls with the fused epilogue ignore it. 1 = per launch     */
                             /*     group, only when the plan skips enough cells to pay for the checkpoint  */
                             /*     steps (> 12 %); 2 = always (testing)                                    */
    /* ---- ABI version 2 ---- */
    uint32_t device_mask;    /* != 0: a MULTI-DEVICE handle over the CUDA ordinals whose bits are set (`device` is ignored). One process,  */
                             /*     one handle, as Mutect2Cpp is one process with T threads (M2/main.cpp:199,570-574): a batch call      */
                             /*     (_compute_batch / _flat / _raw / _normalized) splits its regions over the devices by DP cells with   */
                             /*     the LPT rule -- whole regions first, a region's reads are cut into parts only when the region        */
                             /*     exceeds 1 / (4 x devices) of the batch -- one host worker thread, one pinned arena and two slots per  */
                             /*     device, and every device's results land in the caller's ONE result array (that is the gather;        */
                             /*     there is no collective: pairs are independent, IntelPairHmm.cc:332-351). Single-region calls         */
                             /*     (_compute, _submit) go to the devices in turn.                                                       */
    int32_t batch_window_us; /* > 0: blocking single-region calls (_compute, _compute_raw, _compute_normalized) of ALL handles created   */
                             /*     with it on one device go through one process-wide batcher: regions submitted by different host      */
                             /*     threads while the GPU is busy (or within this many microseconds of the first one) are packed into    */
                             /*     ONE launch group and scattered back per caller -- the counterpart of the reference's cross-thread     */
                             /*     work sharing (computeLikelihoodsNative_concurrent, IntelPairHmm.cc:296-330;                          */
                             /*     PairHMMConcurrentControl.h:15-53; the stealing loop M2/main.cpp:396-417). 0 = every handle launches  */
                             /*     for itself.                                                                                         */
    int32_t reserved2[6];
} pairhmm_cuda_config;

typedef struct pairhmm_cuda_stats {
    uint64_t calls, regions, reads, unique_reads, pairs, unique_pairs, cells;
    uint64_t rescued_pairs;      /* pairs routed to the FP64 kernel                       */
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t kernel_launches;    /* launches of this library's own kernels                */
    double gpu_ms_float, gpu_ms_double, gpu_ms_total; /* CUDA-event times, summed over calls   */
    double host_ms_pack, host_ms_scatter, host_ms_total;
    uint64_t cells_swept;        /* DP cells the FP32 pass swept: `cells` less duplicate reads and,      */
                                 /* with prefix sharing, less the shared haplotype prefixes              */
} pairhmm_cuda_stats;

typedef struct pairhmm_cuda_handle pairhmm_cuda_handle;
typedef int64_t pairhmm_cuda_ticket;

int pairhmm_cuda_abi_version(void);
/* number of usable sm_100 devices (0 if none / no driver) */
int pairhmm_cuda_device_count(void);

void pairhmm_cuda_default_config(pairhmm_cuda_config *cfg);
int pairhmm_cuda_create(const pairhmm_cuda_config *cfg, pairhmm_cuda_handle **out);
int pairhmm_cuda_destroy(pairhmm_cuda_handle *h);

/* PairHMM::initialize(readMaxLength, haplotypeMaxLength): both must be > 0 (PairHMM.cpp:16).
 * Sizes the staging for one region; larger inputs later make the buffers grow, they are not an error. */
int pairhmm_cuda_initialize(pairhmm_cuda_handle *h, int read_max_len, int hap_max_len);

/* One region, blocking. n_reads == 0 returns immediately (VectorLoglessPairHMM.cpp:46-47). */
int pairhmm_cuda_compute(pairhmm_cuda_handle *h, const phmm_read *reads, int n_reads,
                         const phmm_hap *haps, int n_haps, double *log10_out, uint8_t *rescued_out);

/* Same, asynchronous: inputs are consumed before _submit returns; outputs are written by _wait. */
int pairhmm_cuda_submit(pairhmm_cuda_handle *h, const phmm_read *reads, int n_reads,
                        const phmm_hap *haps, int n_haps, double *log10_out, uint8_t *rescued_out,
                        pairhmm_cuda_ticket *ticket);
int pairhmm_cuda_wait(pairhmm_cuda_handle *h, pairhmm_cuda_ticket ticket);

/* Many regions in one call (coalesced into few large launches, packing overlapped with compute). */
int pairhmm_cuda_compute_batch(pairhmm_cuda_handle *h, const phmm_region *regions, int64_t n_regions);

/* Flat form of the same batch for language bindings: five read byte arrays and the haplotype bytes
 * concatenated, per-item offsets/lengths, region r owns reads [reg_read_start[r], reg_read_start[r+1])
 * and haps [reg_hap_start[r], reg_hap_start[r+1]); output region-major then read-major, i.e.
 * region r starts at sum_{q<r} nReads_q * nHaps_q. */
int pairhmm_cuda_compute_flat(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                              const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *gcp,
                              const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                              const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len,
                              int64_t n_haps, const int64_t *reg_read_start, const int64_t *reg_hap_start,
                              int64_t n_regions, double *log10_out, uint8_t *rescued_out);

/* Explicit pair list (computeLikelihoodsNative's testcase vector): out[p] for pair
 * (reads[read_idx[p]], haps[hap_idx[p]]). */
int pairhmm_cuda_compute_pairs(pairhmm_cuda_handle *h, const phmm_read *reads, int n_reads,
                               const phmm_hap *haps, int n_haps, const uint32_t *read_idx,
                               const uint32_t *hap_idx, size_t n_pairs, double *log10_out,
                               uint8_t *rescued_out);

/* Device-resident replay, for measuring the kernels without host<->device traffic: stage a flat batch
 * once (H2D + layout), then run the float+rescue kernels on it any number of times; results stay on
 * the device until _resident_fetch. kernel_ms (nullable) receives the CUDA-event time of the whole run
 * (FP32 forward + finalize + FP64 rescue), float_kernel_ms that of the FP32 forward kernels alone. */
int pairhmm_cuda_resident_load(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                               const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *gcp,
                               const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                               const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len,
                               int64_t n_haps, const int64_t *reg_read_start, const int64_t *reg_hap_start,
                               int64_t n_regions);
int pairhmm_cuda_resident_run(pairhmm_cuda_handle *h, float *kernel_ms, float *float_kernel_ms);
int pairhmm_cuda_resident_fetch(pairhmm_cuda_handle *h, double *log10_out, uint8_t *rescued_out);
/* Raw FP32-pass sums (2^120-scaled, before log10) of the last _resident_run, one per input pair in result
 * order; n must equal the pair count. Test hook for bit-exactness of the FP32 kernel against
 * compute_full_prob_avxs (avx-pairhmm-template.h:204-345). */
int pairhmm_cuda_resident_fetch_raw(pairhmm_cuda_handle *h, float *fres_out, int64_t n);

/* ---- prefix sharing -------------------------------------------------------------------
 * The reference's trie (VectorLoglessPairHMM::initialize picks it when nHaps / nDistinctLens > 3,
 * VectorLoglessPairHMM.cpp:35; buildTreeUtils::buildTreeWithHaplotype_same_height, M2/trie/buildTreeUtils.cpp:155-258;
 * compute_full_prob_with_trie_, avx-pairhmm-template.h:475-721) computes the DP columns of a prefix shared by several
 * SAME-LENGTH haplotypes once per read. With config.reserved[4] != 0 the engine does the same inside its packed FP32
 * pass, at the granularity of its 4-column tiles: a region's haplotypes are swept in (length, bytes) order, a sweep
 * may leave one checkpoint of the lane state at a tile boundary and may start from an earlier one. Values are
 * bit-identical to the per-pair sweep (same operations in the same order), as the reference's trie path is to its
 * per-pair path; the FP64 rescue stays per pair (IntelPairHmm.cc:700-703).
 *
 * pairhmm_cuda_prefix_plan exposes the host-side plan for one region (no device needed; what the packer computes):
 * order[q] = index in haps[] of the haplotype swept q-th; for the q-th sweep t_start[q] = tiles (of 4 columns) taken
 * from a checkpoint, restore_slot[q] = the slot it is read from (-1: fresh sweep, t_start == 0), t_save[q] / save_slot[q]
 * = the tile boundary at which it writes a checkpoint and where (t_save == 0: none). unit_haps = haplotypes per work
 * unit (sharing never crosses a unit; 0 = as the packer would choose for a large batch). shared_cols (nullable) = sum of
 * the columns not swept again. */
int pairhmm_cuda_prefix_plan(const phmm_hap *haps, int n_haps, int unit_haps, uint32_t *order, uint32_t *t_start,
                             int32_t *restore_slot, uint32_t *t_save, uint32_t *save_slot, uint64_t *shared_cols);

/* ---- fused per-read epilogue ---------------------------------------------------------
 * What PairHMMLikelihoodCalculationEngine::computeReadLikelihoods does to the matrix right after the PairHMM
 * (PairHMMLikelihoodCalculationEngine.cpp:93-94) and Mutect2Engine does before genotyping (Mutect2Engine.cpp:224),
 * done on the device on the rows of the result matrix before they are copied back:
 *   1. AlleleLikelihoods::normalizeLikelihoods(likelihood_cap)  (AlleleLikelihoods.h:372-390,150-166): per read, every
 *      value below best + likelihood_cap is raised to it (skipped for one haplotype or an infinite cap);
 *   2. the predicate of filterPoorlyModeledEvidence (AlleleLikelihoods.h:404-419): best < log10MinTrueLikelihood(read)
 *      = min(2, ceil(len * max_error_per_base)) * -4  (PairHMMLikelihoodCalculationEngine.cpp:294-299) -- reported
 *      as a flag, the caller removes the evidence;
 *   3. searchBestAllele(canBeReference = true, no priorities) on the normalised row (AlleleLikelihoods.h:92-148);
 *   4. optionally switchToNaturalLog (AlleleLikelihoods.h:462-476): every value * std::log(10). */
typedef struct pairhmm_cuda_epilogue {
    double likelihood_cap;      /* log10globalReadMismappingRate, < 0 (default -4.5 = phred 45); -inf = no cap      */
    double max_error_per_base;  /* EXPECTED_ERROR_RATE_PER_BASE = 0.02 (PairHMMLikelihoodCalculationEngine.cpp:20)  */
    int32_t natural_log;        /* 1 = matrix, best and second_best returned as natural logs                         */
    int32_t reserved;
} pairhmm_cuda_epilogue;

typedef struct phmm_read_summary {
    double best;                /* likelihood of the best haplotype                                                  */
    double second_best;         /* of the runner-up on the normalised row (-inf with one haplotype)                  */
    int32_t best_hap;           /* index of the best haplotype (first maximum)                                       */
    uint8_t poorly_modeled;     /* 1 = filterPoorlyModeledEvidence would remove this read                            */
    uint8_t pad[3];
} phmm_read_summary;

/* One region, blocking: log10_out receives the NORMALISED matrix, summary_out[n_reads] (nullable) the per-read results. */
int pairhmm_cuda_compute_normalized(pairhmm_cuda_handle *h, const phmm_read *reads, int n_reads,
                                    const phmm_hap *haps, int n_haps, const pairhmm_cuda_epilogue *ep,
                                    double *log10_out, uint8_t *rescued_out, phmm_read_summary *summary_out);
/* Many regions: summary_out[r] (nullable, as the array itself) points at region r's [n_reads] summaries. */
int pairhmm_cuda_compute_batch_normalized(pairhmm_cuda_handle *h, const phmm_region *regions, int64_t n_regions,
                                          const pairhmm_cuda_epilogue *ep, phmm_read_summary *const *summary_out);
/* Flat form: summary_out[n_reads] in input read order. */
int pairhmm_cuda_compute_flat_normalized(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                                         const uint8_t *ins_q, const uint8_t *del_q, const uint8_t *gcp,
                                         const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                                         const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len,
                                         int64_t n_haps, const int64_t *reg_read_start, const int64_t *reg_hap_start,
                                         int64_t n_regions, const pairhmm_cuda_epilogue *ep, double *log10_out,
                                         uint8_t *rescued_out, phmm_read_summary *summary_out);

/* ---- device-side read quality pre-pass ---------------------------------------------------
 * Takes reads as the BAM record holds them and does on the device what
 * PairHMMLikelihoodCalculationEngine::modifyReadQualities does on the host before every PairHMM call
 * (PairHMMLikelihoodCalculationEngine.cpp:123-148): applyPCRErrorModel (:150-158, tandem-repeat scan
 * findTandemRepeatUnits :177-254), capMinimumReadQualities (:256-267: base quality capped by the mapping quality, below
 * base_quality_threshold -> 6; insertion / deletion qualities below 6 -> 6), and the constant gap-continuation penalty of
 * buildGapContinuationPenalties (:284-292). Only bases and base qualities (and BI/BD arrays where a read has them)
 * cross PCIe: 2 bytes per base instead of 5. */
typedef struct phmm_raw_read {
    const uint8_t *bases;
    const uint8_t *base_q;   /* getBaseQualities()                                                    */
    const uint8_t *ins_q;    /* existing base insertion qualities, or NULL = default 45 (ReadUtils.h:40) */
    const uint8_t *del_q;    /* existing base deletion qualities, or NULL = default 45                 */
    int32_t len;
    int32_t mapq;            /* getMappingQuality()                                                    */
} phmm_raw_read;

typedef struct phmm_raw_region {
    const phmm_raw_read *reads;
    int32_t n_reads;
    const phmm_hap *haps;
    int32_t n_haps;
    double *log10_out;       /* [n_reads * n_haps], read-major */
    uint8_t *rescued_out;    /* nullable */
} phmm_raw_region;

typedef struct pairhmm_cuda_read_model {
    int32_t pcr_rate_factor;        /* PCRErrorModel: HOSTILE 1, AGGRESSIVE 2, CONSERVATIVE 3 (the default,            */
                                    /* LikelihoodEngineArgumentCollection.h:30); 0 = no PCR adjustment (GATK's NONE)   */
    int32_t base_quality_threshold; /* BASE_QUALITY_SCORE_THRESHOLD, 18; must be >= 6 (PHLCE.cpp:39)                   */
    int32_t constant_gcp;           /* gcpHMM, 10                                                                      */
    int32_t reserved;
} pairhmm_cuda_read_model;
void pairhmm_cuda_default_read_model(pairhmm_cuda_read_model *m);

int pairhmm_cuda_compute_raw(pairhmm_cuda_handle *h, const phmm_raw_read *reads, int n_reads, const phmm_hap *haps,
                             int n_haps, const pairhmm_cuda_read_model *model, double *log10_out, uint8_t *rescued_out);
int pairhmm_cuda_compute_batch_raw(pairhmm_cuda_handle *h, const phmm_raw_region *regions, int64_t n_regions,
                                   const pairhmm_cuda_read_model *model);
/* Flat form: ins_q / del_q may be NULL (every read takes the default); mapq[n_reads]. */
int pairhmm_cuda_compute_flat_raw(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                                  const uint8_t *ins_q, const uint8_t *del_q, const int32_t *mapq,
                                  const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                                  const uint8_t *hap_bases, const int64_t *hap_off, const int32_t *hap_len,
                                  int64_t n_haps, const int64_t *reg_read_start, const int64_t *reg_hap_start,
                                  int64_t n_regions, const pairhmm_cuda_read_model *model, double *log10_out,
                                  uint8_t *rescued_out);
/* The pre-pass alone (test hook and stand-alone use): processed base / insertion / deletion / gap-continuation
 * qualities of every read, each array laid out like base_q (read_off, read_len). */
int pairhmm_cuda_prepare_reads_flat(pairhmm_cuda_handle *h, const uint8_t *read_bases, const uint8_t *base_q,
                                    const uint8_t *ins_q, const uint8_t *del_q, const int32_t *mapq,
                                    const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                                    const pairhmm_cuda_read_model *model, uint8_t *out_base_q, uint8_t *out_ins_q,
                                    uint8_t *out_del_q, uint8_t *out_gcp);

/* ---- allocation-free entry point ---------------------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it). It calls no cudaMalloc / cudaFree and no
 * allocating library routine, takes no handle, and touches nothing of the engine's state.
 *
 * Explicit pair list on DEVICE arrays the caller owns: pair p = read d_pair_read[p] x haplotype d_pair_hap[p];
 * d_read_planes = five byte planes plane_stride apart (bases | base qualities | insertion | deletion | gap-continuation
 * qualities, as phmm_read), read r at byte d_read_off[r] of every plane, d_read_len[r] bases (any length >= 1; a
 * zero-length read yields -inf, rescued 1); haplotype h = d_hap_bases[d_hap_off[h] .. + d_hap_len[h]). Writes
 * d_log10_out[p] (double) and, unless NULL, d_rescued_out[p] -- same numerics as every other entry point (FP32 pass,
 * < 1e-28f -> FP64, log10), computed by the row-striped kernel (one warp per pair: ~0.8 TCUPS, a quarter of the packed
 * kernels' rate -- the price of needing no per-call layout pass). The launch is asynchronous on cuda_stream (a
 * cudaStream_t; NULL = the default stream); the caller synchronises.
 *
 * Limit: the kernel's scratch -- three hand-over lines per resident warp -- lives in a static __device__ array of 24.1 MiB
 * (+ 98 KiB of lookup tables) that every GPU loading this library holds until exit, sized for haplotypes of at most
 * PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN bases. max_hap_len = the caller's bound on d_hap_len[]; a larger bound is refused
 * with PAIRHMM_ERR_ARG and a message (a haplotype that nevertheless exceeds the limit gets NaN, rescued 255). ALL calls on
 * a device share that array: calls that overlap on the GPU (different streams, or threads) race -- serialise them. The
 * first call on each device performs a one-time set-up (symbol addresses, table upload: a synchronous copy) and must not
 * sit inside a stream capture. error_buf (nullable, error_buf_len bytes) receives the text of an error. */
#define PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN 2048
int pairhmm_cuda_forward_pairs_noalloc(int device, void *cuda_stream, const uint8_t *d_read_planes, uint64_t plane_stride,
                                       const uint32_t *d_read_off, const int32_t *d_read_len,
                                       const uint8_t *d_hap_bases, const uint64_t *d_hap_off, const int32_t *d_hap_len,
                                       const uint32_t *d_pair_read, const uint32_t *d_pair_hap, uint32_t n_pairs,
                                       int32_t max_hap_len, int32_t use_double, double *d_log10_out, uint8_t *d_rescued_out,
                                       char *error_buf, size_t error_buf_len);

/* ---- multi-device plan (host logic, no device needed; what a multi-device handle computes for a batch call) ----
 * Region r has n_reads[r] reads whose lengths sum to read_len_sum[r] and haplotypes whose lengths sum to hap_len_sum[r]
 * (cost = DP cells = the product). Writes at most max_parts parts in (region, first read) order: part p = reads
 * [part_first_read[p], + part_n_reads[p]) of region part_region[p] on device part_device[p] (an index into the set bits
 * of device_mask, lowest first). Regions are split into parts of whole reads only when their cost exceeds
 * 1 / (4 x n_devices) of the batch. Returns PAIRHMM_ERR_ARG if max_parts is too small. */
int pairhmm_cuda_device_plan(const uint64_t *read_len_sum, const uint64_t *hap_len_sum, const int32_t *n_reads,
                             int64_t n_regions, int n_devices, int64_t max_parts, int64_t *part_region,
                             int32_t *part_first_read, int32_t *part_n_reads, int32_t *part_device, int64_t *n_parts);

/* Statistics. On a multi-device handle the counters are summed over the devices; the three gpu_ms_* and host_ms_pack /
 * host_ms_scatter are the MAXIMUM over the devices (they work concurrently), host_ms_total is the caller's wall time. */
int pairhmm_cuda_stats_get(pairhmm_cuda_handle *h, pairhmm_cuda_stats *out);
int pairhmm_cuda_stats_reset(pairhmm_cuda_handle *h);
const char *pairhmm_cuda_last_error(pairhmm_cuda_handle *h);

/* The lookup tables the engine uploaded (ph2pr[128], matchToMatchProb triangle for q<128 = 8256
 * entries), for tests: they must equal Context<float/double>'s (M2/intel/pairhmm/Context.h:75-89). */
#define PAIRHMM_CUDA_MM_COUNT 8256
int pairhmm_cuda_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d);

#ifdef __cplusplus
}
#endif
#endif /* PAIRHMM_CUDA_H */
