/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference PairHMM numerics.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may use it.
 * Parity status: PINNED against the reference's own code compiled here (oracle/_ref, see
 * tests/test_oracle_vs_reference.py and the golden fixtures under tests/golden/ made from it);
 * the reference ships no golden vectors of its own for this path. */
#ifndef PAIRHMM_ORACLE_H
#define PAIRHMM_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORACLE_MM_COUNT 32640 /* ((254+1)*(254+2))>>1, Context.h:52 */

/* Lookup tables exactly as Context<float>/Context<double> build them (Context.h:65-89,137-183). */
void oracle_tables(float *ph2pr_f, float *mm_f, double *ph2pr_d, double *mm_d, int mm_count);

/* One pair. Inputs are the raw bytes the reference's testcase carries (quals are masked &127 here,
 * ReadForPairHMM.cpp:35). Returns the scaled forward sum (2^120 float / 2^1020 double scale). */
float oracle_forward_float(const uint8_t *read, const uint8_t *base_q, const uint8_t *ins_q,
                           const uint8_t *del_q, const uint8_t *gcp, int R, const uint8_t *hap, int H);
double oracle_forward_double(const uint8_t *read, const uint8_t *base_q, const uint8_t *ins_q,
                             const uint8_t *del_q, const uint8_t *gcp, int R, const uint8_t *hap, int H);

/* Flat pair list, same signature as oracle/ref_driver.cpp:ref_pairhmm_compute (dispatch + rescue rule
 * of IntelPairHmm.cc:332-351). Returns wall seconds of the compute loop. */
double oracle_compute(const uint8_t *read_bases, const uint8_t *base_q, const uint8_t *ins_q,
                      const uint8_t *del_q, const uint8_t *gcp, const int64_t *read_off,
                      const int32_t *read_len, int n_reads, const uint8_t *hap_bases,
                      const int64_t *hap_off, const int32_t *hap_len, int n_haps,
                      const uint32_t *pair_read, const uint32_t *pair_hap, int64_t n_pairs,
                      int use_double, int n_threads, int tables_in_timed_region, double *out,
                      uint8_t *rescued);
/* Per-read epilogue on a read-major matrix m[n_reads][n_haps] (in place), restating what the reference's host code
 * does to the PairHMM's result: AlleleLikelihoods::normalizeLikelihoods (AlleleLikelihoods.h:372-390 with
 * normalizeLikelihoodsPerEvidence :150-166 and searchBestAllele :92-148, canBeReference = true, no priorities), the
 * predicate of filterPoorlyModeledEvidence (:404-419) with PairHMMLikelihoodCalculationEngine::log10MinTrueLikelihood
 * (PairHMMLikelihoodCalculationEngine.cpp:294-299), and switchToNaturalLog (:462-476, MathUtils.cpp:16-18).
 * Outputs per read (each nullable): best / second_best value (second on the normalised row), index of the best
 * haplotype, poorly-modelled flag. */
void oracle_epilogue(double *m, int64_t n_reads, int32_t n_haps, const int32_t *read_len, double likelihood_cap,
                     double max_error_per_base, int natural_log, double *best, double *second_best,
                     int32_t *best_hap, uint8_t *poorly_modeled);
/* Quality pre-pass the caller runs on every read before the PairHMM:
 * PairHMMLikelihoodCalculationEngine::modifyReadQualities (PairHMMLikelihoodCalculationEngine.cpp:123-148) =
 * applyPCRErrorModel (:150-158, tandem repeats by findTandemRepeatUnits :177-254 over
 * GATKVariantContextUtils::findNumberOfRepetitions, GATKVariantContextUtils.cpp:59-100, quality table
 * getErrorModelAdjustedQual :45-61) + capMinimumReadQualities (:256-267), and the constant gap-continuation array of
 * buildGapContinuationPenalties (:284-292). ins_q / del_q NULL = ReadUtils' default 45 (ReadUtils.cpp:254-262,
 * ReadUtils.h:40). pcr_rate_factor = the PCRErrorModel enum value (HOSTILE 1, AGGRESSIVE 2, CONSERVATIVE 3); 0 = no PCR
 * adjustment (GATK's NONE; the reference has no such value). PARITY UNPINNED: these functions sit in translation units
 * that need the whole Mutect2Cpp build (htslib, boost), so they are restated, not compiled from the reference. */
int oracle_tandem_repeat_length(const uint8_t *bases, int len, int offset);
void oracle_pcr_table(int pcr_rate_factor, uint8_t table[21]);
void oracle_prepare_read(const uint8_t *bases, const uint8_t *base_q, const uint8_t *ins_q, const uint8_t *del_q,
                         int len, int mapq, int pcr_rate_factor, int base_q_threshold, int constant_gcp,
                         uint8_t *out_base_q, uint8_t *out_ins_q, uint8_t *out_del_q, uint8_t *out_gcp);
#ifdef __cplusplus
}
#endif
#endif
